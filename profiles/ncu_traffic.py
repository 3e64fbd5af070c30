"""profiles/ncu_traffic.json from `ncu --set full` captures: per-launch dram__bytes_read.sum + dram__bytes_write.sum of the
env kernel and the actor kernel.  bench.py reads that file for roofline.traffic.
    python profiles/ncu_traffic.py <envs>:<file.ncu-rep> [<envs>:<file.ncu-rep> ...]"""
import csv
import io
import json
import os
import subprocess
import sys

entries, sources = [], []
for spec in sys.argv[1:]:
    envs, path = spec.split(":", 1)
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head = rows[0]
    ik, ir, iw = head.index("Kernel Name"), head.index("dram__bytes_read.sum"), head.index("dram__bytes_write.sum")
    units = rows[1]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    seen = {}
    for r in rows[2:]:
        kind = ("env" if "env_step_kernel" in r[ik] else "fused" if "actor_kernel<32, 1>" in r[ik]      # kEnv: actor + env step
                else "actor" if "actor_kernel" in r[ik] else None)
        if kind is None:
            continue
        b = float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]]
        seen.setdefault(kind, []).append(b)
    for kind, v in seen.items():
        entries.append({"kernel": kind, "envs": int(envs), "bytes": int(sum(v) / len(v)), "launches": len(v)})
    sources.append(os.path.basename(path))
json.dump({"source": "ncu --set full --clock-control none captures " + ", ".join(sources) + " (summaries under profiles/); cache control = ncu "
                     "default: every replay starts from a flushed L2",
           "note": "dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over the captured launches); bench.py reads this "
                   "file for roofline.traffic.  Regenerate with profiles/ncu_traffic.py after every new --set full capture.",
           "entries": entries}, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json"), "w"), indent=1)
print(json.dumps(entries))
