#!/bin/bash
# Short single-GPU pass (for the last GPU-minutes of a round): parity tests, smoke, the bench line and the reference arm.
# usage (from the repo root, under gpurun): bash profiles/gpu_round_short.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
timeout 240 python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${tag}_smoke.log
timeout 240 python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
timeout 120 python bench.py --impl reference --steps 20 --warmup 20 > gpurun_out/${tag}_bench_ref_n1.json 2>> gpurun_out/${tag}_bench.err; echo "ref exit $?"
