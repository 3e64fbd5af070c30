"""CPU: the C-ABI shared library builds for sm_100a, loads, and exports every symbol include/ois_b200.h
declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ois_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ois_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(built_lib):
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(built_lib, n), f"{n} declared in include/ois_b200.h but not exported"


def test_binding_table_matches_header(built_lib):
    from options_in_simulation_b200 import _lib as L

    assert sorted(L.SYMBOLS) == declared_symbols()
    assert built_lib.ois_abi_version() == 3
    assert b"sm_100a" in built_lib.ois_build_info()


def test_struct_sizes_match_c_layout(built_lib):
    """ctypes mirrors vs the compiler's layout (ois_struct_size is part of the ABI for exactly this)."""
    from options_in_simulation_b200 import _lib as L

    for idx, cls in enumerate((L.EnvParams, L.EnvState, L.EnvIO, L.MlpShape, L.Scaler, L.ActIO, L.UpdateIO, L.StepCtl, L.SampleCtl)):
        assert built_lib.ois_struct_size(idx) == ctypes.sizeof(cls), cls.__name__


def test_argument_errors_do_not_need_a_gpu(built_lib):
    from options_in_simulation_b200 import _lib as L

    rc = built_lib.ois_env_step(None, None, None, 0, None)
    assert rc == -1
    assert b"null" in built_lib.ois_last_error_string()
    rc = built_lib.ois_replay_sample_indices(None, 4, 2, 0, 0, None)
    assert rc == -1


def test_no_cpu_fallback_in_product():
    """The product package never imports the oracle and has no eager fallback path."""
    pkg = os.path.join(ROOT, "options_in_simulation_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert "import oracle" not in src and "from oracle" not in src, f


def test_sass_is_sm100a(built_lib):
    import subprocess

    from options_in_simulation_b200 import LIB_PATH

    out = subprocess.run(["cuobjdump", "-lelf", LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out


def test_abi_version_is_stated_consistently(built_lib):
    """include/ois_b200.h, the ctypes loader, __graft_entry__.build() and INTEGRATION.md must agree."""
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ver = int(re.search(r"#define OIS_ABI_VERSION (\d+)", open(os.path.join(root, "include", "ois_b200.h")).read()).group(1))
    assert built_lib.ois_abi_version() == ver
    for rel in ("__graft_entry__.py", "INTEGRATION.md", os.path.join("options_in_simulation_b200", "_lib.py")):
        text = open(os.path.join(root, rel)).read()
        found = re.findall(r"ois_abi_version\(\) [!=]= (\d+)", text)
        assert found and all(int(v) == ver for v in found), (rel, found)
