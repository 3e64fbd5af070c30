import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def built_lib():
    """The in-tree shared library; compiled on first use (nvcc cross-compiles without a GPU)."""
    import options_in_simulation_b200 as ois

    if not os.path.exists(ois.LIB_PATH):
        ois.build()
    return ois.lib()


def golden_files(prefix="env_"):
    return sorted(f for f in os.listdir(GOLDEN) if f.startswith(prefix) and f.endswith(".npz"))
