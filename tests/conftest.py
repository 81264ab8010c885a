import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    """Build the native libraries / binaries once per session (make is incremental)."""
    res = subprocess.run(["make", "-C", ROOT, "-j", str(os.cpu_count() or 4), "cpu"], stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 0, res.stdout[-4000:]
    if not os.path.exists(os.path.join(ROOT, "mpi_cuda_cnn_b200", "lib", "libb200cnn_cuda.so")):
        res = subprocess.run(["make", "-C", ROOT, "-j", str(os.cpu_count() or 4), "cuda"], stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True)
        assert res.returncode == 0, res.stdout[-4000:]
    yield


@pytest.fixture(scope="session")
def repo_root():
    return ROOT


@pytest.fixture(scope="session")
def dataset_paths(tmp_path_factory):
    from mpi_cuda_cnn_b200.utils import write_synthetic_dataset

    d = tmp_path_factory.mktemp("idx")
    return write_synthetic_dataset(str(d), ntrain=200, ntest=60)
