"""bench.py: the reference arm (runs on the CPU, no GPU needed) prints the JSON line the driver expects and keeps this
repository's package out of its process."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    if not os.path.exists(os.path.join(ROOT, "baseline", "_ref", "cnn")):
        pytest.skip("reference not installed under baseline/_ref (run baseline/install_reference.sh)")
    env = dict(os.environ, PYTHONPATH="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "20", "--warmup", "3"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "train_images_per_sec" and line["unit"] == "images/s"
    assert line["n_gpus"] == 1 and line["higher_is_better"] is True and line["value"] > 0
    assert line["dtype"] == "fp64" and line["data"] == "synthetic" and line["gpu_launches"] == 0
    e2e = line["e2e"]
    assert e2e["value"] > 0 and "h2d_bytes_per_step" in e2e and "d2h_bytes_per_step" in e2e
    assert "cnn" in line["config"]["program"]


def test_reference_arm_does_not_import_the_package():
    """The driver records which of this repository's shared objects a bench process loads: the reference arm must load none.
    (Static check: the function runs before anything of the package is imported.)"""
    src = open(os.path.join(ROOT, "bench.py")).read()
    head = src[:src.index("def parse")]                                        # module level: imports
    ref_fn = src[src.index("def _write_idx"):src.index("class _Arm")]          # _write_idx, _synthetic_idx, run_reference
    assert "mpi_cuda_cnn_b200" not in head.replace("mpi_cuda_cnn_b200/", "") or "import mpi_cuda_cnn_b200" not in head
    assert "import mpi_cuda_cnn_b200" not in head and "from mpi_cuda_cnn_b200" not in head
    assert "mpi_cuda_cnn_b200" not in ref_fn
