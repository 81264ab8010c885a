"""Entry points: same command-line contract as the reference's programs (cnn.c:408-412, :432-518)."""
import os
import re
import subprocess
import sys

import pytest

PY = [sys.executable, "-m", "mpi_cuda_cnn_b200.cli"]


def _run(args, repo_root, timeout=600):
    env = dict(os.environ, PYTHONPATH=repo_root)
    return subprocess.run(args, stderr=subprocess.PIPE, stdout=subprocess.PIPE, text=True, cwd=repo_root, env=env, timeout=timeout)


def test_cli_dispatch_cpu_programs(dataset_paths, repo_root):
    assert _run(PY + ["cnn"], repo_root).returncode == 100
    assert _run(PY + ["nonsense"], repo_root).returncode == 100
    r = _run(PY + ["cnn", "--nepoch", "1"] + dataset_paths, repo_root)
    assert r.returncode == 0 and "ntests=60" in r.stderr
    r = _run(PY + ["cnnmpi", "-np", "2", "--nepoch", "1"] + dataset_paths, repo_root)
    assert r.returncode == 0 and r.stderr.count("param_checksum") == 2


def test_gpu_programs_usage_errors_need_no_gpu(repo_root):
    assert _run(PY + ["cudacnn"], repo_root).returncode == 100
    assert _run(PY + ["cudampi", "a", "b"], repo_root).returncode == 100
    assert _run([os.path.join(repo_root, "bin", "cudacnn")], repo_root).returncode == 100


@pytest.mark.gpu
@pytest.mark.parametrize("prog", ["python", "native"])
def test_cudacnn_trains_and_reports_like_the_reference(prog, dataset_paths, repo_root):
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    cmd = (PY + ["cudacnn"]) if prog == "python" else [os.path.join(repo_root, "bin", "cudacnn")]
    r = _run(cmd + ["--nepoch", "30", "--batch-size", "50", "--rate", "0.3", "--timing"] + dataset_paths, repo_root)
    assert r.returncode == 0, r.stderr
    lines = r.stderr.splitlines()
    assert lines[0] == "training..." and re.match(r"i=0, error=0\.\d{4}$", lines[1]) and "testing..." in lines
    assert "train_images_per_sec=" in r.stderr
    m = re.search(r"ntests=60, ncorrect=(\d+)", lines[-1])
    assert m and int(m.group(1)) >= 50
    bad = _run(cmd + ["a", "b", "c", "d"], repo_root)
    assert bad.returncode == 111


@pytest.mark.gpu
def test_cudampi_two_gpus(dataset_paths, repo_root):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run(PY + ["cudampi", "-np", "2", "--nepoch", "20", "--batch-size", "50", "--rate", "0.3"] + dataset_paths, repo_root)
    assert r.returncode == 0, r.stderr
    sums = re.findall(r"rank \d param_checksum=(\S+)", r.stderr)
    tail = r.stderr[-3000:]  # shown when an assertion fails: which of the checks, and what the ranks printed
    assert len(sums) == 2 and sums[0] == sums[1], tail
    assert re.search(r"^0 0 100$", r.stderr, re.M) and re.search(r"^1 100 200$", r.stderr, re.M), r.stderr[:3000]
    assert int(re.search(r"ncorrect=(\d+)", r.stderr).group(1)) >= 45, tail


@pytest.mark.gpu
def test_native_cudampi_two_gpus(dataset_paths, repo_root):
    """The native launcher: fork per GPU, CUDA-IPC regions, fused NVLink update — no Python, MPI or NCCL involved."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    exe = os.path.join(repo_root, "bin", "cudampi")
    assert _run([exe], repo_root).returncode == 100
    r = _run([exe, "-np", "2", "--nepoch", "20", "--batch-size", "50", "--rate", "0.3", "--timing"] + dataset_paths, repo_root)
    assert r.returncode == 0, r.stderr
    sums = re.findall(r"rank \d param_checksum=(\S+)", r.stderr)
    assert len(sums) == 2 and sums[0] == sums[1]  # replicas bitwise identical
    assert re.search(r"^0 0 100$", r.stderr, re.M) and re.search(r"^1 100 200$", r.stderr, re.M)
    assert int(re.search(r"ncorrect=(\d+)", r.stderr).group(1)) >= 45
    assert "train_images_per_sec=" in r.stderr
