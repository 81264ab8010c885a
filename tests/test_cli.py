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


@pytest.mark.gpu
@pytest.mark.parametrize("update", ["fused", "nccl"])
def test_cudampi_load_reaches_every_rank(dataset_paths, repo_root, tmp_path, update):
    """`cudampi --load F`: the checkpoint is read on rank 0 AFTER the trainer has broadcast the seed initialisation, so
    it must be broadcast again; with zero epochs every rank's checksum equals the checksum of the file's parameters."""
    import numpy as np
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    ck = str(tmp_path / "w.ckpt")
    r = _run(PY + ["cudacnn", "--nepoch", "2", "--batch-size", "50", "--save", ck] + dataset_paths, repo_root)
    assert r.returncode == 0 and os.path.exists(ck), r.stderr
    want = float(np.fromfile(ck, dtype=np.float32).astype(np.float64).sum())
    r = _run(PY + ["cudampi", "-np", "2", "--nepoch", "0", "--batch-size", "50", "--update", update, "--load", ck] + dataset_paths, repo_root)
    assert r.returncode == 0, r.stderr
    sums = [float(v) for v in re.findall(r"rank \d param_checksum=(\S+)", r.stderr)]
    assert len(sums) == 2 and sums[0] == sums[1], r.stderr[-2000:]
    assert abs(sums[0] - want) <= 1e-6 * max(1.0, abs(want)), (sums, want)


@pytest.mark.gpu
def test_native_cudacnn_flags_and_checkpoint_round_trip(dataset_paths, repo_root, tmp_path):
    """bin/cudacnn takes the same optional flags as the Python entrypoint: --pool, --backend, --synthetic-test, and
    --save / --load in the Python CLI's checkpoint format (dense float32), so a model moves between the two."""
    exe = os.path.join(repo_root, "bin", "cudacnn")
    ck = str(tmp_path / "native.ckpt")
    r = _run([exe, "--nepoch", "20", "--batch-size", "50", "--rate", "0.3", "--save", ck] + dataset_paths, repo_root)
    assert r.returncode == 0 and os.path.getsize(ck) == 4 * 360810, r.stderr
    trained = int(re.search(r"ncorrect=(\d+)", r.stderr).group(1))
    # zero epochs from the checkpoint: the test pass must reproduce the trained model's score, natively and in Python
    r = _run([exe, "--nepoch", "0", "--batch-size", "50", "--load", ck] + dataset_paths, repo_root)
    assert r.returncode == 0 and int(re.search(r"ncorrect=(\d+)", r.stderr).group(1)) == trained, r.stderr
    r = _run(PY + ["cudacnn", "--nepoch", "0", "--batch-size", "50", "--load", ck] + dataset_paths, repo_root)
    assert r.returncode == 0 and int(re.search(r"ncorrect=(\d+)", r.stderr).group(1)) == trained, r.stderr
    r = _run([exe, "--nepoch", "0", "--load", str(tmp_path / "missing")] + dataset_paths, repo_root)
    assert r.returncode == 111
    r = _run([exe, "--pool", "--backend", "simt", "--nepoch", "15", "--batch-size", "50", "--rate", "0.3", "--synthetic", "300",
              "--synthetic-test", "80"], repo_root)
    assert r.returncode == 0 and re.search(r"ntests=80, ncorrect=(\d+)", r.stderr), r.stderr
    assert _run([exe, "--backend", "nonsense"] + dataset_paths, repo_root).returncode == 100


@pytest.mark.gpu
def test_native_cudampi_failing_rank_does_not_hang(dataset_paths, repo_root):
    """A rank that leaves early must take the whole job down quickly: bad path (every rank fails, 111) and fewer visible
    GPUs than ranks (the launcher reports the failure instead of waiting for barriers that can never complete)."""
    import time

    exe = os.path.join(repo_root, "bin", "cudampi")
    t0 = time.time()
    r = _run([exe, "-np", "2", "a", "b", "c", "d"], repo_root)
    assert r.returncode == 111, r.stderr
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="0", B200CNN_COMM_TIMEOUT_S="20")
    r = subprocess.run([exe, "-np", "2", "--nepoch", "1"] + dataset_paths, stderr=subprocess.PIPE, text=True, env=env, timeout=120)
    assert r.returncode != 0 and "needs 2 visible GPUs" in r.stderr, r.stderr
    assert time.time() - t0 < 60
