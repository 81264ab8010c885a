"""The CPU engine and both CPU trainers under AddressSanitizer + UndefinedBehaviorSanitizer (`make asan`).

The reference ships no sanitizer build and checks no return code; its IDX reader can be driven
past its buffers by a short file (cnn.c:378-379).  Here every code path of the serial trainer (clean mode, reference
quirks mode, per-sample and mini-batch schedules, pooling topology, checkpoints, a truncated input file) and of the
data-parallel trainer (2 forked ranks over shared memory) must run clean: any report fails the test.
GPU kernels are covered separately (canary tests in test_gpu_guards.py; compute-sanitizer is not usable on the pool).
"""
import os
import subprocess

import pytest


@pytest.fixture(scope="module")
def asan_bins(repo_root):
    probe = subprocess.run(["make", "-C", repo_root, "-s", "--eval", "print-sancxx: ; @echo $(SANCXX)", "print-sancxx"], stdout=subprocess.PIPE, text=True)
    if not probe.stdout.strip():
        pytest.skip("no compiler with the AddressSanitizer runtime on this machine")
    res = subprocess.run(["make", "-C", repo_root, "-j", str(os.cpu_count() or 4), "asan"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 0, res.stdout[-4000:]
    return os.path.join(repo_root, "bin", "cnn_asan"), os.path.join(repo_root, "bin", "cnnmpi_asan")


def _run(cmd, expect=0):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0:exitcode=99", UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1:exitcode=98")
    r = subprocess.run(cmd, stderr=subprocess.PIPE, stdout=subprocess.PIPE, text=True, env=env, timeout=900)
    assert "AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr and "LeakSanitizer" not in r.stderr, r.stderr[-4000:]
    assert r.returncode == expect, (r.returncode, r.stderr[-2000:])
    return r


def test_serial_trainer_clean_under_asan_ubsan(asan_bins, dataset_paths, tmp_path):
    cnn, _ = asan_bins
    ck = str(tmp_path / "w.ckpt")
    _run([cnn, "--nepoch", "1", "--batch-size", "8", "--save", ck] + dataset_paths)           # mini-batch schedule
    _run([cnn, "--nepoch", "1", "--batch-size", "8", "--per-sample", "--load", ck] + dataset_paths)
    _run([cnn, "--nepoch", "1", "--ref-quirks"] + dataset_paths)                             # the reference's exact loop
    _run([cnn, "--nepoch", "1", "--pool", "--batch-size", "8"] + dataset_paths)
    _run([cnn, "--synthetic", "64", "--synthetic-test", "16", "--nepoch", "1"])
    assert _run([cnn], expect=100).returncode == 100


def test_truncated_idx_file_is_an_error_not_an_overrun(asan_bins, dataset_paths, tmp_path):
    cnn, _ = asan_bins
    short = tmp_path / "short-images"
    short.write_bytes(open(dataset_paths[0], "rb").read()[:1000])  # header promises 200 images, payload holds one
    _run([cnn, "--nepoch", "1", str(short)] + dataset_paths[1:], expect=111)
    (tmp_path / "garbage").write_bytes(b"\x00\x00\x08\x03" + b"\xff" * 12)  # absurd dimensions
    _run([cnn, "--nepoch", "1", str(tmp_path / "garbage")] + dataset_paths[1:], expect=111)


def test_data_parallel_trainer_clean_under_asan_ubsan(asan_bins, dataset_paths):
    _, cnnmpi = asan_bins
    r = _run([cnnmpi, "-np", "2", "--nepoch", "1", "--batch-size", "8"] + dataset_paths)
    assert r.stderr.count("param_checksum") == 2
    _run([cnnmpi, "-np", "2", "--nepoch", "1", "--per-sample", "--batch-size", "4"] + dataset_paths)
    _run([cnnmpi, "-np", "2", "a", "b", "c", "d"], expect=111)
