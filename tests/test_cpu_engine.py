"""CPU engine: glibc rand clone, IDX io, layer maths vs PyTorch fp64, quirks mode vs the unmodified
reference binary, CLI contract (Oracle A = cnn.c, Oracle B = PyTorch fp64)."""
import ctypes
import os
import shutil
import subprocess

import numpy as np
import pytest
import torch

from mpi_cuda_cnn_b200.models import RefCNN, dense_grads, load_dense_params
from mpi_cuda_cnn_b200.ops.cpu import QUIRKS_REFERENCE, CpuNet, glibc_rand
from mpi_cuda_cnn_b200.utils import make_synthetic, read_idx, write_idx

REF = "/root/reference/cnn.c"


def test_glibc_rand_matches_libc():
    libc = ctypes.CDLL("libc.so.6")
    for seed in (0, 1, 12345):
        libc.srand(seed)
        want = [libc.rand() for _ in range(1000)]
        assert glibc_rand(seed, 1000) == want


def test_idx_roundtrip(tmp_path):
    x, y = make_synthetic(37, seed=3)
    assert x.shape == (37, 28, 28) and y.shape == (37,) and y.max() < 10
    write_idx(tmp_path / "x", x)
    write_idx(tmp_path / "y", y)
    assert np.array_equal(read_idx(tmp_path / "x"), x)
    assert np.array_equal(read_idx(tmp_path / "y"), y)
    raw = open(tmp_path / "x", "rb").read()
    assert raw[:4] == b"\x00\x00\x08\x03" and raw[4:8] == (37).to_bytes(4, "big")
    with pytest.raises(OSError):
        read_idx(tmp_path / "missing")
    (tmp_path / "short").write_bytes(raw[:100])  # truncated payload must be an error
    with pytest.raises(OSError):
        read_idx(tmp_path / "short")


def test_init_statistics_and_layout():
    net = CpuNet.mnist(seed=0)
    assert net.num_params == 360810 and net.num_layers == 6
    p = net.get_params()
    w1 = p[:144]
    assert abs(p[: 144 + 16 + 4608].std() - 0.1) < 0.02
    assert np.all(p[144:160] == 0)  # biases start at 0
    info = net.layer_info(2)
    assert (info["depth"], info["width"], info["height"], info["nweights"], info["nbiases"]) == (32, 7, 7, 4608, 32)
    # first weight = 0.1 * nrnd() from the first four rand() draws after srand(0)
    r = [v / 2147483647 for v in glibc_rand(0, 4)]
    assert w1[0] == pytest.approx(0.1 * (sum(r) - 2.0) * 1.724, rel=1e-12)


@pytest.mark.parametrize("pool", [False, True])
def test_forward_backward_vs_torch_fp64(pool):
    from mpi_cuda_cnn_b200.engine import mnist_layers, pool_layers

    net = CpuNet.from_layers(pool_layers() if pool else mnist_layers(), seed=0)
    model = load_dense_params(RefCNN(pool=pool), net.get_params())
    x, y = make_synthetic(5, seed=7)
    xs = x.astype(np.float64) / 255.0
    for i in range(5):
        net.set_inputs(xs[i])
        onehot = np.eye(10)[y[i]]
        net.learn_outputs(onehot)
        logits = model(torch.from_numpy(xs[i])[None, None])
        assert np.allclose(net.get_outputs(), torch.softmax(logits, 1)[0].detach().numpy(), atol=1e-12)
        err = net.error_total()
        assert err == pytest.approx(float(RefCNN.mse_metric(logits.detach(), torch.tensor([int(y[i])]))), rel=1e-9)
    loss = model.loss_sum(torch.from_numpy(xs)[:, None], torch.from_numpy(y.astype(np.int64)))
    loss.backward()
    assert np.allclose(net.get_updates(), dense_grads(model), atol=1e-10)
    before = net.get_params()
    net.update(0.1 / 5)
    assert np.allclose(net.get_params(), before - 0.02 * dense_grads(model), atol=1e-12)
    assert np.all(net.get_updates() == 0)


@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("n", [1, 7, 32])
def test_learn_batch_equals_per_sample_calls_and_torch(pool, n):
    """The mini-batch schedule (net_batch.cpp) against the per-sample Layer_* calls and against autograd: gradients,
    per-sample metric, outputs; over several batches with an update in between (stale scratch state shows up on the
    second batch, not the first)."""
    from mpi_cuda_cnn_b200.engine import mnist_layers, pool_layers

    layers = pool_layers() if pool else mnist_layers()
    a, b = CpuNet.from_layers(layers, seed=3), CpuNet.from_layers(layers, seed=3)
    assert b.batch_supported
    for it in range(3):
        x, y = make_synthetic(n, seed=11 + it)
        xs = x.astype(np.float64) / 255.0
        onehot = np.eye(10)[y]
        errs_a = []
        for i in range(n):
            a.set_inputs(xs[i])
            a.learn_outputs(onehot[i])
            errs_a.append(a.error_total())
        errs_b = b.learn_batch(xs, onehot)
        assert np.allclose(errs_b, errs_a, rtol=1e-12, atol=1e-15)
        ua, ub = a.get_updates(), b.get_updates()
        assert np.allclose(ub, ua, rtol=1e-10, atol=1e-12 * max(1.0, np.abs(ua).max()))
        if it == 0:
            model = load_dense_params(RefCNN(pool=pool), b.get_params())
            model.loss_sum(torch.from_numpy(xs)[:, None], torch.from_numpy(y.astype(np.int64))).backward()
            assert np.allclose(ub, dense_grads(model), atol=1e-10)
        assert np.allclose(b.get_outputs(), a.get_outputs(), atol=1e-13)  # single-sample view = the last sample
        assert np.allclose(b.forward_batch(xs)[-1], a.get_outputs(), atol=1e-13)
        a.update(0.1 / n)
        b.update(0.1 / n)
        assert np.allclose(b.get_params(), a.get_params(), rtol=1e-10, atol=1e-12)


def test_learn_batch_falls_back_in_quirks_mode():
    from mpi_cuda_cnn_b200.engine import mnist_layers

    a, b = CpuNet.from_layers(mnist_layers(), seed=1, quirks=7), CpuNet.from_layers(mnist_layers(), seed=1, quirks=7)
    assert not b.batch_supported
    x, y = make_synthetic(3, seed=2)
    xs, onehot = x.astype(np.float64) / 255.0, np.eye(10)[y]
    for i in range(3):
        a.set_inputs(xs[i])
        a.learn_outputs(onehot[i])
    b.learn_batch(xs, onehot)
    assert np.array_equal(a.get_updates(), b.get_updates())  # same code path, bit for bit


def test_quirk_d1_shared_conv_slice():
    clean, quirky = CpuNet.mnist(seed=0), CpuNet.mnist(seed=0, quirks=QUIRKS_REFERENCE)
    x, y = make_synthetic(1, seed=9)
    for n in (clean, quirky):
        n.set_inputs(x[0].astype(np.float64) / 255.0)
        n.learn_outputs(np.eye(10)[y[0]])
    uq = quirky.get_updates()[160:160 + 4608]
    uc = clean.get_updates()[160:160 + 4608]
    assert np.count_nonzero(uc) > 1000
    assert np.all(uq[np.arange(4608) % 144 > 8] == 0)  # only slice [oc][0][:][:] is ever touched


def test_checkpoint_roundtrip(tmp_path):
    a, b = CpuNet.mnist(seed=3), CpuNet.mnist(seed=4)
    assert not np.array_equal(a.get_params(), b.get_params())
    assert a.save(tmp_path / "ck.bin") == 0 and b.load(tmp_path / "ck.bin") == 0
    assert np.array_equal(a.get_params(), b.get_params())
    assert b.load(tmp_path / "nope.bin") == 111


@pytest.fixture(scope="module")
def ref_binary(tmp_path_factory):
    if not os.path.exists(REF) or shutil.which("gcc") is None:
        pytest.skip("reference source or gcc not available")
    out = tmp_path_factory.mktemp("ref") / "cnn_ref"
    subprocess.run(["gcc", "-O2", "-o", str(out), REF, "-lm"], check=True, stderr=subprocess.DEVNULL)
    return str(out)


def test_quirks_mode_is_bit_identical_to_reference(ref_binary, dataset_paths, repo_root):
    """Oracle A: the unmodified cnn.c and our trainer in --ref-quirks mode print the same stderr."""
    ref = subprocess.run([ref_binary] + dataset_paths, stderr=subprocess.PIPE, text=True)
    ours = subprocess.run([os.path.join(repo_root, "bin", "cnn"), "--ref-quirks"] + dataset_paths, stderr=subprocess.PIPE, text=True)
    assert ref.returncode == 0 and ours.returncode == 0
    assert ours.stderr == ref.stderr
    assert "ntests=60" in ours.stderr


def test_cli_contract(dataset_paths, repo_root):
    cnn = os.path.join(repo_root, "bin", "cnn")
    assert subprocess.run([cnn], stderr=subprocess.DEVNULL).returncode == 100      # cnn.c:412
    assert subprocess.run([cnn, "a", "b", "c"], stderr=subprocess.DEVNULL).returncode == 100  # D3 fixed
    assert subprocess.run([cnn, "a", "b", "c", "d"], stderr=subprocess.DEVNULL).returncode == 111  # cnn.c:432
    r = subprocess.run([cnn, "--nepoch", "2", "--batch-size", "16", "--rate", "0.2"] + dataset_paths, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0
    lines = r.stderr.splitlines()
    assert lines[0] == "training..." and lines[1].startswith("i=0, error=") and "testing..." in lines
    ncorrect = int(lines[-1].split("ncorrect=")[1])
    assert lines[-1].startswith("ntests=60") and ncorrect >= 50  # the synthetic set is learnable


def test_serial_synthetic_flag(repo_root):
    r = subprocess.run([os.path.join(repo_root, "bin", "cnn"), "--synthetic", "150", "--nepoch", "1", "--timing"], stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0 and "train_images_per_sec=" in r.stderr
