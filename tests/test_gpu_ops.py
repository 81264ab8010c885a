"""Every CUDA kernel against a plain PyTorch fp32 reference of the same op.

Runs under both backends where the op has two implementations: 'simt' (CUDA cores, fp32 multiply) and
'tcgen05' (tensor cores, TF32 multiply, fp32 accumulate).  Tolerances: fp32 path 2e-4 relative to the
output scale, TF32 path 4e-3 (10-bit mantissa operands).
"""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

TOL = {"simt": 2e-4, "tcgen05": 4e-3, "auto": 4e-3}


@pytest.fixture(autouse=True)
def _cuda_only():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.manual_seed(0)
    # the PyTorch references are plain fp32 whatever test runs first
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    yield
    from mpi_cuda_cnn_b200.ops import functional as F200

    F200.set_backend("auto")


def _close(got, want, tol):
    scale = max(float(want.abs().max()), 1e-6)
    err = float((got.double() - want.double()).abs().max()) / scale
    assert err < tol, f"max rel err {err:.3e} (tol {tol})"


def test_library_is_native_sm100a():
    from mpi_cuda_cnn_b200 import _native as N

    info = N.cuda().b200cnn_build_info().decode()
    assert "sm_100a" in info
    assert torch.cuda.get_device_capability()[0] == 10


@pytest.mark.parametrize("cfg", [
    # n, c, h, w, oc, k, pad, stride, act
    (256, 1, 28, 28, 16, 3, 1, 2, "relu"),   # conv1 of the reference network (cnn.c:419)
    (256, 16, 14, 14, 32, 3, 1, 2, "relu"),  # conv2 (cnn.c:422)
    (7, 3, 11, 9, 5, 3, 0, 1, "none"),
    (4, 16, 14, 14, 32, 3, 1, 1, "relu"),    # pooled variant
    (3, 2, 12, 12, 20, 5, 2, 3, "tanh"),
])
def test_conv_forward_backward(cfg):
    from mpi_cuda_cnn_b200.ops import functional as F200

    n, c, h, w, oc, k, pad, stride, act = cfg
    x = torch.randn(n, c, h, w, device="cuda")
    wt = (0.2 * torch.randn(oc, c, k, k, device="cuda")).requires_grad_()
    b = (0.1 * torch.randn(oc, device="cuda")).requires_grad_()
    torch.backends.cudnn.allow_tf32 = False
    xr = x.clone().requires_grad_()
    x.requires_grad_()
    y = F200.conv2d_act(x, wt, b, stride, pad, act)
    wr, br = wt.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()
    z = F.conv2d(xr, wr, br, stride, pad)
    yr = {"relu": torch.relu, "tanh": torch.tanh, "none": lambda t: t}[act](z)
    _close(y, yr, 2e-4)
    g = torch.randn_like(yr)
    y.backward(g)
    yr.backward(g)
    _close(x.grad, xr.grad, 3e-4)
    _close(wt.grad, wr.grad, 3e-4)
    _close(b.grad, br.grad, 3e-4)


def test_conv_uint8_input_fuses_normalisation():
    from mpi_cuda_cnn_b200.ops import functional as F200

    x8 = torch.randint(0, 256, (32, 1, 28, 28), device="cuda", dtype=torch.uint8)
    wt = 0.2 * torch.randn(16, 1, 3, 3, device="cuda")
    b = 0.1 * torch.randn(16, device="cuda")
    y = F200.conv2d_act(x8, wt, b, 2, 1, "relu")
    yr = torch.relu(F.conv2d(x8.float() / 255.0, wt, b, 2, 1))
    _close(y, yr, 2e-4)


TC_SHAPES = [(256, 200, 1568), (256, 200, 200), (256, 16, 200), (1, 200, 1568), (512, 64, 96), (100, 36, 44), (300, 132, 260)]
ANY_SHAPES = TC_SHAPES + [(256, 10, 200), (33, 17, 70)]


def _linear_case(backend, mnk, act):
    from mpi_cuda_cnn_b200.ops import functional as F200

    F200.set_backend(backend)
    m, n, k = mnk
    x = torch.randn(m, k, device="cuda").requires_grad_()
    w = (torch.randn(n, k, device="cuda") / k ** 0.5).requires_grad_()
    b = (0.1 * torch.randn(n, device="cuda")).requires_grad_()
    xr, wr, br = (t.detach().clone().requires_grad_() for t in (x, w, b))
    y = F200.linear_act(x, w, b, act)
    torch.backends.cuda.matmul.allow_tf32 = False
    z = F.linear(xr, wr, br)
    yr = torch.tanh(z) if act == "tanh" else z
    tol = TOL[backend]
    _close(y, yr, tol)
    g = torch.randn_like(yr)
    y.backward(g)
    yr.backward(g)
    _close(x.grad, xr.grad, tol)
    _close(w.grad, wr.grad, tol)
    _close(b.grad, br.grad, tol)


@pytest.mark.parametrize("mnk", ANY_SHAPES)
@pytest.mark.parametrize("act", ["tanh", "none"])
def test_linear_forward_backward_simt(mnk, act):
    _linear_case("simt", mnk, act)


@pytest.mark.parametrize("mnk", TC_SHAPES)
@pytest.mark.parametrize("act", ["tanh", "none"])
def test_linear_forward_backward_tcgen05(mnk, act):
    """tcgen05 / TMEM / TMA GEMM, all three operand-major combinations (forward K/K, dX K/N, dW M/N)."""
    _linear_case("tcgen05", mnk, act)


def test_linear_auto_falls_back_for_unaligned_shapes():
    _linear_case("auto", (33, 17, 70), "tanh")
    _linear_case("auto", (256, 10, 200), "none")


def test_tcgen05_padded_logits_layout():
    """The engine's last layer: 10 classes in rows of 16 floats (ld = 16) through the C ABI."""
    import ctypes as C
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    m, n, k, ld = 256, 10, 200, 16
    x = torch.randn(m, k, device="cuda")
    w = torch.randn(n, k, device="cuda") / k ** 0.5
    b = torch.randn(n, device="cuda")
    dz = torch.zeros(m, ld, device="cuda")
    dz[:, :n] = torch.randn(m, n, device="cuda")
    y = torch.zeros(m, ld, device="cuda")
    dx, dw, db = torch.empty(m, k, device="cuda"), torch.empty(n, k, device="cuda"), torch.empty(n, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(lib.b200cnn_set_backend(N.BACKEND_TCGEN05))
    try:
        N.check(lib.b200cnn_linear_fwd(p(x), p(w), p(b), p(y), m, n, k, ld, N.ACT_NONE, st), "fwd")
        N.check(lib.b200cnn_linear_bwd_data(p(dz), p(w), p(x), N.ACT_TANH, p(dx), m, n, k, ld, st), "bwd_data")
        N.check(lib.b200cnn_linear_bwd_weight(p(dz), p(x), p(dw), p(db), 0, m, n, k, ld, st), "bwd_weight")
    finally:
        lib.b200cnn_set_backend(N.BACKEND_AUTO)
    torch.backends.cuda.matmul.allow_tf32 = False
    _close(y[:, :n], x @ w.T + b, 4e-3)
    assert float(y[:, n:].abs().max()) == 0.0
    _close(dx, (dz[:, :n] @ w) * (1 - x * x), 4e-3)
    _close(dw, dz[:, :n].T @ x, 4e-3)
    _close(db, dz[:, :n].sum(0), 1e-4)


def test_softmax_xent_and_metrics():
    from mpi_cuda_cnn_b200.ops import functional as F200

    m, n = 300, 10
    logits = (3 * torch.randn(m, n, device="cuda")).requires_grad_()
    labels = torch.randint(0, n, (m,), device="cuda", dtype=torch.uint8)
    lr = logits.detach().clone().requires_grad_()
    loss, probs, metrics = F200.softmax_xent(logits, labels)
    want = F.cross_entropy(lr, labels.long(), reduction="sum")
    _close(loss, want, 1e-4)
    _close(probs, torch.softmax(lr, 1), 1e-5)
    loss.backward()
    want.backward()
    _close(logits.grad, lr.grad, 1e-5)
    p = torch.softmax(lr.detach(), 1)
    mse = ((p - F.one_hot(labels.long(), n)) ** 2).mean(1).sum()
    correct = (p.argmax(1) == labels.long()).sum()
    got = metrics.cpu().numpy()
    assert got[1] == pytest.approx(float(mse), rel=1e-4)
    assert got[2] == float(correct) and got[3] == m


def test_maxpool_forward_backward():
    from mpi_cuda_cnn_b200.ops import functional as F200

    for (k, s, h, w) in [(2, 2, 28, 28), (3, 2, 15, 13), (3, 1, 9, 9)]:
        x = torch.randn(5, 4, h, w, device="cuda").requires_grad_()
        xr = x.detach().clone().requires_grad_()
        y = F200.max_pool2d(x, k, s)
        yr = F.max_pool2d(xr, k, s)
        assert torch.equal(y, yr)
        g = torch.randn_like(yr)
        y.backward(g)
        yr.backward(g)
        _close(x.grad, xr.grad, 1e-6)


def test_sgd_update():
    from mpi_cuda_cnn_b200.ops import functional as F200

    w = torch.randn(360812, device="cuda")
    g = torch.randn_like(w)
    want = w - 0.05 * g
    F200.sgd_update_(w, g, 0.05, zero_grad=True)
    _close(w, want, 1e-6)
    assert float(g.abs().max()) == 0.0


def test_synthetic_and_gather():
    from mpi_cuda_cnn_b200.ops import functional as F200

    x, y = F200.synthetic_batch(512, seed=3)
    x2, y2 = F200.synthetic_batch(256, seed=3, first_index=256)
    assert torch.equal(x[256:], x2) and torch.equal(y[256:], y2)  # counter-based: ranges compose
    assert int(y.max()) < 10 and len(torch.unique(y)) == 10 and float(x.float().mean()) > 5
    bx, by = F200.gather_batch(x, y, 64, seed=1, step=5, shard_begin=128, shard_len=128)
    # every gathered sample is a row of the shard with its own label
    for i in range(0, 64, 7):
        match = (x[128:256] == bx[i]).all(1).nonzero()
        assert match.numel() >= 1 and int(y[128 + int(match[0])]) == int(by[i])


@pytest.mark.parametrize("shape", [(256, 200, 200, 10), (100, 64, 128, 7), (300, 208, 36, 16), (1, 16, 16, 2)])
def test_fused_head_kernel(shape):
    """fc(tanh) -> fc(softmax) + loss gradient + both backward-data GEMMs in one tcgen05 kernel vs PyTorch."""
    import ctypes as C
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    m, n1, n2, nc = shape
    torch.backends.cuda.matmul.allow_tf32 = False
    h1 = torch.tanh(torch.randn(m, n1, device="cuda"))
    w2, b2 = torch.randn(n2, n1, device="cuda") / n1 ** 0.5, 0.1 * torch.randn(n2, device="cuda")
    w3, b3 = torch.randn(nc, n2, device="cuda") / n2 ** 0.5, 0.1 * torch.randn(nc, device="cuda")
    lab = torch.randint(0, nc, (m,), device="cuda", dtype=torch.uint8)
    h2, dz2 = torch.empty(m, n2, device="cuda"), torch.empty(m, n2, device="cuda")
    pr, dz3 = torch.empty(m, 16, device="cuda"), torch.empty(m, 16, device="cuda")
    dz1, met = torch.empty(m, n1, device="cuda"), torch.zeros(4, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(lib.b200cnn_mlp_head(p(h1), p(w2), p(b2), p(w3), p(b3), p(lab), p(h2), p(pr), p(dz3), p(dz2), p(dz1), p(met), m, n1, n2, nc, 16, st), "mlp_head")
    rh2 = torch.tanh(h1 @ w2.T + b2)
    rp = torch.softmax(rh2 @ w3.T + b3, 1)
    rdz3 = rp - F.one_hot(lab.long(), nc)
    rdz2 = (rdz3 @ w3) * (1 - rh2 ** 2)
    rdz1 = (rdz2 @ w2) * (1 - h1 ** 2)
    for got, want in ((h2, rh2), (pr[:, :nc], rp), (dz3[:, :nc], rdz3), (dz2, rdz2), (dz1, rdz1)):
        _close(got, want, 5e-3)
    assert float(pr[:, nc:].abs().max()) == 0.0 if nc < 16 else True
    got = met.cpu().numpy()
    assert got[3] == m and got[2] == float((rp.argmax(1) == lab.long()).sum())
    assert got[0] == pytest.approx(float(-torch.log(rp[torch.arange(m), lab.long()]).sum()), rel=2e-3)


@pytest.mark.parametrize("shape", [(256, 1568, 200, 200, 10), (16, 1568, 200, 200, 10), (37, 1568, 200, 200, 10),
                                   (5, 64, 40, 24, 3), (130, 800, 208, 120, 16), (64, 196, 200, 200, 10)])
@pytest.mark.parametrize("act_x", ["relu", "none"])
def test_fc_chain_kernel(shape, act_x):
    """tc_chain.cu: three FC layers forward, softmax + loss gradient + metrics and the three data-gradient GEMMs in
    one 8-CTA-cluster kernel, against plain PyTorch fp32 (TF32 off), twice in a row (barriers / TMEM are per launch)."""
    import ctypes as C
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    m, k1, n1, n2, nc = shape
    torch.backends.cuda.matmul.allow_tf32 = False
    x = torch.randn(m, k1, device="cuda")
    if act_x == "relu":
        x = torch.relu(x)
    w1, b1 = torch.randn(n1, k1, device="cuda") / k1 ** 0.5, 0.1 * torch.randn(n1, device="cuda")
    w2, b2 = torch.randn(n2, n1, device="cuda") / n1 ** 0.5, 0.1 * torch.randn(n2, device="cuda")
    w3, b3 = torch.randn(nc, n2, device="cuda") / n2 ** 0.5, 0.1 * torch.randn(nc, device="cuda")
    lab = torch.randint(0, nc, (m,), device="cuda", dtype=torch.uint8)
    rh1 = torch.tanh(x @ w1.T + b1)
    rh2 = torch.tanh(rh1 @ w2.T + b2)
    rp = torch.softmax(rh2 @ w3.T + b3, 1)
    rdz3 = rp - F.one_hot(lab.long(), nc)
    rdz2 = (rdz3 @ w3) * (1 - rh2 ** 2)
    rdz1 = (rdz2 @ w2) * (1 - rh1 ** 2)
    rdx = rdz1 @ w1
    if act_x == "relu":
        rdx = rdx * (x > 0)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rep in range(2):
        h1, dz1 = torch.full((m, n1), 9.0, device="cuda"), torch.full((m, n1), 9.0, device="cuda")
        h2, dz2 = torch.full((m, n2), 9.0, device="cuda"), torch.full((m, n2), 9.0, device="cuda")
        pr, dz3 = torch.full((m, 16), 9.0, device="cuda"), torch.full((m, 16), 9.0, device="cuda")
        dx, met = torch.full((m, k1), 9.0, device="cuda"), torch.zeros(4, device="cuda")
        N.check(lib.b200cnn_fc_chain(p(x), N.ACT_RELU if act_x == "relu" else N.ACT_NONE, p(w1), p(b1), p(w2), p(b2), p(w3), p(b3), p(lab),
                                     p(h1), p(h2), p(pr), p(dz3), p(dz2), p(dz1), p(dx), p(met), m, k1, n1, n2, nc, 16, st), "fc_chain")
        torch.cuda.synchronize()
        for name, got, want in (("h1", h1, rh1), ("h2", h2, rh2), ("probs", pr[:, :nc], rp), ("dz3", dz3[:, :nc], rdz3),
                                ("dz2", dz2, rdz2), ("dz1", dz1, rdz1), ("dx", dx, rdx)):
            err = float((got - want).abs().max() / want.abs().max().clamp_min(1e-6))
            assert err < 6e-3, (name, rep, err)
        if nc < 16:
            assert float(pr[:, nc:].abs().max()) == 0.0 and float(dz3[:, nc:].abs().max()) == 0.0
        got = met.cpu().numpy()
        assert got[3] == m
        assert abs(got[2] - float((rp.argmax(1) == lab.long()).sum())) <= 1  # a TF32 near-tie may flip one argmax
        assert got[0] == pytest.approx(float(-torch.log(rp[torch.arange(m), lab.long()]).sum()), rel=3e-3)
        assert got[1] == pytest.approx(float((rdz3 ** 2).sum() / nc), rel=3e-3)


@pytest.mark.parametrize("n", [8, 72, 256, 20])
@pytest.mark.parametrize("backend", ["simt", "auto"])
def test_nhwc_conv2_weight_gradient(n, backend):
    """conv2d_bwd_weight_nhwc (16 -> 32 channels, 14x14 -> 7x7): the tcgen05 split-K implicit GEMM (batch % 8 == 0,
    backend != simt) and the CUDA-core kernel, dz in CHW and NHWC layout, twice in a row on one scratch buffer, with
    and without accumulation, against autograd."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N
    from mpi_cuda_cnn_b200.ops import functional as F200

    lib = N.cuda()
    F200.set_backend(backend)
    p = lambda t: C.c_void_p(t.data_ptr())
    stream = torch.cuda.Stream()
    st = C.c_void_p(stream.cuda_stream)
    dz = torch.randn(n, 32, 7, 7, device="cuda")
    a1 = torch.randn(n, 16, 14, 14, device="cuda")
    dz_nhwc, a1_nhwc = dz.permute(0, 2, 3, 1).contiguous(), a1.permute(0, 2, 3, 1).contiguous()
    w = torch.zeros(32, 16, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(a1, w, None, 2, 1).backward(dz)
    want_w, want_b = w.grad, dz.sum((0, 2, 3))
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp)), dtype=torch.uint8, device="cuda")
    tol = TOL["simt" if backend == "simt" else "tcgen05"]
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        for chw, src, acc in ((1, dz, 0), (0, dz_nhwc, 0), (1, dz, 1)):
            dw = torch.full((32, 16, 3, 3), 7.0, device="cuda")
            db = torch.full((32,), 7.0, device="cuda")
            N.check(lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1_nhwc), 0, p(src), chw, p(dw), p(db), acc, p(scratch), C.byref(shp), st), "bwd_weight")
            stream.synchronize()
            _close(dw, want_w + 7.0 * acc, tol)
            _close(db, want_b + 7.0 * acc, tol)


@pytest.mark.parametrize("n", [1, 3, 255, 256])
@pytest.mark.parametrize("act", ["relu", "none"])
def test_tcgen05_conv3x3_forward_direct(n, act):
    """b200cnn_conv3x3_fwd_nhwc (the engine's tcgen05 implicit GEMM for the second convolution, 16 -> 32 channels,
    14x14 -> 7x7, stride 2) against F.conv2d in fp32 with TF32 off.  NHWC input, CHW-flattened output with row stride
    ldy; guard columns behind every output row must stay untouched."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    x = torch.randn(n, 16, 14, 14, device="cuda")
    w, b = 0.2 * torch.randn(32, 16, 3, 3, device="cuda"), 0.1 * torch.randn(32, device="cuda")
    want = F.conv2d(x, w, b, 2, 1)
    if act == "relu":
        want = torch.relu(want)
    want = want.reshape(n, -1)
    ldy = 32 * 49 + 8
    y = torch.full((n, ldy), 7.0, device="cuda")
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(lib.b200cnn_conv3x3_fwd_nhwc(p(x.permute(0, 2, 3, 1).contiguous()), p(w), p(b), p(y), ldy, C.byref(shp),
                                         N.ACT_RELU if act == "relu" else N.ACT_NONE, st), "conv3x3_fwd_nhwc")
    torch.cuda.synchronize()
    _close(y[:, :32 * 49], want, 4e-3)
    assert bool((y[:, 32 * 49:] == 7.0).all())


@pytest.mark.parametrize("n", [1, 3, 255, 256])
@pytest.mark.parametrize("dz_is_chw", [0, 1])
@pytest.mark.parametrize("mask", [True, False])
def test_tcgen05_conv3x3s2_backward_data_direct(n, dz_is_chw, mask):
    """b200cnn_conv3x3s2_bwd_data_nhwc (32 -> 16 channels, 7x7 -> 14x14) against conv_transpose2d in fp32 (TF32 off),
    both gradient layouts, with and without the ReLU mask of the layer below."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dz = torch.randn(n, 32, 7, 7, device="cuda")
    w = 0.2 * torch.randn(32, 16, 3, 3, device="cuda")
    a_prev = torch.relu(torch.randn(n, 16, 14, 14, device="cuda"))
    want = F.conv_transpose2d(dz, w, None, 2, 1, output_padding=1)
    if mask:
        want = want * (a_prev > 0)
    dx = torch.full((n, 14, 14, 16), 7.0, device="cuda")
    src = dz.contiguous() if dz_is_chw else dz.permute(0, 2, 3, 1).contiguous()
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ap = a_prev.permute(0, 2, 3, 1).contiguous()
    N.check(lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(src), dz_is_chw, p(w), p(ap) if mask else None, N.ACT_RELU if mask else N.ACT_NONE, p(dx),
                                                C.byref(shp), st), "conv3x3s2_bwd_data_nhwc")
    torch.cuda.synchronize()
    _close(dx.permute(0, 3, 1, 2), want, 4e-3)


@pytest.mark.parametrize("n", [1, 2, 7, 64, 255, 256])
@pytest.mark.parametrize("tc", ["1", "0"])
def test_conv1_weight_gradient_direct(n, tc, monkeypatch):
    """Weight / bias gradient of the first convolution (1 -> 16 channels, uint8 image) through
    b200cnn_conv2d_bwd_weight_nhwc: the tcgen05 kernel (default) and the CUDA-core kernel (B200CNN_TC_C1=0 is read once
    per process, so the second variant runs in a child process), against autograd; canaries around dw / db."""
    import ctypes as C
    import subprocess
    import sys

    if tc == "0":
        code = ("import os,sys; os.environ['B200CNN_TC_C1']='0'; sys.path.insert(0, %r); import pytest; "
                "sys.exit(pytest.main(['-q','-x','-m','gpu','-p','no:cacheprovider','%s::test_conv1_weight_gradient_direct[1-%d]']))"
                % (ROOT, __file__, n))
        if os.environ.get("B200CNN_TC_C1") == "0":
            pytest.skip("already the CUDA-core child")
        res = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert res.returncode == 0, res.stdout[-2000:]
        return
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    torch.backends.cudnn.allow_tf32 = False
    x = torch.randint(0, 256, (n, 1, 28, 28), device="cuda", dtype=torch.uint8)
    dz = torch.randn(n, 16, 14, 14, device="cuda")
    w = torch.zeros(16, 1, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(x.float() / 255.0, w, None, 2, 1).backward(dz)
    want_w, want_b = w.grad, dz.sum((0, 2, 3))
    shp = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp)), dtype=torch.uint8, device="cuda")
    buf = torch.full((16 * 9 + 16 + 64,), 7.0, device="cuda")
    dw, db = buf[16:16 + 144], buf[16 + 144 + 16:16 + 144 + 32]
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rep in range(2):  # the completion counter in the scratch header must reset itself
        N.check(lib.b200cnn_conv2d_bwd_weight_nhwc(p(x), 1, p(dz.permute(0, 2, 3, 1).contiguous()), 0, p(dw), p(db), 0, p(scratch),
                                                   C.byref(shp), st), "conv2d_bwd_weight_nhwc")
        torch.cuda.synchronize()
        _close(dw.reshape(16, 1, 3, 3), want_w, 4e-3)
        _close(db, want_b, 4e-3)
    assert bool((buf[:16] == 7.0).all()) and bool((buf[16 + 144:16 + 144 + 16] == 7.0).all()) and bool((buf[16 + 144 + 32:] == 7.0).all())


@pytest.mark.parametrize("shape", [(256, 200, 200, 10), (16, 200, 200, 10), (37, 200, 200, 10), (5, 40, 24, 3), (130, 208, 120, 16), (300, 64, 200, 7)])
def test_fc_mid_kernel(shape):
    """The middle of the FC chain as its own cluster kernel (b200cnn_fc_mid: fc -> softmax/loss -> two data-gradient
    GEMMs from h1) against PyTorch fp32 (TF32 off), twice in a row."""
    import ctypes as C
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    m, n1, n2, nc = shape
    torch.backends.cuda.matmul.allow_tf32 = False
    h1 = torch.tanh(torch.randn(m, n1, device="cuda"))
    w2, b2 = torch.randn(n2, n1, device="cuda") / n1 ** 0.5, 0.1 * torch.randn(n2, device="cuda")
    w3, b3 = torch.randn(nc, n2, device="cuda") / n2 ** 0.5, 0.1 * torch.randn(nc, device="cuda")
    lab = torch.randint(0, nc, (m,), device="cuda", dtype=torch.uint8)
    rh2 = torch.tanh(h1 @ w2.T + b2)
    rp = torch.softmax(rh2 @ w3.T + b3, 1)
    rdz3 = rp - F.one_hot(lab.long(), nc)
    rdz2 = (rdz3 @ w3) * (1 - rh2 ** 2)
    rdz1 = (rdz2 @ w2) * (1 - h1 ** 2)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rep in range(2):
        h2, dz2 = torch.full((m, n2), 9.0, device="cuda"), torch.full((m, n2), 9.0, device="cuda")
        pr, dz3 = torch.full((m, 16), 9.0, device="cuda"), torch.full((m, 16), 9.0, device="cuda")
        dz1, met = torch.full((m, n1), 9.0, device="cuda"), torch.zeros(4, device="cuda")
        N.check(lib.b200cnn_fc_mid(p(h1), p(w2), p(b2), p(w3), p(b3), p(lab), p(h2), p(pr), p(dz3), p(dz2), p(dz1), p(met), m, n1, n2, nc, 16, st), "fc_mid")
        torch.cuda.synchronize()
        for name, got, want in (("h2", h2, rh2), ("probs", pr[:, :nc], rp), ("dz3", dz3[:, :nc], rdz3), ("dz2", dz2, rdz2), ("dz1", dz1, rdz1)):
            err = float((got - want).abs().max() / want.abs().max().clamp_min(1e-6))
            assert err < 6e-3, (name, rep, err)
        got = met.cpu().numpy()
        assert got[3] == m and abs(got[2] - float((rp.argmax(1) == lab.long()).sum())) <= 1
        assert got[0] == pytest.approx(float(-torch.log(rp[torch.arange(m), lab.long()]).sum()), rel=3e-3)


@pytest.mark.parametrize("n", [1, 3, 64])
@pytest.mark.parametrize("act", ["relu", "none"])
def test_tcgen05_conv3x3_stride1_banded_nhwc_out(n, act):
    """The pooled topology's second convolution (16 -> 32 channels at 14x14, stride 1): 196 output pixels per image are
    cut into two 7-row bands of the tcgen05 implicit GEMM; output NHWC, then also the CHW-flattened form."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    torch.backends.cudnn.allow_tf32 = False
    x = torch.randn(n, 16, 14, 14, device="cuda")
    w, b = 0.2 * torch.randn(32, 16, 3, 3, device="cuda"), 0.1 * torch.randn(32, device="cuda")
    want = F.conv2d(x, w, b, 1, 1)
    if act == "relu":
        want = torch.relu(want)
    shp = N.ConvShape(n, 16, 14, 14, 32, 14, 14, 3, 1, 1)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    code = N.ACT_RELU if act == "relu" else N.ACT_NONE
    xn = x.permute(0, 2, 3, 1).contiguous()
    # (the one-load kernel computes 16 columns per image row and two rows past the image: none of that may be stored)
    ybuf = torch.full((n * 14 * 14 * 32 + 512,), 7.0, device="cuda")
    y = ybuf[:n * 14 * 14 * 32].view(n, 14, 14, 32)
    N.check(lib.b200cnn_conv3x3_fwd_nhwc_to_nhwc(p(xn), p(w), p(b), p(y), C.byref(shp), code, st), "conv3x3_fwd_nhwc_to_nhwc")
    torch.cuda.synchronize()
    _close(y.permute(0, 3, 1, 2), want, 4e-3)
    assert bool((ybuf[n * 14 * 14 * 32:] == 7.0).all())
    ldy = 32 * 196 + 16
    y2 = torch.full((n, ldy), 7.0, device="cuda")
    N.check(lib.b200cnn_conv3x3_fwd_nhwc(p(xn), p(w), p(b), p(y2), ldy, C.byref(shp), code, st), "conv3x3_fwd_nhwc")
    torch.cuda.synchronize()
    _close(y2[:, :32 * 196].reshape(n, 32, 14, 14), want, 4e-3)
    assert bool((y2[:, 32 * 196:] == 7.0).all())


@pytest.mark.parametrize("n", [1, 3, 64])
@pytest.mark.parametrize("mask", [True, False])
def test_tcgen05_conv3x3_stride1_backward_data(n, mask):
    """b200cnn_conv3x3s1_bwd_data_nhwc (32 -> 16 channels at 14x14, stride 1) against conv_transpose2d in fp32."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    torch.backends.cudnn.allow_tf32 = False
    dz = torch.randn(n, 32, 14, 14, device="cuda")
    w = 0.2 * torch.randn(32, 16, 3, 3, device="cuda")
    a_prev = torch.relu(torch.randn(n, 16, 14, 14, device="cuda"))
    want = F.conv_transpose2d(dz, w, None, 1, 1)
    if mask:
        want = want * (a_prev > 0)
    dxbuf = torch.full((n * 14 * 14 * 16 + 512,), 7.0, device="cuda")
    dx = dxbuf[:n * 14 * 14 * 16].view(n, 14, 14, 16)
    shp = N.ConvShape(n, 16, 14, 14, 32, 14, 14, 3, 1, 1)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ap = a_prev.permute(0, 2, 3, 1).contiguous()
    N.check(lib.b200cnn_conv3x3s1_bwd_data_nhwc(p(dz.permute(0, 2, 3, 1).contiguous()), p(w), p(ap) if mask else None,
                                                N.ACT_RELU if mask else N.ACT_NONE, p(dx), C.byref(shp), st), "conv3x3s1_bwd_data_nhwc")
    torch.cuda.synchronize()
    _close(dx.permute(0, 3, 1, 2), want, 4e-3)
    assert bool((dxbuf[n * 14 * 14 * 16:] == 7.0).all())  # nothing stored for the padded columns / rows past the image


@pytest.mark.parametrize("shape", [(5, 16, 28, 28), (3, 32, 14, 14)])
@pytest.mark.parametrize("chw", [0, 1])
@pytest.mark.parametrize("ks", [(2, 2), (3, 2)])
def test_maxpool_nhwc_forward_backward(shape, chw, ks):
    """NHWC max pooling: NHWC or CHW-flattened output / incoming gradient, ReLU mask of the layer below.  2x2 / stride 2
    takes the per-sample and scatter kernels, 3x3 / stride 2 (overlapping windows) the general gather kernels."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    n, c, h, w = shape
    x = torch.relu(torch.randn(n, c, h, w, device="cuda")).requires_grad_(True)
    k, stride = ks
    ref = F.max_pool2d(x, k, stride)
    dy = torch.randn_like(ref)
    ref.backward(dy)
    want_dx = x.grad * (x > 0)
    xn = x.detach().permute(0, 2, 3, 1).contiguous()
    oh, ow = (h - k) // stride + 1, (w - k) // stride + 1
    y = torch.empty(n * c * oh * ow, device="cuda")
    arg = torch.empty(n * oh * ow * c, dtype=torch.int32, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(lib.b200cnn_maxpool_nhwc_fwd(p(xn), p(y), p(arg), n, c, h, w, k, stride, chw, st), "maxpool_nhwc_fwd")
    got = y.reshape(n, c, oh, ow) if chw else y.reshape(n, oh, ow, c).permute(0, 3, 1, 2)
    assert torch.equal(got, ref.detach())
    dyl = dy.contiguous() if chw else dy.permute(0, 2, 3, 1).contiguous()
    dx = torch.empty(n, h, w, c, device="cuda")
    N.check(lib.b200cnn_maxpool_nhwc_bwd(p(dyl), chw, p(arg), p(xn), N.ACT_RELU, p(dx), n, c, h, w, k, stride, st), "maxpool_nhwc_bwd")
    torch.cuda.synchronize()
    assert float((dx.permute(0, 3, 1, 2) - want_dx).abs().max()) < 1e-6


@pytest.mark.parametrize("n", [1, 7, 256, 300])
@pytest.mark.parametrize("u8", [1, 0])
def test_conv1_pool_fused_forward_and_weight_gradient(n, u8):
    """First layer of the pooled topology as one kernel (conv 1 -> 16, 3x3 / s1 / p1 + ReLU + 2x2 max pool) and its weight
    gradient taken from the POOLED gradient, against PyTorch fp32 autograd through conv -> relu -> max_pool."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    torch.backends.cudnn.allow_tf32 = False
    xi = torch.randint(0, 256, (n, 1, 28, 28), device="cuda", dtype=torch.uint8)
    xf = xi.float() / 255.0
    x_dev = xi if u8 else xf.contiguous()
    w = (0.3 * torch.randn(16, 1, 3, 3, device="cuda")).requires_grad_(True)
    b = (0.1 * torch.randn(16, device="cuda")).requires_grad_(True)
    a = torch.relu(F.conv2d(xf, w, b, 1, 1))
    yp_ref = F.max_pool2d(a, 2, 2)
    dyp = torch.randn_like(yp_ref)
    yp_ref.backward(dyp)
    shp = N.ConvShape(n, 1, 28, 28, 16, 28, 28, 3, 1, 1)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for keep in (1, 0):
        a_full = torch.full((n, 28, 28, 16), 9.0, device="cuda")
        yp = torch.full((n, 14, 14, 16), 9.0, device="cuda")
        arg = torch.full((n, 14, 14, 16), -5, dtype=torch.int32, device="cuda")
        N.check(lib.b200cnn_conv1_pool_fwd(p(x_dev), u8, p(w.detach()), p(b.detach()), p(a_full) if keep else None, p(yp), p(arg),
                                           C.byref(shp), N.ACT_RELU, st), "conv1_pool_fwd")
        torch.cuda.synchronize()
        _close(yp.permute(0, 3, 1, 2), yp_ref.detach(), 1e-5)
        if keep:
            _close(a_full.permute(0, 3, 1, 2), a.detach(), 1e-5)
            # the recorded winner holds the pooled value and lies inside its own window
            win = torch.gather(a_full.reshape(n, 784, 16), 1, arg.reshape(n, 196, 16).long())
            assert torch.equal(win, yp.reshape(n, 196, 16))
            q = torch.arange(196, device="cuda").reshape(1, 196, 1)
            ay, ax = arg.reshape(n, 196, 16) // 28, arg.reshape(n, 196, 16) % 28
            assert bool(((ay // 2) == (q // 14)).all()) and bool(((ax // 2) == (q % 14)).all())
        else:
            assert bool((a_full == 9.0).all())
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp)), dtype=torch.uint8, device="cuda")
    buf = torch.full((16 * 9 + 16 + 64,), 7.0, device="cuda")
    dw, db = buf[16:16 + 144], buf[16 + 144 + 16:16 + 144 + 32]
    dyn = dyp.permute(0, 2, 3, 1).contiguous()
    for rep in range(2):  # the completion counter in the scratch header must reset itself
        N.check(lib.b200cnn_conv1_pool_bwd_weight(p(x_dev), u8, p(dyn), p(yp), p(arg), N.ACT_RELU, p(dw), p(db), 0, p(scratch),
                                                  C.byref(shp), st), "conv1_pool_bwd_weight")
        torch.cuda.synchronize()
        _close(dw.reshape(16, 1, 3, 3), w.grad, 1e-4)
        _close(db, b.grad, 1e-4)
    assert bool((buf[:16] == 7.0).all()) and bool((buf[16 + 144:16 + 144 + 16] == 7.0).all()) and bool((buf[16 + 144 + 32:] == 7.0).all())


@pytest.mark.parametrize("n", [1, 3, 4, 66, 256])
@pytest.mark.parametrize("backend", ["simt", "auto"])
def test_nhwc_conv2_stride1_weight_gradient(n, backend):
    """conv2d_bwd_weight_nhwc for the pooled topology's second convolution (16 -> 32 channels at 14x14, stride 1): the
    tcgen05 split-K implicit GEMM over even / odd pixel pairs (any batch: images past the end are TMA zero fill) and the
    CUDA-core kernel, twice on one scratch buffer, with and without accumulation, against autograd."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N
    from mpi_cuda_cnn_b200.ops import functional as F200

    lib = N.cuda()
    F200.set_backend(backend)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.backends.cudnn.allow_tf32 = False
    dz = torch.randn(n, 32, 14, 14, device="cuda")
    a1 = torch.randn(n, 16, 14, 14, device="cuda")
    dz_nhwc, a1_nhwc = dz.permute(0, 2, 3, 1).contiguous(), a1.permute(0, 2, 3, 1).contiguous()
    w = torch.zeros(32, 16, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(a1, w, None, 1, 1).backward(dz)
    want_w, want_b = w.grad, dz.sum((0, 2, 3))
    shp = N.ConvShape(n, 16, 14, 14, 32, 14, 14, 3, 1, 1)
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp)), dtype=torch.uint8, device="cuda")
    tol = TOL["simt" if backend == "simt" else "tcgen05"]
    try:
        for acc in (0, 0, 1):
            buf = torch.full((64 + 4608 + 64 + 32 + 64,), 7.0, device="cuda")  # canaries in front of, between and behind dw / db
            dw, db = buf[64:64 + 4608].view(32, 16, 3, 3), buf[64 + 4608 + 64:64 + 4608 + 64 + 32]
            N.check(lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1_nhwc), 0, p(dz_nhwc), 0, p(dw), p(db), acc, p(scratch), C.byref(shp), st), "bwd_weight")
            torch.cuda.synchronize()
            _close(dw, want_w + 7.0 * acc, tol)
            _close(db, want_b + 7.0 * acc, tol)
            assert bool((buf[:64] == 7.0).all()) and bool((buf[64 + 4608:64 + 4608 + 64] == 7.0).all()) and bool((buf[-64:] == 7.0).all())
    finally:
        F200.set_backend("auto")


@pytest.mark.parametrize("n", [1, 3, 255, 256])
def test_conv12_fused_forward_direct(n):
    """b200cnn_conv12_fwd: both convolutions of the reference topology in one kernel (conv1 on the CUDA cores, written to
    global memory AND into the tcgen05 operand tiles of conv2) against F.conv2d in fp32; guard columns behind every
    output row and behind a1 must stay untouched."""
    import ctypes as C

    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    x = torch.randint(0, 256, (n, 1, 28, 28), device="cuda", dtype=torch.uint8)
    w1, b1 = 0.3 * torch.randn(16, 1, 3, 3, device="cuda"), 0.1 * torch.randn(16, device="cuda")
    w2, b2 = 0.2 * torch.randn(32, 16, 3, 3, device="cuda"), 0.1 * torch.randn(32, device="cuda")
    a1_ref = torch.relu(F.conv2d(x.float() / 255.0, w1, b1, 2, 1))
    a2_ref = torch.relu(F.conv2d(a1_ref, w2, b2, 2, 1))
    s1 = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
    s2 = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    ldy = 32 * 49 + 16
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rep in range(2):
        a1 = torch.full((n * 14 * 14 * 16 + 64,), 7.0, device="cuda")
        y = torch.full((n, ldy), 7.0, device="cuda")
        N.check(lib.b200cnn_conv12_fwd(p(x), p(w1), p(b1), p(a1), p(w2), p(b2), p(y), ldy, C.byref(s1), C.byref(s2), st), "conv12_fwd")
        torch.cuda.synchronize()
        _close(a1[:n * 3136].reshape(n, 14, 14, 16).permute(0, 3, 1, 2), a1_ref, 1e-5)
        assert bool((a1[n * 3136:] == 7.0).all())
        _close(y[:, :1568].reshape(n, 32, 7, 7), a2_ref, 4e-3)
        assert bool((y[:, 1568:] == 7.0).all())
