"""Out-of-bounds canaries.  compute-sanitizer is closed on the GPU pool, so every tensor-core op is run with
its outputs embedded between guard regions and with awkward (non-multiple-of-tile) sizes; a stray store shows up
as a changed guard word."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
GUARD, CANARY = 256, 12345.0


@pytest.fixture(autouse=True)
def _cuda_only():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def guarded(*shape):
    n = 1
    for s in shape:
        n *= s
    buf = torch.full((n + 2 * GUARD,), CANARY, device="cuda")
    return buf, buf[GUARD:GUARD + n].view(*shape)


def intact(buf):
    return bool((buf[:GUARD] == CANARY).all()) and bool((buf[-GUARD:] == CANARY).all())


def p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


@pytest.mark.parametrize("mnk", [(256, 200, 1568), (130, 36, 44), (1, 200, 1568), (257, 16, 200), (100, 132, 260)])
def test_linear_ops_stay_in_bounds(mnk):
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    m, n, k = mnk
    x, w, b, dz = torch.randn(m, k, device="cuda"), torch.randn(n, k, device="cuda"), torch.randn(n, device="cuda"), torch.randn(m, n, device="cuda")
    by, y = guarded(m, n)
    bdx, dx = guarded(m, k)
    bdw, dw = guarded(n, k)
    bdb, db = guarded(n)
    N.check(lib.b200cnn_set_backend(N.BACKEND_TCGEN05))
    try:
        N.check(lib.b200cnn_linear_fwd(p(x), p(w), p(b), p(y), m, n, k, n, N.ACT_TANH, st), "fwd")
        N.check(lib.b200cnn_linear_bwd_data(p(dz), p(w), p(x), N.ACT_TANH, p(dx), m, n, k, n, st), "bwd_data")
        N.check(lib.b200cnn_linear_bwd_weight(p(dz), p(x), p(dw), p(db), 0, m, n, k, n, st), "bwd_weight")
    finally:
        lib.b200cnn_set_backend(N.BACKEND_AUTO)
    torch.cuda.synchronize()
    for buf, view in ((by, y), (bdx, dx), (bdw, dw), (bdb, db)):
        assert intact(buf)
        assert not bool((view == CANARY).any())  # and every output word was written


@pytest.mark.parametrize("n", [1, 2, 3, 255, 256])
def test_conv_chain_ops_stay_in_bounds(n):
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    s1 = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
    s2 = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    x8 = torch.randint(0, 256, (n, 784), device="cuda", dtype=torch.uint8)
    a1, w2, b2 = torch.randn(n, 14, 14, 16, device="cuda"), torch.randn(32, 16, 3, 3, device="cuda"), torch.randn(32, device="cuda")
    dz2 = torch.randn(n, 1568, device="cuda")
    ba2, a2 = guarded(n, 1568)
    bdz1, dz1 = guarded(n, 14, 14, 16)
    bdw2, dw2 = guarded(32, 144)
    bdb2, db2 = guarded(32)
    bdw1, dw1 = guarded(16, 9)
    bdb1, db1 = guarded(16)
    scr = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(s2)), dtype=torch.uint8, device="cuda")
    N.check(lib.b200cnn_conv3x3_fwd_nhwc(p(a1), p(w2), p(b2), p(a2), 1568, C.byref(s2), N.ACT_RELU, st), "conv fwd")
    N.check(lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz2), 1, p(w2), p(a1), N.ACT_RELU, p(dz1), C.byref(s2), st), "conv bwd data")
    N.check(lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1), 0, p(dz2), 1, p(dw2), p(db2), 0, p(scr), C.byref(s2), st), "conv2 bwd weight")
    dz1c = dz1.contiguous()
    N.check(lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1c), 0, p(dw1), p(db1), 0, p(scr), C.byref(s1), st), "conv1 bwd weight")
    torch.cuda.synchronize()
    for buf, view in ((ba2, a2), (bdz1, dz1), (bdw2, dw2), (bdb2, db2), (bdw1, dw1), (bdb1, db1)):
        assert intact(buf)
        assert not bool((view == CANARY).any())
