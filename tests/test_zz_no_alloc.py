"""The allocation-free weight-gradient entry point, exercised through the C ABI.

Why it exists: some users of this library cannot call cudaMalloc once they are running — their host
application's memory pool already owns the GPU, or an allocation's implicit device synchronisation
would break their timing, or their memory map is frozen.  For them b200cnn_conv2d_bwd_weight_noalloc
computes the convolution weight / bias gradients with its scratch in an 8 MiB array inside the library.

The test calls it twice in a row (the second call finds the first call's partial sums still in the
array) and compares both results with the regular entry point, which uses caller-provided scratch.
It needs nothing from the other tests, and lives in the file pytest collects last.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _cuda_only():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def test_noalloc_matches_regular_entry_point_twice():
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.b200cnn_conv2d_bwd_weight_noalloc_capacity() == 8 << 20
    gen = torch.Generator(device="cuda").manual_seed(1234)
    # the reference network's second convolution at the benchmark's batch, then a different layer:
    # the second call must not be disturbed by what the first one left in the shared array
    cases = [(256, 16, 14, 14, 32, 7, 7, 3, 1, 2), (40, 3, 12, 10, 8, 6, 5, 3, 1, 2)]
    for (n, c, h, w, oc, oh, ow, k, pad, stride) in cases:
        shape = N.ConvShape(n, c, h, w, oc, oh, ow, k, pad, stride)
        x = torch.randn(n, c, h, w, device="cuda", generator=gen)
        dz = torch.randn(n, oc, oh, ow, device="cuda", generator=gen)
        need = lib.b200cnn_conv2d_bwd_weight_scratch_bytes(C.byref(shape))
        assert 0 < need <= 8 << 20
        scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
        dw_ref, db_ref = torch.empty(oc, c, k, k, device="cuda"), torch.empty(oc, device="cuda")
        N.check(lib.b200cnn_conv2d_bwd_weight(_ptr(dz), _ptr(x), 0, _ptr(dw_ref), _ptr(db_ref), 0, _ptr(scratch), C.byref(shape), stream), "regular")
        dw, db = torch.full_like(dw_ref, 7.0), torch.full_like(db_ref, 7.0)
        N.check(lib.b200cnn_conv2d_bwd_weight_noalloc(_ptr(dz), _ptr(x), 0, _ptr(dw), _ptr(db), 0, C.byref(shape), stream), "noalloc")
        torch.cuda.synchronize()
        for got, want in ((dw, dw_ref), (db, db_ref)):
            scale = float(want.abs().max())
            assert float((got - want).abs().max()) / scale < 3e-4
        # and both agree with PyTorch
        wr = torch.zeros(oc, c, k, k, device="cuda", requires_grad=True)
        torch.backends.cudnn.allow_tf32 = False
        torch.nn.functional.conv2d(x, wr, None, stride, pad)[:, :, :oh, :ow].backward(dz)
        assert float((dw - wr.grad).abs().max()) / float(wr.grad.abs().max()) < 3e-4


def test_noalloc_rejects_oversized_layers():
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    shape = N.ConvShape(4096, 64, 8, 8, 64, 8, 8, 3, 1, 1)  # 296 chunks x 36,928 floats = 43.7 MB of partial sums
    assert lib.b200cnn_conv2d_bwd_weight_scratch_bytes(C.byref(shape)) > 8 << 20
    rc = lib.b200cnn_conv2d_bwd_weight_noalloc(None, None, 0, None, None, 0, C.byref(shape), None)
    assert rc != 0
    msg = lib.b200cnn_last_error().decode()
    assert "built-in buffer holds 8388608" in msg and "needs" in msg
