"""Runs the conv-chain kernels a few times (for `ncu -k regex:...` captures)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda(); B = 256
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
R = lambda *s: torch.randn(*s, device="cuda")
s1 = N.ConvShape(B, 1, 28, 28, 16, 14, 14, 3, 1, 2); s2 = N.ConvShape(B, 16, 14, 14, 32, 7, 7, 3, 1, 2)
x8 = torch.randint(0, 256, (B, 784), device="cuda", dtype=torch.uint8)
a1, w2, b2, a2, dz2, dz1 = R(B, 14, 14, 16), R(32, 16, 3, 3), R(32), R(B, 1568), R(B, 1568), R(B, 14, 14, 16)
dw1, db1, dw2, db2 = R(16, 9), R(16), R(32, 144), R(32)
scr = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(s2)), dtype=torch.uint8, device="cuda")
for it in range(4):
    lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz2), 1, p(w2), p(a1), 2, p(dz1), C.byref(s2), st)
    lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1), 0, p(dz2), 1, p(dw2), p(db2), 0, p(scr), C.byref(s2), st)
    lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1), 0, p(dw1), p(db1), 0, p(scr), C.byref(s1), st)
torch.cuda.synchronize(); print("ok")
