"""Runs each tensor-core op of the flagship step a few times (for `ncu -k regex:...` captures)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda(); B = 256
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
R = lambda *s: torch.randn(*s, device="cuda")
x, w, b, y = R(B, 1568), R(200, 1568), R(200), R(B, 200)
x2, w3, b3, lab, pr, dz3, met = R(B, 200), R(10, 200), R(10), torch.randint(0, 10, (B,), device="cuda", dtype=torch.uint8), R(B, 16), R(B, 16), torch.zeros(4, device="cuda")
dz, dx, dw, db = R(B, 200), R(B, 1568), R(200, 1568), R(200)
s2 = N.ConvShape(B, 16, 14, 14, 32, 7, 7, 3, 1, 2)
a1, w2, b2, a2, dz2, dz1 = R(B, 14, 14, 16), R(32, 16, 3, 3), R(32), R(B, 1568), R(B, 1568), R(B, 14, 14, 16)
for it in range(4):
    lib.b200cnn_linear_fwd(p(x), p(w), p(b), p(y), B, 200, 1568, 200, 1, st)
    lib.b200cnn_linear_softmax_xent(p(x2), p(w3), p(b3), p(lab), p(pr), p(dz3), p(met), B, 10, 200, 16, st)
    lib.b200cnn_linear_bwd_data(p(dz), p(w), p(x), 2, p(dx), B, 200, 1568, 200, st)
    lib.b200cnn_linear_bwd_weight(p(dz), p(x), p(dw), p(db), 0, B, 200, 1568, 200, st)
    lib.b200cnn_conv3x3_fwd_nhwc(p(a1), p(w2), p(b2), p(a2), 1568, C.byref(s2), 2, st)
    lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz2), 1, p(w2), p(a1), 2, p(dz1), C.byref(s2), st)
torch.cuda.synchronize()
print("ok")
