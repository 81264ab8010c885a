"""In-kernel clock stamps of the first convolution's weight-gradient kernel (block 0 and the final-reduction block)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda()
p = lambda t: C.c_void_p(t.data_ptr())
stream = torch.cuda.Stream(); st = C.c_void_p(stream.cuda_stream)
n = 256
x8 = torch.randint(0, 256, (n, 784), device="cuda", dtype=torch.uint8)
dz1 = torch.randn(n, 14, 14, 16, device="cuda")
s1 = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
scr = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(s1)), dtype=torch.uint8, device="cuda")
dw = torch.zeros(16, 9, device="cuda"); db = torch.zeros(16, device="cuda")
lib.b200cnn_debug_c1_enable(1)
with torch.cuda.stream(stream):
    for it in range(5):
        lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1), 0, p(dw), p(db), 0, p(scr), C.byref(s1), st)
        torch.cuda.synchronize()
buf = (C.c_longlong * 16)()
lib.b200cnn_debug_c1_stamps(buf)
us = lambda a, b: (buf[a] - buf[b]) / 1.965e3
print(f"block 0: pdl_wait {us(1,0):.2f}  loads {us(2,1):.2f}  compute+partial {us(3,2):.2f}  fence+atomic {us(4,3):.2f}")
print(f"last block: loads {us(9,8):.2f}  final {us(10,9):.2f}")
