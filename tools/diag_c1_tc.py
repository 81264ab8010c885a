"""First thing to run on a GPU next round: the opt-in tensor-core conv1 weight gradient (B200CNN_TC_C1=1) against
autograd and against the default CUDA-core kernel, then its in-graph time.
    B200CNN_TC_C1=1 python tools/diag_c1_tc.py      (and once without the variable for the baseline numbers)"""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda()
p = lambda t: C.c_void_p(t.data_ptr())
stream = torch.cuda.Stream(); st = C.c_void_p(stream.cuda_stream)
print("B200CNN_TC_C1 =", os.environ.get("B200CNN_TC_C1"))
for n in (256, 2, 7, 64):
    torch.manual_seed(n)
    x8 = torch.randint(0, 256, (n, 1, 28, 28), device="cuda", dtype=torch.uint8)
    dz1 = torch.randn(n, 16, 14, 14, device="cuda"); dz1_nhwc = dz1.permute(0, 2, 3, 1).contiguous()
    w1 = torch.zeros(16, 1, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(x8.float() / 255.0, w1, None, 2, 1).backward(dz1)
    s1 = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
    scr = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(s1)), dtype=torch.uint8, device="cuda")
    with torch.cuda.stream(stream):
        for rep in range(2):  # twice: the completion counter must have reset itself
            dw = torch.full((16, 1, 3, 3), 7.0, device="cuda"); db = torch.full((16,), 7.0, device="cuda")
            rc = lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1_nhwc), 0, p(dw), p(db), 0, p(scr), C.byref(s1), st)
            torch.cuda.synchronize()
            ew = ((dw - w1.grad).abs().max() / w1.grad.abs().max()).item()
            eb = ((db - dz1.sum((0, 2, 3))).abs().max() / dz1.sum((0, 2, 3)).abs().max()).item()
            print(f"n={n:3d} rep={rep} rc={rc} rel err dW {ew:.3e}  db {eb:.3e}")
n = 256
x8 = torch.randint(0, 256, (n, 784), device="cuda", dtype=torch.uint8); dz1 = torch.randn(n, 14, 14, 16, device="cuda")
s1 = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
scr = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(s1)), dtype=torch.uint8, device="cuda")
dw = torch.zeros(16, 9, device="cuda"); db = torch.zeros(16, device="cuda")
with torch.cuda.stream(stream):
    g = torch.cuda.CUDAGraph()
    lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1), 0, p(dw), p(db), 0, p(scr), C.byref(s1), st)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=stream):
        for _ in range(20):
            lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1), 0, p(dw), p(db), 0, p(scr), C.byref(s1), st)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10):
        g.replay()
    e1.record(stream)
    torch.cuda.synchronize()
    print(f"in-graph time per launch: {e0.elapsed_time(e1) * 1000 / 200:.2f} us")
