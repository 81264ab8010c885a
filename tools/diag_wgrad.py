"""GPU diagnostic for the tcgen05 conv2 weight-gradient kernel (tc_conv_wgrad.cu): error statistics vs autograd."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda()
torch.backends.cudnn.allow_tf32 = False
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
stream = torch.cuda.Stream()
st = C.c_void_p(stream.cuda_stream)
def stat(name, got, want):
    torch.cuda.synchronize()
    err = (got - want).abs().max().item() / max(want.abs().max().item(), 1e-9)
    bad = ((got - want).abs() > 1e-2 * want.abs().max()).float().mean().item()
    print(f"{name:40s} rel_err={err:.3e} bad_frac={bad:.4f} got={[round(v,3) for v in got.flatten()[:5].tolist()]} want={[round(v,3) for v in want.flatten()[:5].tolist()]}", flush=True)
for n in (256, 72, 8):
    torch.manual_seed(1)
    dz = torch.randn(n, 32, 7, 7, device="cuda"); a1 = torch.randn(n, 16, 14, 14, device="cuda")
    dz_nhwc = dz.permute(0, 2, 3, 1).contiguous(); a1_nhwc = a1.permute(0, 2, 3, 1).contiguous()
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp)), dtype=torch.uint8, device="cuda")
    wr = torch.zeros(32, 16, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(a1, wr, None, 2, 1).backward(dz)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        for chw, src in ((1, dz), (0, dz_nhwc), (1, dz)):
            dw = torch.full((32, 16, 3, 3), 7.0, device="cuda"); db = torch.full((32,), 7.0, device="cuda")
            torch.cuda.synchronize()
            print(f"launch n={n} chw={chw}", flush=True)
            rc = lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1_nhwc), 0, p(src), chw, p(dw), p(db), 0, p(scratch), C.byref(shp), st)
            stat(f"bwd_weight n{n} dz_chw={chw} rc={rc}", dw, wr.grad)
            stat(f"bwd_bias   n{n} dz_chw={chw}", db, dz.sum((0, 2, 3)))
            if chw == 0:
                for kx in range(3):
                    stat(f"   kx={kx}", dw[:, :, :, kx], wr.grad[:, :, :, kx])
if os.environ.get("B200CNN_WG_DEBUG") == "1":
    n = 256
    dz = torch.randn(n, 32, 7, 7, device="cuda"); a1_nhwc = torch.randn(n, 14, 14, 16, device="cuda")
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp)), dtype=torch.uint8, device="cuda")
    dw = torch.zeros(32, 16, 3, 3, device="cuda"); db = torch.zeros(32, device="cuda")
    with torch.cuda.stream(stream):
        for it in range(5):
            lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1_nhwc), 0, p(dz), 1, p(dw), p(db), 0, p(scratch), C.byref(shp), st)
            torch.cuda.synchronize()
    buf = (C.c_longlong * (12 * 16))()
    lib.b200cnn_debug_wgrad_stamps.argtypes = [C.c_void_p, C.c_int]
    print("stamps rc", lib.b200cnn_debug_wgrad_stamps(buf, 12 * 16))
    names = ["entry", "pdl", "tma0", "b0", "mma0done", "tma1", "b1", "acc", "stored"]
    for L in range(12):
        row = [buf[L * 16 + i] for i in range(9)]
        print(f"leader {L:2d}: " + " ".join(f"{names[i]}={(row[i] - row[0]) / 1.965e3:6.2f}" if row[i] else f"{names[i]}=   -- " for i in range(1, 9)))
