"""GPU diagnostic for the tcgen05 implicit-GEMM convolutions (prints error statistics)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda()
torch.backends.cudnn.allow_tf32 = False
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
st = C.c_void_p(0)
def stat(name, got, want):
    torch.cuda.synchronize()
    err = (got - want).abs().max().item() / max(want.abs().max().item(), 1e-9)
    bad = ((got - want).abs() > 1e-2 * want.abs().max()).float().mean().item()
    print(f"{name:44s} rel_err={err:.3e} bad_frac={bad:.4f} got={[round(v,4) for v in got.flatten()[:4].tolist()]} want={[round(v,4) for v in want.flatten()[:4].tolist()]}")
for (n, h, oc, s, pad) in [(2, 14, 32, 2, 1), (256, 14, 32, 2, 1), (5, 14, 16, 2, 1), (3, 9, 32, 1, 1), (4, 7, 64, 1, 0)]:
    torch.manual_seed(0)
    oh = (h + 2 * pad - 3) // s + 1
    x = torch.randn(n, 16, h, h, device="cuda"); w = torch.randn(oc, 16, 3, 3, device="cuda") * 0.1; b = torch.randn(oc, device="cuda")
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    ld = oc * oh * oh + 4
    y = torch.full((n, ld), 7.0, device="cuda")
    shp = N.ConvShape(n, 16, h, h, oc, oh, oh, 3, pad, s)
    rc = lib.b200cnn_conv3x3_fwd_nhwc(p(x_nhwc), p(w), p(b), p(y), ld, C.byref(shp), N.ACT_RELU, st)
    torch.cuda.synchronize()
    want = torch.relu(F.conv2d(x, w, b, s, pad)).reshape(n, -1)
    stat(f"fwd n{n} h{h} oc{oc} s{s} p{pad} rc={rc} {lib.b200cnn_last_error().decode()[:30]}", y[:, :oc * oh * oh], want)
for n in (2, 256, 7):
    torch.manual_seed(1)
    dz = torch.randn(n, 32, 7, 7, device="cuda"); w = torch.randn(32, 16, 3, 3, device="cuda") * 0.1
    a1 = torch.randn(n, 16, 14, 14, device="cuda")
    dz_nhwc = dz.permute(0, 2, 3, 1).contiguous(); a1_nhwc = a1.permute(0, 2, 3, 1).contiguous()
    dx = torch.full((n, 14, 14, 16), 7.0, device="cuda")
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    rc = lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz_nhwc), p(w), p(a1_nhwc), N.ACT_RELU, p(dx), C.byref(shp), st)
    torch.cuda.synchronize()
    want = F.conv_transpose2d(dz, w, None, 2, 1, output_padding=1) * (a1 > 0)
    stat(f"bwd_data n{n} rc={rc} {lib.b200cnn_last_error().decode()[:30]}", dx.permute(0, 3, 1, 2), want)
