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
for n in (2, 256, 7, 72):
    torch.manual_seed(1)
    dz = torch.randn(n, 32, 7, 7, device="cuda"); w = torch.randn(32, 16, 3, 3, device="cuda") * 0.1
    a1 = torch.randn(n, 16, 14, 14, device="cuda")
    dz_nhwc = dz.permute(0, 2, 3, 1).contiguous(); a1_nhwc = a1.permute(0, 2, 3, 1).contiguous()
    dx = torch.full((n, 14, 14, 16), 7.0, device="cuda")
    shp = N.ConvShape(n, 16, 14, 14, 32, 7, 7, 3, 1, 2)
    rc = lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz_nhwc), 0, p(w), p(a1_nhwc), N.ACT_RELU, p(dx), C.byref(shp), st)
    torch.cuda.synchronize()
    want = F.conv_transpose2d(dz, w, None, 2, 1, output_padding=1) * (a1 > 0)
    stat(f"bwd_data n{n} rc={rc} {lib.b200cnn_last_error().decode()[:30]}", dx.permute(0, 3, 1, 2), want)
    dx2 = torch.full((n, 14, 14, 16), 7.0, device="cuda")
    rc = lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz), 1, p(w), p(a1_nhwc), N.ACT_RELU, p(dx2), C.byref(shp), st)
    stat(f"bwd_data(chw dz) n{n} rc={rc}", dx2.permute(0, 3, 1, 2), want)
    # backward-weight, NHWC x + CHW / NHWC dz
    nbytes = lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp))
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    dzm = dz * 1.0
    wr = w.clone().requires_grad_(); a1r = a1.clone()
    y = F.conv2d(a1r, wr, None, 2, 1); y.backward(dzm)
    for chw, src in ((1, dz), (0, dz_nhwc)):
        dw = torch.full_like(w, 7.0); db = torch.full((32,), 7.0, device="cuda")
        rc = lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1_nhwc), 0, p(src), chw, p(dw), p(db), 0, p(scratch), C.byref(shp), st)
        stat(f"bwd_weight nhwc dz_chw={chw} n{n} rc={rc}", dw, wr.grad)
        stat(f"bwd_bias   nhwc dz_chw={chw} n{n}", db, dz.sum((0, 2, 3)))
    # first conv: 1 input channel (uint8 image), dz NHWC [n][14][14][16]
    x8 = torch.randint(0, 256, (n, 1, 28, 28), device="cuda", dtype=torch.uint8)
    dz1 = torch.randn(n, 16, 14, 14, device="cuda"); dz1_nhwc = dz1.permute(0, 2, 3, 1).contiguous()
    shp1 = N.ConvShape(n, 1, 28, 28, 16, 14, 14, 3, 1, 2)
    w1 = torch.zeros(16, 1, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(x8.float() / 255.0, w1, None, 2, 1).backward(dz1)
    scratch = torch.zeros(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(shp1)), dtype=torch.uint8, device="cuda")
    dw1 = torch.full((16, 1, 3, 3), 7.0, device="cuda"); db1 = torch.full((16,), 7.0, device="cuda")
    rc = lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1_nhwc), 0, p(dw1), p(db1), 0, p(scratch), C.byref(shp1), st)
    stat(f"bwd_weight c1 u8 n{n} rc={rc}", dw1, w1.grad)
    stat(f"bwd_bias   c1 u8 n{n}", db1, dz1.sum((0, 2, 3)))
