"""GPU diagnostic for the tcgen05 GEMM paths (prints error statistics, never asserts)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda()
torch.backends.cuda.matmul.allow_tf32 = False
p = lambda t: C.c_void_p(t.data_ptr())
st = C.c_void_p(0)
def stat(name, got, want):
    torch.cuda.synchronize()
    err = (got - want).abs().max().item() / max(want.abs().max().item(), 1e-9)
    nz = (got != 0).float().mean().item()
    print(f"{name:40s} rel_err={err:.3e} nonzero_frac={nz:.3f} got[0,:4]={got.flatten()[:4].tolist()} want={want.flatten()[:4].tolist()}")
lib.b200cnn_set_backend(N.BACKEND_TCGEN05)
for (m, n, k) in [(128, 32, 32), (128, 64, 64), (256, 200, 200), (256, 1568, 200), (2048, 512, 64)]:
    torch.manual_seed(0)
    x = torch.randn(m, k, device="cuda"); w = torch.randn(n, k, device="cuda"); b = torch.zeros(n, device="cuda")
    y = torch.full((m, n), 7.0, device="cuda")
    rc = lib.b200cnn_linear_fwd(p(x), p(w), p(b), p(y), m, n, k, n, 0, st)
    stat(f"fwd m{m} n{n} k{k} rc={rc}", y, x @ w.T)
    dz = torch.randn(m, n, device="cuda")
    dx = torch.full((m, k), 7.0, device="cuda")
    rc = lib.b200cnn_linear_bwd_data(p(dz), p(w), None, 0, p(dx), m, n, k, n, st)
    stat(f"bwd_data m{m} n{n} k{k} rc={rc}", dx, dz @ w)
    dw = torch.full((n, k), 7.0, device="cuda"); db = torch.full((n,), 7.0, device="cuda")
    rc = lib.b200cnn_linear_bwd_weight(p(dz), p(x), p(dw), p(db), 0, m, n, k, n, st)
    stat(f"bwd_weight m{m} n{n} k{k} rc={rc}", dw, dz.T @ x)
    stat(f"bwd_bias   m{m} n{n} k{k}", db, dz.sum(0))
print(lib.b200cnn_last_error())
