"""GPU diagnostic for the fused head kernel vs PyTorch."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda(); torch.backends.cuda.matmul.allow_tf32 = False
p = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(0)
def stat(name, got, want):
    err = (got - want).abs().max().item() / max(want.abs().max().item(), 1e-9)
    print(f"{name:28s} rel_err={err:.3e} got={[round(x,4) for x in got.flatten()[:3].tolist()]} want={[round(x,4) for x in want.flatten()[:3].tolist()]}")
for (m, n1, n2, nc) in [(256, 200, 200, 10), (100, 64, 128, 7), (300, 208, 36, 16)]:
    torch.manual_seed(0)
    h1 = torch.tanh(torch.randn(m, n1, device="cuda")); w2 = torch.randn(n2, n1, device="cuda") / n1 ** 0.5; b2 = 0.1 * torch.randn(n2, device="cuda")
    w3 = torch.randn(nc, n2, device="cuda") / n2 ** 0.5; b3 = 0.1 * torch.randn(nc, device="cuda")
    lab = torch.randint(0, nc, (m,), device="cuda", dtype=torch.uint8)
    h2, pr, dz3, dz2, dz1, met = (torch.full(s, 7.0, device="cuda") for s in ((m, n2), (m, 16), (m, 16), (m, n2), (m, n1), (4,)))
    met.zero_()
    rc = lib.b200cnn_mlp_head(p(h1), p(w2), p(b2), p(w3), p(b3), p(lab), p(h2), p(pr), p(dz3), p(dz2), p(dz1), p(met), m, n1, n2, nc, 16, st)
    torch.cuda.synchronize()
    print(f"--- m={m} n1={n1} n2={n2} nc={nc} rc={rc} {lib.b200cnn_last_error().decode()}")
    rh2 = torch.tanh(h1 @ w2.T + b2); rp = torch.softmax(rh2 @ w3.T + b3, 1)
    rdz3 = rp - torch.nn.functional.one_hot(lab.long(), nc); rdz2 = (rdz3 @ w3) * (1 - rh2 ** 2); rdz1 = (rdz2 @ w2) * (1 - h1 ** 2)
    stat("h2", h2, rh2); stat("probs", pr[:, :nc], rp); stat("dz3", dz3[:, :nc], rdz3); stat("dz2", dz2, rdz2); stat("dz1", dz1, rdz1)
    ce = -torch.log(rp[torch.arange(m), lab.long()]).sum()
    print("metrics", met.tolist(), "want ce", ce.item(), "mse", ((rdz3 ** 2).mean(1)).sum().item(), "count", m)
