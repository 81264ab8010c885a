"""Diagnostic for csrc/cuda/tc_chain.cu: per-tensor error of the fused FC chain kernel against PyTorch fp32, and its
time inside a CUDA graph next to the seven separate kernels it replaces.  Usage: python tools/diag_chain.py [m]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from mpi_cuda_cnn_b200 import _native as N

lib = N.cuda()
m = int(sys.argv[1]) if len(sys.argv) > 1 else 256
k1, n1, n2, nc = 1568, 200, 200, 10
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(0)
x = torch.relu(torch.randn(m, k1, device="cuda"))
w1, b1 = torch.randn(n1, k1, device="cuda") / k1 ** 0.5, 0.1 * torch.randn(n1, device="cuda")
w2, b2 = torch.randn(n2, n1, device="cuda") / n1 ** 0.5, 0.1 * torch.randn(n2, device="cuda")
w3, b3 = torch.randn(nc, n2, device="cuda") / n2 ** 0.5, 0.1 * torch.randn(nc, device="cuda")
lab = torch.randint(0, nc, (m,), device="cuda", dtype=torch.uint8)
rh1 = torch.tanh(x @ w1.T + b1)
rh2 = torch.tanh(rh1 @ w2.T + b2)
rp = torch.softmax(rh2 @ w3.T + b3, 1)
rdz3 = rp - F.one_hot(lab.long(), nc)
rdz2 = (rdz3 @ w3) * (1 - rh2 ** 2)
rdz1 = (rdz2 @ w2) * (1 - rh1 ** 2)
rdx = (rdz1 @ w1) * (x > 0)
p = lambda t: C.c_void_p(t.data_ptr())
stream = torch.cuda.Stream()
st = C.c_void_p(stream.cuda_stream)
h1, dz1 = torch.full((m, n1), 9.0, device="cuda"), torch.full((m, n1), 9.0, device="cuda")
h2, dz2 = torch.full((m, n2), 9.0, device="cuda"), torch.full((m, n2), 9.0, device="cuda")
pr, dz3 = torch.full((m, 16), 9.0, device="cuda"), torch.full((m, 16), 9.0, device="cuda")
dx, met = torch.full((m, k1), 9.0, device="cuda"), torch.zeros(4, device="cuda")
torch.cuda.synchronize()


def run():
    N.check(lib.b200cnn_fc_chain(p(x), N.ACT_RELU, p(w1), p(b1), p(w2), p(b2), p(w3), p(b3), p(lab), p(h1), p(h2), p(pr), p(dz3),
                                 p(dz2), p(dz1), p(dx), p(met), m, k1, n1, n2, nc, 16, st), "fc_chain")


with torch.cuda.stream(stream):
    run()
    stream.synchronize()
for name, got, want in (("h1", h1, rh1), ("h2", h2, rh2), ("probs", pr[:, :nc], rp), ("dz3", dz3[:, :nc], rdz3), ("dz2", dz2, rdz2),
                        ("dz1", dz1, rdz1), ("dx", dx, rdx)):
    d = (got - want).abs()
    print(f"{name:6s} max rel err {float(d.max() / want.abs().max()):.3e}   untouched (== 9.0): {int((got == 9.0).sum())}   "
          f"nan: {int(torch.isnan(got).sum())}   worst at {tuple(int(v) for v in torch.nonzero(d == d.max())[0])}")
print("metrics", met.tolist(), "want ce", float(-torch.log(rp[torch.arange(m), lab.long()]).sum()), "mse", float((rdz3 ** 2).sum() / nc),
      "correct", int((rp.argmax(1) == lab.long()).sum()))
lib.b200cnn_fc_chain_max_active_clusters.restype = C.c_int
print("max active clusters of the chain kernel on this device:", lib.b200cnn_fc_chain_max_active_clusters())
# in-kernel milestones (SM clock of CTA 0 of cluster 0), one launch with warm caches
stamps = torch.zeros(32, dtype=torch.int64, device="cuda")
lib.b200cnn_fc_chain_debug_stamps.restype = None
lib.b200cnn_fc_chain_debug_stamps.argtypes = [C.c_void_p]
lib.b200cnn_fc_chain_debug_stamps(p(stamps))
with torch.cuda.stream(stream):
    run()
    stream.synchronize()
lib.b200cnn_fc_chain_debug_stamps(C.c_void_p(0))
t = stamps.tolist()
names = {0: "start", 1: "producer: GEMM-1 loads issued", 2: "producer: GEMM-6 loads issued", 3: "GEMM-1 MMAs issued",
         4: "boundary 1 passed (h1 tile)", 5: "GEMM-2 issued", 6: "boundary 2 passed (h2 tile)", 7: "GEMM-3 issued", 8: "dz3 tile ready",
         9: "boundary 3 passed (dz2 tile)", 10: "GEMM-5 issued", 11: "boundary 4 passed (dz1 tile)", 12: "GEMM-6 issued",
         13: "acc 1 ready", 14: "partials of all CTAs here", 15: "h1 delivered", 16: "acc 2 ready", 17: "h2 delivered",
         18: "logits ready", 19: "softmax done", 20: "acc 4 ready", 21: "dz2 delivered", 22: "acc 5 ready", 23: "dz1 delivered",
         24: "acc 6 ready", 25: "epilogue done"}
order = [0, 1, 3, 13, 14, 15, 4, 5, 16, 17, 6, 7, 18, 19, 8, 20, 21, 9, 10, 22, 23, 11, 2, 12, 24, 25]
for i in order:
    if t[i]:
        print(f"  stamp {i:2d} {names[i]:32s} {(t[i] - t[0]) / 1.965e3:8.2f} us")
# time in a graph: 20 launches
with torch.cuda.stream(stream):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for _ in range(20):
            run()
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10):
        g.replay()
    e1.record(stream)
    stream.synchronize()
    print(f"fc_chain in a graph: {e0.elapsed_time(e1) * 1e3 / 200:.2f} us per launch (m = {m})")
