"""Per-op device time (CUDA events, warm, back-to-back launches of the SAME op) for the engine's shapes."""
import ctypes as C, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_cuda_cnn_b200 import _native as N
lib = N.cuda()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
stream = torch.cuda.Stream(); st = C.c_void_p(stream.cuda_stream)
def timeit(name, fn, iters=20, reps=20):
    """GPU time per call: `iters` calls captured into one CUDA graph (no host launch cost), replayed `reps` times."""
    with torch.cuda.stream(stream):
        for _ in range(3): fn()
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for _ in range(iters): fn()
        g.replay(); stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps): g.replay()
        e1.record(stream); stream.synchronize()
    us = e0.elapsed_time(e1) / (iters * reps) * 1e3
    print(f"{name:46s} {us:8.2f} us"); return us
R = lambda *s: torch.randn(*s, device="cuda")
out = {}
for backend, bname in ((N.BACKEND_SIMT, "simt"), (N.BACKEND_AUTO, "tc")):
    lib.b200cnn_set_backend(backend)
    for (n, k, ld, tag) in ((200, 1568, 200, "fc1"), (200, 200, 200, "fc2"), (10, 200, 16, "out")):
        x, w, b, y, dz, dx, dw, db = R(B, k), R(n, k), R(n), R(B, ld), R(B, ld), R(B, k), R(n, k), R(n)
        out[f"{bname}.{tag}.fwd"] = timeit(f"{bname} {tag} fwd m{B} n{n} k{k}", lambda: lib.b200cnn_linear_fwd(p(x), p(w), p(b), p(y), B, n, k, ld, 1, st))
        out[f"{bname}.{tag}.bwd_data"] = timeit(f"{bname} {tag} bwd_data", lambda: lib.b200cnn_linear_bwd_data(p(dz), p(w), p(x), 1, p(dx), B, n, k, ld, st))
        out[f"{bname}.{tag}.bwd_weight"] = timeit(f"{bname} {tag} bwd_weight", lambda: lib.b200cnn_linear_bwd_weight(p(dz), p(x), p(dw), p(db), 0, B, n, k, ld, st))
    x, w, b, lab, pr, dz, met = R(B, 200), R(10, 200), R(10), torch.randint(0, 10, (B,), device="cuda", dtype=torch.uint8), R(B, 16), R(B, 16), torch.zeros(4, device="cuda")
    out[f"{bname}.out.softmax"] = timeit(f"{bname} out fwd+softmax+xent", lambda: lib.b200cnn_linear_softmax_xent(p(x), p(w), p(b), p(lab), p(pr), p(dz), p(met), B, 10, 200, 16, st))
lib.b200cnn_set_backend(N.BACKEND_AUTO)
h1, w2, b2, w3, b3 = torch.tanh(R(B, 200)), R(200, 200) / 14, R(200), R(10, 200) / 14, R(10)
lab = torch.randint(0, 10, (B,), device="cuda", dtype=torch.uint8)
h2, pr, dz3, dz2, dzh1, met = R(B, 200), R(B, 16), R(B, 16), R(B, 200), R(B, 200), torch.zeros(4, device="cuda")
out["tc.head"] = timeit("tc fused head (fc2+out+softmax+2 dX)", lambda: lib.b200cnn_mlp_head(p(h1), p(w2), p(b2), p(w3), p(b3), p(lab), p(h2), p(pr), p(dz3), p(dz2), p(dzh1), p(met), B, 200, 200, 10, 16, st))
s1 = N.ConvShape(B, 1, 28, 28, 16, 14, 14, 3, 1, 2); s2 = N.ConvShape(B, 16, 14, 14, 32, 7, 7, 3, 1, 2)
x8 = torch.randint(0, 256, (B, 784), device="cuda", dtype=torch.uint8)
w1, b1, a1, dz1 = R(16, 1, 3, 3), R(16), R(B, 14, 14, 16), R(B, 14, 14, 16)
w2, b2, a2, dz2 = R(32, 16, 3, 3), R(32), R(B, 1568), R(B, 1568)
dw1, db1, dw2, db2 = R(16, 9), R(16), R(32, 144), R(32)
a1n, y1 = R(B, 16, 14, 14), R(B, 16, 14, 14)
scr = torch.zeros(max(lib.b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(C.byref(s2)), lib.b200cnn_conv2d_bwd_weight_scratch_bytes(C.byref(s2))), dtype=torch.uint8, device="cuda")
out["simt.conv1.fwd"] = timeit("simt conv1 fwd (u8 -> NCHW)", lambda: lib.b200cnn_conv2d_fwd(p(x8), 1, p(w1), p(b1), p(y1), C.byref(s1), 2, st))
out["simt.conv2.fwd"] = timeit("simt conv2 fwd NCHW", lambda: lib.b200cnn_conv2d_fwd(p(a1n), 0, p(w2), p(b2), p(a2), C.byref(s2), 2, st))
out["tc.conv2.fwd"] = timeit("tc   conv2 fwd NHWC implicit GEMM", lambda: lib.b200cnn_conv3x3_fwd_nhwc(p(a1), p(w2), p(b2), p(a2), 1568, C.byref(s2), 2, st))
out["simt.conv2.bwd_data"] = timeit("simt conv2 bwd_data NCHW", lambda: lib.b200cnn_conv2d_bwd_data(p(dz2), p(w2), p(a1n), 2, p(y1), C.byref(s2), st))
out["tc.conv2.bwd_data"] = timeit("tc   conv2 bwd_data (chw dz)", lambda: lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz2), 1, p(w2), p(a1), 2, p(dz1), C.byref(s2), st))
out["tc.conv2.bwd_data_tma"] = timeit("tc   conv2 bwd_data (nhwc dz, TMA)", lambda: lib.b200cnn_conv3x3s2_bwd_data_nhwc(p(dz2), 0, p(w2), p(a1), 2, p(dz1), C.byref(s2), st))
out["simt.conv2.bwd_weight"] = timeit("simt conv2 bwd_weight NCHW (2 kernels)", lambda: lib.b200cnn_conv2d_bwd_weight(p(dz2), p(a1n), 0, p(dw2), p(db2), 0, p(scr), C.byref(s2), st))
out["nhwc.conv2.bwd_weight"] = timeit("nhwc conv2 bwd_weight (2 kernels)", lambda: lib.b200cnn_conv2d_bwd_weight_nhwc(p(a1), 0, p(dz2), 1, p(dw2), p(db2), 0, p(scr), C.byref(s2), st))
out["nhwc.conv1.bwd_weight"] = timeit("nhwc conv1 bwd_weight u8 (2 kernels)", lambda: lib.b200cnn_conv2d_bwd_weight_nhwc(p(x8), 1, p(dz1), 0, p(dw1), p(db1), 0, p(scr), C.byref(s1), st))
w = R(360832); g = R(360832)
out["sgd"] = timeit("sgd 360,832 floats", lambda: lib.b200cnn_sgd_update(p(w), p(g), 360832, 0.01, 0, st))
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "bench_kernels.json"), "w"), indent=1)
