"""Loss / metric curve of the GPU engine vs the native fp64 CPU engine on IDENTICAL mini-batches.

North-star check: "final loss matching the serial cnn.c reference within tolerance".  The CPU engine in quirks mode is
byte-identical to cnn.c (tests/test_cpu_engine.py); here its clean mode (same maths, correct conv indexing, true
mini-batches) is the fp64 oracle and the GPU engine runs the same batches in TF32 (and fp32 with --backend simt)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mpi_cuda_cnn_b200 import Engine
from mpi_cuda_cnn_b200.ops import functional as F200
from mpi_cuda_cnn_b200.utils import make_synthetic

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=150)
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--rate", type=float, default=0.1)
args = ap.parse_args()
B, S = args.batch, args.steps
x, y = make_synthetic(B * S, seed=1)
xt, yt = make_synthetic(1000, seed=2)
for backend in ("simt", "auto"):
    F200.set_backend(backend)
    eng = Engine(max_batch=max(B, 250))
    cpu = eng.init_params(seed=0)
    dx, dy = torch.from_numpy(x.reshape(-1, 784)).cuda(), torch.from_numpy(y).cuda()
    xs, ys = torch.empty((B, 784), dtype=torch.uint8, device="cuda"), torch.empty((B,), dtype=torch.uint8, device="cuda")
    gpu_curve, cpu_curve = [], []
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for s in range(S):
            xs.copy_(dx[s * B:(s + 1) * B]); ys.copy_(dy[s * B:(s + 1) * B])
            eng.reset_metrics()
            eng.train_step(xs, ys, args.rate / B, use_graph=True)
            gpu_curve.append(eng.read_metrics()["mse_sum"] / B)
    for s in range(S):
        e = 0.0
        for i in range(s * B, (s + 1) * B):
            cpu.set_inputs(x[i].astype(np.float64) / 255.0)
            cpu.learn_outputs(np.eye(10)[y[i]])
            e += cpu.error_total()
        cpu.update(args.rate / B)
        cpu_curve.append(e / B)
    g, c = np.array(gpu_curve), np.array(cpu_curve)
    pg = eng.predict(torch.from_numpy(xt.reshape(-1, 784)[:250]).cuda()).argmax(1)
    acc_gpu = float((pg == yt[:250]).mean())
    correct = 0
    for i in range(250):
        cpu.set_inputs(xt[i].astype(np.float64) / 255.0)
        correct += int(np.argmax(cpu.get_outputs()) == yt[i])
    print(f"backend={backend:5s} steps={S} batch={B}: reference 'error' metric  first gpu/cpu {g[0]:.5f}/{c[0]:.5f}  last {g[-1]:.5f}/{c[-1]:.5f}  "
          f"max|gpu-cpu| over the curve {np.abs(g - c).max():.2e}  mean {np.abs(g - c).mean():.2e};  "
          f"param max rel diff {np.abs(eng.get_params() - cpu.get_params()).max() / np.abs(cpu.get_params()).max():.2e};  "
          f"test accuracy gpu {acc_gpu:.3f} cpu {correct / 250:.3f}")
    eng.close()
