"""Per-tensor activation / gradient error of the engine vs the fp64 PyTorch model, for each backend."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mpi_cuda_cnn_b200 import Engine, mnist_layers, pool_layers
from mpi_cuda_cnn_b200.models import RefCNN, dense_grads, load_dense_params
from mpi_cuda_cnn_b200.ops import functional as F200
torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False
def rel(a, b): return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))
POOL = "--pool" in sys.argv
_pos = [a for a in sys.argv[1:] if not a.startswith("--")]
B = int(_pos[0]) if _pos else 64
names = ["conv1.w", "conv1.b", "conv2.w", "conv2.b", "fc1.w", "fc1.b", "fc2.w", "fc2.b", "out.w", "out.b"]
for backend in ("simt", "auto"):
    F200.set_backend(backend)
    eng = Engine(pool_layers() if POOL else mnist_layers(), max_batch=B); cpu = eng.init_params(seed=0)
    model = load_dense_params(RefCNN(pool=POOL, dtype=torch.float64), cpu.get_params()).cuda()
    x, y = F200.synthetic_batch(B, seed=11)
    eng.forward(x, y); eng.backward(); torch.cuda.synchronize()
    xr = (x.double() / 255.0).reshape(B, 1, 28, 28)
    feats = model.features(xr); model.loss_sum(xr, y).backward()
    L = len(eng.layers)
    acts = {L - 5 if not POOL else 2: feats[0].flatten(1), L - 4: feats[1].flatten(1), L - 3: feats[2], L - 2: feats[3], L - 1: torch.softmax(feats[4], 1)}
    print(backend, eng.conv_path, "activations:", {k: f"{rel(eng.layer_output(k, B), v.detach().cpu().numpy()):.2e}" for k, v in acts.items()})
    g, gr = eng.get_grads(), dense_grads(model)
    o = 0; out = {}
    for nme, p in zip(names, model.param_list()):
        n = p.numel(); out[nme] = f"{rel(g[o:o+n], gr[o:o+n]):.2e}"; o += n
    print(backend, "grads:", out)
    eng.close()
