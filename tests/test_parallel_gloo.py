"""Host-side logic of the Python data-parallel path on CPU (gloo, world_size 2): the launcher's rendezvous and the
shard / sum / average / update algebra the GPU trainers implement (plumbing tests need no GPU)."""
import numpy as np
import pytest
import torch


def _rank_main(rank, world, out_dir):
    import torch.distributed as dist

    from mpi_cuda_cnn_b200.engine import mnist_layers
    from mpi_cuda_cnn_b200.ops.cpu import CpuNet
    from mpi_cuda_cnn_b200.parallel.launcher import env_rank_world, init_distributed
    from mpi_cuda_cnn_b200.utils import make_synthetic

    assert env_rank_world() == (rank, world, rank)
    r, w, device = init_distributed(backend="gloo")
    assert (r, w) == (rank, world) and dist.get_backend() == "gloo"
    # every rank seeds identically (the reference seeds per rank and never synchronises: SURVEY D7)
    net = CpuNet.from_layers(mnist_layers(), seed=5)
    x, y = make_synthetic(16, seed=9)
    xs, onehot = x.astype(np.float64) / 255.0, np.eye(10)[y]
    shard = len(xs) // world  # contiguous shard per rank (cnnmpi.c:457-458)
    lo, hi = shard * rank, shard * (rank + 1)
    for step in range(2):
        net.learn_batch(xs[lo:hi], onehot[lo:hi])
        g = torch.from_numpy(net.get_updates())
        dist.all_reduce(g, op=dist.ReduceOp.SUM)  # the intended collective of cnnmpi.c:487-499
        # apply: w -= rate / (local batch * world) * sum of gradients
        p = net.get_params() - (0.1 / (shard * world)) * g.numpy()
        net.set_params(p)
        net.update(0.0)  # clears the local accumulators
    np.save(f"{out_dir}/params_{rank}.npy", net.get_params())


@pytest.mark.timeout(300)
def test_gloo_two_ranks_match_one_process_on_the_global_batch(tmp_path):
    from mpi_cuda_cnn_b200.engine import mnist_layers
    from mpi_cuda_cnn_b200.ops.cpu import CpuNet
    from mpi_cuda_cnn_b200.parallel.launcher import spawn
    from mpi_cuda_cnn_b200.utils import make_synthetic

    spawn(_rank_main, 2, args=(str(tmp_path),))
    p0, p1 = np.load(tmp_path / "params_0.npy"), np.load(tmp_path / "params_1.npy")
    assert np.array_equal(p0, p1)  # replicas bitwise identical
    net = CpuNet.from_layers(mnist_layers(), seed=5)
    x, y = make_synthetic(16, seed=9)
    xs, onehot = x.astype(np.float64) / 255.0, np.eye(10)[y]
    for step in range(2):
        net.learn_batch(xs, onehot)
        net.update(0.1 / 16)
    assert np.allclose(p0, net.get_params(), rtol=1e-9, atol=1e-12)
