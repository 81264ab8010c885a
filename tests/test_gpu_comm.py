"""Fused all-reduce + SGD: protocol / numerics on ONE GPU (virtual ranks inside one launch), and the
real multi-GPU path when the box has >= 2 GPUs."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _cuda_only():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("algo", ["one_shot", "two_shot"])
def test_emulated_ranks_one_gpu(world, algo):
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    flat = 40000
    nbytes = lib.b200cnn_comm_region_bytes(flat)
    poff, goff = lib.b200cnn_comm_params_offset(flat), [lib.b200cnn_comm_grads_offset(flat, p) for p in (0, 1)]
    regions = [torch.zeros(nbytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
    gen = torch.Generator(device="cuda").manual_seed(1)
    w0 = torch.randn(flat, device="cuda", generator=gen)
    views = []
    for r in range(world):
        p = regions[r][poff:poff + 4 * flat].view(torch.float32)
        g = [regions[r][o:o + 4 * flat].view(torch.float32) for o in goff]
        p.copy_(w0)
        views.append((p, g))
    seq = torch.zeros(world * 128 + 8, dtype=torch.int32, device="cuda")
    ptrs = (C.c_void_p * world)(*[C.c_void_p(t.data_ptr()) for t in regions])
    code = {"one_shot": N.ALGO_ONE_SHOT, "two_shot": N.ALGO_TWO_SHOT}[algo]
    want = w0.double().clone()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for step in range(4):  # several steps: sequence numbers grow, parity alternates
        par = step % 2
        grads = [torch.randn(flat, device="cuda", generator=gen) for _ in range(world)]
        for r in range(world):
            views[r][1][par].copy_(grads[r])
        off, cnt = (0, flat) if step % 2 == 0 else (1024, 20000)
        slot = (0, 64, 0, 64 + 61)[step]  # both flag channels of one region, interleaved; the last one ends at slot 127
        N.check(lib.b200cnn_comm_emulate_allreduce_sgd_slot(world, ptrs, flat, off, cnt, 0.5, code, par, 3, slot,
                                                            C.c_void_p(seq.data_ptr()), st), "emulate")
        torch.cuda.synchronize()
        assert int(seq[world * 128]) == 0, "a virtual rank timed out waiting for a peer"
        total = torch.zeros(flat, dtype=torch.float64, device="cuda")
        for r in range(world):
            total += grads[r].double()
        want[off:off + cnt] -= (0.5 / world) * total[off:off + cnt]
        for r in range(world):
            assert torch.equal(views[r][0], views[0][0]), f"replica {r} differs from replica 0"
        err = float((views[0][0].double() - want).abs().max())
        assert err < 1e-5, err
        want = views[0][0].double().clone()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("two_shot", [0, 1])
def test_emulated_push_protocol_one_gpu(world, two_shot):
    """Push (LL) all-reduce + SGD with `world` virtual ranks in one launch: tags, buffer parity, both channels, the
    deferred weight wait, replicas bitwise identical, values against an fp64 sum."""
    from mpi_cuda_cnn_b200 import _native as N

    lib = N.cuda()
    flat = 40000
    nbytes = lib.b200cnn_comm_region_bytes(flat)
    poff, goff = lib.b200cnn_comm_params_offset(flat), [lib.b200cnn_comm_grads_offset(flat, p) for p in (0, 1)]
    regions = [torch.zeros(nbytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
    gen = torch.Generator(device="cuda").manual_seed(2)
    w0 = torch.randn(flat, device="cuda", generator=gen)
    views = []
    for r in range(world):
        p = regions[r][poff:poff + 4 * flat].view(torch.float32)
        g = [regions[r][o:o + 4 * flat].view(torch.float32) for o in goff]
        p.copy_(w0)
        views.append((p, g))
    sbytes = lib.b200cnn_comm_emulate_ll_scratch_bytes(world)
    scratch = torch.zeros(sbytes, dtype=torch.uint8, device="cuda")
    ptrs = (C.c_void_p * world)(*[C.c_void_p(t.data_ptr()) for t in regions])
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    want = w0.double().clone()
    for step in range(6):  # tags grow, inbox buffers alternate, slice / grid / channel change between calls
        par = step % 2
        grads = [torch.randn(flat, device="cuda", generator=gen) for _ in range(world)]
        for r in range(world):
            views[r][1][par].copy_(grads[r])
        off, cnt = [(0, flat), (1024, 20000), (0, 4800), (36, 160), (0, flat), (8, 12000)][step]
        blocks = [4, 2, 3, 1, 5, 2][step]
        channel = 0 if two_shot or cnt * 4 > 128 * 1024 else step % 2
        update = 0 if step == 3 else 1
        N.check(lib.b200cnn_comm_emulate_ll(world, ptrs, flat, off, cnt, 0.5, two_shot, par, blocks, channel, update,
                                            C.c_void_p(scratch.data_ptr()), st), "emulate_ll")
        torch.cuda.synchronize()
        assert int(scratch[sbytes - 16:sbytes - 12].view(torch.int32)[0]) == 0, "a virtual rank timed out waiting for a peer"
        total = torch.zeros(flat, dtype=torch.float64, device="cuda")
        for r in range(world):
            total += grads[r].double()
        if update:
            want[off:off + cnt] -= (0.5 / world) * total[off:off + cnt]
        else:  # plain all-reduce: the sum lands in every rank's gradient buffer, parameters untouched
            for r in range(world):
                assert torch.equal(views[r][1][par][off:off + cnt], views[0][1][par][off:off + cnt])
            assert float((views[0][1][par][off:off + cnt].double() - total[off:off + cnt]).abs().max()) < 1e-4
        for r in range(world):
            assert torch.equal(views[r][0], views[0][0]), f"replica {r} differs from replica 0 after step {step}"
        err = float((views[0][0].double() - want).abs().max())
        assert err < 1e-5, (step, err)
        want = views[0][0].double().clone()


def _dp_worker(rank, world, mode, algo, steps, out_dir, pool=False):
    import os
    import torch
    import torch.distributed as dist
    from mpi_cuda_cnn_b200 import mnist_layers, pool_layers
    from mpi_cuda_cnn_b200.ops import functional as F200
    from mpi_cuda_cnn_b200.parallel import DataParallelTrainer, init_distributed

    rank, world, device = init_distributed()
    B = 64
    tr = DataParallelTrainer(rank, world, device, layers=pool_layers() if pool else mnist_layers(), batch=B, mode=mode, algo=algo)
    if pool:
        assert tr.engine.conv_path == "tcgen05_pool"
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for s in range(steps):
            x, y = F200.synthetic_batch(B, seed=7, first_index=(s * world + rank) * B)
            tr.train_step(x, y, 0.1, use_graph=(mode == "fused"))
        stream.synchronize()
    torch.cuda.synchronize()
    if tr.comm is not None:
        assert tr.comm.error() == 0
    params = tr.engine.get_params()
    np.save(os.path.join(out_dir, f"{mode}_{algo}_{rank}.npy"), params)
    dist.barrier()
    tr.close()


@pytest.mark.parametrize("algo", ["auto", "one_shot", "two_shot", "ll_one_shot", "ll_two_shot", "nvls"])
def test_multi_gpu_fused_equals_single_gpu_big_batch(algo, tmp_path):
    """W GPUs x batch B with the fused update == 1 GPU with batch W*B (same samples), replicas bitwise equal."""
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    from mpi_cuda_cnn_b200 import Engine
    from mpi_cuda_cnn_b200.ops import functional as F200
    from mpi_cuda_cnn_b200.parallel import spawn

    steps, B = 4, 64  # both step parities captured AND replayed
    spawn(_dp_worker, world, args=("fused", algo, steps, str(tmp_path)))
    reps = [np.load(tmp_path / f"fused_{algo}_{r}.npy") for r in range(world)]
    for r in range(1, world):
        assert np.array_equal(reps[0], reps[r])
    eng = Engine(max_batch=B * world)
    eng.init_params(seed=0)
    for s in range(steps):
        x, y = F200.synthetic_batch(B * world, seed=7, first_index=s * world * B)
        eng.train_step(x, y, 0.1 / (B * world), use_graph=False)
    torch.cuda.synchronize()
    ref = eng.get_params()
    # (TF32 products of differently split batches: ~1e-3 per step of drift relative to the largest parameter)
    assert float(np.abs(reps[0] - ref).max() / np.abs(ref).max()) < 4e-3
    eng.close()


def test_multi_gpu_pooled_topology_fused_equals_single_gpu(tmp_path):
    """The conv -> pool -> conv -> pool topology on its tensor-core path, data parallel with the fused update: replicas
    bitwise equal and equal (to TF32 drift) to one GPU training on the concatenated batch."""
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    from mpi_cuda_cnn_b200 import Engine, pool_layers
    from mpi_cuda_cnn_b200.ops import functional as F200
    from mpi_cuda_cnn_b200.parallel import spawn

    steps, B = 4, 64
    spawn(_dp_worker, world, args=("fused", "auto", steps, str(tmp_path), True))
    reps = [np.load(tmp_path / f"fused_auto_{r}.npy") for r in range(world)]
    for r in range(1, world):
        assert np.array_equal(reps[0], reps[r])
    eng = Engine(pool_layers(), max_batch=B * world)
    eng.init_params(seed=0)
    for s in range(steps):
        x, y = F200.synthetic_batch(B * world, seed=7, first_index=s * world * B)
        eng.train_step(x, y, 0.1 / (B * world), use_graph=False)
    torch.cuda.synchronize()
    ref = eng.get_params()
    eng.close()
    # TF32 products of differently split batches, conv gradients that sum 4x the pixels of the reference topology:
    # measured < 4e-3 at 2 GPUs, 7.2e-3 at 8 GPUs after these 4 steps (against an fp32 CUDA-core run of the concatenated
    # batch BOTH tensor-core runs sit at 1.3e-2: the TF32 truncation bias is common to them).
    err = float(np.abs(reps[0] - ref).max() / np.abs(ref).max())
    print(f"pooled topology, {world} GPUs: max parameter difference to the single-GPU run {err:.2e}")
    assert err < 1e-2


def test_multi_gpu_nccl_baseline_matches_fused(tmp_path):
    world = min(torch.cuda.device_count(), 2)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    from mpi_cuda_cnn_b200.parallel import spawn

    # same per-rank kernels, same samples: only the reduction differs, so this is the tight check of the protocol
    spawn(_dp_worker, world, args=("nccl", "auto", 6, str(tmp_path)))
    spawn(_dp_worker, world, args=("fused", "auto", 6, str(tmp_path)))
    a, b = np.load(tmp_path / "nccl_auto_0.npy"), np.load(tmp_path / "fused_auto_0.npy")
    assert float(np.abs(a - b).max() / np.abs(a).max()) < 1e-4
