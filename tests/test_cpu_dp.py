"""CPU data-parallel trainer (configuration 2) over the shared-memory communicator."""
import os
import re
import subprocess


def _run(repo_root, args, env=None):
    return subprocess.run([os.path.join(repo_root, "bin", "cnnmpi")] + args, stderr=subprocess.PIPE, text=True, env=env, timeout=300)


def test_two_ranks_stay_identical_and_learn(dataset_paths, repo_root):
    r = _run(repo_root, ["-np", "2", "--nepoch", "3", "--batch-size", "8"] + dataset_paths)
    assert r.returncode == 0, r.stderr
    sums = re.findall(r"rank (\d) param_checksum=(\S+)", r.stderr)
    assert len(sums) == 2 and sums[0][1] == sums[1][1]  # replicas bitwise identical
    assert re.search(r"^0 0 100$", r.stderr, re.M) and re.search(r"^1 100 200$", r.stderr, re.M)  # cnnmpi.c:457-459
    assert "epoch = 2" in r.stderr
    ncorrect = int(re.search(r"ncorrect=(\d+)", r.stderr).group(1))
    assert ncorrect >= 45


def test_world2_equals_world1_with_same_global_batches(dataset_paths, repo_root, tmp_path):
    """2 ranks x batch B over contiguous shards == 1 rank x batch 2B over the SAME global mini-batches.

    World 2 trains, in step b, on samples [bB, bB+B) of rank 0's shard and [100+bB, 100+bB+B) of rank 1's.  The
    single-rank run gets a copy of the data set in which exactly those 2B samples are adjacent, so its step b sees the
    same samples; sums are taken in a different order, so parameters agree to rounding (fp64), not bitwise."""
    import numpy as np

    from mpi_cuda_cnn_b200.utils import read_idx, write_idx

    B = 5
    x, y = read_idx(dataset_paths[0]), read_idx(dataset_paths[1])
    n, shard = len(x), len(x) // 2
    order = []
    for b in range(shard // B):
        order += list(range(b * B, (b + 1) * B)) + list(range(shard + b * B, shard + (b + 1) * B))
    assert sorted(order) == list(range(n))
    px, py = str(tmp_path / "train-images-perm"), str(tmp_path / "train-labels-perm")
    write_idx(px, x[order])
    write_idx(py, y[order])
    two = _run(repo_root, ["-np", "2", "--nepoch", "2", "--batch-size", str(B)] + dataset_paths)
    one = _run(repo_root, ["-np", "1", "--nepoch", "2", "--batch-size", str(2 * B), px, py] + dataset_paths[2:])
    assert two.returncode == 0 and one.returncode == 0, two.stderr + one.stderr
    c2 = [float(v) for v in re.findall(r"param_checksum=(\S+)", two.stderr)]
    c1 = [float(v) for v in re.findall(r"param_checksum=(\S+)", one.stderr)]
    assert len(c2) == 2 and c2[0] == c2[1] and len(c1) == 1
    assert abs(c2[0] - c1[0]) <= 1e-9 * max(1.0, abs(c1[0])), (c2[0], c1[0])
    assert re.search(r"ncorrect=(\d+)", two.stderr).group(1) == re.search(r"ncorrect=(\d+)", one.stderr).group(1)
    # and the plumbing is deterministic: the same world twice gives the same bits
    again = _run(repo_root, ["-np", "2", "--nepoch", "2", "--batch-size", str(B)] + dataset_paths)
    assert re.findall(r"param_checksum=(\S+)", again.stderr) == re.findall(r"param_checksum=(\S+)", two.stderr)


def test_real_mpi_backend_builds_and_matches_shm(dataset_paths, repo_root):
    """The -DB200CNN_HAVE_MPI transport (five MPI calls, csrc/cpu/mpi_comm.hpp), built against the in-repo five-call
    mpi.h stand-in because the image has no MPI: two ranks give the same parameters as the shared-memory transport."""
    res = subprocess.run(["make", "-C", repo_root, "cnnmpi_mpishim"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 0, res.stdout[-3000:]
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_PORT="29878")
        procs.append(subprocess.Popen([os.path.join(repo_root, "bin", "cnnmpi_mpishim"), "--nepoch", "1", "--batch-size", "8"] + dataset_paths,
                                      stderr=subprocess.PIPE, text=True, env=env))
    outs = [p.communicate(timeout=300)[1] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    mpi = [re.search(r"param_checksum=(\S+)", o).group(1) for o in outs]
    shm = _run(repo_root, ["-np", "2", "--nepoch", "1", "--batch-size", "8"] + dataset_paths)
    assert mpi[0] == mpi[1] == re.findall(r"param_checksum=(\S+)", shm.stderr)[0]


def test_bad_checkpoint_ends_every_rank(dataset_paths, repo_root, tmp_path):
    r = _run(repo_root, ["-np", "2", "--nepoch", "1", "--load", str(tmp_path / "missing.ckpt")] + dataset_paths)
    assert r.returncode == 111 and "param_checksum" not in r.stderr


def test_rank_that_dies_releases_its_peers(dataset_paths, repo_root, tmp_path):
    """External launch of a 2-rank world where rank 1 never shows up in the second barrier: rank 0 must not spin forever."""
    env = dict(os.environ, RANK="0", WORLD_SIZE="2", MASTER_PORT="29879", B200CNN_COMM_TIMEOUT_S="3")
    p = subprocess.Popen([os.path.join(repo_root, "bin", "cnnmpi"), "--nepoch", "1"] + dataset_paths, stderr=subprocess.PIPE, text=True, env=env)
    err = p.communicate(timeout=120)[1]
    assert p.returncode in (111, 112), (p.returncode, err)  # no segment partner / no barrier partner within 3 s


def test_external_launcher_env(dataset_paths, repo_root):
    """RANK / WORLD_SIZE in the environment (torchrun-style) instead of -np."""
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_PORT="29877")
        procs.append(subprocess.Popen([os.path.join(repo_root, "bin", "cnnmpi"), "--nepoch", "1"] + dataset_paths, stderr=subprocess.PIPE, text=True, env=env))
    outs = [p.communicate(timeout=300)[1] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    sums = [re.search(r"param_checksum=(\S+)", o).group(1) for o in outs]
    assert sums[0] == sums[1]


def test_missing_file_does_not_hang(repo_root):
    r = _run(repo_root, ["-np", "2", "a", "b", "c", "d"])
    assert r.returncode == 111
