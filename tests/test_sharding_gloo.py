"""CPU, world_size 2 over gloo: the host-side sharding of the batched robots.  The kernels are
not run here (no GPU); a deterministic per-robot function stands in for `qpmpc_solve_batched`, which -- like the real
path -- depends only on the GLOBAL robot id (Philox subsequence = robot id), never on the rank layout."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from quadruped_pympc_tamols_b200.sharding import gather_to_rank0, shard, shard_bounds


def _fake_solve(robot_ids, state):
    """Stand-in with the real path's dependency structure: row r is a function of (global robot id, its own state)."""
    rows = [np.random.default_rng(1000 + int(r)).normal(size=5) for r in robot_ids]
    noise = np.stack(rows) if rows else np.zeros((0, 5))
    return noise.astype(np.float32) + state[:, :5].astype(np.float32)


def _worker(rank, world, port, num_robots, out_path):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    states = np.random.default_rng(7).normal(size=(num_robots, 24))
    lo, hi = shard_bounds(num_robots, world, rank)
    local = _fake_solve(np.arange(lo, hi), shard(states, world, rank))
    full = gather_to_rank0(local, num_robots, dist)
    if rank == 0:
        np.save(out_path, full)
    else:
        assert full is None
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("num_robots", [8, 7, 1])
def test_two_rank_sharding_equals_single_process(tmp_path, num_robots):
    out = str(tmp_path / "full.npy")
    mp.spawn(_worker, args=(2, _free_port(), num_robots, out), nprocs=2, join=True)
    got = np.load(out)
    states = np.random.default_rng(7).normal(size=(num_robots, 24))
    want = _fake_solve(np.arange(num_robots), states)
    assert got.shape == want.shape and np.array_equal(got, want)


def test_shard_bounds_cover_without_overlap():
    for n in (0, 1, 5, 8192, 8191):
        for world in (1, 2, 4, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[k][1] == spans[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)
