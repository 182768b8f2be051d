"""CPU: BatchedHostSolver has no CPU fallback -- without a CUDA device its constructor fails loudly -- and its chunking
is the contiguous robot split of sharding.shard_bounds."""
import pytest
import torch

from quadruped_pympc_tamols_b200 import _cabi
from quadruped_pympc_tamols_b200.batched_host import BatchedHostSolver
from quadruped_pympc_tamols_b200.sharding import shard_bounds


def test_constructor_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_cabi.QpmpcError, match="no CPU fallback"):
        BatchedHostSolver(8, mass=10.0)


def test_chunk_bounds_cover_the_robots_once():
    for robots, chunks in ((8192, 8), (37, 4), (5, 1), (1024, 3)):
        b = [shard_bounds(robots, chunks, k) for k in range(chunks)]
        assert b[0][0] == 0 and b[-1][1] == robots
        assert all(b[k][1] == b[k + 1][0] for k in range(chunks - 1))
        assert max(hi - lo for lo, hi in b) - min(hi - lo for lo, hi in b) <= 1
