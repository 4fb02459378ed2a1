"""Property tests of the sharding / reassembly rules (FSDP dim 0, train_ffns.py:265-266 and :284-287; MP W1 dim 0 /
W2 dim 1, :316-319 and :335-338) and of the seed split (:182)."""
import torch
from hypothesis import given, settings, strategies as st

from distributed_llm_code_samples_b200.parallel.base import fsdp_shard, tp_shard, tp_shard_dim
from distributed_llm_code_samples_b200.utils.data import split_seeds


@settings(max_examples=30, deadline=None)
@given(world=st.sampled_from([1, 2, 4, 8]), d=st.integers(1, 4), f=st.integers(1, 4))
def test_fsdp_shards_concatenate_back(world, d, f):
    D, F = 8 * world * d, 8 * world * f
    W1, W2 = torch.randn(F, D), torch.randn(D, F)
    for W in (W1, W2):
        shards = [fsdp_shard(W, r, world) for r in range(world)]
        assert all(s.shape[0] == W.shape[0] // world for s in shards)
        assert torch.equal(torch.cat(shards, dim=0), W)


@settings(max_examples=30, deadline=None)
@given(world=st.sampled_from([1, 2, 4, 8]), d=st.integers(1, 4), f=st.integers(1, 4))
def test_tp_shards_reassemble_and_sum_to_the_full_block(world, d, f):
    D, F = 8 * d, 8 * world * f
    W1, W2, b1, b2 = torch.randn(F, D), torch.randn(D, F), torch.randn(F), torch.randn(D)
    parts = [[tp_shard(p, j, r, world) for j, p in enumerate((W1, W2, b1, b2))] for r in range(world)]
    assert torch.equal(torch.cat([p[0] for p in parts], dim=0), W1)          # column-parallel: rows of W1
    assert torch.equal(torch.cat([p[1] for p in parts], dim=1), W2)          # row-parallel: columns of W2
    assert torch.equal(torch.cat([p[2] for p in parts], dim=0), b1)
    assert all(torch.equal(p[3], b2) for p in parts) and tp_shard_dim(3) is None
    # the Megatron identity the MP trainer relies on: sum_r relu(x W1_r^T) W2_r^T == relu(x W1^T) W2^T
    x = torch.randn(5, D)
    full = torch.relu(x @ W1.T) @ W2.T
    part = sum(torch.relu(x @ p[0].T) @ p[1].T for p in parts)
    assert torch.allclose(full, part, atol=1e-4)


@settings(max_examples=30, deadline=None)
@given(world=st.sampled_from([1, 2, 4, 8]), steps=st.integers(1, 6))
def test_seed_split_is_a_strided_partition(world, steps):
    seeds = torch.arange(world * steps)
    parts = split_seeds(seeds, world)
    assert [p.tolist() for p in parts] == [seeds[r::world].tolist() for r in range(world)]
