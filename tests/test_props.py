"""Property-based checks (hypothesis) of the host-side invariants the device path relies on: the block-cyclic shard
map, the flat arena layout, the per-worker data partition (reference: worker.py:166-179) and the staleness weight
(reference: server.py:173-178)."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from dps_b200.models.resnet import build_model
from dps_b200.rpc.servicer import staleness_weight
from dps_b200.store.shardmap import build_layout
from dps_b200.utils.data import partition_range

_MODEL = build_model(100, seed=0)


@settings(max_examples=12, deadline=None)
@given(world=st.integers(1, 32), block=st.sampled_from([256, 1024, 2048]))
def test_layout_and_shard_map(world, block):
    L = build_layout(_MODEL, world, block)
    assert L.total % (world * block) == 0 and L.shard_elems * world == L.total
    # entries are in forward order, disjoint, inside the arena, and cover all 62 trainable tensors
    assert len(L.order) == 62
    end = 0
    for n in L.order:
        e = L.entries[n]
        assert e.offset >= end and e.span >= e.numel
        end = e.offset + e.span
    assert end <= L.total
    # shard map: to_shards / from_shards are inverse permutations; owner_of agrees with them
    flat = torch.arange(L.total, dtype=torch.float32)
    sh = L.to_shards(flat)
    assert sh.shape == (world, L.shard_elems)
    assert torch.equal(L.from_shards(sh), flat)
    rng = np.random.default_rng(world * 7 + block)
    for g in rng.integers(0, L.total, size=16):
        o, l = L.owner_of(int(g))
        assert 0 <= o < world and float(sh[o, l]) == float(g)
    # a pull streams the arena front to back from all owners at once: consecutive blocks rotate over the owners
    assert [L.owner_of(k * block)[0] for k in range(min(2 * world, L.total // block))] == \
        [k % world for k in range(min(2 * world, L.total // block))]


@settings(max_examples=10, deadline=None)
@given(world=st.integers(1, 8))
def test_state_dict_views_roundtrip(world):
    L = build_layout(_MODEL, world)
    flat = torch.zeros(L.total)
    sd = {k: v.detach() for k, v in _MODEL.named_parameters()}
    L.flatten_into(flat, sd)
    back = L.unflatten(flat)
    assert set(back) == set(sd)
    for k in sd:
        assert back[k].shape == sd[k].shape and torch.equal(back[k], sd[k]), k


@given(n=st.integers(0, 100_000), w=st.integers(1, 32))
def test_partition_is_disjoint_contiguous_complete(n, w):
    prev_end = 0
    for i in range(w):
        s, e = partition_range(i, w, n)
        assert s == prev_end and e >= s
        if i < w - 1:
            assert e - s == n // w
        prev_end = e
    assert prev_end == n                                  # the last worker takes the remainder
    assert partition_range(w, w, n) == (n, n)             # ids >= W own nothing
    assert partition_range(-1, w, n) == (n, n)


@given(s=st.integers(0, 1000), alpha=st.floats(0.01, 1.0), floor=st.floats(0.0, 0.5))
def test_staleness_weight_properties(s, alpha, floor):
    w = staleness_weight(s, alpha, floor)
    assert floor <= w <= 1.0 or (s == 0 and w == 1.0)
    assert staleness_weight(0, alpha, floor) == pytest.approx(max(floor, 1.0))
    assert staleness_weight(s + 1, alpha, floor) <= w + 1e-12          # monotone non-increasing
    if s <= 5:
        assert staleness_weight(s) == pytest.approx(max(0.1, 1.0 / (1.0 + 0.1 * s)))   # reference constants
