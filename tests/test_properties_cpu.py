"""Property tests (hypothesis) of the host-side arithmetic the GPU paths rely on: sampler sharding against torch's own
DistributedSampler, the block sampler's coverage, gradient-bucket plans against torch DDP's, cron matching against a
minute-by-minute walk, and the parallel checkpoint archive against torch.load for arbitrary nested states."""
import datetime as dt
import os

import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from ray_torch_distributed_checkpoint_b200 import data
from ray_torch_distributed_checkpoint_b200.data.loader import BlockSampler
from ray_torch_distributed_checkpoint_b200.flow.core import CronSchedule

_FAST = settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])


@_FAST
@given(n=st.integers(1, 300), world=st.integers(1, 9), shuffle=st.booleans(), drop_last=st.booleans(), seed=st.integers(0, 1000),
       epoch=st.integers(0, 5), data_=st.data())
def test_distributed_sampler_equals_torch(n, world, shuffle, drop_last, seed, epoch, data_):
    rank = data_.draw(st.integers(0, world - 1))
    ours = data.DistributedSampler(n, world, rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
    ours.set_epoch(epoch)
    ref = torch.utils.data.distributed.DistributedSampler(range(n), num_replicas=world, rank=rank, shuffle=shuffle, seed=seed,
                                                         drop_last=drop_last)
    ref.set_epoch(epoch)
    want = list(ref)
    assert ours.indices() == want and len(ours) == len(ref)
    assert ours.indices_tensor().tolist() == want


@_FAST
@given(n=st.integers(1, 5000), block=st.integers(1, 700), world=st.integers(1, 8), shuffle=st.booleans(), epoch=st.integers(0, 3))
def test_block_sampler_covers_every_row_with_equal_shards(n, block, world, shuffle, epoch):
    shards = []
    for r in range(world):
        s = BlockSampler(n, block, world, r, shuffle=shuffle, seed=7)
        s.set_epoch(epoch)
        starts = s.block_starts()
        assert all(0 <= b and b + s.block <= n for b in starts)              # every block lies inside the dataset
        assert len(s) == len(starts) * s.block
        shards.append(starts)
    assert len({len(x) for x in shards}) == 1                                # same number of rows on every rank
    covered = set()
    for starts in shards:
        for b in starts:
            covered.update(range(b, b + min(block, n)))
    assert covered == set(range(n))                                          # nothing is left out of an epoch


@_FAST
@given(sizes=st.lists(st.integers(1, 3_000_000), min_size=1, max_size=40), cap_mb=st.sampled_from([1, 5, 25]))
def test_bucket_plan_equals_torch_ddp(sizes, cap_mb):
    """sizes are in reduction order (reverse registration order), as both planners receive them."""
    import torch.distributed as dist
    from ray_torch_distributed_checkpoint_b200.parallel import buckets
    tensors = [torch.empty(s, device="meta") for s in sizes]
    first, cap = 1024 * 1024, cap_mb * 1024 * 1024
    ref, _ = dist._compute_bucket_assignment_by_size(tensors, [first, cap])
    ours = buckets.assign_buckets([t.numel() * 4 for t in tensors], first, cap)
    assert ours == [list(b) for b in ref]


_FIELD = lambda lo, hi: st.one_of(
    st.just("*"),
    st.integers(lo, hi).map(str),
    st.tuples(st.integers(lo, hi), st.integers(lo, hi)).map(lambda ab: f"{min(ab)}-{max(ab)}"),
    st.integers(1, max(1, (hi - lo) // 2)).map(lambda s: f"*/{s}"),
    st.lists(st.integers(lo, hi), min_size=2, max_size=3).map(lambda xs: ",".join(map(str, xs))))


@_FAST
@given(mi=_FIELD(0, 59), ho=_FIELD(0, 23), dom=_FIELD(1, 28), mo=st.sampled_from(["*", "*/2", "1-6", "7,8,12"]), dow=_FIELD(0, 7),
       start=st.integers(0, 500_000))
def test_cron_next_after_is_the_first_matching_minute(mi, ho, dom, mo, dow, start):
    c = CronSchedule(f"{mi} {ho} {dom} {mo} {dow}")
    t0 = dt.datetime(2026, 1, 1) + dt.timedelta(minutes=start)
    nxt = c.next_after(t0)
    assert nxt > t0 and c.matches(nxt) and nxt.second == 0
    # nothing earlier matches (walk at most two days minute by minute; longer gaps are checked at the day level)
    t = t0.replace(second=0, microsecond=0) + dt.timedelta(minutes=1)
    limit = min(nxt, t + dt.timedelta(days=2))
    while t < limit:
        assert not c.matches(t), (c.expr, t)
        t += dt.timedelta(minutes=1)


def test_cron_leap_day_is_found_years_ahead():
    c = CronSchedule("0 0 29 2 *")
    assert c.next_after(dt.datetime(2025, 3, 1)) == dt.datetime(2028, 2, 29)
    assert c.next_after(dt.datetime(1896, 3, 1)) == dt.datetime(1904, 2, 29)


_LEAF = st.one_of(st.integers(-5, 5), st.floats(allow_nan=False, allow_infinity=False, width=32), st.text(max_size=5), st.none(),
                  st.tuples(st.integers(0, 40), st.sampled_from([torch.float32, torch.bfloat16, torch.int64, torch.uint8, torch.float16])))


def _build(node):
    if isinstance(node, tuple) and len(node) == 2 and isinstance(node[1], torch.dtype):
        n, dtype = node
        return (torch.arange(n) % 251).to(dtype)
    if isinstance(node, dict):
        return {k: _build(v) for k, v in node.items()}
    if isinstance(node, list):
        return [_build(v) for v in node]
    return node


def _same(a, b):
    if isinstance(a, torch.Tensor):
        return isinstance(b, torch.Tensor) and a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b)
    if isinstance(a, dict):
        return isinstance(b, dict) and list(a) == list(b) and all(_same(a[k], b[k]) for k in a)
    if isinstance(a, list):
        return isinstance(b, list) and len(a) == len(b) and all(_same(x, y) for x, y in zip(a, b))
    return a == b


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(tree=st.recursive(_LEAF, lambda kids: st.one_of(st.lists(kids, max_size=4), st.dictionaries(st.text(min_size=1, max_size=4), kids, max_size=4)),
                         max_leaves=12), threads=st.integers(1, 4), chunk=st.sampled_from([16, 64, 1 << 20]))
def test_parallel_archive_round_trips_arbitrary_states(tmp_path, tree, threads, chunk):
    import zipfile
    from ray_torch_distributed_checkpoint_b200.ckpt import fast_zip
    state = {"root": _build(tree)}
    path = str(tmp_path / "latest_model.pt")
    fast_zip.save(state, path, threads=threads, chunk_bytes=chunk)
    assert zipfile.ZipFile(path).testzip() is None
    assert _same(torch.load(path, weights_only=True), state)
    os.remove(path)
