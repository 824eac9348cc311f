"""Bucket layout of the bucketed p2p parameter plane (parallel/bucketed.py): pure index arithmetic, no GPU."""
import torch

from bigdl_b200.models.inception import Inception_v1_NoAuxClassifier
from bigdl_b200.models.lenet import LeNet5
from bigdl_b200.parallel.bucketed import BucketedP2P, build_buckets, plan_units


def _check_tiling(buckets, ptot, P, world):
    pos = 0
    for b in buckets:
        assert b.off == pos and b.length > 0 and b.off % 8 == 0 and b.length % 8 == 0
        cover = 0
        for r in range(world):
            s, l = b.slices[r]
            assert s == b.off + cover and s % 8 == 0 and l % 8 == 0
            cover += l
        assert cover == b.length                       # the ranks' sub-slices tile the bucket
        pos += b.length
    assert pos == P
    for r in range(world):                             # partition offsets are the running sums of the sub-slice lengths
        run = 0
        for b in buckets:
            assert b.poff[r] == run
            run += b.slices[r][1]
        assert run == ptot[r]
    assert sum(ptot) == P


def test_units_follow_top_level_layers_and_tile_the_vector():
    m = Inception_v1_NoAuxClassifier(1000)
    m.getParameters()
    fl = m._flat
    units = plan_units(m, fl)
    assert len(units) > 10 and units[0][1] == 0
    assert sum(l for _, _, l in units) == fl.numel
    # an Inception block (Concat: branches on side streams) is never split
    from bigdl_b200.nn.layers_table import Concat
    assert any(isinstance(u[0], Concat) or any(isinstance(c, Concat) for c in u[0].all_modules()) for u in units)
    for world in (2, 3, 8):
        buckets, ptot = build_buckets(units, world, cap=512 * 1024, first_cap=64 * 1024)
        assert len(buckets) >= 4 and buckets[0].length <= 2 * 1024 * 1024
        _check_tiling(buckets, ptot, fl.numel, world)


def test_method_ranges_force_bucket_borders():
    m = LeNet5(10)
    m.getParameters()
    fl = m._flat
    units = plan_units(m, fl)
    P = fl.numel
    cut = units[len(units) // 2][1]
    ranges = [(0, cut), (cut, P - cut)]
    buckets, ptot = build_buckets(units, 4, cap=1 << 30, first_cap=1 << 30, method_ranges=ranges)
    _check_tiling(buckets, ptot, P, 4)
    assert [b.off for b in buckets] == [0, cut] and [b.method for b in buckets] == [0, 1]


def test_subtract_ranges():
    f = BucketedP2P._subtract
    assert f(0, 100, []) == [(0, 100)]
    assert f(0, 100, [(10, 20)]) == [(0, 10), (30, 70)]
    assert f(16, 64, [(0, 32), (64, 100)]) == [(32, 32)]
    assert f(0, 8, [(0, 8)]) == []
