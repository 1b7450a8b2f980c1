"""pbrtgpu_intersect_closest_noalloc (include/pbrtgpu.h): the closest-hit entry point whose device scratch is a
fixed file-scope array instead of cudaMalloc'ed buffers.  It has no counterpart in the reference,
so the oracle here is the existing entry point, pbrtgpu_intersect_closest, which
the tier-1 tests pin to the reference's BVHAccel::Intersect.  Tolerance: the one those tests use -- bit-exact IDs,
t and barycentrics.

Collected last on purpose.  The test passes alone in a fresh process and after
the other GPU tests: it shares their session context and depends on nothing they leave behind.
"""
import numpy as np
import pytest

import util
from conftest import scene_path

pytestmark = pytest.mark.gpu


def _same(a, b):
    assert np.array_equal(a[0], b[0])
    assert np.array_equal(util.bits(a[1]), util.bits(b[1]))
    assert np.array_equal(util.bits(a[2]), util.bits(b[2]))


def test_noalloc_entry_point_matches_the_allocating_one_twice_in_a_row(pkg, gpu):
    hs = pkg.HostScene(scene_path("tri_soup.pbrt"))
    gpu.upload(hs)
    first = np.concatenate([util.rays_incoherent_closest(hs.view, 12000, seed=101),
                            util.rays_toward_box(12000, [-10, -10, -10], [10, 10, 10], seed=102),
                            util.rays_edge_cases(hs.view, 4000, seed=103)])
    assert len(first) <= pkg.NOALLOC_MAX_RAYS
    # second call: fewer rays of another kind -- what the first call left in the shared array must not show through
    second = util.rays_toward_box(5000, [-10, -10, -10], [10, 10, 10], seed=104)
    for rays in (first, second):
        want = gpu.intersect_closest(rays)
        got = gpu.intersect_closest_noalloc(rays)
        assert (want[0] >= 0).sum() > len(rays) // 10  # a real mix of hits and misses
        _same(got, want)
    # and the very same call again: the queue cursor is cleared by every call
    _same(gpu.intersect_closest_noalloc(second), gpu.intersect_closest(second))


def test_noalloc_entry_point_limits(pkg, gpu):
    hs = pkg.HostScene(scene_path("tri_soup.pbrt"))
    gpu.upload(hs)
    full = util.rays_toward_box(pkg.NOALLOC_MAX_RAYS, [-10, -10, -10], [10, 10, 10], seed=105)
    _same(gpu.intersect_closest_noalloc(full), gpu.intersect_closest(full))  # exactly the maximum is served
    with pytest.raises(pkg.PbrtGpuError, match="PBRTGPU_NOALLOC_MAX_RAYS"):
        gpu.intersect_closest_noalloc(np.concatenate([full, full[:1]]))
    ids, t, b = gpu.intersect_closest_noalloc(np.zeros((0, 8), np.float32))
    assert len(ids) == 0
