"""srt_trace_rays_static -- the allocation-free variant of srt_trace_rays -- against srt_trace_rays, through the C-ABI.

No counterpart in the reference: the entry point was added for callers that cannot
call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Its device staging (rays in, hit
index and distance out) is a file-scope `static __device__` array in csrc/srt_kernels.cu, reached through
cudaGetSymbolAddress once per device; every call shares it and it keeps the last call's contents.  So the checks are made on
TWO CALLS IN A ROW with different batches (a long one, then a shorter one whose results must not be the first call's
left-overs), bit for bit -- the tolerance of this repo's integer / index tests is zero, and the distances come from the same
kernel, so they are compared as bit patterns too.

This file sorts last on purpose (a failure here hides nothing after it) and passes both alone
in a fresh process and after the rest of the suite (nothing it needs is left over from other tests: own context, own scene).
"""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def make_rays(pkg, n, seed):
    """Rays at the Rubik model: from points around it towards points inside its bounds, every fifth with a finite
    intersection_distance, a few aimed away (misses)."""
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, pkg.RAY_DT)
    origin = rng.uniform(-40.0, 40.0, (n, 3)).astype(np.float32)
    origin[:, 1] = np.abs(origin[:, 1])
    target = np.stack([rng.uniform(-9.0, 9.0, n), rng.uniform(0.0, 18.0, n), rng.uniform(-2.5, 9.0, n)], 1).astype(np.float32)
    d = target - origin
    d[::17] *= -1.0  # these look away from the model
    rays["origin"] = origin
    rays["direction"] = d
    rays["dist"] = np.float32(1e30)
    rays["dist"][::5] = rng.uniform(0.2, 1.2, len(rays["dist"][::5])).astype(np.float32)  # in units of |direction|
    return rays


def test_static_scratch_variant_equals_the_allocating_one_on_two_calls_in_a_row(pkg):
    ctx = pkg.Context(0)
    try:
        ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
        ctx.upload_scene(scenes.rubik_scene())
        first, second = make_rays(pkg, 50000, 1), make_rays(pkg, 1237, 2)
        for rays in (first, second):  # the second batch is shorter and different: stale results would show
            want_hit, want_t = ctx.trace_rays(rays)
            got_hit, got_t = ctx.trace_rays_static(rays)
            assert np.array_equal(got_hit, want_hit)
            assert np.array_equal(got_t.view(np.uint32), want_t.view(np.uint32))
            assert (want_hit != pkg.NO_HIT).mean() > 0.2 and (want_hit == pkg.NO_HIT).any()
        assert not np.array_equal(ctx.trace_rays_static(first)[0][:len(second)], ctx.trace_rays_static(second)[0])
        # the ordered traversal goes through the same scratch
        ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_ORDERED)
        want_hit, want_t = ctx.trace_rays(second)
        got_hit, got_t = ctx.trace_rays_static(second)
        assert np.array_equal(got_hit, want_hit) and np.array_equal(got_t.view(np.uint32), want_t.view(np.uint32))
    finally:
        ctx.close()


def test_static_scratch_variant_states_its_limit(pkg):
    ctx = pkg.Context(0)
    try:
        ctx.upload_scene(scenes.rubik_scene())
        at_limit = make_rays(pkg, pkg.TRACE_STATIC_MAX_RAYS, 3)
        hit, t = ctx.trace_rays_static(at_limit)
        want_hit, want_t = ctx.trace_rays(at_limit)
        assert np.array_equal(hit, want_hit) and np.array_equal(t.view(np.uint32), want_t.view(np.uint32))
        with pytest.raises(pkg.SrtError, match="65537 rays exceed the static scratch's limit of 65536"):
            ctx.trace_rays_static(make_rays(pkg, pkg.TRACE_STATIC_MAX_RAYS + 1, 4))
        hit, _ = ctx.trace_rays_static(at_limit[:10])  # and the context still works after the refusal
        assert np.array_equal(hit, want_hit[:10])
        assert len(ctx.trace_rays_static(at_limit[:0])[0]) == 0
    finally:
        ctx.close()


def test_static_scratch_variant_needs_a_scene(pkg):
    ctx = pkg.Context(0)
    try:
        with pytest.raises(pkg.SrtError, match="no scene uploaded"):
            ctx.trace_rays_static(make_rays(pkg, 8, 5))
    finally:
        ctx.close()
