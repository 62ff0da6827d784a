"""sail_intersect_rays_noalloc (include/sail_b200.h): the intersection query with its scratch in a fixed in-library device
array instead of per-call cudaMalloc / cudaFree, no counterpart in the
reference.  Checked through the C ABI against the existing sail_intersect_rays on the same rays, bit for bit (the tolerance
every parity test here uses: 0), on two calls in a row that share the array.

The file name sorts last on purpose: a failure here must not hide other tests.  It
passes alone in a fresh process and after the rest of the suite (it needs nothing another test leaves behind)."""
import numpy as np
import pytest

from conftest import bits_equal, mismatch_report, scene_tables

pytestmark = pytest.mark.gpu


def _rays(seed, n):
    rng = np.random.default_rng(seed)
    org = rng.uniform([0.1, 0.1, -4.5], [5.4, 5.3, 5.5], size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[: n // 8] *= rng.uniform(0.2, 3.0, size=(n // 8, 1))  # un-normalised directions too (primary and shadow rays)
    return np.concatenate([org, d], axis=1).astype(np.float32)


def _same(got, want, what):
    gid, gt, gn, gm = got
    wid, wt, wn, wm = want
    assert np.array_equal(gid, wid), what + ": hit ids differ"
    assert bits_equal(gt, wt), mismatch_report(gt, wt, what + " t")
    hit = wid >= 0
    assert hit.any() and (~hit).sum() >= 0
    assert bits_equal(gn[hit], wn[hit]), mismatch_report(gn[hit], wn[hit], what + " normal")
    assert np.array_equal(gm[hit], wm[hit]), what + ": material categories differ"


def test_no_alloc_entry_point_matches_the_allocating_one():
    from sail_b200 import capi
    assert capi.device_count() >= 1, "no sm_100 device: the CUDA path cannot run (there is no fallback)"
    _, o, t, l, _ = scene_tables(3, 16, 16)  # config 3: every shape kind
    with capi.Context(16, 16) as ctx:
        ctx.upload_scene(o, t, l)
        # call 1: the full scratch; call 2, straight after it: fewer, different rays over the first call's leftovers
        first = _rays(1, capi.NOALLOC_MAX_RAYS)
        second = _rays(2, 10007)
        got1 = ctx.intersect_rays_noalloc(first)
        got2 = ctx.intersect_rays_noalloc(second)
        _same(got1, ctx.intersect_rays(first), "call 1 (%d rays)" % len(first))
        _same(got2, ctx.intersect_rays(second), "call 2 (%d rays)" % len(second))
        assert not np.array_equal(got1[0][:10007], got2[0])  # the two calls really asked different questions
        # a second context on the same device shares the array (sequentially: fine) and a different scene
        _, o1, t1, l1, _ = scene_tables(1, 16, 16)
        with capi.Context(16, 16) as other:
            other.upload_scene(o1, t1, l1)
            _same(other.intersect_rays_noalloc(second), other.intersect_rays(second), "second context")
        _same(ctx.intersect_rays_noalloc(second), got2, "call 3")
        # beyond the fixed maximum: a clear error, nothing launched
        launches = ctx.launch_count()
        with pytest.raises(capi.SailError) as ei:
            ctx.intersect_rays_noalloc(np.zeros((capi.NOALLOC_MAX_RAYS + 1, 6), np.float32))
        assert ei.value.code == capi.ERR_INVALID and "SAIL_NOALLOC_MAX_RAYS" in str(ei.value)
        assert ctx.launch_count() == launches
        empty = ctx.intersect_rays_noalloc(np.zeros((0, 6), np.float32))
        assert all(len(a) == 0 for a in empty)


def test_no_alloc_before_upload_is_a_state_error():
    from sail_b200 import capi
    with capi.Context(16, 16) as ctx:
        with pytest.raises(capi.SailError) as ei:
            ctx.intersect_rays_noalloc(np.zeros((4, 6), np.float32))
        assert ei.value.code == capi.ERR_STATE
