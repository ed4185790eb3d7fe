"""sptx_intersect_batch_noalloc (include/sptx.h): the ray-batch entry point whose device scratch
is a fixed file-scope device array instead of cudaMalloc'd buffers.  Checked through the C ABI
against the allocating entry point (sptx_intersect_batch), bit for bit, on calls in a row that
share the scratch.  Collected last on purpose."""
import numpy as np
import pytest

from conftest import blob_path, golden

pytestmark = pytest.mark.gpu


def _same(a, b):
    for f in ("prim", "instance", "t", "b0", "b1", "b2"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


def test_noalloc_ray_batch_matches_the_allocating_entry_point(capi):
    g = golden("cornell64")
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path("cornell64"))
    ctx.upload(sf)
    sf.close()
    rays = np.ascontiguousarray(g["rays"]).view(capi.RAY_DTYPE).reshape(-1)
    assert 0 < len(rays) <= capi.NOALLOC_MAX_RAYS
    # call 1: closest hit on the whole batch
    want = ctx.intersect(rays)
    got = ctx.intersect_noalloc(rays)
    _same(got, want)
    assert np.array_equal(got["prim"], g["hit_prim"])  # and therefore the reference's ids
    # call 2, right behind it on the same scratch: any hit, on a shorter, reversed batch (the
    # scratch still holds call 1's rays and hits beyond this batch's length)
    sub = np.ascontiguousarray(rays[::-1][: len(rays) // 3])
    want_any = ctx.intersect(sub, any_hit=True)
    got_any = ctx.intersect_noalloc(sub, any_hit=True)
    assert np.array_equal(got_any["prim"], want_any["prim"])
    assert np.array_equal(got_any["t"].view(np.uint32), want_any["t"].view(np.uint32))
    # call 3: closest hit again; nothing of call 2 leaks into it
    _same(ctx.intersect_noalloc(rays), want)
    # a full-size batch (the scratch's capacity) and one ray more
    reps = -(-capi.NOALLOC_MAX_RAYS // len(rays))
    full = np.ascontiguousarray(np.tile(rays, reps)[: capi.NOALLOC_MAX_RAYS])
    _same(ctx.intersect_noalloc(full), ctx.intersect(full))
    over = np.ascontiguousarray(np.tile(rays, reps + 1)[: capi.NOALLOC_MAX_RAYS + 1])
    with pytest.raises(capi.SptxError, match="SPTX_NOALLOC_MAX_RAYS"):
        ctx.intersect_noalloc(over)
    # empty batch: nothing to do, no error
    assert len(ctx.intersect_noalloc(rays[:0])) == 0
    ctx.close()
