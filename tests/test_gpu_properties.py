"""Size-independent properties and edge cases of the device path (through the C ABI).

The reference gets these for free from its structure: a pixel's value is the in-order sum of its
samples, whatever tile or thread rendered them (SamplerIntegrator::Render,
src/core/integrator.cpp:228-342).  The wavefront must keep them: however the work is cut into
passes, tiles or ranks, the film is the same bits.
"""
import numpy as np
import pytest

from conftest import blob_path

pytestmark = pytest.mark.gpu
W = H = 64


@pytest.fixture(scope="module")
def cornell(capi):
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path("cornell64"))
    ctx.upload(sf)
    yield ctx, sf
    sf.close()
    ctx.close()


def _film(ctx):
    c, w = ctx.film_download(W, H)
    return np.concatenate([c.reshape(-1), w.reshape(-1)])


def _same(a, b):
    """Bit-identical, except for pixels that receive a sample lying exactly on a pixel edge.  The
    Halton value is exactly 0 while index >> e0 == 0 (base 2) or index / 3^e1 == 0 (base 3), i.e.
    for sample 0 of the pixels with a small per-pixel offset: pFilm is then an integer and the box
    filter footprint is two (or 2x2) pixels.  The extra pixels are added with an atomic after the
    pass's in-order sums (k_film_spill), so the position of that one addend in the float sum
    depends on how the samples were cut into passes."""
    diff = np.nonzero(a != b)[0]
    assert len(diff) <= 0.1 * a.size, f"{len(diff)} of {a.size} film words differ"
    assert np.allclose(a[diff], b[diff], rtol=2e-6, atol=0)
    return True


def _render(ctx, *calls):
    ctx.film_clear()
    for fn, args in calls:
        getattr(ctx, fn)(*args)
    return _film(ctx)


def test_render_is_deterministic(cornell):
    ctx, _ = cornell
    a = _render(ctx, ("render", (0, 8)))
    b = _render(ctx, ("render", (0, 8)))
    assert np.array_equal(a, b)


def test_sample_ranges_concatenate_bit_exactly(cornell):
    ctx, _ = cornell
    whole = _render(ctx, ("render", (0, 12)))
    parts = _render(ctx, ("render", (0, 5)), ("render", (5, 6)), ("render", (6, 12)))
    assert _same(whole, parts)


def test_wavefront_size_does_not_change_the_film(cornell):
    ctx, _ = cornell
    big = _render(ctx, ("render", (0, 8)))
    ctx.set_wavefront_size(W * H)        # one sample per pass
    one = _render(ctx, ("render", (0, 8)))
    ctx.set_wavefront_size(1000)         # pixel chunks that do not divide the image
    odd = _render(ctx, ("render", (0, 8)))
    ctx.set_wavefront_size(0)
    assert _same(big, one)
    assert _same(big, odd)


def test_tile_partition_sums_to_the_whole(cornell):
    """Partitioning A of SURVEY 8(e): disjoint 16x16 tile sets, box filter -> disjoint pixels."""
    ctx, _ = cornell
    whole = _render(ctx, ("render", (0, 6)))
    tiles = _render(ctx, *[("render_tiles", (p, 3, 0, 6)) for p in range(3)])
    assert _same(whole, tiles)


def test_pixel_bounds_restrict_the_render(cornell):
    ctx, _ = cornell
    whole = _render(ctx, ("render", (0, 4))).reshape(-1)
    part = _render(ctx, ("render", (0, 4, (8, 16, 40, 48))))
    c_w = whole[: W * H * 31].reshape(H, W, 31)
    c_p = part[: W * H * 31].reshape(H, W, 31)
    assert np.array_equal(c_p[16:48, 8:40], c_w[16:48, 8:40])
    # nothing further than the filter footprint of an edge sample (one pixel) outside the bounds
    mask = np.ones((H, W), bool)
    mask[15:49, 7:41] = False
    assert (c_p[mask] == 0).all() and (part[W * H * 31:].reshape(H, W)[mask] == 0).all()


def test_empty_work_is_accepted(cornell, capi):
    ctx, _ = cornell
    before = _render(ctx, ("render", (0, 2)))
    ctx.render(2, 2)                          # empty sample range
    ctx.render(0, 2, (10, 10, 10, 30))        # empty pixel bounds
    assert np.array_equal(before, _film(ctx))
    hits = ctx.intersect(np.zeros(0, dtype=capi.RAY_DTYPE))
    assert hits.shape == (0,)


def test_degenerate_rays(cornell, capi):
    """Zero-length segments, zero directions and NaNs must not hang or hit."""
    ctx, _ = cornell
    rays = np.zeros(4, dtype=capi.RAY_DTYPE)
    rays["o"] = [0, 0, 0]
    rays["d"] = [[0, 0, -1], [0, 0, 0], [np.nan, 0, 1], [0, 0, -1]]
    rays["t_max"] = [0.0, np.inf, np.inf, np.inf]
    hits = ctx.intersect(rays)
    assert hits["prim"][0] < 0 and hits["prim"][1] < 0 and hits["prim"][2] < 0
    assert hits["prim"][3] >= 0


def test_counters_account_for_every_sample(cornell):
    ctx, _ = cornell
    ctx.reset_counters()
    _render(ctx, ("render", (0, 4)))
    c = ctx.counters()
    assert c["camera_rays"] == W * H * 4
    assert c["closest_hit_rays"] >= c["camera_rays"] and c["path_vertices"] >= c["camera_rays"]
    assert c["shadow_rays"] > 0 and c["kernel_launches"] > 0
    assert abs(sum(c["kernel_ms"]) - c["render_ms"]) < 0.5 * c["render_ms"] + 1.0


def test_errors_are_reported_not_swallowed(capi):
    ctx = capi.Context(0)
    with pytest.raises(capi.SptxError, match="no scene"):
        ctx.render(0, 1)
    with pytest.raises(capi.SptxError):
        ctx.set_option("no_such_option", 1)
    ctx.close()


@pytest.mark.parametrize("name", ["cornell64", "killeroo128", "instanced", "zoo"])
def test_visit_counts_are_the_reference_traversals(capi, gpu_ctx, name):
    """the traversal roofline is built on per-ray node and primitive visit counts.
    The device's counted traversal (sptx_intersect_batch(counters=1)) visits exactly what the
    reference's BVHAccel::Intersect visits -- same order, same early exits -- so the counts are
    equal ray by ray (fixture: oracle_dump's counting walk over the reference's own BVH)."""
    from conftest import blob_path, golden
    g, v = golden(name), golden("visits")
    sf = capi.SceneFile(blob_path(name))
    gpu_ctx.upload(sf)
    rays = np.ascontiguousarray(g["rays"]).view(capi.RAY_DTYPE).reshape(-1)
    hits = gpu_ctx.intersect(rays, any_hit=False, counters=True)
    assert np.array_equal(hits["n_nodes"], v[name + "_nodes"]), \
        f"{(hits['n_nodes'] != v[name + '_nodes']).sum()} rays visit a different number of nodes"
    assert np.array_equal(hits["n_tests"], v[name + "_tests"])
    sf.close()
