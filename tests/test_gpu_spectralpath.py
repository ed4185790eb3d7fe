"""Integrator "spectralpath" (src/integrators/spectralpath.cpp:198-342) on the
device: per camera sample one path per wavelength band, every band drawing from the sample's one
Halton stream where the previous band stopped, each band's bins taken from its own path.

Fixtures (tests/golden/make_golden.py: BAND_SCENES) are the unmodified reference's render of the
same scene file and the reference camera's rays at every band's wavelength.
"""
import numpy as np
import pytest

from conftest import blob_path, golden

pytestmark = pytest.mark.gpu

BAND_SCENES = ["cornell_lens_ca", "cornell_bands3", "cornell_bands1", "texbox_bands"]


@pytest.fixture(scope="module", params=BAND_SCENES)
def band_scene(request, capi, gpu_ctx):
    name = request.param
    sf = capi.SceneFile(blob_path(name))
    gpu_ctx.upload(sf)
    yield name, golden(name), gpu_ctx
    sf.close()


def test_camera_rays_at_band_wavelengths(band_scene):
    """RealisticCamera with chromaticAberrationEnabled refracts by wavelength
    (cameras/realistic.cpp:352-358); the perspective camera ignores it."""
    name, g, ctx = band_scene
    differs = False
    for lam, ref in zip(g["wavelengths"], g["cam_rays"]):
        rays, w = ctx.camera_rays(g["cam_samples"], wavelength=float(lam))
        assert np.array_equal(w > 0, ref[:, 7] > 0), f"{name} at {lam} nm: vignetting differs"
        ok = ref[:, 7] > 0
        np.testing.assert_allclose(w[ok], ref[ok, 7], rtol=2e-6)
        np.testing.assert_allclose(rays["o"][ok], ref[ok, 0:3], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(rays["d"][ok], ref[ok, 4:7], rtol=1e-6, atol=1e-7)
        differs |= not np.array_equal(ref[:, 4:7], g["cam_rays"][0][:, 4:7])
    if name == "cornell_lens_ca":
        assert differs, "the fixture's bands should leave the lens in different directions"


def test_render_matches_reference_dat(band_scene):
    name, g, ctx = band_scene
    w, h = [int(v) for v in g["res"]]
    spp = int(g["spp"])
    ctx.film_clear()
    ctx.render(0, spp)
    planes = ctx.film_pack_dat(w, h)
    ref = g["dat"].astype(np.float64)
    assert planes.shape == ref.shape
    # same bar as test_gpu_parity.test_render_matches_reference_dat: identical streams, so only
    # samples sitting exactly on a discrete threshold may differ (at most 0.05 % of the pixels)
    tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
    flipped = np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * ref.shape[0])
    assert flipped.sum() <= max(1, int(5e-4 * flipped.size)), f"{name}: {flipped.sum()} pixels differ"
    keep = ~flipped
    ref_mean = ref[:, keep].mean(axis=1)
    rel_rmse = np.sqrt((((planes - ref) ** 2)[:, keep]).mean(axis=1)) / np.maximum(ref_mean, 1e-12)
    assert rel_rmse[ref_mean > 0].max() < 0.005, f"{name}: per-bin relRMSE {rel_rmse.max():.4%}"
    # bins no band owns stay black (3 bands: round(31 / 3) = 10 bins each, bin 30 unassigned)
    assert np.array_equal(planes[ref_mean == 0], ref[ref_mean == 0])
    if name == "cornell_bands3":
        assert (ref_mean == 0).sum() == 1 and ref_mean[30] == 0


def test_bands_share_one_sampler_stream(band_scene):
    """Rendering the samples in two calls, or with a small wavefront, gives the same film: a band's
    starting dimension is carried per path, not per launch."""
    name, g, ctx = band_scene
    w, h = [int(v) for v in g["res"]]
    spp = int(g["spp"])
    ctx.film_clear()
    ctx.render(0, spp)
    whole = ctx.film_download(w, h)[0].copy()
    ctx.film_clear()
    ctx.set_wavefront_size(3000)
    ctx.render(0, spp // 2)
    ctx.render(spp // 2, spp)
    ctx.set_wavefront_size(0)
    split = ctx.film_download(w, h)[0]
    np.testing.assert_allclose(split, whole, rtol=1e-5, atol=1e-7)
