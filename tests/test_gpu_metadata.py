"""Integrator "metadata" (src/integrators/metadata.cpp): the standard sample loop
with an Li that returns the first hit's depth, materialId, instanceId or world position.  Fixtures
are the unmodified reference's renders (tests/golden/make_golden.py: META_SCENES)."""
import numpy as np
import pytest

from conftest import blob_path, golden

pytestmark = pytest.mark.gpu

META_SCENES = ["meta_depth", "meta_mesh", "meta_material", "meta_coordinates"]


@pytest.mark.parametrize("name", META_SCENES)
def test_metadata_render_matches_reference(capi, gpu_ctx, name):
    g = golden(name)
    sf = capi.SceneFile(blob_path(name))
    gpu_ctx.upload(sf)
    w, h = [int(v) for v in g["res"]]
    gpu_ctx.film_clear()
    gpu_ctx.render(0, int(g["spp"]))
    planes = gpu_ctx.film_pack_dat(w, h)
    ref = g["dat"].astype(np.float64)
    assert planes.shape == ref.shape
    # hit ids are bit-exact, so every pixel sees the same primitives; what is left is float rounding
    # of the hit point (fixtures are float32)
    scale = np.abs(ref).max()
    np.testing.assert_allclose(planes, ref, rtol=2e-6, atol=2e-6 * scale)
    if name in ("meta_mesh", "meta_material"):
        assert np.array_equal(planes.astype(np.float32), g["dat"])  # ids: exact
    if name != "meta_coordinates":
        assert np.array_equal(planes[0], planes[30])  # Spectrum(v): every bin
    else:
        assert not planes[3:].any()
    sf.close()


def test_metadata_needs_no_sampler_dimensions_beyond_the_camera(capi, gpu_ctx):
    """Sample ranges add up (the film is linear in the samples) also for this integrator."""
    g = golden("meta_depth")
    sf = capi.SceneFile(blob_path("meta_depth"))
    gpu_ctx.upload(sf)
    w, h = [int(v) for v in g["res"]]
    spp = int(g["spp"])
    gpu_ctx.film_clear()
    gpu_ctx.render(0, spp)
    whole = gpu_ctx.film_download(w, h)[0].copy()
    gpu_ctx.film_clear()
    gpu_ctx.render(0, 1)
    gpu_ctx.render(1, spp)
    np.testing.assert_allclose(gpu_ctx.film_download(w, h)[0], whole, rtol=1e-6, atol=1e-6)
    sf.close()
