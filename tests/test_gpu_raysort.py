"""Ray binning in sptx_intersect_batch (set_option "ray_sort"): the walk order changes, nothing else --
hit ids, t, barycentrics and the per-ray visit counts are those of the unsorted walk, bit for bit, and
the reference's ids (tests/golden) with them."""
import numpy as np
import pytest

from conftest import SCENES, blob_path, golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["cornell64", "instanced", "killeroo128", "zoo"])
def scene(request, capi, gpu_ctx):
    name = request.param
    sf = capi.SceneFile(blob_path(name))
    gpu_ctx.upload(sf)
    yield name, golden(name), gpu_ctx
    sf.close()

FIELDS = ("prim", "instance", "t", "b0", "b1", "b2", "n_nodes", "n_tests")


def ctx_ray_dtype():
    import importlib
    return importlib.import_module("pbrt-v3-spectral_b200.capi").RAY_DTYPE


def _same(a, b, any_hit):
    for f in (("prim", "t", "n_nodes", "n_tests") if any_hit else FIELDS):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


@pytest.mark.parametrize("any_hit", [False, True])
def test_forced_sort_gives_identical_hits(scene, any_hit):
    name, g, ctx = scene
    rays = np.ascontiguousarray(g["rays"]).view(ctx_ray_dtype()).reshape(-1)
    try:
        ctx.set_option("ray_sort", 0)
        plain = ctx.intersect(rays, any_hit=any_hit, counters=True)
        assert not ctx.last_batch_sorted()
        ctx.set_option("ray_sort", 1)
        binned = ctx.intersect(rays, any_hit=any_hit, counters=True)
        assert ctx.last_batch_sorted()
    finally:
        ctx.set_option("ray_sort", 2)
    _same(plain, binned, any_hit)
    if not any_hit:
        assert np.array_equal(binned["prim"], g["hit_prim"])
        assert np.array_equal(binned["instance"], g["hit_inst"])


def test_auto_mode_sorts_large_batches_only(capi):
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path("instanced"))
    ctx.upload(sf)
    sf.close()
    g = golden("instanced")
    rng = np.random.RandomState(5)
    n = 1 << 17
    gr = np.ascontiguousarray(g["rays"]).view(capi.RAY_DTYPE).reshape(-1)
    lo, hi = gr["o"].min(axis=0), gr["o"].max(axis=0)
    rays = np.zeros(n, dtype=capi.RAY_DTYPE)
    rays["o"] = (lo + rng.rand(n, 3) * (hi - lo)).astype(np.float32)
    d = rng.randn(n, 3).astype(np.float32)
    rays["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays["t_max"] = np.inf
    rays["t_max"][::3] = np.float32(0.5)
    rays["o"][5] = np.nan                   # a broken ray must not break the binning of the others
    auto = ctx.intersect(rays, counters=True)
    assert ctx.last_batch_sorted()
    ctx.set_option("ray_sort", 0)
    plain = ctx.intersect(rays, counters=True)
    assert not ctx.last_batch_sorted()
    ctx.set_option("ray_sort", 2)
    _same(plain, auto, False)
    small = ctx.intersect(rays[:1000], counters=True)     # below the size threshold: caller's order
    assert not ctx.last_batch_sorted()
    _same(plain[:1000], small, False)
    # the no-allocation entry point never sorts (the sort needs cudaMalloc'd scratch)
    na = ctx.intersect_noalloc(rays[:capi.NOALLOC_MAX_RAYS])
    assert not ctx.last_batch_sorted()
    assert np.array_equal(na["prim"], plain["prim"][:capi.NOALLOC_MAX_RAYS])
    ctx.close()
