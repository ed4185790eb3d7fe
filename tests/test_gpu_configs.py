"""Parity at the full sizes of BASELINE.json's configurations (C1 .. C5).

The fixtures are reference outputs reduced to a committable size by tools/make_configs.py
(block means of the reference's .dat; reference hits of a seeded sub-batch).  The two large scene
blobs (C1, C4) are committed xz-compressed and unpacked on first use (capi.unpack_blob); no test
here skips.
"""
import os
import sys

import numpy as np
import pytest

from conftest import GEN, ROOT, golden

sys.path.insert(0, os.path.join(ROOT, "tools"))
import config_rays  # noqa: E402

pytestmark = pytest.mark.gpu


def _ctx(capi, name):
    p = capi.unpack_blob(os.path.join(GEN, name + ".sptx"))
    assert p, f"{name}.sptx(.xz) is missing from scenes/_gen (tools/make_configs.py makes it)"
    ctx = capi.Context(0)
    sf = capi.SceneFile(p)
    ctx.upload(sf)
    sf.close()
    return ctx


def _block_means(planes, k):
    b, h, w = planes.shape
    return planes.reshape(b, h // k, k, w // k, k).mean(axis=(2, 4))


def _check_render(ctx, g, tol):
    w, h = [int(v) for v in g["res"]]
    spp, k = int(g["spp"]), int(g["block"])
    ctx.film_clear()
    ctx.render(0, spp)
    planes = ctx.film_pack_dat(w, h)
    # whole-image energy per bin: float32 film sums against the reference's
    sums = planes.sum(axis=(1, 2))
    assert np.allclose(sums, g["bin_sums"], rtol=1e-4), np.abs(sums / g["bin_sums"] - 1).max()
    mine, ref = _block_means(planes, k), g["means"].astype(np.float64)
    rel_rmse = np.sqrt(((mine - ref) ** 2).mean(axis=(1, 2))) / np.maximum(ref.mean(axis=(1, 2)), 1e-12)
    assert rel_rmse.max() < tol, f"per-bin relRMSE of block means {rel_rmse.max():.4%}"
    return planes


def test_c1_killeroo_400x400_64spp(capi):
    """BASELINE config 1: the reference's own CPU-runnable case, at its full size."""
    ctx = _ctx(capi, "killeroo400")
    _check_render(ctx, golden("c1_killeroo400"), 0.005)
    ctx.close()


def test_c2_cornell_1024(capi):
    """BASELINE config 2 (the bench workload) at its resolution, first 4 sample numbers of the same
    Halton streams as the reference."""
    ctx = _ctx(capi, "cornell1024")
    _check_render(ctx, golden("c2_cornell1024"), 0.005)
    ctx.close()


def test_c5_cornell_4k_sample_partitions_sum_to_the_whole(capi):
    """BASELINE config 5 geometry (3840x2160): the reference's 2-spp render, and partitioning B of
    SURVEY 8(e) on one device -- two 'ranks' render one sample number each into their own film,
    the films' sum is the whole render (what sptx_film_merge_peers adds up across GPUs)."""
    ctx = _ctx(capi, "cornell3840x2160")
    g = golden("c5_cornell4k")
    w, h = [int(v) for v in g["res"]]
    whole = _check_render(ctx, g, 0.005)
    parts = []
    for r in range(2):
        ctx.film_clear()
        ctx.render(r, r + 1)
        parts.append(ctx.film_pack_dat(w, h))
    np.testing.assert_allclose(parts[0] + parts[1], whole, rtol=1e-5, atol=1e-7)
    ctx.close()


def test_c3_realistic_camera_2048(capi):
    """BASELINE config 3 geometry (2048x2048 double-Gauss) at 2 spp: same Halton streams as the reference."""
    ctx = _ctx(capi, "cornell_lens2048")
    _check_render(ctx, golden("c3_cornell_lens2048"), 0.005)
    ctx.close()


def test_c4_instanced_20m_triangles(capi):
    """BASELINE config 4: hit ids of the seeded incoherent batch are bit-exact on the sub-batch the
    reference traced; over the full 2^22-ray batch closest-hit and any-hit agree (size-independent)."""
    ctx = _ctx(capi, "instanced_c4")
    g = golden("c4_instanced")
    n = int(g["n"])
    rays = config_rays.incoherent_rays(n, g["lo"], g["hi"])
    hits = ctx.intersect(rays, any_hit=False)
    assert np.array_equal(hits["prim"], g["hit_prim"]), f"{(hits['prim'] != g['hit_prim']).sum()} ids differ"
    assert np.array_equal(hits["instance"], g["hit_inst"])
    hit = g["hit_prim"] >= 0
    rel = np.abs(hits["t"][hit] - g["hit_t"][hit]) / np.abs(g["hit_t"][hit])
    assert rel.max() <= 1e-5
    anyh = ctx.intersect(rays, any_hit=True)
    assert np.array_equal((anyh["prim"] >= 0).astype(np.int8), g["anyhit"])
    # full-size property: tMax = inf, so "some hit" == "closest hit exists"
    big = config_rays.incoherent_rays(1 << 22, g["lo"], g["hi"])
    a = ctx.intersect(big, any_hit=False)
    b = ctx.intersect(big, any_hit=True)
    assert np.array_equal(a["prim"] >= 0, b["prim"] >= 0)
    # the first rays of the big batch are the sub-batch
    assert np.array_equal(a["prim"][:n], g["hit_prim"])
    ctx.close()
