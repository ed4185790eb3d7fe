"""The reference's own edge-case vectors for traversal (src/tests/shapes.cpp), run on the device:
Triangle.BadCases (:544-559), Triangle.Watertight (:28-129) and Triangle.Reintersect (:154-205).
Inputs are rebuilt by tests/golden/make_refvectors.py's generators (PCG32 as in src/core/rng.h), the
expected answers are the reference's (oracle_dump), kept in tests/golden/refvectors.npz."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLD, blob_path, golden

sys.path.insert(0, GOLD)
import make_refvectors as rv  # noqa: E402  (generators only; nothing of the oracle runs here)

pytestmark = pytest.mark.gpu


def _ctx(capi, name):
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path(name))
    ctx.upload(sf)
    sf.close()
    return ctx


def test_badcases_ray_misses_the_needle_triangle(capi):
    g = golden("refvectors")
    ctx = _ctx(capi, "rv_badcase")
    r = np.zeros(1, dtype=capi.RAY_DTYPE)
    r["o"], r["d"], r["t_max"] = rv.BADCASE_RAY
    hit = ctx.intersect(r)
    assert hit["prim"][0] == -1 == g["badcase_prim"][0]          # EXPECT_FALSE(mesh[0]->Intersect(...))
    assert hit["t"][0] == np.float32(0.9999)                     # tMax untouched
    # Triangle::IntersectP has no degenerate-triangle rejection: the reference reports a hit there
    anyh = ctx.intersect(r, any_hit=True)
    assert int(anyh["prim"][0] >= 0) == int(g["badcase_any"][0])
    ctx.close()


def test_watertight_mesh_never_leaks(capi):
    g = golden("refvectors")
    ctx = _ctx(capi, "rv_watertight")
    verts = g["watertight_verts"]
    rays = rv.watertight_rays(verts).view(capi.RAY_DTYPE)
    assert len(rays) == 200000
    hits = ctx.intersect(rays)
    assert (hits["prim"] >= 0).all(), f"{(hits['prim'] < 0).sum()} rays leak through the mesh"
    assert np.array_equal(hits["prim"], g["watertight_prim"])    # and the same triangle as the reference
    assert np.array_equal(hits["t"].view(np.uint32), g["watertight_t"].view(np.uint32))
    anyh = ctx.intersect(rays, any_hit=True)
    assert (anyh["prim"] >= 0).all()
    ctx.close()


def test_spawned_rays_never_rehit_their_triangle(capi):
    g = golden("refvectors")
    ctx = _ctx(capi, "rv_watertight")
    rays = np.ascontiguousarray(g["spawn_rays"]).view(capi.RAY_DTYPE).reshape(-1)
    parent = g["spawn_parent"]
    hits = ctx.intersect(rays)
    assert not (hits["prim"] == parent).any(), "a spawned ray re-hits the triangle it left"
    assert np.array_equal(hits["prim"], g["spawn_prim"])
    assert np.array_equal(hits["t"].view(np.uint32), g["spawn_t"].view(np.uint32))
    anyh = ctx.intersect(rays, any_hit=True)
    assert np.array_equal((anyh["prim"] >= 0).astype(np.int8), g["spawn_any"])
    ctx.close()
