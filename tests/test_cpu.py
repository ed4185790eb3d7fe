"""CPU-side tests (no GPU): the library loads and exports the whole C ABI, scene
files round-trip, and the golden vectors recorded from the reference satisfy the
reference's own known answers."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import GEN, ROOT, SCENES, golden


def test_library_exports_every_declared_symbol(capi):
    header = open(os.path.join(ROOT, "include", "sptx.h")).read()
    declared = set(re.findall(r"\b(sptx_[a-z_0-9]+)\s*\(", header))
    assert declared, "no declarations found in include/sptx.h"
    assert declared == set(capi.SYMBOLS)
    lib = capi.lib()
    for sym in sorted(declared):
        assert hasattr(lib, sym), f"libsptx.so does not export {sym}"
    assert lib.sptx_abi_version() == 1


def test_no_cpu_fallback_without_device(capi):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.SptxError, match="no CPU fallback|no CUDA"):
        capi.Context(0)


@pytest.mark.parametrize("name", SCENES)
def test_scene_file_roundtrip(capi, name, tmp_path):
    p = os.path.join(GEN, name + ".sptx")
    if not os.path.exists(p):
        pytest.skip("scene blob not generated here")
    sf = capi.SceneFile(p)
    out = str(tmp_path / "copy.sptx")
    assert capi.lib().sptx_scene_save(out.encode(), sf.handle) == 0
    assert open(out, "rb").read() == open(p, "rb").read()
    sf.close()


def test_scene_file_rejects_garbage(capi, tmp_path):
    p = tmp_path / "bad.sptx"
    p.write_bytes(b"not a scene" * 100)
    with pytest.raises(capi.SptxError):
        capi.SceneFile(str(p))


@pytest.mark.parametrize("name", ["furnace_point", "furnace_area"])
def test_golden_furnace_known_answer(name):
    """The oracle itself reproduces src/tests/analytic_scenes.cpp:54-66 (mean 1 +- 0.02)."""
    g = golden(name)
    spp = int(g["spp"])
    mean = g["dat"].mean() / spp
    assert abs(mean - 1.0) < 0.02
    assert abs(g["li_v"][:, :31].mean() - 1.0) < 0.05


@pytest.mark.parametrize("name", SCENES)
def test_golden_is_self_consistent(name):
    g = golden(name)
    assert g["rays"].shape[0] == g["hit_prim"].shape[0] == g["anyhit"].shape[0]
    hit = g["hit_prim"] >= 0
    # any-hit agrees with closest-hit for unbounded rays
    rays = g["rays"]
    tmax = rays.view(np.float32).reshape(len(rays), -1)[:, 3]
    unbounded = np.isinf(tmax)
    assert np.array_equal(hit[unbounded], g["anyhit"][unbounded].astype(bool))
    v = g["halton_v"]
    assert ((v >= 0) & (v <= 1)).all()
    assert np.isfinite(g["dat"]).all()


def test_write_dat_matches_reference_layout(capi, tmp_path):
    planes = np.arange(31 * 3 * 2, dtype=np.float64).reshape(31, 3, 2)
    p = str(tmp_path / "x.dat")
    capi.write_dat(p, planes)
    raw = open(p, "rb").read()
    assert raw.startswith(b"2 3 31\nv3 \n")
    assert len(raw) == len(b"2 3 31\nv3 \n") + 31 * 3 * 2 * 8
    assert np.array_equal(capi.read_dat(p), planes)
