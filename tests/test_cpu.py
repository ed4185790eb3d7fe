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


# ---- Integrator "spectralpath"  ----------------------------------------

BAND_SCENES = ["cornell_lens_ca", "cornell_bands3", "cornell_bands1"]


@pytest.mark.parametrize("name", BAND_SCENES)
def test_band_fixture_follows_the_reference_band_layout(name):
    """integrators/spectralpath.cpp:240-318: deltaIndex = round(31 / numCABands) bins per band,
    ray.wavelength = 395 + 10 * deltaIndex * (s + 1/2); bins past the last band stay black."""
    g = golden(name)
    n = int(g["bands"])
    di = int(np.floor(31 / n + 0.5))
    lam = [395 + 10 * di * s + 10 * di / 2 for s in range(n)]
    assert np.allclose(g["wavelengths"], lam)
    owned = np.zeros(31, bool)
    for s in range(n):
        owned[di * s:min(di * (s + 1), 31)] = True
    means = g["dat"].mean(axis=(1, 2))
    assert (means[owned] > 0).all() and (means[~owned] == 0).all()
    assert g["cam_rays"].shape == (n, g["cam_samples"].shape[0], 8)


def test_one_band_is_the_path_integrator():
    """numCABands = 1 traces one 550 nm path per sample from the same Halton stream: the
    reference's two integrators then write the same image."""
    a, b = golden("cornell_bands1"), golden("cornell_bands1_path")
    assert np.array_equal(a["dat"], b["dat"])


def test_flattener_rejects_bands_with_random_sampler(tmp_path):
    """No silent approximation: the tile-serial RandomSampler path does not run bands."""
    import subprocess
    import sys
    exe = os.path.join(ROOT, "host", "_build", "pbrt_b200")
    if not os.path.exists(exe):
        pytest.skip("host binary is built only where the reference sources are")
    scene = str(tmp_path / "s.pbrt")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", scene,
                           "res=16", "spp=1", "sampler=random", "bands=2"])
    env = dict(os.environ, SPTX_DUMP_SCENE=str(tmp_path / "s.sptx"), SPTX_DUMP_ONLY="1")
    r = subprocess.run([exe, "--quiet", scene], cwd=str(tmp_path), env=env, capture_output=True, text=True)
    assert r.returncode != 0 and "halton and sobol samplers only" in (r.stdout + r.stderr)
    assert not os.path.exists(str(tmp_path / "s.sptx"))
