"""The drop-in route end to end: host/_build/pbrt_b200 = the reference's own main(), parser, scene
construction and Film::WriteImage with GpuPathIntegrator bound for Integrator "path" /
"spectralpath" (host/gpu_path_integrator.cpp).  `pbrt_b200 scene.pbrt` must write
the .dat the reference's `pbrt scene.pbrt` writes (fixtures: tests/golden/make_golden.py).

The binary is built where the reference's sources are (the build container) and travels to the
GPU box with the tree; scene files are regenerated there by scenes/gen_scenes.py.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, golden

pytestmark = pytest.mark.gpu

EXE = os.path.join(ROOT, "host", "_build", "pbrt_b200")

CASES = {
    "cornell64": ["cornell", "res=64", "spp=16"],
    "cornell_lens_ca": ["cornell_lens", "res=48", "spp=8", "bands=4", "ca=true"],
    "zoo_gauss": ["zoo", "res=64x48", "spp=4", "strategy=uniform", "filter=gaussian"],
    "lobes_box": ["lobes_box", "res=64", "spp=16"],
    "meta_depth": ["instanced", "res=64", "spp=4", "grid=48", "inst=3", "metadata=depth"],
    # SURVEY 8(f) rank 3 through the reference's own front end: image textures (PFM read by the host,
    # MIPMap built by the reference's code, evaluated on the device), a mapped infinite light, and the
    # many-light scene whose spatial light tables are sparse
    "texbox_uber": ["texbox", "res=56x40", "spp=8", "floor=uber", "strategy=power"],
    "envbox_map": ["envbox", "res=64x48", "spp=16", "env=map", "strategy=power"],
    "manylights": ["manylights", "res=48x36", "spp=8"],
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_pbrt_b200_writes_the_reference_dat(capi, name, tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("host/_build/pbrt_b200 not built in this tree (needs the reference's sources)")
    scene = str(tmp_path / (name + ".pbrt"))
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), CASES[name][0], scene] +
                          CASES[name][1:])
    r = subprocess.run([EXE, "--quiet", scene], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "GpuPathIntegrator:" in r.stderr  # the device path ran (its summary line)
    planes = capi.read_dat(str(tmp_path / (name + ".dat")))
    ref = golden(name)["dat"].astype(np.float64)
    assert planes.shape == ref.shape
    tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
    flipped = np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * ref.shape[0])
    assert flipped.sum() <= max(1, int(5e-4 * flipped.size)), f"{name}: {flipped.sum()} pixels differ"
    keep = ~flipped
    ref_mean = ref[:, keep].mean(axis=1)
    rel_rmse = np.sqrt((((planes - ref) ** 2)[:, keep]).mean(axis=1)) / np.maximum(ref_mean, 1e-12)
    assert rel_rmse[ref_mean > 0].max() < 0.005, f"{name}: per-bin relRMSE {rel_rmse.max():.4%}"


def test_pbrt_b200_on_several_contexts_writes_the_single_context_dat(capi, tmp_path):
    """Multi-GPU through the reference's own entry point: SPTX_DEVICES=0,0,0 makes
    GpuPathIntegrator::Render drive three contexts (all on cuda:0 here; one per GPU on a multi-GPU
    box) from one host thread each, partition the sample numbers, and sum the films through peer
    access (sptx_film_merge_contexts) before the reference's Film::WriteImage.  C5's geometry
    (16:9 spectral Cornell box) at a reduced size; the .dat equals the single-context one up to
    the float summation order of the partial films."""
    if not os.path.exists(EXE):
        pytest.skip("host/_build/pbrt_b200 not built in this tree (needs the reference's sources)")
    scene = str(tmp_path / "c5small.pbrt")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", scene,
                           "res=480x270", "spp=10"])
    dats = {}
    for tag, devs in (("one", "0"), ("three", "0,0,0")):
        env = dict(os.environ, SPTX_DEVICES=devs)
        r = subprocess.run([EXE, "--quiet", scene], cwd=str(tmp_path), capture_output=True, text=True,
                           timeout=300, env=env)
        assert r.returncode == 0, r.stdout + r.stderr
        assert f"{len(devs.split(','))} GPU context(s)" in r.stderr
        dats[tag] = capi.read_dat(str(tmp_path / "c5small.dat"))
        os.remove(str(tmp_path / "c5small.dat"))
    np.testing.assert_allclose(dats["three"], dats["one"], rtol=1e-5, atol=1e-5 * dats["one"].mean())


def test_pbrt_b200_reports_unsupported_scenes(tmp_path):
    """features outside the hot path are an error, never a silent approximation."""
    if not os.path.exists(EXE):
        pytest.skip("host/_build/pbrt_b200 not built in this tree")
    scene = str(tmp_path / "s.pbrt")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", scene,
                           "res=16", "spp=1", "sampler=random", "bands=2"])
    r = subprocess.run([EXE, "--quiet", scene], cwd=str(tmp_path), capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "outside the GPU hot path" in (r.stdout + r.stderr)
    assert not os.path.exists(str(tmp_path / "s.dat"))
