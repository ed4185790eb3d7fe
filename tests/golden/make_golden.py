#!/usr/bin/env python3
"""Generates the golden fixtures of tests/golden/ from the REFERENCE itself.

Runs only where /root/reference exists (the build container).  For every scene
it (1) flattens the scene with the product host binary (host/_build/pbrt_b200,
SPTX_DUMP_SCENE) into scenes/_gen/<name>.sptx -- git-ignored, travels to the GPU
box -- and (2) drives the unmodified reference through oracle/_ref/oracle_dump
and oracle/_ref/pbrt to record, for seeded inputs:
    hits     Scene::Intersect  -> (primitive, instance, t)   [bit-exact target]
    anyhit   Scene::IntersectP
    halton   HaltonSampler::SampleDimension                   [bit-exact target]
    camrays  Camera::GenerateRayDifferential main ray + weight
    li       PathIntegrator::Li per (pixel, sample)
    dat      the reference's own render (.dat planes, float32)
into tests/golden/<name>.npz (committed).

    python tests/golden/make_golden.py [scene ...]
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
GEN = os.path.join(ROOT, "scenes", "_gen")
GOLD = os.path.join(ROOT, "tests", "golden")
PBRT_REF = os.path.join(ROOT, "oracle", "_ref", "pbrt")
ORACLE = os.path.join(ROOT, "oracle", "_ref", "oracle_dump")
PBRT_B200 = os.path.join(ROOT, "host", "_build", "pbrt_b200")
REF_SCENES = "/root/reference/scenes"

RAY = np.dtype([("o", "<f4", 3), ("t_max", "<f4"), ("d", "<f4", 3), ("time", "<f4")])
HIT = np.dtype([("prim", "<i4"), ("instance", "<i4"), ("t", "<f4"), ("b0", "<f4"), ("b1", "<f4"),
                ("b2", "<f4"), ("n_nodes", "<i4"), ("n_tests", "<i4")])

# name -> (how to get the .pbrt, render resolution, spp)
SCENES = {
    "cornell64": dict(gen=["cornell", "res=64", "spp=16"], res=(64, 64), spp=16),
    "cornell_lens48": dict(gen=["cornell_lens", "res=48", "spp=8"], res=(48, 48), spp=8),
    "instanced": dict(gen=["instanced", "res=64", "spp=4", "grid=48", "inst=3"], res=(64, 64), spp=4),
    "furnace_point": dict(gen=["furnace_point", "res=10", "spp=256", "maxdepth=8"], res=(10, 10), spp=256),
    "furnace_area": dict(gen=["furnace_area", "res=10", "spp=256", "maxdepth=8"], res=(10, 10), spp=256),
    # Material "uber" (SURVEY 8f rank 4): every lobe combination src/materials/uber.cpp builds
    "uber_box": dict(gen=["uber_box", "res=64", "spp=16"], res=(64, 64), spp=16),
    # "substrate", "translucent", rough "glass": FresnelBlend, Lambertian / microfacet transmission
    "lobes_box": dict(gen=["lobes_box", "res=64", "spp=16"], res=(64, 64), spp=16),
    # Sampler "sobol" (SURVEY 8f rank 4); non-power-of-two resolution, so the 64x64 Sobol' grid overhangs
    "cornell_sobol": dict(gen=["cornell", "res=56x40", "spp=8", "sampler=sobol"], res=(56, 40), spp=8),
    # Light "infinite" (SURVEY 8f rank 3): a constant sky, and a lat-long map with a sun (12 x 6,
    # resampled by the MIPMap), next to an area light; light selection by the spatial / power tables
    "envbox_const": dict(gen=["envbox", "res=64x48", "spp=16", "env=const"], res=(64, 48), spp=16),
    "texbox": dict(gen=["texbox", "res=64x48", "spp=16"], res=(64, 48), spp=16),
    "texbox_lens": dict(gen=["texbox", "res=64x48", "spp=16", "lens=.06", "strategy=power"], res=(64, 48), spp=16),
    "texbox_uber": dict(gen=["texbox", "res=56x40", "spp=8", "floor=uber", "strategy=power"], res=(56, 40), spp=8),
    "texbox_real": dict(gen=["texbox", "res=48x48", "spp=8", "camera=realistic", "strategy=uniform"], res=(48, 48), spp=8),
    "manylights": dict(gen=["manylights", "res=48x36", "spp=8"], res=(48, 36), spp=8),
    "envbox_map": dict(gen=["envbox", "res=64x48", "spp=16", "env=map", "strategy=power"], res=(64, 48), spp=16),
    "killeroo128": dict(ref="killeroo-simple.pbrt", res=(128, 128), spp=8),
    # crop = croppedPixelBounds (ceil(res * cropwindow), core/film.cpp:57-63): the film / .dat size
    "zoo": dict(gen=["zoo", "res=96x64", "spp=8", "strategy=power"], res=(96, 64), spp=8, crop=(10, 4, 92, 58)),
    "zoo_gauss": dict(gen=["zoo", "res=64x48", "spp=4", "strategy=uniform", "filter=gaussian"], res=(64, 48),
                      spp=4, crop=(7, 3, 61, 44)),
}


def run(cmd, env=None, cwd=None):
    e = dict(os.environ)
    if env:
        e.update(env)
    r = subprocess.run(cmd, env=e, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"{' '.join(cmd)} failed:\n{r.stdout}")
    return r.stdout


def scene_file(name, spec, work):
    """Writes <work>/<name>.pbrt (and side files) and returns its path."""
    path = os.path.join(work, name + ".pbrt")
    if "gen" in spec:
        run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), spec["gen"][0], path] + spec["gen"][1:])
        return path
    # A reference scene: same directives, resolution / spp overridden (BASELINE config 1 style).
    src = open(os.path.join(REF_SCENES, spec["ref"])).read()
    w, h = spec["res"]
    import re
    src = re.sub(r'"integer xresolution" \[\d+\]', f'"integer xresolution" [{w}]', src, count=1)
    src = re.sub(r'"integer yresolution" \[\d+\]', f'"integer yresolution" [{h}]', src, count=1)
    src = re.sub(r'"integer pixelsamples" \[\d+\]', f'"integer pixelsamples" [{spec["spp"]}]', src, count=1)
    src = re.sub(r'"string filename" "[^"]*"', f'"string filename" "{name}.exr"', src, count=1)
    src = src.replace('Include "geometry/', f'Include "{REF_SCENES}/geometry/')
    open(path, "w").write(src)
    return path


def seeded_rays(rng, bounds_min, bounds_max, cam_rays, n_random, n_shadow):
    """Camera rays (exact reference floats) + incoherent rays + shadow-style segments."""
    lo, hi = np.asarray(bounds_min, np.float32), np.asarray(bounds_max, np.float32)
    ext = hi - lo
    rays = np.zeros(len(cam_rays) + n_random + n_shadow, dtype=RAY)
    k = len(cam_rays)
    rays[:k] = cam_rays
    o = (lo - 0.1 * ext + rng.rand(n_random, 3).astype(np.float32) * 1.2 * ext).astype(np.float32)
    u = rng.rand(n_random, 2).astype(np.float32)
    z = 1 - 2 * u[:, 0]
    r = np.sqrt(np.maximum(0, 1 - z * z))
    phi = 2 * np.pi * u[:, 1]
    d = np.stack([r * np.cos(phi), r * np.sin(phi), z], -1).astype(np.float32)
    rays["o"][k:k + n_random] = o
    rays["d"][k:k + n_random] = d
    rays["t_max"][k:k + n_random] = np.inf
    a = (lo + rng.rand(n_shadow, 3).astype(np.float32) * ext).astype(np.float32)
    b = (lo + rng.rand(n_shadow, 3).astype(np.float32) * ext).astype(np.float32)
    rays["o"][k + n_random:] = a
    rays["d"][k + n_random:] = b - a
    rays["t_max"][k + n_random:] = np.float32(1 - 0.0001)
    return rays


def make(name):
    spec = SCENES[name]
    os.makedirs(GEN, exist_ok=True)
    rng = np.random.RandomState(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)))
    with tempfile.TemporaryDirectory() as work:
        pbrt = scene_file(name, spec, work)
        blob = os.path.join(GEN, name + ".sptx")
        run([PBRT_B200, "--quiet", pbrt], env={"SPTX_DUMP_SCENE": blob, "SPTX_DUMP_ONLY": "1"}, cwd=work)
        w, h = spec["res"]
        x0, y0, x1, y1 = spec.get("crop", (0, 0, w, h))
        spp = spec["spp"]
        # camera samples -> reference camera rays
        n_cam = 1024
        cs = np.zeros((n_cam, 5), np.float32)
        cs[:, 0] = rng.rand(n_cam) * w
        cs[:, 1] = rng.rand(n_cam) * h
        cs[:, 2:] = rng.rand(n_cam, 3)
        cs.tofile(os.path.join(work, "cs.bin"))
        # halton queries
        n_h = 4096
        hq = np.stack([rng.randint(x0, x1, n_h), rng.randint(y0, y1, n_h), rng.randint(0, max(spp, 64), n_h),
                       rng.randint(0, 45, n_h)], -1).astype(np.int32)
        hq[:64, 3] = np.arange(64) % 2  # make sure dims 0/1 are covered
        hq.tofile(os.path.join(work, "hq.bin"))
        # Li queries
        n_li = 2048
        lq = np.stack([rng.randint(x0, x1, n_li), rng.randint(y0, y1, n_li), rng.randint(0, spp, n_li)], -1).astype(np.int32)
        lq.tofile(os.path.join(work, "lq.bin"))
        env = {"ORACLE_CAMSAMPLES_IN": "cs.bin", "ORACLE_CAMRAYS_OUT": "cr.bin",
               "ORACLE_HALTON_IN": "hq.bin", "ORACLE_HALTON_OUT": "hv.bin",
               "ORACLE_LI_IN": "lq.bin", "ORACLE_LI_OUT": "lv.bin"}
        run([ORACLE, "--quiet", "--nthreads", "8", pbrt], env=env, cwd=work)
        cr = np.fromfile(os.path.join(work, "cr.bin"), np.float32).reshape(-1, 8)
        hv = np.fromfile(os.path.join(work, "hv.bin"), np.float32)
        lv = np.fromfile(os.path.join(work, "lv.bin"), np.float32).reshape(-1, 34)
        # rays: the camera rays that passed + random + shadow-style
        ok = cr[:, 7] > 0
        cam_rays = np.zeros(int(ok.sum()), dtype=RAY)
        cam_rays["o"] = cr[ok, 0:3]
        cam_rays["t_max"] = cr[ok, 3]
        cam_rays["d"] = cr[ok, 4:7]
        # scene bounds from the blob's root node
        sys.path.insert(0, ROOT)
        import importlib
        capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
        root = read_root_bounds(blob)
        rays = seeded_rays(rng, root[0], root[1], cam_rays, 4096, 2048)
        rays.tofile(os.path.join(work, "rays.bin"))
        env = {"ORACLE_RAYS_IN": "rays.bin", "ORACLE_HITS_OUT": "hits.bin", "ORACLE_ANYHIT_OUT": "any.bin"}
        run([ORACLE, "--quiet", "--nthreads", "8", pbrt], env=env, cwd=work)
        hits = np.fromfile(os.path.join(work, "hits.bin"), dtype=HIT)
        anyh = np.fromfile(os.path.join(work, "any.bin"), np.int32)
        # secondary rays: leave from the hit points of the first batch in fresh directions (the
        # rays a path tracer really traces: origins on and inside shapes, grazing directions)
        hp = np.nonzero(hits["prim"] >= 0)[0]
        if len(hp):
            pick = hp[rng.randint(0, len(hp), 4096)]
            p = rays["o"][pick] + rays["d"][pick] * hits["t"][pick][:, None]
            u = rng.rand(4096, 2).astype(np.float32)
            z = 1 - 2 * u[:, 0]
            r = np.sqrt(np.maximum(0, 1 - z * z))
            phi = 2 * np.pi * u[:, 1]
            d2 = np.stack([r * np.cos(phi), r * np.sin(phi), z], -1).astype(np.float32)
            sec = np.zeros(4096, dtype=RAY)
            sec["o"] = (p + np.float32(1e-3) * d2).astype(np.float32)
            sec["d"] = d2
            sec["t_max"] = np.inf
            sec["t_max"][::4] = np.float32(0.75)   # some bounded segments
            sec.tofile(os.path.join(work, "rays2.bin"))
            env = {"ORACLE_RAYS_IN": "rays2.bin", "ORACLE_HITS_OUT": "hits2.bin", "ORACLE_ANYHIT_OUT": "any2.bin"}
            run([ORACLE, "--quiet", "--nthreads", "8", pbrt], env=env, cwd=work)
            rays = np.concatenate([rays, sec])
            hits = np.concatenate([hits, np.fromfile(os.path.join(work, "hits2.bin"), dtype=HIT)])
            anyh = np.concatenate([anyh, np.fromfile(os.path.join(work, "any2.bin"), np.int32)])
        # the reference's own render
        run([PBRT_REF, "--quiet", "--nthreads", "8", pbrt], cwd=work)
        dat = capi.read_dat(os.path.join(work, name + ".dat")).astype(np.float32)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), res=np.array([x1 - x0, y1 - y0]), spp=np.array(spp),
                            cam_samples=cs, cam_rays=cr, halton_q=hq, halton_v=hv, li_q=lq, li_v=lv,
                            rays=rays, hit_prim=hits["prim"], hit_inst=hits["instance"], hit_t=hits["t"],
                            anyhit=anyh.astype(np.int8), dat=dat)
        print(f"{name}: {len(rays)} rays ({(hits['prim'] >= 0).mean():.2f} hit), "
              f"{int(ok.sum())}/{n_cam} camera rays, Li mean {lv[:, :31].mean():.4f}, dat mean {dat.mean():.4f}")


# Samplers fed by one RNG stream per tile: name -> (gen_scenes arguments, resolution, spp)
TILE_SERIAL = {
    "cornell_random": (["cornell", "res=48x40", "spp=4", "sampler=random"], (48, 40), 4),
    # ZeroTwoSequenceSampler; 6 spp is rounded up to 8 by the sampler, and the
    # path leaves the 4 sampled dimensions after the first bounce (RNG draws from then on)
    "cornell_02seq": (["cornell", "res=48x40", "spp=6", "sampler=02sequence"], (48, 40), 8),
}


def make_random(name="cornell_random"):
    """RandomSampler / ZeroTwoSequenceSampler fixtures: the reference's render only (their values are
    not a function of (pixel, sample, dimension), so there is nothing else to query)."""
    gen, res, spp = TILE_SERIAL[name]
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, name + ".pbrt")
        run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), gen[0], pbrt] + gen[1:])
        blob = os.path.join(GEN, name + ".sptx")
        run([PBRT_B200, "--quiet", pbrt], env={"SPTX_DUMP_SCENE": blob, "SPTX_DUMP_ONLY": "1"}, cwd=work)
        # one thread: the merge order of the tiles is then fixed too (the streams do not depend on it)
        run([PBRT_REF, "--quiet", "--nthreads", "1", pbrt], cwd=work)
        sys.path.insert(0, ROOT)
        import importlib
        capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
        dat = capi.read_dat(os.path.join(work, name + ".dat")).astype(np.float32)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), res=np.array(res), spp=np.array(spp), dat=dat)
    print(f"{name}: dat mean {dat.mean():.4f}")


# SpectralPathIntegrator fixtures: name -> gen_scenes arguments
BAND_SCENES = {
    # double-Gauss lens with wavelength-dependent refraction, 4 bands (the fork's default)
    "cornell_lens_ca": dict(gen=["cornell_lens", "res=48", "spp=8", "bands=4", "ca=true"], res=(48, 48), spp=8,
                            bands=4),
    # perspective camera, 3 bands: round(31/3) = 10 bins per band leaves bin 30 unassigned (black)
    "cornell_bands3": dict(gen=["cornell", "res=40x32", "spp=4", "bands=3"], res=(40, 32), spp=4, bands=3),
    # one band at 550 nm == Integrator "path" (cornell_bands1_path: the same file with "path")
    "cornell_bands1": dict(gen=["cornell", "res=40x32", "spp=4", "bands=1"], res=(40, 32), spp=4, bands=1),
    # image textures under the band loop: the textured variant recomputes the camera ray's differentials
    # per band from the sample's CameraSample (the bands share it), through the lens with aberration
    "texbox_bands": dict(gen=["texbox", "res=40x32", "spp=4", "bands=4", "camera=realistic", "ca=true"], res=(40, 32),
                         spp=4, bands=4),
}


# MetadataIntegrator fixtures: the reference's render only
META_SCENES = {
    "meta_depth": dict(gen=["instanced", "res=64", "spp=4", "grid=48", "inst=3", "metadata=depth"], res=(64, 64), spp=4),
    "meta_mesh": dict(gen=["instanced", "res=64", "spp=4", "grid=48", "inst=3", "metadata=mesh"], res=(64, 64), spp=4),
    "meta_material": dict(gen=["cornell", "res=48", "spp=2", "metadata=material"], res=(48, 48), spp=2),
    "meta_coordinates": dict(gen=["cornell_lens", "res=48", "spp=2", "metadata=coordinates"], res=(48, 48), spp=2),
}


def make_meta(name):
    spec = META_SCENES[name]
    gen = spec["gen"]
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, name + ".pbrt")
        run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), gen[0], pbrt] + gen[1:])
        blob = os.path.join(GEN, name + ".sptx")
        run([PBRT_B200, "--quiet", pbrt], env={"SPTX_DUMP_SCENE": blob, "SPTX_DUMP_ONLY": "1"}, cwd=work)
        run([PBRT_REF, "--quiet", "--nthreads", "8", pbrt], cwd=work)
        sys.path.insert(0, ROOT)
        import importlib
        capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
        dat = capi.read_dat(os.path.join(work, name + ".dat")).astype(np.float32)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), res=np.array(spec["res"]), spp=np.array(spec["spp"]),
                        dat=dat)
    print(f"{name}: dat mean {dat.mean():.4f}, bins 0..3 {np.array2string(dat.mean(axis=(1, 2))[:4], precision=4)}, "
          f"{len(np.unique(dat[0]))} distinct values in bin 0")


def band_wavelengths(n_bands):
    """ray.wavelength of each band, integrators/spectralpath.cpp:240-268 (float arithmetic)."""
    delta_index = int(np.floor(np.float32(31) / np.float32(n_bands) + 0.5))  # round(), positive argument
    dw_ca = np.float32((705 - 395) // 31) * np.float32(delta_index)
    return [float(np.float32(395) + dw_ca * np.float32(s) + dw_ca / np.float32(2)) for s in range(n_bands)]


def make_bands(name):
    """Integrator "spectralpath": the reference's render, and the reference camera's rays at every
    band's wavelength (through a sibling Integrator "path" file with the same camera, which is
    what oracle_dump hooks)."""
    spec = BAND_SCENES[name]
    gen = spec["gen"]
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, name + ".pbrt")
        run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), gen[0], pbrt] + gen[1:])
        blob = os.path.join(GEN, name + ".sptx")
        run([PBRT_B200, "--quiet", pbrt], env={"SPTX_DUMP_SCENE": blob, "SPTX_DUMP_ONLY": "1"}, cwd=work)
        run([PBRT_REF, "--quiet", "--nthreads", "8", pbrt], cwd=work)
        sys.path.insert(0, ROOT)
        import importlib
        capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
        dat = capi.read_dat(os.path.join(work, name + ".dat")).astype(np.float32)
        sib = os.path.join(work, name + "_path.pbrt")
        run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), gen[0], sib] +
            [a for a in gen[1:] if not a.startswith("bands=")])
        rng = np.random.RandomState(sum(map(ord, name)))
        w, h = spec["res"]
        n_cam = 1024
        cs = np.zeros((n_cam, 5), np.float32)
        cs[:, 0] = rng.rand(n_cam) * w
        cs[:, 1] = rng.rand(n_cam) * h
        cs[:, 2:] = rng.rand(n_cam, 3)
        cs.tofile(os.path.join(work, "cs.bin"))
        if name == "cornell_bands1":
            run([PBRT_REF, "--quiet", "--nthreads", "8", sib], cwd=work)
            dat_path = capi.read_dat(os.path.join(work, name + "_path.dat")).astype(np.float32)
            np.savez_compressed(os.path.join(GOLD, name + "_path.npz"), res=np.array(spec["res"]),
                                spp=np.array(spec["spp"]), dat=dat_path)
        lams = band_wavelengths(spec["bands"])
        crs = []
        for lam in lams:
            run([ORACLE, "--quiet", "--nthreads", "8", sib], cwd=work,
                env={"ORACLE_CAMSAMPLES_IN": "cs.bin", "ORACLE_CAMRAYS_OUT": "cr.bin", "ORACLE_WAVELENGTH": repr(lam)})
            crs.append(np.fromfile(os.path.join(work, "cr.bin"), np.float32).reshape(-1, 8))
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), res=np.array(spec["res"]), spp=np.array(spec["spp"]),
                        bands=np.array(spec["bands"]), wavelengths=np.array(lams, np.float32), cam_samples=cs,
                        cam_rays=np.stack(crs), dat=dat)
    print(f"{name}: bands at {lams} nm, dat mean {dat.mean():.4f}, per-bin means "
          f"{np.array2string(dat.mean(axis=(1, 2))[[0, 7, 8, 15, 16, 23, 24, 29, 30]], precision=4)}")


def make_halton_large():
    """HaltonSampler values at the bench resolution (1024^2: sample stride 2 239 488) for sample
    numbers up to 8191 -- indices on both sides of 2^32 and of 2^32 / base, where the device's
    interleaved digit loops and its scalar 64-bit path take over from one another."""
    rng = np.random.RandomState(20261105)
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, "cornell1024.pbrt")
        run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", pbrt, "res=1024", "spp=1024"])
        n = 1 << 15
        hq = np.stack([rng.randint(0, 1024, n), rng.randint(0, 1024, n), rng.randint(0, 8192, n),
                       rng.randint(0, 60, n)], -1).astype(np.int32)
        hq[:4096, 2] = rng.randint(1800, 2000, 4096)     # around index = 2^32 (sample number ~1918)
        hq[4096:8192, 2] = rng.randint(0, 64, 4096)
        hq.tofile(os.path.join(work, "hq.bin"))
        run([ORACLE, "--quiet", "--nthreads", "8", pbrt],
            env={"ORACLE_HALTON_IN": "hq.bin", "ORACLE_HALTON_OUT": "hv.bin"}, cwd=work)
        hv = np.fromfile(os.path.join(work, "hv.bin"), np.float32)
        np.savez_compressed(os.path.join(GOLD, "halton_large.npz"), halton_q=hq, halton_v=hv)
        print(f"halton_large: {n} values, mean {hv.mean():.4f}")


VISIT_SCENES = ["cornell64", "killeroo128", "instanced", "zoo"]


def make_visits():
    """Per-ray node / primitive visit counts of the reference's traversal (oracle_dump's counting
    walk over the reference's own BVH, checked there against Scene::Intersect) for the rays the
    scene fixtures already hold -> tests/golden/visits.npz."""
    out = {}
    for name in VISIT_SCENES:
        spec = SCENES[name]
        g = np.load(os.path.join(GOLD, name + ".npz"))
        with tempfile.TemporaryDirectory() as work:
            pbrt = scene_file(name, spec, work)
            np.ascontiguousarray(g["rays"]).tofile(os.path.join(work, "rays.bin"))
            run([ORACLE, "--quiet", "--nthreads", "8", pbrt], cwd=work,
                env={"ORACLE_RAYS_IN": "rays.bin", "ORACLE_HITS_OUT": "hits.bin", "ORACLE_COUNT_VISITS": "1"})
            hits = np.fromfile(os.path.join(work, "hits.bin"), dtype=HIT)
        assert np.array_equal(hits["prim"], g["hit_prim"]) and np.array_equal(hits["t"], g["hit_t"])
        out[name + "_nodes"] = hits["n_nodes"].astype(np.int32)
        out[name + "_tests"] = hits["n_tests"].astype(np.int32)
        print(f"{name}: {hits['n_nodes'].mean():.2f} nodes, {hits['n_tests'].mean():.2f} primitive tests per ray")
    np.savez_compressed(os.path.join(GOLD, "visits.npz"), **out)


def read_root_bounds(blob):
    """Root node bounds of the scene aggregate, read straight from the scene file."""
    import ctypes as C
    import importlib
    capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
    sf = capi.SceneFile(blob)
    # sptx_scene_desc: find `nodes` via a tiny C view: n_nodes (int64) precedes the pointer.
    # Rather than mirror the whole struct here, scan the file: the first array after
    # the header is nodes (csrc/sceneio.cpp), and node 0 is the root.
    sf.close()
    raw = open(blob, "rb").read()
    header_bytes = int(np.frombuffer(raw[8:16], "<u8")[0])
    pos = 16 + header_bytes
    pos += (-pos) % 16
    b = np.frombuffer(raw[pos:pos + 24], "<f4")
    return b[:3].copy(), b[3:6].copy()


if __name__ == "__main__":
    names = sys.argv[1:] or list(SCENES) + list(TILE_SERIAL) + list(BAND_SCENES) + list(META_SCENES) + ["visits", "halton_large"]
    for n in names:
        if n == "visits":
            make_visits()
        elif n == "halton_large":
            make_halton_large()
        elif n in TILE_SERIAL:
            make_random(n)
        elif n in BAND_SCENES:
            make_bands(n)
        elif n in META_SCENES:
            make_meta(n)
        else:
            make(n)
