#!/usr/bin/env python3
"""north_star's render bar at its stated sample count: the spectral Cornell box at 4096 spp,
GPU against the reference's own CPU render of the same scene file, per-bin relative RMSE.

Runs on the GPU box: the scene file is generated there, flattened by host/_build/pbrt_b200
(SPTX_DUMP_ONLY), rendered by oracle/_ref/pbrt on the host cores and by libsptx.so on the device.

    python tools/converged_parity.py [res=256] [spp=4096] > profiles/<round>_converged_parity.json
"""
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
opts = dict(a.split("=") for a in sys.argv[1:])
res, spp = int(opts.get("res", 256)), int(opts.get("spp", 4096))
scene_kind = opts.get("scene", "cornell")
exe_ref = os.path.join(ROOT, "oracle", "_ref", "pbrt")
exe_b200 = os.path.join(ROOT, "host", "_build", "pbrt_b200")
with tempfile.TemporaryDirectory() as work:
    pbrt = os.path.join(work, "s.pbrt")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), scene_kind, pbrt,
                           f"res={res}", f"spp={spp}"])
    blob = os.path.join(work, "s.sptx")
    subprocess.check_call([exe_b200, "--quiet", pbrt], cwd=work,
                          env=dict(os.environ, SPTX_DUMP_SCENE=blob, SPTX_DUMP_ONLY="1"),
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t0 = time.perf_counter()
    subprocess.check_call([exe_ref, "--quiet", "--nthreads", str(os.cpu_count()), pbrt], cwd=work,
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    cpu_s = time.perf_counter() - t0
    ref = capi.read_dat(os.path.join(work, "s.dat"))
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob)
    ctx.upload(sf)
    ctx.render(0, spp)
    gpu_ms = ctx.counters()["render_ms"]
    mine = ctx.film_pack_dat(res, res)
rel_rmse = np.sqrt(((mine - ref) ** 2).mean(axis=(1, 2))) / np.maximum(ref.mean(axis=(1, 2)), 1e-30)
tot_ref, tot_mine = ref.sum(axis=0), mine.sum(axis=0)
differing = int((np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * 31)).sum())
print(json.dumps({
    "scene": f"{scene_kind} {res}x{res}, Halton, {spp} spp, maxdepth 5", "samples": res * res * spp,
    "per_bin_rel_rmse_max": float(rel_rmse.max()), "per_bin_rel_rmse_mean": float(rel_rmse.mean()),
    "tolerance": 0.005, "pixels_with_a_differing_sample": differing, "pixels": res * res,
    "energy_ratio": float(mine.sum() / ref.sum()),
    "reference_cpu_seconds": cpu_s, "cores": os.cpu_count(), "gpu_render_ms": gpu_ms}, indent=1))
