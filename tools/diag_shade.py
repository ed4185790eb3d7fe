#!/usr/bin/env python3
"""Per-sample and per-pixel differences against the golden fixtures, for comparing builds of K3
(SPTX_LIB=libsptx_fast.so python tools/diag_shade.py [scene ...]).  Prints one JSON line per scene."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import SCENES, blob_path, golden  # noqa: E402

capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
ctx = capi.Context(0)
for name in sys.argv[1:] or SCENES:
    g = golden(name)
    sf = capi.SceneFile(blob_path(name))
    ctx.upload(sf)
    sf.close()
    out = {"scene": name, "lib": os.environ.get("SPTX_LIB", "libsptx.so")}
    try:
        q = g["li_q"]
        L, pw = ctx.debug_radiance(q[:, 0], q[:, 1], q[:, 2])
        ref = g["li_v"]
        scale = np.maximum(np.abs(ref[:, :31]).max(axis=1), 1e-3)
        err = np.abs(L - ref[:, :31]).max(axis=1) / scale
        out["li_samples"] = int(len(err))
        for t in (1e-6, 1e-5, 1e-4, 1e-3, 1e-2):
            out[f"li_frac_gt_{t:g}"] = float((err > t).mean())
        out["li_err_median"] = float(np.median(err))
        out["li_mean_ratio"] = float(L.mean() / max(ref[:, :31].mean(), 1e-30))
    except Exception as exc:
        out["li_error"] = str(exc)[:200]
    try:
        w, h = [int(v) for v in g["res"]]
        ctx.film_clear()
        ctx.render(0, int(g["spp"]))
        planes = ctx.film_pack_dat(w, h)
        ref = g["dat"].astype(np.float64)
        tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
        rel = np.abs(tot_mine - tot_ref) / np.maximum(tot_ref, ref.mean() * ref.shape[0])
        out["pixels"] = int(rel.size)
        for t in (1e-5, 1e-4, 1e-3, 1e-2):
            out[f"px_gt_{t:g}"] = int((rel > t).sum())
        out["rel_rmse_all_max"] = float((np.sqrt(((planes - ref) ** 2).mean(axis=(1, 2))) /
                                         np.maximum(ref.mean(axis=(1, 2)), 1e-12)).max())
        out["energy_ratio"] = float(planes.sum() / ref.sum())
    except Exception as exc:
        out["render_error"] = str(exc)[:200]
    print(json.dumps(out), flush=True)
ctx.close()
