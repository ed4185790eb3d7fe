#!/usr/bin/env python3
"""Device-timed numbers for the other BASELINE.json configurations (bench.py measures C2):

    C1  killeroo-simple 400x400, 64 spp          Msamples/s, kernel shares
    C3  RealisticCamera 2048x2048, 256 spp       Msamples/s + camera-ray generation alone
    C4  2.0e7 instanced triangles, 2^24-ray batches (coherent primary / incoherent), closest and
        any hit                                  Mrays/s and traversal roofline from visit counts

    python tools/bench_configs.py [c1 c3 c4] > profiles/<round>_configs.json   (on the GPU box)
"""
import ctypes as C
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import config_rays  # noqa: E402

capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
GEN = os.path.join(ROOT, "scenes", "_gen")
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0) \
    if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def load(name):
    ctx = capi.Context(0)
    sf = capi.SceneFile(os.path.join(GEN, name + ".sptx"))
    ctx.upload(sf)
    sf.close()
    return ctx


def render_rate(name, w, h, spp, warm_spp):
    ctx = load(name)
    ctx.render(0, warm_spp)
    ctx.film_clear()
    ctx.reset_counters()
    ctx.render(0, spp)
    c = ctx.counters()
    ctx.close()
    ms = c["render_ms"]
    k = {n: round(c["kernel_ms"][i], 2) for i, n in enumerate(capi.KERNEL_NAMES[:6])}
    rays = c["closest_hit_rays"] + c["shadow_rays"]
    return {"scene": name, "res": [w, h], "spp": spp, "render_ms": ms,
            "msamples_per_s": w * h * spp / ms * 1e-3, "kernel_ms": k,
            "rays_per_sample": rays / max(c["camera_rays"], 1),
            "mrays_per_s_traversal": rays / max(c["traverse_ms"], 1e-9) * 1e-3}


def c1():
    return render_rate("killeroo400", 400, 400, 64, 8)


def c3():
    out = render_rate("cornell_lens2048", 2048, 2048, 256, 4)
    # camera-ray generation alone (K1 with lens tracing): share of the render
    out["generate_ms"] = out["kernel_ms"]["generate"]
    out["camera_mrays_per_s"] = 2048 * 2048 * 256 / out["generate_ms"] * 1e-3
    return out


def c4():
    import torch
    ctx = load("instanced_c4")
    n = 1 << 24
    blob = open(os.path.join(GEN, "instanced_c4.sptx"), "rb").read()
    hb = int(np.frombuffer(blob[8:16], "<u8")[0])
    pos = 16 + hb
    pos += (-pos) % 16
    b = np.frombuffer(blob[pos:pos + 24], "<f4")
    lo, hi = b[:3], b[3:6]
    res = {"scene": "instanced_c4", "rays_per_batch": n, "batches": {}}
    # coherent primary rays: the scene's own camera over a 4096x4096 sample grid
    g = np.random.Generator(np.random.Philox(key=7))
    cs = np.zeros((n, 5), np.float32)
    side = 4096
    yy, xx = np.divmod(np.arange(n), side)
    cs[:, 0] = (xx + 0.5) * (1024.0 / side)
    cs[:, 1] = (yy + 0.5) * (1024.0 / side)
    cs[:, 2:] = 0.5
    prim, w = ctx.camera_rays(cs)
    batches = {"coherent_primary": prim, "incoherent": config_rays.incoherent_rays(n, lo, hi)}
    d_hits = torch.empty(n * 32, dtype=torch.uint8, device="cuda")
    for name, rays in batches.items():
        d_rays = torch.from_numpy(np.ascontiguousarray(rays).view(np.uint8).reshape(-1)).cuda()
        for any_hit in (False, True):
            ctx.intersect_device(d_rays.data_ptr(), n, d_hits.data_ptr(), any_hit=any_hit)  # warm-up
            ms = min(ctx.intersect_device(d_rays.data_ptr(), n, d_hits.data_ptr(), any_hit=any_hit)
                     for _ in range(3))
            # visit counts from a counted run (same traversal order as the reference)
            ctx.intersect_device(d_rays.data_ptr(), n, d_hits.data_ptr(), any_hit=any_hit, counters=True)
            hits = d_hits.cpu().numpy().view(capi.HIT_DTYPE)
            nodes, tests = float(hits["n_nodes"].mean()), float(hits["n_tests"].mean())
            bytes_per_ray = 48.0 + 32.0 * nodes + 48.0 * tests
            gbs = bytes_per_ray * n / (ms * 1e-3) * 1e-9
            res["batches"][f"{name}/{'any' if any_hit else 'closest'}"] = {
                "ms": ms, "mrays_per_s": n / ms * 1e-3, "hit_rate": float((hits["prim"] >= 0).mean()),
                "nodes_per_ray": nodes, "tests_per_ray": tests, "algorithmic_bytes_per_ray": bytes_per_ray,
                "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / PEAK}
    ctx.close()
    return res


if __name__ == "__main__":
    todo = sys.argv[1:] or ["c1", "c3", "c4"]
    out = {"hbm_peak_gbs": PEAK}
    for t in todo:
        out[t] = {"c1": c1, "c3": c3, "c4": c4}[t]()
    print(json.dumps(out, indent=1))
