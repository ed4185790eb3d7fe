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
L2_PEAK = None   # measured by the run itself (sptx_measure_l2, a 32 MB buffer): see l2_peak()


def l2_peak():
    global L2_PEAK
    if L2_PEAK is None:
        ctx = capi.Context(0)
        L2_PEAK = max(ctx.measure_l2(32 << 20, 200) for _ in range(3))
        ctx.close()
    return L2_PEAK


def load(name, device_bvh=False):
    ctx = capi.Context(0)
    if device_bvh:
        ctx.set_option("device_bvh", 1)
    sf = capi.SceneFile(capi.unpack_blob(os.path.join(GEN, name + ".sptx")))
    ctx.upload(sf)
    sf.close()
    return ctx


def render_rate(name, w, h, spp, warm_spp, device_bvh=False):
    ctx = load(name, device_bvh)
    build_ms = ctx.bvh_build_ms()
    ctx.render(0, warm_spp)
    ctx.film_clear()
    ctx.reset_counters()
    ctx.render(0, spp)
    c = ctx.counters()
    ctx.close()
    ms = c["render_ms"]
    k = {n: round(c["kernel_ms"][i], 2) for i, n in enumerate(capi.KERNEL_NAMES[:6])}
    rays = c["closest_hit_rays"] + c["shadow_rays"]
    return {"scene": name, "res": [w, h], "spp": spp, "render_ms": ms, "bvh": "device LBVH" if device_bvh else "reference SAH (host)",
            "bvh_build_ms": build_ms,
            "msamples_per_s": w * h * spp / ms * 1e-3, "kernel_ms": k,
            "rays_per_sample": rays / max(c["camera_rays"], 1),
            "mrays_per_s_traversal": rays / max(c["traverse_ms"], 1e-9) * 1e-3}


def reference_rate(gen, w, h, spp_lo, spp_hi):
    """The reference's own CPU renderer (oracle/_ref/pbrt, all host threads) on a procedural scene:
    samples / (t(spp_hi) - t(spp_lo)), which removes parsing, BVH build and the .dat write."""
    import subprocess
    import tempfile
    import time
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/pbrt was not built"}
    t = []
    with tempfile.TemporaryDirectory() as work:
        for spp in (spp_lo, spp_hi):
            pbrt = os.path.join(work, f"s{spp}.pbrt")
            subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), gen[0], pbrt] +
                                  gen[1:] + [f"res={w}x{h}", f"spp={spp}"])
            t0 = time.perf_counter()
            subprocess.run([exe, "--quiet", "--nthreads", str(os.cpu_count()), pbrt], cwd=work, check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            t.append(time.perf_counter() - t0)
    dt = max(t[1] - t[0], 1e-6)
    return {"msamples_per_s": w * h * (spp_hi - spp_lo) / dt * 1e-6, "cores": os.cpu_count(),
            "sample": f"{spp_hi - spp_lo} spp (difference of a {spp_hi}-spp and a {spp_lo}-spp run, {dt:.1f} s)"}


def c1():
    return render_rate("killeroo400", 400, 400, 64, 8)


def c3(spp=256, cpu=True):
    out = render_rate("cornell_lens2048", 2048, 2048, spp, 4)
    # camera-ray generation alone (K1 with lens tracing): share of the render
    out["generate_ms"] = out["kernel_ms"]["generate"]
    out["camera_mrays_per_s"] = 2048 * 2048 * spp / out["generate_ms"] * 1e-3
    if cpu:
        out["cpu_reference"] = reference_rate(["cornell_lens"], 2048, 2048, 1, 3)
    return out


def c1_device_bvh():
    """C1 through a BVH built on the device (csrc/lbvh.cuh) instead of the reference's SAH tree."""
    return render_rate("killeroo400", 400, 400, 64, 8, device_bvh=True)


def c4(n=1 << 24, n_cpu=1 << 20, device_bvh=False):
    import torch
    ctx = load("instanced_c4", device_bvh)
    blob = open(os.path.join(GEN, "instanced_c4.sptx"), "rb").read()
    hb = int(np.frombuffer(blob[8:16], "<u8")[0])
    pos = 16 + hb
    pos += (-pos) % 16
    b = np.frombuffer(blob[pos:pos + 24], "<f4")
    lo, hi = b[:3], b[3:6]
    res = {"scene": "instanced_c4", "rays_per_batch": n, "batches": {},
           "bvh": "device LBVH" if device_bvh else "reference SAH (host)", "bvh_build_ms": ctx.bvh_build_ms()}
    # coherent primary rays: the scene's own camera over a 4096x4096 sample grid
    g = np.random.Generator(np.random.Philox(key=7))
    cs = np.zeros((n, 5), np.float32)
    side = int(round(n ** 0.5))
    assert side * side == n, "n must be a square (the coherent batch is a side x side sample grid)"
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
            # The instanced BVH (11 MB) and the top level stay in L2 (126 MB): the node / triangle bytes
            # are L2 traffic, only the rays and hit records (48 B per ray) come from HBM -- so the
            # algorithmic rate is set against the L2 bandwidth, with the HBM share beside it.
            binned = ctx.last_batch_sorted()
            # the other walk order, for the record: forced on where the library left it off and vice versa
            ctx.set_option("ray_sort", 0 if binned else 1)
            ctx.intersect_device(d_rays.data_ptr(), n, d_hits.data_ptr(), any_hit=any_hit)
            ms_other = min(ctx.intersect_device(d_rays.data_ptr(), n, d_hits.data_ptr(), any_hit=any_hit)
                           for _ in range(3))
            ctx.set_option("ray_sort", 2)
            res["batches"][f"{name}/{'any' if any_hit else 'closest'}"] = {
                "ms": ms, "rays_binned": binned, "ms_unbinned" if binned else "ms_binned": ms_other, "mrays_per_s": n / ms * 1e-3, "hit_rate": float((hits["prim"] >= 0).mean()),
                "nodes_per_ray": nodes, "tests_per_ray": tests, "algorithmic_bytes_per_ray": bytes_per_ray,
                "achieved_gbs": gbs, "bound": "l2", "l2_peak_gbs": l2_peak(), "frac_of_l2_peak": gbs / l2_peak(),
                "hbm_gbs_rays_and_hits": 48.0 * n / (ms * 1e-3) * 1e-9,
                "frac_of_hbm_peak_rays_and_hits": 48.0 * n / (ms * 1e-3) * 1e-9 / PEAK}
    ctx.close()
    # the reference's BVHAccel::Intersect / IntersectP on the box's host cores, timed by the oracle
    # (BASELINE.md 3.3): 2^20 rays of each batch, every host thread
    if n_cpu:
        res["cpu_reference"] = c4_cpu({k: v[:n_cpu] for k, v in batches.items()})
    return res


def c4_cpu(batches):
    import re
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "oracle_dump")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/oracle_dump was not built"}
    out = {"cores": os.cpu_count(), "rays_per_batch": int(len(next(iter(batches.values()))))}
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, "instanced_c4.pbrt")
        subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "instanced", pbrt,
                               "res=1024", "spp=4", "grid=317", "inst=10"])
        for name, rays in batches.items():
            np.ascontiguousarray(rays).tofile(os.path.join(work, "rays.bin"))
            env = dict(os.environ, ORACLE_RAYS_IN="rays.bin", ORACLE_TIME_RAYS="4")
            r = subprocess.run([exe, "--quiet", "--nthreads", str(os.cpu_count()), pbrt], cwd=work, env=env,
                               capture_output=True, text=True)
            for kind, rate in re.findall(r"ORACLE_TIMING (\w+) .*?mrays_per_s=([0-9.]+)", r.stdout + r.stderr):
                out[f"{name}/{kind}"] = {"mrays_per_s": float(rate)}
    return out


if __name__ == "__main__":
    todo = sys.argv[1:] or ["c1", "c3", "c4"]
    out = {"hbm_peak_gbs": PEAK}
    for t in todo:
        out[t] = {"c1": c1, "c3": c3, "c4": c4, "c1_lbvh": c1_device_bvh,
                  "c4_lbvh": lambda: c4(n=1 << 22, n_cpu=0, device_bvh=True),
                  "c4_small": lambda: c4(n=1 << 22, n_cpu=0)}[t]()
    print(json.dumps(out, indent=1))
