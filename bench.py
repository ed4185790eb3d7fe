#!/usr/bin/env python3
"""bench.py -- spectral Msamples/s of the PathIntegrator hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): the procedural spectral Cornell box (diffuse + glass +
metal, 31 bins), 1024x1024, Halton, maxdepth 5, 1024 spp.  A "step" renders SPP_STEP = 64
consecutive sample numbers of every pixel (67.1 M camera samples); the default 16 steps are the
whole 1024-spp render.  `value` is camera samples / device time (CUDA events on the render
stream, scene resident in HBM); `e2e` is the same step through the C ABI from HOST buffers:
sptx_upload_scene(host arrays) + sptx_render + sptx_film_download(host) per step, wall clock.

N > 1 (torchrun, one rank per GPU): the render is partitioned by sample number -- rank r renders
its own SPP_STEP samples per step of the full image (weak scaling) -- and the per-GPU films are
summed onto rank 0 with one NCCL reduce inside the timed region (the reference's MergeFilmTile,
src/core/film.cpp:124-142).

--impl reference times the reference's own CPU renderer (oracle/_ref/pbrt, the unmodified
reference compiled in place) on the box's host cores on a bounded sample of the same workload.
"""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GEN = os.path.join(ROOT, "scenes", "_gen")
RES_W = RES_H = 1024   # --workload c2 (default); c5 = 3840 x 2160, 4096 spp (set_workload)
SPP_TOTAL = 1024
SPP_STEP = 64
REF_SPP_STEP = 8       # reference arm: sample numbers per step (bounded sample; a step is seconds)
CPU_BASELINE_SPP = 8   # cpu_baseline leg of our arm
METRIC = "spectral_msamples_per_s"
UNIT = "Msamples/s"
WORKLOAD = ""


def set_workload(name):
    """c2 = BASELINE.json configs[1] (the metric's single-GPU configuration, the default);
    c5 = configs[4], the 4K render at 4096 spp partitioned across GPUs."""
    global RES_W, RES_H, SPP_TOTAL, WORKLOAD
    RES_W, RES_H, SPP_TOTAL = {"c2": (1024, 1024, 1024), "c5": (3840, 2160, 4096)}[name]
    WORKLOAD = (f"procedural spectral Cornell box (diffuse+glass+metal), 31 bins, {RES_W}x{RES_H}, Halton, "
                f"maxdepth 5, {SPP_TOTAL} spp")


def res_arg():
    return f"res={RES_W}x{RES_H}"


def blob_name():
    return f"cornell{RES_W}" if RES_W == RES_H else f"cornell{RES_W}x{RES_H}"


set_workload("c2")


def emit(text):  # rebound by main() to the process's original stdout
    print(text)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def scene_pbrt(path, spp):
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", path,
                           res_arg(), f"spp={spp}"])


def scene_blob():
    blob = os.path.join(GEN, blob_name() + ".sptx")
    if not os.path.exists(blob):
        entry = importlib.import_module("__graft_entry__")
        entry.make_scene_blob(blob_name(), ["cornell", res_arg(), f"spp={SPP_TOTAL}"], blob)
    return blob


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "250"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 6 and r[2 + k] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def reference_rate(spp_lo, spp_hi, threads=None):
    """Reference throughput from two runs (removes parse / .dat-write constants):
    (spp_hi - spp_lo) * pixels / (t_hi - t_lo)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    cores = threads or os.cpu_count() or 1
    times = []
    with tempfile.TemporaryDirectory() as work:
        for spp in (spp_lo, spp_hi):
            pbrt = os.path.join(work, f"cornell_ref_{spp}.pbrt")
            scene_pbrt(pbrt, spp)
            t0 = time.perf_counter()
            out = subprocess.run([exe, "--quiet", "--nthreads", str(cores), pbrt], cwd=work,
                                 stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            times.append(time.perf_counter() - t0)
            if out.returncode != 0:
                raise RuntimeError("reference renderer failed:\n" + out.stdout[-2000:])
    dt = max(times[1] - times[0], 1e-6)
    return RES_W * RES_H * (spp_hi - spp_lo) / dt * 1e-6, dt, cores, times


class CudaArray:
    """Minimal __cuda_array_interface__ view so torch can wrap the device film for NCCL."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def bench_ours(args):
    import torch
    capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
    part = importlib.import_module("pbrt-v3-spectral_b200.partition")
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this benchmark has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    blob = scene_blob() if rank == 0 else None
    if world > 1:
        dist.barrier()
        blob = os.path.join(GEN, blob_name() + ".sptx")
    ctx = capi.Context(local)
    sf = capi.SceneFile(blob)
    ctx.upload(sf)
    film = torch.as_tensor(CudaArray(ctx.film_ptr, ctx.film_floats), device=f"cuda:{local}")
    K, W = args.steps, args.warmup
    # the exchange step at N > 1: "reduce_scatter" (default) / "gather" = the two peer-memory merge
    # kernels, "nccl" = a plain NCCL reduce of the film tensor, for comparison
    merge_shape = os.environ.get("SPTX_BENCH_MERGE", "nccl" if os.environ.get("SPTX_BENCH_NCCL_REDUCE") else
                                 "reduce_scatter")

    # SPTX_BENCH_EMULATE="r,N": a single GPU renders the sample numbers rank r of N would get
    # (used to separate "larger Halton indices" from "more GPUs" in the scaling numbers)
    emu = os.environ.get("SPTX_BENCH_EMULATE")
    emu_rank, emu_world = [int(v) for v in emu.split(",")] if (emu and world == 1) else (rank, world)

    def step_range(k):
        # consecutive sample numbers; with N ranks, rank r takes the r-th block of each step
        return part.sample_block(k, emu_rank, emu_world, SPP_STEP, SPP_TOTAL)

    for k in range(W):
        ctx.render(*step_range(k))
    ctx.film_clear()
    ctx.reset_counters()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    clocks = ClockSampler(local)   # every rank samples its own GPU
    clocks.start()
    t0 = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for k in range(K):
        ctx.render(*step_range(k))
    c = ctx.counters()
    dev_ms = c["render_ms"]
    reduce_ms = 0.0
    reduce_wall_ms = 0.0
    if dist:
        # the one exchange step of the path: per-rank films -> rank 0, summed in rank order by one
        # kernel reading the peers over NVLink (sptx_film_merge_peers); NCCL reduce if asked for
        f0 = ctx.counters()["film_ms"]
        t_red = time.perf_counter()
        if merge_shape == "nccl":
            ev0.record()
            part.reduce_film(film, dist, dst=0)
            ev1.record()
            torch.cuda.synchronize()
            reduce_ms = ev0.elapsed_time(ev1)
        else:
            part.merge_films_over_peers(ctx, dist, dst=0, shape=merge_shape)
            reduce_ms = ctx.counters()["film_ms"] - f0   # this rank's merge kernel (max over ranks below)
        reduce_wall_ms = (time.perf_counter() - t_red) * 1e3
        dev_ms += reduce_ms
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clk = clocks.stop()
    clk_all = [clk]
    if dist:
        clk_all = [None] * world
        dist.all_gather_object(clk_all, clk)
    if dist:
        t = torch.tensor([dev_ms, wall_ms, reduce_ms, reduce_wall_ms], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms, reduce_ms, reduce_wall_ms = [float(v) for v in t]
    samples_total = float(RES_W * RES_H * SPP_STEP * K * world)
    value = samples_total / (dev_ms * 1e-3) * 1e-6

    # ---- e2e through the C ABI with host buffers (every rank; max over ranks) ----
    w, h = RES_W, RES_H
    e2e_steps = max(1, min(K, 4))
    h2d = os.path.getsize(blob)
    d2h = w * h * 32 * 4
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    # caller-owned result buffers, page-locked (what a host integration registers once per film)
    host_film = (torch.zeros((h, w, 31), dtype=torch.float32).pin_memory().numpy(),
                 torch.zeros((h, w), dtype=torch.float32).pin_memory().numpy())
    # one untimed round trip first (first touch of the page-locked buffers, allocator warm-up)
    ctx.upload(sf)
    ctx.render(0, 1)
    ctx.film_download(w, h, out=host_film)
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    e2e_parts = [0.0, 0.0, 0.0, 0.0]
    for k in range(e2e_steps):
        ta = time.perf_counter()
        ctx.upload(sf)                      # host arrays -> HBM (also a fresh, cleared film)
        tb = time.perf_counter()
        ctx.render(*step_range(k))
        tc = time.perf_counter()
        if dist:
            # the product of a multi-GPU step is the MERGED film on the host: handle exchange,
            # barriers, the peer-memory merge kernel on rank 0 -- all inside the timed region
            if merge_shape == "nccl":
                part.reduce_film(film, dist, dst=0)
                torch.cuda.synchronize()
            else:
                part.merge_films_over_peers(ctx, dist, dst=0, shape=merge_shape)
        tm = time.perf_counter()
        if rank == 0:
            contrib, weight = ctx.film_download(w, h, out=host_film)  # device film -> host
        td = time.perf_counter()
        e2e_parts = [e2e_parts[0] + tb - ta, e2e_parts[1] + tc - tb, e2e_parts[2] + tm - tc,
                     e2e_parts[3] + td - tm]
    if dist:
        dist.barrier()
    e2e_s = time.perf_counter() - t0
    if dist:
        t = torch.tensor([e2e_s], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t[0])
    e2e_value = RES_W * RES_H * SPP_STEP * e2e_steps * world / e2e_s * 1e-6

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel (one counted re-run of a step for visit counts) ----
        kms, kcnt = c["kernel_ms"], c["kernel_count"]
        names = capi.KERNEL_NAMES
        dom = int(np.argmax(kms[:6]))
        ctx.upload(sf)
        ctx.set_option("count_visits", 1)
        ctx.reset_counters()
        ctx.render(*step_range(0))
        cc = ctx.counters()
        ctx.set_option("count_visits", 0)
        steps_counted = 1
        rays_int = cc["path_vertices"]
        rays_res = cc["shadow_rays"] + (cc["closest_hit_rays"] - cc["path_vertices"])
        launches_per_step = {n: kcnt[i] / K for i, n in enumerate(names) if i < 6}
        # B_ray = 32 (ray in) + 32 * nodes + 48 * triangle tests + 16 (hit out)
        bytes_int = 48.0 * rays_int + 32.0 * cc["nodes_visited"][0] + 48.0 * cc["prims_tested"][0]
        bytes_res = 48.0 * rays_res + 32.0 * cc["nodes_visited"][1] + 48.0 * cc["prims_tested"][1]
        # shading: B_vertex = 2 * (124 + 124 + 64) state + 124 * 2 spectra + 124 light
        bytes_shade = cc["path_vertices"] * (2 * (124 + 124 + 64) + 124 * 2 + 124.0)
        algo = {"intersect": bytes_int, "resolve": bytes_res, "shade": bytes_shade}
        peak, peak_src = peaks()
        name = names[dom]
        # per kernel: algorithmic bytes of one launch / its average launch time (CUDA events on the
        # render stream).  K3's "launch" is one bounce = its per-class launches together.
        rooflines = {}
        for kn in ("intersect", "shade", "resolve"):
            i = names.index(kn)
            ms1 = kms[i] / max(kcnt[i], 1)
            b1 = algo[kn] / steps_counted / max(launches_per_step[kn], 1)
            ach = b1 / (ms1 * 1e-3) * 1e-9
            rooflines[kn] = {"bound": "hbm", "kernel": "k_" + kn, "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "algorithmic_bytes_per_launch": b1, "avg_launch_ms": ms1}
        # the traversal working set of this scene (100 KB of nodes and leaf blocks) is L2 / L1
        # resident: only 48 B per ray (ray in, hit out) are HBM traffic -- say so beside the number
        for kn, nrays in (("intersect", rays_int), ("resolve", rays_res)):
            rooflines[kn]["note"] = ("node / primitive bytes are cache hits for this scene; HBM carries the "
                                     "rays and hit records only")
            rooflines[kn]["hbm_bytes_per_launch_rays_and_hits"] = 48.0 * nrays / steps_counted / max(
                launches_per_step[kn], 1)
        roof = None
        if name in rooflines:
            roof = dict(rooflines[name])
            traffic, traffic_src = None, None
            tpath = os.path.join(ROOT, "profiles", f"k_{name}_traffic.json")
            if os.path.exists(tpath):   # dram bytes per launch from a committed ncu --set full capture
                tj = json.load(open(tpath))
                traffic = tj.get("dram_bytes_per_launch")
                traffic_src = "static capture, not measured by this run: " + tj.get("source", tpath)
            roof.update({"traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "note": "algorithmic bytes per SURVEY.md 8(d); K3 launch = one bounce (all shading classes)"})
        # SURVEY 8(d) whole-sample roofline: B_sample = sum of B_ray + n_vertices * B_vertex + B_film
        samples_step = float(RES_W * RES_H * SPP_STEP)
        b_sample = (bytes_int + bytes_res + bytes_shade) / steps_counted / samples_step + 256.0
        ach_sample = b_sample * samples_step / (dev_ms / K * 1e-3) * 1e-9
        roofline_sample = {"bound": "hbm", "bytes_per_sample": b_sample, "achieved": ach_sample, "peak": peak,
                           "unit": "GB/s", "frac": ach_sample / peak,
                           "msamples_per_s_at_peak": peak * 1e9 / b_sample * 1e-6,
                           "rays_per_sample": (rays_int + rays_res) / steps_counted / samples_step,
                           "vertices_per_sample": cc["path_vertices"] / steps_counted / samples_step,
                           "note": "SURVEY.md 8(d): traversal bytes from the counted visit counts of this run, "
                                   "996 B per path vertex, 256 B of film per sample"}
        kernels = {names[i]: {"ms_per_step": kms[i] / K, "launches_per_step": kcnt[i] / K,
                              "share": kms[i] / max(sum(kms[:6]), 1e-9)} for i in range(6)}
        # ---- CPU baseline: the reference renderer on this box's host cores, bounded sample ----
        try:
            cpu_v, cpu_dt, cores, _ = reference_rate(2, 2 + CPU_BASELINE_SPP)
            cpu = {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "reference",
                   "sample": f"same scene and resolution, {CPU_BASELINE_SPP} spp "
                             f"(difference of a {2 + CPU_BASELINE_SPP}-spp and a 2-spp run, {cpu_dt:.1f} s)"}
        except Exception as exc:  # the baseline is reported, never a gate
            cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                   "sample": f"unavailable: {exc}"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "samples_per_step": RES_W * RES_H * SPP_STEP,
                       "spp_per_step": SPP_STEP, "partition": "sample numbers across ranks + film merge on rank 0"
                       if world > 1 else "single GPU",
                       "l2": "path state of a pass (4.2 M paths, ~3.7 GB) exceeds L2 between launches",
                       "shade_arithmetic": ctx.shade_arithmetic,
                       "last_segment": "not traced after a non-specular event at bounces == maxdepth - 1: the "
                                       "reference intersects it and uses nothing of it (path.cpp:86-102); "
                                       "films are bit-identical",
                       "sample_numbers_of_rank": [emu_rank, emu_world] if emu else None},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps,
                    "ms_per_step": {n: 1e3 * v / e2e_steps
                                    for n, v in zip(("upload", "render", "film_merge", "film_download"), e2e_parts)},
                    "product": "merged film on rank 0's host" if world > 1 else "film on the host"},
            "gpu_launches": int(c["kernel_launches"]),
            "clocks": clk, "clocks_per_rank": clk_all if world > 1 else None, "roofline": roof, "rooflines": rooflines, "roofline_sample": roofline_sample,
            "cpu_baseline": cpu, "kernels": kernels, "configs": other_configs(args) if world == 1 else None,
            "wall_ms_per_step": wall_ms / K, "film_reduce_ms": reduce_ms,
            "film_reduce_wall_ms": reduce_wall_ms if world > 1 else None,
            "film_reduce": {"nccl": "NCCL reduce of the film tensor (torch.distributed)",
                            "gather": "rank 0 sums the peers' films over NVLink peer memory (sptx_film_merge_peers)",
                            "reduce_scatter": "every rank sums its slice of all films over NVLink peer memory "
                                              "into rank 0's film (sptx_film_reduce_scatter)"}[merge_shape]
            if world > 1 else None,
            "rays_per_sample": (c["closest_hit_rays"] + c["shadow_rays"]) / max(c["camera_rays"], 1),
            "mrays_per_s": (c["closest_hit_rays"] + c["shadow_rays"]) / (dev_ms * 1e-3) * 1e-6,
        }
    sf.close()
    ctx.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    if line:
        emit(json.dumps(line))


def other_configs(args):
    """The other single-GPU configurations of BASELINE.json, measured by this run (bounded sizes):
    C1 killeroo 400x400x64 and C3 RealisticCamera 2048x2048 (16 of its 256 spp) in Msamples/s, C4
    (2.0e7 instanced triangles) in Mrays/s for 2^22-ray batches with the reference's BVHAccel loop
    on the host cores beside it.  Skipped with --no-configs; a missing blob is reported, not fatal."""
    if args.no_configs:
        return None
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    out = {}
    try:
        bc = importlib.import_module("bench_configs")
        out["l2_read_gbs_measured"] = bc.l2_peak()
        for key, fn in (("c1", lambda: bc.c1()), ("c3", lambda: bc.c3(spp=16, cpu=False)),
                        ("c4", lambda: bc.c4(n=1 << 22, n_cpu=1 << 18))):
            try:
                out[key] = fn()
            except Exception as exc:   # a configuration is reported, never a gate for the bench line
                out[key] = {"unavailable": str(exc)[:300]}
    except Exception as exc:
        out["unavailable"] = str(exc)[:300]
    return out


def bench_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    K, W = args.steps, args.warmup
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    if not os.path.exists(exe):
        emit(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/pbrt was not built"}))
        return
    cores = os.cpu_count() or 1
    secs = []
    for k in range(W + K):
        v, dt, cores, _ = reference_rate(1, 1 + REF_SPP_STEP)
        if k >= W:
            secs.append(dt)
    total = sum(secs)
    samples = RES_W * RES_H * REF_SPP_STEP * K
    value = samples / total * 1e-6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": total / K * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "samples_per_step": RES_W * RES_H * REF_SPP_STEP,
                       "spp_per_step": REF_SPP_STEP},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                             "sample": f"{REF_SPP_STEP} spp of the same scene per step (difference of a "
                                       f"{1 + REF_SPP_STEP}-spp and a 1-spp run of oracle/_ref/pbrt)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=SPP_TOTAL // SPP_STEP)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C4 legs of the line")
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"],
                    help="c2: 1024x1024 at 1024 spp (default, the metric's configuration); "
                         "c5: 3840x2160 at 4096 spp, for --gpus 2/4/8")
    args = ap.parse_args()
    set_workload(args.workload)
    # stdout carries exactly ONE JSON line: anything a library prints there while the benchmark
    # runs (NCCL's version banner, for one) is sent to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    global emit
    emit = lambda text: os.write(real_stdout, (text + "\n").encode())
    if args.impl == "reference":
        bench_reference(args)
    else:
        bench_ours(args)


if __name__ == "__main__":
    main()
