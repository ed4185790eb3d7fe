#!/usr/bin/env python3
"""One wavefront pass of the bench workload (for ncu): cornell 1024x1024, samples [a, b).
Usage: profile_step.py [scene] [spp per pass] [wavefront paths]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
scene = sys.argv[1] if len(sys.argv) > 1 else "cornell1024"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
ctx = capi.Context(0)
if len(sys.argv) > 3:
    capi.lib().sptx_set_wavefront_size(ctx.h, int(sys.argv[3]))
sf = capi.SceneFile(os.path.join(ROOT, "scenes", "_gen", scene + ".sptx"))
ctx.upload(sf)
ctx.render(0, spp)      # warm-up pass (also allocates the path arrays)
ctx.reset_counters()
ctx.render(spp, 2 * spp)
c = ctx.counters()
print({k: c[k] for k in ("render_ms", "kernel_launches", "camera_rays", "closest_hit_rays", "shadow_rays")})
print({n: round(ms, 3) for n, ms in zip(capi.KERNEL_NAMES, c["kernel_ms"])})
