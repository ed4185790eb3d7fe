#!/usr/bin/env python3
"""Wall time of the e2e legs around sptx_render: upload (SPTX_TRACE_UPLOAD=1 prints its phases)
and film download, on the bench scene."""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
scene = sys.argv[1] if len(sys.argv) > 1 else "cornell1024"
ctx = capi.Context(0)
sf = capi.SceneFile(os.path.join(ROOT, "scenes", "_gen", scene + ".sptx"))
ctx.upload(sf)
ctx.render(0, 4)
w = h = 1024
out = (np.zeros((h, w, 31), np.float32), np.zeros((h, w), np.float32))
for k in range(4):
    t0 = time.perf_counter()
    ctx.upload(sf)
    t1 = time.perf_counter()
    ctx.render(0, 4)
    t2 = time.perf_counter()
    ctx.film_download(w, h, out=out)
    t3 = time.perf_counter()
    print("upload %.2f ms  render(4 spp) %.2f ms  download %.2f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
