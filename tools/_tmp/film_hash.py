"""Film hashes of the golden scenes with the library SPTX_LIB names (experiments: two builds must agree)."""
import hashlib, importlib, os, sys, time
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
os.environ.setdefault("SPTX_EXACT", "1")
capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
from conftest import SCENES, blob_path, golden
import bench
out = []
for name in SCENES:
    try:
        g = golden(name)
    except Exception as e:
        continue
    ctx = capi.Context(0)
    sf = capi.SceneFile(blob_path(name))
    t = time.perf_counter(); ctx.upload(sf); up = (time.perf_counter() - t) * 1e3
    w, h = [int(v) for v in g["res"]]
    try:
        ctx.render(0, int(g["spp"]))
        planes = ctx.film_pack_dat(w, h)
        out.append((name, hashlib.sha1(np.ascontiguousarray(planes).tobytes()).hexdigest()[:16], round(up, 2)))
    except Exception as e:
        out.append((name, "ERR " + str(e)[:40], 0))
    sf.close(); ctx.close()
ctx = capi.Context(0); sf = capi.SceneFile(bench.scene_blob())
ctx.upload(sf)
t = time.perf_counter(); ctx.upload(sf); up = (time.perf_counter() - t) * 1e3
ctx.render(0, 2)
c, wgt = ctx.film_download(bench.RES_W, bench.RES_H)
out.append(("bench", hashlib.sha1(np.ascontiguousarray(c).tobytes()).hexdigest()[:16], round(up, 2)))
for o in out: print(*o)
