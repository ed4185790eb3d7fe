import importlib,os,sys,time
sys.path.insert(0,'/root/repo')
capi=importlib.import_module("pbrt-v3-spectral_b200.capi")
ctx=capi.Context(0); sf=capi.SceneFile('/root/repo/scenes/_gen/cornell1024.sptx'); ctx.upload(sf)
ctx.render(0,4)
for spp in (4,16,64):
    ctx.reset_counters(); t0=time.perf_counter(); ctx.render(0,spp); w=(time.perf_counter()-t0)*1e3
    c=ctx.counters(); print(spp,"wall %.1f ms device %.1f ms kernels %.1f ms"%(w,c["render_ms"],sum(c["kernel_ms"])))
import numpy as np
buf=(np.zeros((1024,1024,31),np.float32),np.zeros((1024,1024),np.float32))
for k in range(3):
    t0=time.perf_counter(); ctx.upload(sf); t1=time.perf_counter(); ctx.render(0,4); t2=time.perf_counter(); a,b=ctx.film_download(1024,1024,out=buf); t3=time.perf_counter()
    print("upload %.1f ms render4 %.1f ms download %.1f ms"%((t1-t0)*1e3,(t2-t1)*1e3,(t3-t2)*1e3))
