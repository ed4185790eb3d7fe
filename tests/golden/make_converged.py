#!/usr/bin/env python3
"""north_star's render criterion as committed fixtures: the REFERENCE's own CPU render at 4096 spp
(Halton sample numbers 0..4095 of every pixel) of
    converged_cornell   the spectral Cornell box (diffuse + glass + metal), 128x128
    converged_lens      the same scene through the double-Gauss RealisticCamera, 96x96
kept per pixel and per bin (float32 planes of the reference's .dat) in tests/golden/<name>.npz,
with the flattened scene in scenes/_gen/<name>.sptx.  tests/test_gpu_converged.py renders the same
4096 samples on the device and asserts the per-bin relative RMSE < 0.5 %.

Runs only where /root/reference exists (needs oracle/_ref/pbrt and host/_build/pbrt_b200).
    python tests/golden/make_converged.py [converged_cornell] [converged_lens]
"""
import importlib
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402

sys.path.insert(0, mg.ROOT)
capi = importlib.import_module("pbrt-v3-spectral_b200.capi")

SCENES = {
    "converged_cornell": dict(gen="cornell", res=128, spp=4096),
    "converged_lens": dict(gen="cornell_lens", res=96, spp=4096),
}


def make(name):
    spec = SCENES[name]
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, name + ".pbrt")
        mg.run([sys.executable, os.path.join(mg.ROOT, "scenes", "gen_scenes.py"), spec["gen"], pbrt,
                f"res={spec['res']}", f"spp={spec['spp']}"])
        mg.run([mg.PBRT_B200, "--quiet", pbrt], cwd=work,
               env={"SPTX_DUMP_SCENE": os.path.join(mg.GEN, name + ".sptx"), "SPTX_DUMP_ONLY": "1"})
        t0 = time.perf_counter()
        mg.run([mg.PBRT_REF, "--quiet", "--nthreads", str(os.cpu_count() or 8), pbrt], cwd=work)
        secs = time.perf_counter() - t0
        dat = capi.read_dat(os.path.join(work, name + ".dat"))
    np.savez_compressed(os.path.join(mg.GOLD, name + ".npz"), res=np.array([spec["res"], spec["res"]]),
                        spp=np.array(spec["spp"]), dat=dat.astype(np.float32),
                        reference_cpu_seconds=np.array(secs), cores=np.array(os.cpu_count() or 0))
    print(f"{name}: {spec['res']}^2 x {spec['spp']} spp, reference {secs:.1f} s, mean {dat.mean():.5f}")


if __name__ == "__main__":
    for n in sys.argv[1:] or list(SCENES):
        make(n)
