#!/usr/bin/env python3
"""Scene blobs + reference fixtures for the full-size configurations of BASELINE.json.

    C1  killeroo400       scenes/killeroo-simple.pbrt, 400x400, 64 spp (the reference's own case)
    C2  cornell1024       procedural spectral Cornell box, 1024x1024 (the bench workload)
    C5  cornell3840x2160  the same scene at 4K (multi-GPU configuration)
    C3  cornell_lens2048  RealisticCamera double-Gauss, 2048x2048, 256 spp (lens-tracing path)
    C4  instanced_c4      100 instances of a 199,712-triangle mesh (2.0e7 instanced triangles)

Runs only in the build container (needs /root/reference, host/_build/pbrt_b200, oracle/_ref).
Blobs go to scenes/_gen/ (the two large ones, C1 and C4, are committed xz-compressed and unpacked
on first use by capi.unpack_blob).  Fixtures are reference outputs reduced to a committable size:
    c1: the reference's 64-spp .dat as 8x8 block means          -> tests/golden/c1_killeroo400.npz
    c3: the reference's 2-spp .dat as 16x16 block means         -> tests/golden/c3_cornell_lens2048.npz
    c2: the reference's 4-spp .dat as 8x8 block means           -> tests/golden/c2_cornell1024.npz
    c5: the reference's 2-spp .dat as 16x16 block means         -> tests/golden/c5_cornell4k.npz
    c4: Scene::Intersect / IntersectP of the first 2^17 rays of the seeded incoherent batch
        (tools/config_rays.py)                                  -> tests/golden/c4_instanced.npz
"""
import importlib
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_golden as mg  # noqa: E402
import config_rays  # noqa: E402

capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
NT = str(os.cpu_count() or 8)


def block_means(planes, k):
    b, h, w = planes.shape
    return planes.reshape(b, h // k, k, w // k, k).mean(axis=(2, 4)).astype(np.float32)


def flatten(pbrt, blob, work, pack=False):
    mg.run([mg.PBRT_B200, "--quiet", pbrt], env={"SPTX_DUMP_SCENE": blob, "SPTX_DUMP_ONLY": "1"}, cwd=work)
    if pack:  # the committed form of the large blobs (capi.unpack_blob restores them)
        import lzma
        with open(blob, "rb") as src, lzma.open(blob + ".xz", "wb", preset=9) as dst:
            dst.write(src.read())


def c1():
    name = "killeroo400"
    spec = dict(ref="killeroo-simple.pbrt", res=(400, 400), spp=64)
    with tempfile.TemporaryDirectory() as work:
        pbrt = mg.scene_file(name, spec, work)
        flatten(pbrt, os.path.join(mg.GEN, name + ".sptx"), work, pack=True)
        mg.run([mg.PBRT_REF, "--quiet", "--nthreads", NT, pbrt], cwd=work)
        dat = capi.read_dat(os.path.join(work, name + ".dat"))
    np.savez_compressed(os.path.join(mg.GOLD, "c1_killeroo400.npz"), res=np.array([400, 400]), spp=np.array(64),
                        block=np.array(8), means=block_means(dat, 8), bin_sums=dat.sum(axis=(1, 2)))
    print("c1: dat mean %.5f" % dat.mean())


def c3():
    name = "cornell_lens2048"
    with tempfile.TemporaryDirectory() as work:
        # blob for the timed run: 256 spp
        pbrt = os.path.join(work, name + ".pbrt")
        mg.run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell_lens", pbrt,
                "res=2048", "spp=256"])
        flatten(pbrt, os.path.join(mg.GEN, name + ".sptx"), work)
        # reference render at 2 spp (sample numbers 0,1 of the same Halton streams)
        pbrt2 = os.path.join(work, name + "_2spp.pbrt")
        mg.run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell_lens", pbrt2,
                "res=2048", "spp=2"])
        mg.run([mg.PBRT_REF, "--quiet", "--nthreads", NT, pbrt2], cwd=work)
        dat = capi.read_dat(os.path.join(work, name + "_2spp.dat"))
    np.savez_compressed(os.path.join(mg.GOLD, "c3_cornell_lens2048.npz"), res=np.array([2048, 2048]),
                        spp=np.array(2), block=np.array(16), means=block_means(dat, 16),
                        bin_sums=dat.sum(axis=(1, 2)))
    print("c3: dat mean %.5f" % dat.mean())


def c4():
    name = "instanced_c4"
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, name + ".pbrt")
        mg.run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "instanced", pbrt,
                "res=1024", "spp=4", "grid=317", "inst=10"])
        blob = os.path.join(mg.GEN, name + ".sptx")
        flatten(pbrt, blob, work, pack=True)
        lo, hi = mg.read_root_bounds(blob)
        n_sub = 1 << 17
        rays = config_rays.incoherent_rays(n_sub, lo, hi, first=0)
        rays.tofile(os.path.join(work, "rays.bin"))
        env = {"ORACLE_RAYS_IN": "rays.bin", "ORACLE_HITS_OUT": "hits.bin", "ORACLE_ANYHIT_OUT": "any.bin"}
        mg.run([mg.ORACLE, "--quiet", "--nthreads", NT, pbrt], env=env, cwd=work)
        hits = np.fromfile(os.path.join(work, "hits.bin"), dtype=mg.HIT)
        anyh = np.fromfile(os.path.join(work, "any.bin"), np.int32)
    np.savez_compressed(os.path.join(mg.GOLD, "c4_instanced.npz"), n=np.array(n_sub), lo=lo, hi=hi,
                        hit_prim=hits["prim"], hit_inst=hits["instance"], hit_t=hits["t"],
                        n_nodes=hits["n_nodes"], n_tests=hits["n_tests"], anyhit=anyh.astype(np.int8))
    print("c4: %d rays, %.3f hit, %.1f nodes and %.1f tests per ray" %
          (n_sub, (hits["prim"] >= 0).mean(), hits["n_nodes"].mean(), hits["n_tests"].mean()))


def cornell_config(tag, name, res, spp_full, spp_ref, block):
    """C2 / C5: the procedural Cornell box at the configuration's resolution.  The blob carries the
    configuration's spp; the reference renders the first `spp_ref` sample numbers of the same
    Halton streams (the whole job would take the host cores of this container hours)."""
    w, h = res
    with tempfile.TemporaryDirectory() as work:
        pbrt = os.path.join(work, name + ".pbrt")
        mg.run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", pbrt,
                f"res={w}x{h}", f"spp={spp_full}"])
        flatten(pbrt, os.path.join(mg.GEN, name + ".sptx"), work)
        pbrt2 = os.path.join(work, name + "_ref.pbrt")
        mg.run([sys.executable, os.path.join(ROOT, "scenes", "gen_scenes.py"), "cornell", pbrt2,
                f"res={w}x{h}", f"spp={spp_ref}"])
        mg.run([mg.PBRT_REF, "--quiet", "--nthreads", NT, pbrt2], cwd=work)
        dat = capi.read_dat(os.path.join(work, name + "_ref.dat"))
    np.savez_compressed(os.path.join(mg.GOLD, tag + ".npz"), res=np.array([w, h]), spp=np.array(spp_ref),
                        block=np.array(block), means=block_means(dat, block), bin_sums=dat.sum(axis=(1, 2)))
    print("%s: dat mean %.5f" % (tag, dat.mean()))


def c2():
    cornell_config("c2_cornell1024", "cornell1024", (1024, 1024), 1024, 4, 8)


def c5():
    cornell_config("c5_cornell4k", "cornell3840x2160", (3840, 2160), 4096, 2, 16)


if __name__ == "__main__":
    todo = sys.argv[1:] or ["c1", "c2", "c3", "c4", "c5"]
    for t in todo:
        {"c1": c1, "c2": c2, "c3": c3, "c4": c4, "c5": c5}[t]()
