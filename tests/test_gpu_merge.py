"""The multi-GPU film merge (sptx_film_merge_peers) with two processes.

Both ranks use cuda:0 here (the test box has one GPU; CUDA IPC works between processes on one
device as it does across NVLink peers): rank r renders its block of sample numbers, rank 0 maps
the other film and merges + packs in one kernel.  The handle exchange runs over gloo.
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pytestmark = pytest.mark.gpu
W = H = 64
SPP_EACH = 4


def _worker(rank, world, port, q, shape):
    import torch.distributed as dist
    capi = importlib.import_module("pbrt-v3-spectral_b200.capi")
    part = importlib.import_module("pbrt-v3-spectral_b200.partition")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ctx = capi.Context(0)
    sf = capi.SceneFile(os.path.join(ROOT, "scenes", "_gen", "cornell64.sptx"))
    ctx.upload(sf)
    ctx.render(*part.sample_block(0, rank, world, SPP_EACH))
    planes = part.merge_films_over_peers(ctx, dist, dst=0, pack=(W, H), shape=shape)
    if shape == "reduce_scatter":   # a second merge reuses the mapped peer films (and adds the peers again)
        ctx.film_clear()
        dist.barrier()
        ctx.render(*part.sample_block(0, rank, world, SPP_EACH))
        planes = part.merge_films_over_peers(ctx, dist, dst=0, pack=(W, H), shape=shape)
    if rank == 0:
        merged = ctx.film_download(W, H)
        q.put((planes, merged[0], merged[1], ctx.film_pack_dat(W, H)))
    dist.barrier()
    sf.close()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("shape", ["gather", "reduce_scatter"])
def test_two_rank_merge_equals_single_render(capi, shape):
    import torch.multiprocessing as mp
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_worker, args=(r, world, port, q, shape)) for r in range(world)]
    for p in procs:
        p.start()
    planes, contrib, weight, packed = q.get(timeout=300)
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    # the fused pack is the pack of the merged film, bit for bit
    assert np.array_equal(planes, packed)
    # and the merged film is the single-process render of all the samples (float sum order differs)
    ctx = capi.Context(0)
    sf = capi.SceneFile(os.path.join(ROOT, "scenes", "_gen", "cornell64.sptx"))
    ctx.upload(sf)
    ctx.render(0, world * SPP_EACH)
    c1, w1 = ctx.film_download(W, H)
    assert np.array_equal(weight, w1)
    assert np.allclose(contrib, c1, rtol=1e-5, atol=1e-7)
    assert weight.min() >= world * SPP_EACH
    sf.close()
    ctx.close()


def test_contexts_of_one_process_merge(capi):
    """sptx_film_merge_contexts: what GpuPathIntegrator::Render does with SPTX_DEVICES -- one
    context per GPU in one process (both on cuda:0 here), sample numbers split between them, films
    summed through peer access into the first context's film."""
    part = importlib.import_module("pbrt-v3-spectral_b200.partition")
    blob = os.path.join(ROOT, "scenes", "_gen", "cornell64.sptx")
    ctxs = [capi.Context(0) for _ in range(3)]
    for c in ctxs:
        sf = capi.SceneFile(blob)
        c.upload(sf)
        sf.close()
    spp = 10
    for c, (s0, s1) in zip(ctxs, part.split_samples(spp, len(ctxs))):
        c.render(s0, s1)
    capi.merge_contexts(ctxs, dst=0)
    merged, wsum = ctxs[0].film_download(W, H)
    one = capi.Context(0)
    sf = capi.SceneFile(blob)
    one.upload(sf)
    sf.close()
    one.render(0, spp)
    c1, w1 = one.film_download(W, H)
    assert np.array_equal(wsum, w1)
    assert np.allclose(merged, c1, rtol=1e-5, atol=1e-7)
    for c in ctxs + [one]:
        c.close()


def test_nccl_film_reduce_single_rank(capi):
    """sptx_film_reduce with a one-rank communicator made with ncclCommInitAll: the
    reduce is an in-place copy, the film is unchanged and the call goes through the context's
    stream.  (The multi-rank form is timed by bench.py with SPTX_BENCH_NCCL_REDUCE=1.)"""
    import ctypes as C
    try:
        import torch  # noqa: F401  (loads the bundled libnccl.so.2 into the process)
        nccl = C.CDLL("libnccl.so.2")
    except OSError:
        pytest.skip("libnccl.so.2 not loadable on this box")
    comm = C.c_void_p()
    devs = (C.c_int * 1)(0)
    assert nccl.ncclCommInitAll(C.byref(comm), 1, devs) == 0
    ctx = capi.Context(0)
    sf = capi.SceneFile(os.path.join(ROOT, "scenes", "_gen", "cornell64.sptx"))
    ctx.upload(sf)
    sf.close()
    ctx.render(0, 4)
    before = ctx.film_download(W, H)[0].copy()
    ctx.film_reduce_nccl(comm.value, root=0)
    after = ctx.film_download(W, H)[0]
    assert np.array_equal(before, after)
    ctx.close()
    nccl.ncclCommDestroy(comm)
