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


def _worker(rank, world, port, q):
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
    planes = part.merge_films_over_peers(ctx, dist, dst=0, pack=(W, H))
    if rank == 0:
        merged = ctx.film_download(W, H)
        q.put((planes, merged[0], merged[1], ctx.film_pack_dat(W, H)))
    dist.barrier()
    sf.close()
    ctx.close()
    dist.destroy_process_group()


def test_two_rank_merge_equals_single_render(capi):
    import torch.multiprocessing as mp
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
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
