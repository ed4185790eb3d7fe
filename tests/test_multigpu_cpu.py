"""Host-side multi-GPU logic on CPU: sample partition + film reduce with gloo, world_size 2.

The device work of a rank is replaced by a deterministic stand-in film (a function of the sample
numbers it was given), so what is checked is the plumbing bench.py uses at N > 1: every sample
number is rendered exactly once across ranks, and the reduced film on rank 0 equals the
single-process film (the property Film::MergeFilmTile gives the reference,
src/core/film.cpp:124-142).
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
part = importlib.import_module("pbrt-v3-spectral_b200.partition")

H, W, SPP_STEP, STEPS = 6, 5, 4, 3


def fake_film(sample_numbers):
    """[H][W][32] stand-in: each sample adds an integer pattern (exact in float32) + weight 1."""
    film = np.zeros((H, W, 32), np.float32)
    yy, xx = np.mgrid[0:H, 0:W]
    for s in sample_numbers:
        for b in range(31):
            film[:, :, b] += ((s * 7 + b * 3 + yy * 5 + xx) % 11).astype(np.float32)
        film[:, :, 31] += 1.0
    return film


def test_sample_blocks_cover_each_number_once():
    for world in (1, 2, 4, 8):
        seen = []
        for k in range(STEPS):
            for r in range(world):
                a, b = part.sample_block(k, r, world, SPP_STEP)
                assert b - a == SPP_STEP
                seen += list(range(a, b))
        assert sorted(seen) == list(range(STEPS * world * SPP_STEP))


def test_sample_block_wraps_and_validates():
    assert part.sample_block(16, 0, 1, 64, 1024) == (0, 64)
    assert part.sample_block(3, 1, 2, 64, 1024) == (448, 512)
    with pytest.raises(ValueError):
        part.sample_block(0, 2, 2, 64)
    with pytest.raises(ValueError):
        part.sample_block(0, 0, 1, 0)


def test_split_samples():
    assert part.split_samples(10, 4) == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert part.split_samples(4096, 8)[-1] == (3584, 4096)
    assert part.split_samples(0, 2) == [(0, 0), (0, 0)]
    with pytest.raises(ValueError):
        part.split_samples(4, 0)


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # the film-handle exchange of merge_films_over_peers: 64 opaque bytes per rank, rank order
    handles = part.gather_handles(bytes([rank * 16 + (i % 16) for i in range(64)]), dist)
    assert len(handles) == world
    for r in range(world):
        assert handles[r] == bytes([r * 16 + (i % 16) for i in range(64)])
    mine = []
    for k in range(STEPS):
        a, b = part.sample_block(k, rank, world, SPP_STEP)
        mine += list(range(a, b))
    film = torch.from_numpy(fake_film(mine))
    part.reduce_film(film, dist, dst=0)
    # timing convention of bench.py: max over ranks
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        q.put((film.numpy().copy(), float(t[0])))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_film_reduce_matches_single_process():
    import torch.multiprocessing as mp
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    film, tmax = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    expect = fake_film(range(STEPS * world * SPP_STEP))
    assert np.array_equal(film, expect)
    assert film[:, :, 31].min() == STEPS * world * SPP_STEP
    assert tmax == world


def test_reduce_film_single_process_is_identity():
    import torch
    f = torch.ones(4, 4, 32)
    assert part.reduce_film(f) is f
