"""Multi-GPU host logic: how a render is split across ranks and how the films meet again.

The reference splits the image into 16x16 tiles over threads and merges them with
Film::MergeFilmTile (src/core/integrator.cpp:235-249, src/core/film.cpp:124-142).  Across GPUs
the natural split is by Halton sample number instead: every sample of a pixel is independent
(HaltonSampler::GetIndexForSample is a pure function of the sample number,
src/samplers/halton.cpp:96-115), the scene is replicated, and each rank renders ALL pixels for
its own block of sample numbers.  The only exchange step is the sum of the per-rank films.

No CUDA in this file: it is exercised on CPU with the gloo backend (tests/test_cpu.py) and used
with NCCL device tensors by bench.py.
"""


def sample_block(step, rank, world, spp_step, spp_total=None):
    """Half-open range of sample numbers rank `rank` renders in step `step`.

    Step k covers the `world * spp_step` consecutive sample numbers starting at
    k * world * spp_step; rank r takes the r-th block of spp_step.  With spp_total given the
    numbering wraps (benchmark loops longer than one render reuse sample numbers).
    """
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    if spp_step <= 0:
        raise ValueError("spp_step must be positive")
    s0 = (step * world + rank) * spp_step
    if spp_total is not None:
        s0 %= max(spp_total * world, spp_step)
    return s0, s0 + spp_step


def split_samples(spp, world):
    """Partition sample numbers [0, spp) into `world` contiguous blocks (sizes differ by <= 1)."""
    if world <= 0:
        raise ValueError("world must be positive")
    base, extra = divmod(spp, world)
    out, s = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((s, s + n))
        s += n
    return out


def reduce_film(film, dist=None, dst=0):
    """Sum the per-rank films onto rank `dst` (in place), the multi-GPU MergeFilmTile.

    `film` is a torch tensor viewing [H][W][32] float32 (31 bins of contribSum + filterWeightSum);
    both parts are plain sums, so one SUM reduce merges them.  With dist None (single process)
    this is a no-op.  Returns the tensor.
    """
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def gather_handles(handle, dist=None):
    """Every rank's film handle (bytes), indexed by rank, on every rank.  The handles are 64-byte
    CUDA IPC tokens: a host-side exchange of a few bytes, so any backend does (a uint8 all_gather)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [bytes(handle)]
    import torch
    world = dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    mine = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    return [bytes(t.cpu().numpy().tobytes()) for t in out]


def merge_films_over_peers(ctx, dist=None, dst=0, pack=None, shape="reduce_scatter"):
    """The exchange step of a sample-partitioned render: every rank exports its device film and
    the films are summed in rank order over NVLink peer memory into rank `dst`'s film.
    shape "reduce_scatter" (default): every rank sums its own 1/N slice of the film over all ranks
    and writes it into dst's film (sptx_film_reduce_scatter) -- each link carries 1/N of a film;
    shape "gather": dst alone maps the peers and sums whole films with one kernel, fused with the
    .dat pack (sptx_film_merge_peers) -- dst's ingress carries N - 1 films.
    Both give the same bits.  Barriers keep the films alive and quiescent around the merge.
    Returns the .dat planes on `dst` when `pack` = (width, height) is given (else None)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return ctx.film_pack_dat(*pack) if pack is not None else None
    # film_export drains this rank's render stream, and the all_gather returns only once every rank
    # has contributed its handle: when it completes, all films are complete (no separate barrier)
    handles = gather_handles(ctx.film_export(), dist)
    planes = None
    if shape == "gather":
        if dist.get_rank() == dst:
            planes = ctx.film_merge_peers(handles, dst, pack)
        dist.barrier()
        return planes
    ctx.film_reduce_scatter(handles, dist.get_rank(), dst)
    dist.barrier()
    if dist.get_rank() == dst and pack is not None:
        planes = ctx.film_pack_dat(*pack)
    return planes
