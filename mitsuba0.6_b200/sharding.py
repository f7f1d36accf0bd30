"""Sample-index sharding of one render over the ranks of a torch.distributed job.

The path shards naturally: camera samples are iid and the film is a linear accumulator of (w*L, w)
(include/mitsuba/render/imageblock.h:176-186), so any partition of the sample index set is valid.  Rank r of N renders
the FULL frame with the global sample indices r, r+N, r+2N, ... and the float4 films are summed once (NCCL reduce over
NVLink on GPUs; gloo in the CPU tests).  The reference's image-space block split (src/librender/renderproc.cpp) is
deliberately not mirrored: tiles would unbalance (sky vs. geometry) and need a filter-border exchange.
"""


def shard_samples(spp_total, rank, world):
    """(spp of this rank, sample_offset, sample_stride): the remainder goes to the low ranks."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    spp = spp_total // world + (1 if rank < spp_total % world else 0)
    return spp, rank, world


def render_sharded(path_tracer, spp_total, rank, world, seed=0, **kw):
    """Render this rank's share; returns the film (H, W, 4) of the rank (not yet reduced)."""
    spp, off, stride = shard_samples(spp_total, rank, world)
    return path_tracer.render(spp=spp, seed=seed, sample_offset=off, sample_stride=stride, **kw)


def reduce_film(film_tensor, dst=0):
    """Sum the per-rank films onto `dst` (the only exchange step of the path)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return film_tensor
