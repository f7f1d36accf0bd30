"""Sample-index sharding of one render over the ranks of a torch.distributed job.

The path shards naturally: camera samples are iid and the film is a linear accumulator of (w*L, w)
(include/mitsuba/render/imageblock.h:176-186), so any partition of the sample index set is valid.  Rank r of N renders
the FULL frame with the global sample indices r, r+N, r+2N, ... and the float4 films are summed once (NCCL reduce over
NVLink on GPUs; gloo in the CPU tests).  The reference's image-space block split (src/librender/renderproc.cpp) is
deliberately not mirrored: tiles would unbalance (sky vs. geometry) and need a filter-border exchange.
"""


import numpy as np


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


def gather_scene(flat, rank, world, device="cuda", min_bytes=1 << 16):
    """End-to-end leg with N ranks: every rank copies 1/N of each large scene array from pinned host memory to its GPU
    over its own PCIe link, and one NCCL all-gather per array over NVLink gives every GPU the whole scene -- each host byte
    crosses PCIe once per JOB instead of once per rank.  Returns a shallow copy of `flat` whose large arrays are
    abi.DeviceArray views of the gathered buffers (libmtscuda takes them like host buffers: cudaMemcpyDefault), and the
    bytes this rank copied host to device.  `device="cpu"` runs the same slicing / all-gather over gloo (the CPU tests:
    the gathered buffers are then host memory, which every library of this ABI accepts)."""
    import copy
    import torch
    import torch.distributed as dist
    from . import abi
    h2d = 0
    keep = []

    def gather(a):
        nonlocal h2d
        if not isinstance(a, np.ndarray) or a.nbytes < min_bytes:
            return a
        assert a.dtype.itemsize == 4 and a.flags["C_CONTIGUOUS"]
        flat32 = a.reshape(-1).view(np.int32)                      # float32 / uint32 travel as 32-bit words
        n = flat32.size
        chunk = (n + world - 1) // world
        lo, hi = min(rank * chunk, n), min((rank + 1) * chunk, n)
        full = torch.empty(world * chunk, dtype=torch.int32, device=device)
        mine = full[rank * chunk:(rank + 1) * chunk]
        if hi > lo:
            mine[:hi - lo].copy_(torch.from_numpy(flat32[lo:hi]), non_blocking=True)   # pinned host -> this rank's slice
            h2d += 4 * (hi - lo)
        dist.all_gather_into_tensor(full, mine)                    # in place: every rank contributes its own slice
        keep.append(full)
        return abi.DeviceArray(full.data_ptr(), a.shape, a.dtype, owner=full)

    f2 = copy.copy(flat)
    f2.textures = [(d, gather(t)) for d, t in flat.textures]
    f2.emitters = [(d, gather(t) if t is not None else None) for d, t in flat.emitters]
    shapes = []
    for s in flat.shapes:
        if s[0] == "mesh":
            s = (s[0], {k: gather(v) for k, v in s[1].items()}) + tuple(s[2:])
        shapes.append(s)
    f2.shapes = shapes
    f2._gathered = keep
    if device != "cpu":
        torch.cuda.current_stream().synchronize()    # the library reads the buffers on its own stream
    return f2, h2d
