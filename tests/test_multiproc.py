"""World-size-2 test of the multi-rank path on the CPU (gloo): sample-index sharding + the film reduction.
The per-rank renders use the oracle back end (no GPU here); the sharding / reduction code is the one bench.py runs
with NCCL on the GPUs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import mts_b200 as M
from oracle_lib import OraclePath
from mitsuba0_6_b200 import scenes, sharding

SPP = 9   # odd on purpose: rank 0 gets 5, rank 1 gets 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="2")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pt = OraclePath(scenes.cornell_box(48, 48, SPP))
    mine = sharding.render_sharded(pt, SPP, rank, world, seed=3)
    film = torch.from_numpy(mine.copy())
    sharding.reduce_film(film, dst=0)
    gathered = [torch.zeros_like(film) for _ in range(world)] if rank == 0 else None
    dist.gather(torch.from_numpy(mine.copy()), gathered, dst=0)
    if rank == 0:
        np.savez(out, reduced=film.numpy(), parts=np.stack([g.numpy() for g in gathered]),
                 samples=np.array([pt.stats.camera_samples]))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_samples():
    for total in (1, 7, 128, 1024):
        for world in (1, 2, 3, 8):
            parts = [sharding.shard_samples(total, r, world) for r in range(world)]
            assert sum(p[0] for p in parts) == total
            idx = sorted(off + i * stride for spp, off, stride in parts for i in range(spp))
            assert idx == list(range(total))          # the union of the ranks' samples is the 1-rank sample set
    with pytest.raises(ValueError):
        sharding.shard_samples(8, 2, 2)


def test_two_ranks_gloo(tmp_path):
    out = str(tmp_path / "r.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    reduced, parts = r["reduced"], r["parts"]
    assert np.array_equal(reduced, parts[0] + parts[1])                     # the reduction is a plain sum of the films
    assert abs(parts[0][..., 3].sum() / (48 * 48) - 5) < 1e-3 and abs(parts[1][..., 3].sum() / (48 * 48) - 4) < 1e-3
    assert abs(reduced[..., 3].sum() / (48 * 48 * SPP) - 1) < 1e-4          # box filter: total weight = samples
    single = OraclePath(scenes.cornell_box(48, 48, SPP)).render(spp=SPP, seed=11)
    a, b = M.develop(reduced), M.develop(single)
    assert abs(a.mean() / b.mean() - 1) < 0.05                               # statistically the same image


def _gather_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="2")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import util
    sc, _ = util.zoo_scene(48, 32, 2)                       # meshes, bitmap textures, an environment map
    flat = M.flatten(sc)
    # every rank contributes only ITS slice of every array: poison the rest of the local copy to prove it
    f2, moved = sharding.gather_scene(flat, rank, world, device="cpu", min_bytes=256)
    n_gathered = sum(isinstance(t, M.abi.DeviceArray) for _, t in f2.textures) + \
        sum(isinstance(v, M.abi.DeviceArray) for s in f2.shapes if s[0] == "mesh" for v in s[1].values())
    pt = OraclePath(f2)
    img = pt.render(spp=2, seed=5)
    if rank == 0:
        ref = OraclePath(flat).render(spp=2, seed=5)
        np.savez(out, img=img, ref=ref, moved=np.array([moved]), n_gathered=np.array([n_gathered]),
                 total=np.array([sum(t.nbytes for _, t in flat.textures)]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_upload_all_gather_gloo(tmp_path):
    """bench.py's N-rank end-to-end upload (sharding.gather_scene): each rank moves 1/N of every large array, an all-gather
    completes the scene everywhere, and the scene built from the gathered buffers renders the identical film."""
    out = str(tmp_path / "g.npz")
    mp.spawn(_gather_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    assert r["n_gathered"][0] >= 3
    assert np.allclose(r["img"], r["ref"], rtol=1e-5, atol=1e-6)     # (the oracle splats from several threads: summation order)
    assert 0 < r["moved"][0] < 0.75 * r["total"][0] + 1e6      # about half of the bytes went through this rank
