#!/usr/bin/env python
"""bench.py -- headline benchmark of the cudapath hot path (BASELINE.json metric: Msamples/s and Mrays/s).

    python bench.py --gpus 1 --steps 3 --warmup 3            # one B200
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W                  # N B200s, one rank per GPU
    python bench.py --impl reference --gpus 1 --steps 1 --warmup 0   # reference arm: CPU path integrator

A step = one full render of the workload (every pixel x spp camera samples through raygen, intersect,
shade, shadow, splat).  With N ranks every rank renders the full frame with spp/N of the samples
(global sample index = rank + N*i) and the float4 films are summed with one NCCL reduce (strong
scaling: total work fixed).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (generator, kwargs, description)
    "terrain": ("terrain_scene", dict(width=1920, height=1080, spp=128, max_depth=8),
                "configs[2]: procedural 10M-triangle terrain (9,999,392 tris), four 4096^2 RGB bitmap textures (default "
                "filterType ewa) + 4096^2 bump map, bumpmap over roughplastic and over diffuse, 1920x1080, 128 spp, "
                "maxDepth 8, gaussian rfilter, area light + envmap"),
    "spheres": ("material_spheres", dict(width=1024, height=1024, spp=256, max_depth=8),
                "configs[1]: 25 spheres roughconductor/roughplastic/dielectric/coating over envmap, 1024x1024, 256 spp"),
    "cornell": ("cornell_box", dict(width=512, height=512, spp=64, max_depth=5),
                "configs[0]: Cornell box 512x512, 64 spp, maxDepth 5, box rfilter"),
    "caustics": ("caustic_scene", dict(width=1024, height=1024, spp=512, max_depth=16),
                 "configs[3]: blendbsdf + conductor + dielectric caustic scene, 1024x1024, 512 spp, maxDepth 16"),
    "terrain4k": ("terrain_scene", dict(width=3840, height=2160, spp=1024, max_depth=16),
                  "configs[4]: the cfg3 scene at 3840x2160, 1024 spp, maxDepth 16 (sample-split over the ranks, NCCL film reduce)"),
    # reduced variant for quick functional checks of this script (NOT a bench configuration)
    "terrain_small": ("terrain_scene", dict(width=480, height=270, spp=8, max_depth=8, grid=300, tex_size=512),
                      "debug: 180k-triangle terrain"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--workload", default=os.environ.get("MTS_BENCH_WORKLOAD", "terrain"), choices=list(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--ref-spp", type=int, default=0, help="spp of the bounded CPU sample (reference arm / cpu_baseline); 0 = per-workload default")
    ap.add_argument("--ref-threads", type=int, default=0, help="worker threads of the reference arm (0 = all host cores, `-p 1` = per-core rate)")
    ap.add_argument("--no-other-configs", dest="other_configs", action="store_false",
                    help="skip the short cfg1/cfg2/cfg4 measurements appended to the N=1 line")
    return ap.parse_args()


def build_scene(name, spp_override=0):
    import mts_b200 as M
    from mitsuba0_6_b200 import scenes
    gen, kw, desc = WORKLOADS[name]
    kw = dict(kw)
    if spp_override:
        kw["spp"] = spp_override
    t0 = time.time()
    sc = getattr(scenes, gen)(**kw)
    flat = M.flatten(sc)
    return sc, flat, kw, desc, time.time() - t0


def scene_bytes(flat):
    n = 0
    for d, tex in flat.textures:
        n += tex.nbytes
    n += 4 * len(flat.pool) + 144 * len(flat.nodes)
    for s in flat.shapes:
        if s[0] == "mesh":
            n += sum(v.nbytes for v in s[1].values() if isinstance(v, np.ndarray))
        else:
            n += 64
    for d, tex in flat.emitters:
        if tex is not None:
            n += tex.nbytes
    return int(n)


def pin_scene(flat):
    """Re-homes the large host arrays of the flattened scene (vertex/index/tangent buffers, texels) in pinned
    (page-locked) host memory, so that the H2D copies of the e2e leg run from pinned memory as the bench contract
    asks.  The arrays stay ordinary numpy views; libmtscuda sees plain host pointers either way."""
    import torch
    keep = []

    def pin(a):
        if not isinstance(a, np.ndarray) or a.nbytes < (1 << 16):
            return a
        a = np.ascontiguousarray(a)
        t = torch.from_numpy(a).pin_memory()
        keep.append(t)
        return t.numpy()

    flat.textures = [(d, pin(t)) for d, t in flat.textures]
    flat.emitters = [(d, pin(t) if t is not None else None) for d, t in flat.emitters]
    shapes = []
    for s in flat.shapes:
        if s[0] == "mesh":
            s = (s[0], {k: pin(v) for k, v in s[1].items()}) + tuple(s[2:])
        shapes.append(s)
    flat.shapes = shapes
    flat._pinned = keep   # owns the page-locked storage
    return flat


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.idx = device_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                p = [x.strip() for x in line.split(",")]
                if len(p) < 9:
                    continue
                try:
                    sm.append(float(p[1]))
                    mx.append(float(p[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


# ---- algorithmic bytes (SURVEY.md 8d / DESIGN.md "Roofline"): per ray 32 B in + 16 B out (4 B for shadow rays)
#      + node_bytes per BVH node popped + 48 B per primitive tested -------------------------------------------
RAY_IN, HIT_OUT, SHADOW_OUT, PRIM_BYTES = 32, 16, 4, 48


REF_SPP = {"terrain": 8, "terrain4k": 4, "spheres": 16, "cornell": 64, "caustics": 16, "terrain_small": 8}   # bounded CPU samples (reference arm, per step)
# the cpu_baseline leg of the CUDA arm also supplies the reference IMAGE of the relMSE figure: one render, more samples
REF_SPP_IMAGE = {"terrain": 32, "terrain4k": 8, "spheres": 32, "cornell": 64, "caustics": 32, "terrain_small": 16}


def reference_built():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libmtsref.so"))


def block_mean(img, k):
    """Mean over k x k pixel blocks (the frame is cropped to a multiple of k)."""
    h, w = (img.shape[0] // k) * k, (img.shape[1] // k) * k
    return img[:h, :w].reshape(h // k, k, w // k, k, -1).mean(axis=(1, 3))


def image_agreement(gpu_film, ref_film, gpu_spp, ref_spp, block=8):
    """relMSE between the GPU render and the reference's own CPU render of the same frame (BASELINE.json's metric
    includes relMSE).  Neither image is converged at bench sample counts, so the per-pixel figure is the Monte-Carlo noise
    floor of the two estimates; the test of *distributional* agreement is the relMSE of `block` x `block` pixel means
    (variance / block^2) and the ratio of the frame means."""
    import mts_b200 as M
    a, b = M.develop(gpu_film), M.develop(ref_film)
    return {"value": M.rel_mse(block_mean(a, block), block_mean(b, block)), "block": block,
            "per_pixel": M.rel_mse(a, b), "mean_ratio": float(a.mean() / b.mean()),
            "gpu_spp": int(gpu_spp), "ref_spp": int(ref_spp),
            "definition": "rel_mse (eps 1e-2, top 0.1%% trimmed) of %dx%d-pixel block means, GPU render vs the reference's CPU `path` render of the same frame" % (block, block)}


class CpuArm:
    """The CPU implementation of the path on the host cores, on a bounded sample of the workload (the full frame at a
    reduced spp).  kind "reference" = the reference's OWN code (oracle/_ref: unchanged Mitsuba 0.6 sources, its
    Scheduler with one LocalWorker per core, 32x32 blocks); kind "port" = the oracle restatement (OpenMP), used only
    when oracle/_ref has not been built."""

    def __init__(self, name, sc, flat, kw, spp=0, threads=0):
        import copy
        self.kw, self.flat = kw, flat
        self.spp = int(spp or min(kw["spp"], REF_SPP.get(name, 8)))
        self.film = None
        t0 = time.time()
        if reference_built():
            sys.path.insert(0, os.path.join(ROOT, "oracle", "refbuild"))
            import mtsref
            self.kind = "reference"
            sc2 = copy.copy(sc)
            sc2.children = list(sc.children)
            for i, (n, c) in enumerate(sc2.children):       # same scene, sampleCount = the bounded spp
                if c.kind == "sensor":
                    s2 = copy.copy(c)
                    s2.children = [(m, (copy.copy(d) if d.kind == "sampler" else d)) for m, d in c.children]
                    for m, d in s2.children:
                        if d.kind == "sampler":
                            d.props = dict(d.props, sampleCount=self.spp)
                    sc2.children[i] = (n, s2)
            self.ref = mtsref.RefScene(sc2, threads=threads)
            self.cores = mtsref.load().mtsref_thread_count()
            self.build_s = self.ref.kd_build_s
        else:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            from oracle_lib import OraclePath   # CPU baseline leg only
            self.kind = "port"
            self.cores = int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1))
            self.cpu = OraclePath(flat)
            self.cpu.preprocess()
            self.build_s = self.cpu.ctx.get_stats().build_ms / 1e3
        self.setup_s = time.time() - t0

    def step(self, seed=1):
        W, H = self.flat.film_size
        if self.kind == "reference":
            self.film = self.ref.render()
            s = self.ref.render_s
            rays = self.ref.stats.get("normal_rays", 0) + self.ref.stats.get("shadow_rays", 0)
        else:
            self.film = self.cpu.render(spp=self.spp, seed=seed)
            st = self.cpu.stats
            s, rays = st.render_ms / 1e3, st.normal_rays + st.shadow_rays
        samples = W * H * self.spp
        return dict(value=samples / s / 1e6, unit="Msamples/s", cores=self.cores, kind=self.kind,
                    sample="full %dx%d frame at %d spp (of %d): %.1f s of CPU render on %d threads; kd-tree build %.1f s and scene "
                           "loading excluded" % (W, H, self.spp, self.kw["spp"], s, self.cores, self.build_s),
                    mrays_per_s=rays / s / 1e6, rays_per_sample=rays / samples, ref_spp=self.spp, ms=s * 1e3)


def config_of(args, desc, flat, kw, world):
    """The `config` object: identical on the CUDA arm and on the reference arm (same workload, same keys)."""
    W, H = flat.film_size
    spp_total = kw["spp"]
    return {"workload": desc, "name": args.workload, "width": W, "height": H, "spp": spp_total,
            "spp_per_rank": spp_total // world + (1 if spp_total % world else 0), "max_depth": kw["max_depth"],
            "parallelism": "sample-split x%d" % world,
            "l2_policy": "inputs larger than L2 (per-step working set: scene + wavefront state >> 126 MB)"}


def run_reference(args):
    """--impl reference: the CPU path integrator on the host cores, same workload/metric, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, flat, kw, desc, gen_s = build_scene(args.workload, args.spp)
    arm = CpuArm(args.workload, sc, flat, kw, args.ref_spp, threads=args.ref_threads)
    vals = []
    base = None
    for i in range(args.warmup + args.steps):
        base = arm.step(seed=i)
        if i >= args.warmup:
            vals.append(base)
    v = float(np.mean([b["value"] for b in vals]))
    ms = float(np.mean([b["ms"] for b in vals]))
    line = {"impl": "reference", "metric": "Msamples/s", "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(args, desc, flat, kw, max(1, args.gpus)),
            "ref_spp": arm.spp, "cores": arm.cores,
            "mrays_per_s": float(np.mean([b["mrays_per_s"] for b in vals])), "rays_per_sample": base["rays_per_sample"],
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    line["cpu_baseline"]["value"] = v
    print(json.dumps(line), flush=True)


# ---- per-kernel counters that cannot be measured live (DRAM bytes, executed instructions): read from the committed
#      `ncu --set full` capture of the same workload (profiles/kernel_counters.json, written by tools/ncu_counters.py).
#      A number taken under a profiler is never a bench value; these only annotate the live figures. -----------------
def kernel_counters():
    p = os.path.join(ROOT, "profiles", "kernel_counters.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return json.load(f)
        except Exception:
            pass
    return {}


B_HANDOVER = 320   # staged shade (scenes whose leaf materials read bitmaps): one 160-byte surface record written by
                   # k_shade_surface and read by k_shade_scatter per hit
B_SHADE = 436   # algorithmic bytes of one shaded surface hit (ray 32 + hit 16 + state 48 + geometry 108 +
                # material 64 + texels 32(+32 bump) + next ray 32 + shadow ray 48 + state 48 + queue indices 8)


def kernel_roofline(kernel, units, ms, launches, bytes_per_unit, peak, peak_src, sm_mhz, num_sms, kc):
    """HBM roofline (algorithmic bytes / live launch time) + the warp-issue roofline of SURVEY 8(d)(ii) for one kernel."""
    t = max(ms, 1e-9) / 1e3
    achieved = units * bytes_per_unit / t / 1e9
    r = {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
         "peak_source": peak_src, "bytes_per_unit": bytes_per_unit, "units": units, "units_per_launch": units / max(launches, 1),
         "avg_launch_ms": ms / max(launches, 1), "launches": launches, "traffic": None}
    c = kc.get(kernel)
    if c:
        r["traffic"] = c["dram_bytes_per_unit"] * units / max(launches, 1)
        r["traffic_source"] = "%s: %.0f DRAM B/unit (ncu dram__bytes_read+write) x units_per_launch" % (kc.get("source", "profiles/kernel_counters.json"), c["dram_bytes_per_unit"])
        r["dram_frac"] = c["dram_bytes_per_unit"] * units / t / 1e9 / peak
        if sm_mhz and c.get("warp_inst_per_unit"):
            issue_peak = num_sms * 4 * sm_mhz * 1e6          # warp instructions / s: 4 schedulers per SM, 1 issue per clock
            r["issue"] = {"achieved_ginst_s": c["warp_inst_per_unit"] * units / t / 1e9, "peak_ginst_s": issue_peak / 1e9,
                          "frac": c["warp_inst_per_unit"] * units / t / issue_peak, "lanes_per_inst": c.get("lanes_per_inst"),
                          "warp_inst_per_unit": c["warp_inst_per_unit"],
                          "note": "instructions per unit from the committed ncu capture x live units / live time; peak = SMs x 4 x SM clock"}
    return r


def run_cuda(args):
    import torch
    import torch.distributed as dist
    import mts_b200 as M

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libmtscuda has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc, flat, kw, desc, gen_s = build_scene(args.workload, args.spp)
    W, H = flat.film_size
    spp_total = kw["spp"]
    # sample-index sharding: rank r renders global samples r, r+N, r+2N, ...
    spp_rank, _, _ = M.sharding.shard_samples(spp_total, rank, world)

    gpu = M.CudaPath(flat, device=local)
    t0 = time.time()
    gpu.preprocess()
    upload_s = time.time() - t0
    build_ms = gpu.ctx.get_stats().build_ms
    film = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    film_sum = torch.zeros_like(film)          # all timed steps accumulated (different seeds): the image relMSE is computed on
    gpu.ctx.bind_film(film.data_ptr())
    stream = torch.cuda.current_stream()
    gpu.ctx.set_stream(stream.cuda_stream)

    def step(seed):
        gpu.render(spp=spp_rank, read_back=False, seed=seed, sample_offset=rank, sample_stride=world)
        if world > 1:
            M.sharding.reduce_film(film)   # NCCL sum of the float4 film over NVLink (the path's only exchange step)
        return gpu.stats

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(100 + i)
    sync()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    keys_ms = ("intersect_ms", "shade_ms", "shadow_ms", "splat_ms", "raygen_ms", "sort_ms", "miss_ms", "render_ms")
    keys_n = ("normal_rays", "shadow_rays", "kernel_launches", "camera_samples", "intersect_launches", "shade_launches",
              "shadow_launches", "shaded_paths", "miss_paths")
    acc = {k: 0.0 for k in keys_ms + keys_n}
    t_wall = time.time()
    e0.record(stream)
    for i in range(args.steps):
        st = step(i)
        film_sum += film
        for k in keys_ms + keys_n:
            acc[k] += getattr(st, k)
    e1.record(stream)
    sync()
    wall_ms = (time.time() - t_wall) * 1e3
    dev_ms = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([acc["camera_samples"], acc["normal_rays"], acc["shadow_rays"], acc["kernel_launches"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_ms, wall_ms = [float(x) for x in t.tolist()]
    samples, normal, shadow, launches = [float(x) for x in cnt.tolist()]
    ms_per_step = dev_ms / args.steps
    msamples = samples / (dev_ms / 1e3) / 1e6
    mrays = (normal + shadow) / (dev_ms / 1e3) / 1e6
    gpu_film = film_sum.cpu().numpy() if rank == 0 else None     # rank 0 holds the reduced film of every step
    num_sms = torch.cuda.get_device_properties(local).multi_processor_count

    # ---- end to end through the plugin-facing call sequence with HOST buffers: scene upload (H2D) + BVH build +
    #      render + film read-back (D2H), every step, exactly what cudapath::render() does per `mitsuba scene.xml`.
    #      With N ranks the films are summed on the devices (NCCL) and only rank 0 reads the result back.
    e2e = None
    if not args.no_e2e:
        h2d = scene_bytes(flat)
        d2h = W * H * 16
        gpu.close()
        del gpu
        torch.cuda.synchronize()
        pin_scene(flat)
        host_film = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
        e2e_ms, parts = [], {}
        for i in range(1 + max(1, min(args.steps, 3))):
            sync()
            t0 = time.time()
            if world > 1:     # sharded upload + NVLink all-gather (see gather_scene)
                flat_e2e, h2d_rank = M.sharding.gather_scene(flat, rank, world)
                t_gather = time.time()
            else:
                flat_e2e, h2d_rank, t_gather = flat, h2d, t0
            g2 = M.CudaPath(flat_e2e, device=local)
            g2.preprocess()
            t1 = time.time()
            if world > 1:
                g2.ctx.bind_film(film.data_ptr())
                g2.render(spp=spp_rank, read_back=False, seed=i, sample_offset=rank, sample_stride=world)
                t2 = time.time()
                dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)
                if rank == 0:
                    host_film.copy_(film, non_blocking=False)
            else:
                g2.render(spp=spp_rank, read_back=False, seed=i)
                t2 = time.time()
                g2.ctx.read_film()
            sync()
            t3 = time.time()
            if i > 0:
                e2e_ms.append((t3 - t0) * 1e3)
                parts = {"upload_build_ms": (t1 - t0) * 1e3, "render_ms": (t2 - t1) * 1e3, "reduce_readback_ms": (t3 - t2) * 1e3}
                if world > 1:
                    parts["sharded_h2d_allgather_ms"] = (t_gather - t0) * 1e3
            e2e_build_ms = g2.ctx.get_stats().build_ms
            g2.close()
            del g2
        tt = torch.tensor([float(np.mean(e2e_ms))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_t = float(tt.item())
        e2e = {"value": W * H * spp_total / (e2e_t / 1e3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
               "h2d_bytes_per_step_per_rank": h2d_rank,
               "upload": "whole scene from pinned host memory" if world == 1 else
                         "each rank copies 1/%d of every large array over its own PCIe link, NCCL all-gather over NVLink" % world,
               "d2h_bytes_per_step": d2h, "ms_per_step": e2e_t, "bvh_build_ms": e2e_build_ms, "rank0_breakdown_ms": parts,
               "includes": "context + scene upload from pinned host memory + GPU BVH build + render + device-side film reduce + film read-back on rank 0"}

    if rank == 0:
        # ---- rooflines from an instrumented, untimed run (node / primitive counts per ray) + the live kernel times ----
        g3 = M.CudaPath(flat, device=local)
        g3.preprocess()
        ispp = max(1, min(spp_rank, 4))
        g3.render(spp=ispp, read_back=False, seed=0, instrument=True)
        si = g3.stats
        node_bytes = 80   # one eight-wide compressed node (WideNode, scene.cuh)
        nodes_per_ray = si.bvh_nodes_visited / max(si.normal_rays, 1)
        prims_per_ray = si.bvh_prims_tested / max(si.normal_rays, 1)
        nodes_per_sray = si.bvh_nodes_visited_shadow / max(si.shadow_rays, 1)
        prims_per_sray = si.bvh_prims_tested_shadow / max(si.shadow_rays, 1)
        g3.close()
        b_ray = RAY_IN + HIT_OUT + node_bytes * nodes_per_ray + PRIM_BYTES * prims_per_ray
        b_sray = RAY_IN + SHADOW_OUT + node_bytes * nodes_per_sray + PRIM_BYTES * prims_per_sray
        staged = any(r.tex >= 0 for n in flat.nodes for r in (n.t0, n.t1, n.t2))     # what mtscuda_build decides (leafTextured)
        b_shade = B_SHADE + (B_HANDOVER if staged else 0)
        # this rank's own kernel times (CUDA events around every launch on the launching stream)
        times = {k: acc[k] for k in ("intersect_ms", "shade_ms", "shadow_ms", "splat_ms", "raygen_ms", "sort_ms", "miss_ms")}
        peak, peak_src = measured_peak()
        kc = kernel_counters()
        sm_mhz = (clk or {}).get("sm_mhz")
        rl = {
            "k_intersect": kernel_roofline("k_intersect", acc["normal_rays"], acc["intersect_ms"], acc["intersect_launches"], b_ray, peak, peak_src, sm_mhz, num_sms, kc),
            "k_shade": kernel_roofline("k_shade", acc["shaded_paths"], acc["shade_ms"], acc["shade_launches"], b_shade, peak, peak_src, sm_mhz, num_sms, kc),
            "k_shadow": kernel_roofline("k_shadow", acc["shadow_rays"], acc["shadow_ms"], acc["shadow_launches"], b_sray, peak, peak_src, sm_mhz, num_sms, kc),
        }
        rl["k_intersect"].update(nodes_per_ray=nodes_per_ray, prims_per_ray=prims_per_ray)
        rl["k_shadow"].update(nodes_per_ray=nodes_per_sray, prims_per_ray=prims_per_sray)
        dominant = max(rl, key=lambda k: times[{"k_intersect": "intersect_ms", "k_shade": "shade_ms", "k_shadow": "shadow_ms"}[k]])
        roofline = dict(rl[dominant])
        roofline["kernel_ms_share"] = {k: v / max(acc["render_ms"], 1e-9) for k, v in times.items()}
        roofline["dominant_by_time"] = dominant
        roofline["other_kernels"] = {k: v for k, v in rl.items() if k != dominant}
        # whole pipeline, SURVEY 8(d): (rays x B_ray + shadow rays x B_sray + shaded hits x B_shade + samples x B_film) / time
        b_film = 24 + 32 * (1 if flat.film.filter_radius < 1 else 16)
        pipe_bytes = acc["normal_rays"] * b_ray + acc["shadow_rays"] * b_sray + acc["shaded_paths"] * b_shade + acc["camera_samples"] * b_film
        roofline["pipeline"] = {"bytes_per_sample": pipe_bytes / max(acc["camera_samples"], 1),
                                "achieved": pipe_bytes / (acc["render_ms"] / 1e3) / 1e9, "frac": pipe_bytes / (acc["render_ms"] / 1e3) / 1e9 / peak}
        line = {"metric": "Msamples/s", "value": msamples, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config_of(args, desc, flat, kw, world),
                "mrays_per_s": mrays, "rays_per_sample": (normal + shadow) / max(samples, 1),
                "wall_ms_per_step": wall_ms / args.steps, "bvh_build_ms": build_ms, "upload_s": upload_s,
                "scene_gen_s": gen_s, "gpu_launches": int(launches), "clocks": clk, "roofline": roofline}
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu_baseline:
            arm = CpuArm(args.workload, sc, flat, kw, args.ref_spp or REF_SPP_IMAGE.get(args.workload, 0), threads=args.ref_threads)
            cb = arm.step()
            line["cpu_baseline"] = {k: v for k, v in cb.items() if k != "ms"}
            line["relmse"] = image_agreement(gpu_film, arm.film, spp_total * args.steps, arm.spp)
            line["rays_per_sample_reference"] = cb["rays_per_sample"]
            del arm
        if world == 1 and not args.no_cpu_baseline and e2e and reference_built():
            # Cross-check of `e2e` through the drop-in plugin itself: the unchanged reference host (PluginManager, Scene,
            # RenderQueue / RenderJob, HDRFilm) renders the same scene at full spp with `cudapath.so` as its integrator --
            # pageable TriMesh / Bitmap buffers, flattening inside render(), film through Film::put.
            try:
                import mtsref
                t_p = time.time()
                rp = mtsref.RefScene(sc, integrator="cudapath")
                rp.render()                                   # first job: context creation, lazy module load
                rp.render()
                line["e2e"]["plugin"] = {"value": W * H * spp_total / rp.render_s / 1e6, "unit": "Msamples/s", "ms_per_step": rp.render_s * 1e3,
                                         "what": "RenderJob of the reference host with integrator `cudapath` (second of two jobs), "
                                                 "scene flattening + upload + build + render + Film::put inside the job",
                                         "setup_s": time.time() - t_p}
                del rp
            except Exception as e:   # a secondary measurement must never take the headline line down
                line["e2e"]["plugin"] = {"error": "%s: %s" % (type(e).__name__, e)}
        if world == 1 and args.other_configs and not args.no_cpu_baseline:
            line["other_configs"] = other_configs(args, local)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def other_configs(args, device):
    """BASELINE.json's remaining single-GPU configurations (cfg1 Cornell, cfg2 material spheres, cfg4 caustics) at their
    full sizes: device-timed Msamples/s and Mrays/s, the reference's CPU `path` on a bounded sample, image agreement.
    Short runs next to the headline workload (1 warm-up + 2 timed renders each); they are reported, not the `value`."""
    import torch
    import mts_b200 as M
    out = {}
    for name in [w for w in ("cornell", "spheres", "caustics") if w != args.workload]:
        try:
            sc, flat, kw, desc, gen_s = build_scene(name)
            W, H = flat.film_size
            g = M.CudaPath(flat, device=device)
            g.preprocess()
            g.render(read_back=False, seed=50)
            ms, rays, film = 0.0, 0, None
            for i in range(2):
                f = g.render(seed=i)
                film = f if film is None else film + f
                ms += g.stats.render_ms
                rays += g.stats.normal_rays + g.stats.shadow_rays
            samples = 2.0 * W * H * kw["spp"]
            arm = CpuArm(name, sc, flat, kw, REF_SPP_IMAGE.get(name, 0))
            cb = arm.step()
            out[name] = {"workload": desc, "msamples_per_s": samples / ms / 1e3, "mrays_per_s": rays / ms / 1e3,
                         "ms_per_frame": ms / 2, "rays_per_sample": rays / samples, "bvh_build_ms": g.ctx.get_stats().build_ms,
                         "cpu_baseline": {k: v for k, v in cb.items() if k != "ms"}, "speedup_vs_cpu": samples / ms / 1e3 / cb["value"],
                         "relmse": image_agreement(film, arm.film, 2 * kw["spp"], arm.spp)}
            g.close()
            del g, arm
            torch.cuda.synchronize()
        except Exception as e:   # a secondary measurement must never take the headline line down
            out[name] = {"error": "%s: %s" % (type(e).__name__, e)}
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
