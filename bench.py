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
                "configs[2]: procedural 10M-triangle terrain (9,999,392 tris) with bitmap textures + bumpmap, "
                "1920x1080, 128 spp, maxDepth 8, gaussian rfilter, area light + envmap"),
    "spheres": ("material_spheres", dict(width=1024, height=1024, spp=256, max_depth=8),
                "configs[1]: 25 spheres roughconductor/roughplastic/dielectric/coating over envmap, 1024x1024, 256 spp"),
    "cornell": ("cornell_box", dict(width=512, height=512, spp=64, max_depth=5),
                "configs[0]: Cornell box 512x512, 64 spp, maxDepth 5, box rfilter"),
    "caustics": ("caustic_scene", dict(width=1024, height=1024, spp=512, max_depth=16),
                 "configs[3]: blendbsdf + conductor + dielectric caustic scene, 1024x1024, 512 spp, maxDepth 16"),
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


REF_SPP = {"terrain": 8, "spheres": 16, "cornell": 64, "caustics": 16, "terrain_small": 8}   # bounded CPU samples


def reference_built():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libmtsref.so"))


class CpuArm:
    """The CPU implementation of the path on the host cores, on a bounded sample of the workload (the full frame at a
    reduced spp).  kind "reference" = the reference's OWN code (oracle/_ref: unchanged Mitsuba 0.6 sources, its
    Scheduler with one LocalWorker per core, 32x32 blocks); kind "port" = the oracle restatement (OpenMP), used only
    when oracle/_ref has not been built."""

    def __init__(self, name, sc, flat, kw, spp=0):
        import copy
        self.kw, self.flat = kw, flat
        self.spp = int(spp or min(kw["spp"], REF_SPP.get(name, 8)))
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
            self.ref = mtsref.RefScene(sc2)
            self.cores = mtsref.load().mtsref_thread_count()
            self.build_s = self.ref.kd_build_s
        else:
            import mts_b200 as M
            self.kind = "port"
            self.cores = int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1))
            self.cpu = M.CudaPath(flat, backend="oracle")
            self.cpu.preprocess()
            self.build_s = self.cpu.ctx.get_stats().build_ms / 1e3
        self.setup_s = time.time() - t0

    def step(self, seed=1):
        W, H = self.flat.film_size
        if self.kind == "reference":
            self.ref.render()
            s = self.ref.render_s
            rays = self.ref.stats.get("normal_rays", 0) + self.ref.stats.get("shadow_rays", 0)
        else:
            self.cpu.render(spp=self.spp, read_back=False, seed=seed)
            st = self.cpu.stats
            s, rays = st.render_ms / 1e3, st.normal_rays + st.shadow_rays
        samples = W * H * self.spp
        return dict(value=samples / s / 1e6, unit="Msamples/s", cores=self.cores, kind=self.kind,
                    sample="full %dx%d frame at %d spp (of %d): %.1f s of CPU render on %d threads; kd-tree build %.1f s and scene "
                           "loading excluded" % (W, H, self.spp, self.kw["spp"], s, self.cores, self.build_s),
                    mrays_per_s=rays / s / 1e6, ms=s * 1e3)


def run_reference(args):
    """--impl reference: the CPU path integrator on the host cores, same workload/metric, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, flat, kw, desc, gen_s = build_scene(args.workload, args.spp)
    arm = CpuArm(args.workload, sc, flat, kw, args.ref_spp)
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
            "config": {"workload": desc, "name": args.workload},
            "mrays_per_s": float(np.mean([b["mrays_per_s"] for b in vals])),
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    line["cpu_baseline"]["value"] = v
    print(json.dumps(line), flush=True)


def run_cuda(args):
    import torch
    import torch.distributed as dist
    import mts_b200 as M
    from mitsuba0_6_b200 import scene as mscene

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

    gpu = M.CudaPath(flat, backend="cuda", device=local)
    t0 = time.time()
    gpu.preprocess()
    upload_s = time.time() - t0
    build_ms = gpu.ctx.get_stats().build_ms
    film = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
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
    acc = dict(normal=0, shadow=0, launches=0, intersect_ms=0.0, shade_ms=0.0, shadow_ms=0.0, splat_ms=0.0, raygen_ms=0.0,
               intersect_launches=0, render_ms=0.0, samples=0)
    t_wall = time.time()
    e0.record(stream)
    for i in range(args.steps):
        st = step(i)
        acc["normal"] += st.normal_rays; acc["shadow"] += st.shadow_rays; acc["launches"] += st.kernel_launches
        acc["samples"] += st.camera_samples
        for k in ("intersect_ms", "shade_ms", "shadow_ms", "splat_ms", "raygen_ms", "render_ms"):
            acc[k] += getattr(st, k)
        acc["intersect_launches"] += st.intersect_launches
    e1.record(stream)
    sync()
    wall_ms = (time.time() - t_wall) * 1e3
    dev_ms = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([acc["samples"], acc["normal"], acc["shadow"], acc["launches"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_ms, wall_ms = [float(x) for x in t.tolist()]
    samples, normal, shadow, launches = [float(x) for x in cnt.tolist()]
    ms_per_step = dev_ms / args.steps
    msamples = samples / (dev_ms / 1e3) / 1e6
    mrays = (normal + shadow) / (dev_ms / 1e3) / 1e6

    # ---- end to end through the plugin-facing call sequence with HOST buffers: scene upload (H2D) + BVH build +
    #      render + film read-back (D2H), every step, exactly what cudapath::render() does per `mitsuba scene.xml`
    e2e = None
    if not args.no_e2e:
        h2d = scene_bytes(flat)
        d2h = W * H * 16
        gpu.close()
        del gpu
        torch.cuda.synchronize()
        pin_scene(flat)
        e2e_ms = []
        for i in range(1 + max(1, min(args.steps, 2))):
            sync()
            t0 = time.time()
            g2 = M.CudaPath(flat, backend="cuda", device=local)
            g2.preprocess()
            img = g2.render(spp=spp_rank, seed=i, sample_offset=rank, sample_stride=world)   # includes read_film -> host
            if world > 1:
                ft = torch.from_numpy(img).cuda()
                dist.reduce(ft, dst=0, op=dist.ReduceOp.SUM)
                img = ft.cpu().numpy()
            sync()
            if i > 0:
                e2e_ms.append((time.time() - t0) * 1e3)
            e2e_build_ms = g2.ctx.get_stats().build_ms
            g2.close()
            del g2
        tt = torch.tensor([float(np.mean(e2e_ms))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_t = float(tt.item())
        e2e = {"value": W * H * spp_total / (e2e_t / 1e3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": e2e_t, "bvh_build_ms": e2e_build_ms,
               "includes": "scene upload + GPU BVH build + render + film read-back"}

    if rank == 0:
        # ---- roofline of the dominant kernel (closest-hit traversal) from an instrumented, untimed run ----
        g3 = M.CudaPath(flat, backend="cuda", device=local)
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
        # this rank's own kernel times (CUDA events around every launch on the launching stream)
        times = {k: acc[k] for k in ("intersect_ms", "shade_ms", "shadow_ms", "splat_ms", "raygen_ms")}
        dominant = max(times, key=times.get)
        peak, peak_src = measured_peak()
        rays_rank = acc["normal"]
        launches_i = max(acc["intersect_launches"], 1)
        achieved = (rays_rank * b_ray) / (acc["intersect_ms"] / 1e3) / 1e9 if acc["intersect_ms"] > 0 else 0.0
        # DRAM traffic of the same kernel from the committed `ncu --set full` capture (bytes per ray x this run's rays
        # per launch): it cannot be measured live, a number taken under a profiler is never a bench value
        traffic, traffic_src = None, None
        tpath = os.path.join(ROOT, "profiles", "r1f_traffic_k_intersect.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f)
            traffic = tj["dram_bytes_per_ray"] * rays_rank / launches_i
            traffic_src = "profiles/r1f_traffic_k_intersect.json: %.0f DRAM B/ray (ncu dram__bytes_read+write) x rays_per_launch" % tj["dram_bytes_per_ray"]
        roofline = {"bound": "hbm", "kernel": "k_intersect", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "bytes_per_ray": b_ray, "nodes_per_ray": nodes_per_ray, "prims_per_ray": prims_per_ray,
                    "bytes_per_shadow_ray": b_sray, "nodes_per_shadow_ray": nodes_per_sray,
                    "rays_per_launch": rays_rank / launches_i,
                    "avg_launch_ms": acc["intersect_ms"] / launches_i, "launches": launches_i,
                    "kernel_ms_share": {k: v / max(acc["render_ms"], 1e-9) for k, v in times.items()},
                    "dominant_by_time": dominant}
        line = {"metric": "Msamples/s", "value": msamples, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": desc, "name": args.workload, "width": W, "height": H, "spp": spp_total,
                           "spp_per_rank": spp_rank, "max_depth": kw["max_depth"], "parallelism": "sample-split x%d" % world,
                           "l2_policy": "inputs larger than L2 (per-step working set: scene + wavefront state >> 126 MB)"},
                "mrays_per_s": mrays, "rays_per_sample": (normal + shadow) / max(samples, 1),
                "wall_ms_per_step": wall_ms / args.steps, "bvh_build_ms": build_ms, "upload_s": upload_s,
                "scene_gen_s": gen_s, "gpu_launches": int(launches), "clocks": clk, "roofline": roofline}
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = {k: v for k, v in CpuArm(args.workload, sc, flat, kw, args.ref_spp).step().items() if k != "ms"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
