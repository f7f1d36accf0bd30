"""GPU box with N GPUs: one PROCESS rendering a bench workload on 1, 2, ... N devices through mtscuda_create_multi (the path behind
the plugin's `devices` property): wall clock of upload + build (scene replicated GPU to GPU) and of render + NVLink film reduce."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import mts_b200 as M

name = sys.argv[1] if len(sys.argv) > 1 else "terrain"
nmax = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sc, flat, kw, desc, gen_s = bench.build_scene(name)
W, H = flat.film_size
n = int(sys.argv[3]) if len(sys.argv) > 3 else 1      # first device count (one count per process avoids pool effects between contexts)
while n <= nmax:
    t0 = time.time()
    g = M.CudaPath(flat, devices=list(range(n)))
    g.preprocess()
    t_up = time.time() - t0
    g.render(read_back=False, seed=9)           # warm-up
    ts = []
    for i in range(2):
        t1 = time.time()
        film = g.render(seed=i)                 # render on n devices + film reduce + read-back
        ts.append(time.time() - t1)
    t = min(ts)
    print("%d device(s): upload + build %.0f ms, frame %.1f ms wall (device %.1f ms) -> %.1f Msamples/s, film mean %.5f"
          % (n, t_up * 1e3, t * 1e3, g.stats.render_ms, W * H * kw["spp"] / t / 1e6, M.develop(film).mean()), flush=True)
    g.close()
    n *= 2
