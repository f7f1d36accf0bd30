"""GPU box: wall-clock of one RenderJob of the reference host with the cudapath plugin on a bench workload (stage timing of the
plugin on stderr with MTS_PLUGIN_TIMING=1)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle", "refbuild"))
import bench
import mtsref

name = sys.argv[1] if len(sys.argv) > 1 else "terrain"
sc, flat, kw, desc, gen_s = bench.build_scene(name)
W, H = flat.film_size
t0 = time.time()
rp = mtsref.RefScene(sc, integrator="cudapath")
print("reference scene set-up %.1f s (kd-tree %.1f s)" % (time.time() - t0, rp.kd_build_s), flush=True)
for i in range(3):
    rp.render()
    print("job %d: %.0f ms -> %.1f Msamples/s" % (i, rp.render_s * 1e3, W * H * kw["spp"] / rp.render_s / 1e6), flush=True)
