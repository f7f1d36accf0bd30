"""GPU box: relMSE of the reduced image scenes against the committed reference renders at growing GPU sample counts."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import mts_b200 as M
import golden_io as G

gold = G.load("images.npz")
for name in ("spheres", "caustics", "zoo"):
    g = M.CudaPath(G.IMAGE_SCENES[name]())
    ref = gold["image/%s/rgb" % name].astype(np.float32)
    acc, n = None, 0
    for k, spp in enumerate((8192, 8192, 16384, 32768, 65536)):
        f = g.render(spp=spp, seed=100 + k)
        acc = f.astype(np.float64) if acc is None else acc + f.astype(np.float64)   # films are weighted sums: they add
        n += spp
        img = M.develop(acc.astype(np.float32))
        print("%-9s %7d spp: relMSE %.3e  mean ratio %.4f  (%.0f ms)" % (name, n, M.rel_mse(img, ref), img.mean() / ref.mean(), g.stats.render_ms), flush=True)
    g.close()
