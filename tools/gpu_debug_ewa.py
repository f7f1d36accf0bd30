"""Debug helper (GPU box): dumps the 'ewa' texture lookups whose filtered value differs from the reference golden."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import mts_b200 as M
import golden_io as G

sc, tex, mats = G.mip(spp=4)
gpu = M.CudaPath(sc); gpu.preprocess()
gold = G.load("mip.npz")
out = {}
for name in sorted(tex):
    tid = gpu.flat._tex_ids[id(tex[name])][0]
    uv, duv = G.texture_inputs(name)
    v = gpu.ctx.texture_eval(tid, uv, duv)
    e = np.abs(v - gold["tex/%s/filtered" % name]).max(axis=1)
    bad = np.nonzero(e > 2e-3)[0]
    print(name, len(e), (e > 2e-3).mean(), e.max(), bad[:20])
    out[name + "/gpu"] = v
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "ewa_debug.npz"), **out)
