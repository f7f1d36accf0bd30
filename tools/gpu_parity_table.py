"""GPU box: per-model parity table of the CUDA library against the reference-generated goldens (the numbers behind the
percentile bars of tests/golden_check.py): relative error of BSDF eval and pdf on identical inputs -- p99, p99.9, max --
and the agreement of the sampled directions."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import mts_b200 as M
import golden_io as G
import util


def table(title, scene_z, gold, inputs_of):
    sc, z = scene_z
    gpu = M.CudaPath(sc)
    gpu.preprocess()
    flat = gpu.flat
    print("\n### %s\n" % title)
    print("| model | eval p99 | eval p99.9 | eval max | pdf p99 | pdf p99.9 | pdf max | sampled lobe equal | sampled direction p99 |")
    print("|---|---|---|---|---|---|---|---|---|")
    for name in z:
        b = flat._bsdf_ids[id(z[name])]
        surf, wi, wo, u = inputs_of(name)
        ref_s = gold["bsdf/%s/sample" % name]
        got_s = gpu.ctx.bsdf_sample(b, surf, wi, u)
        same = (got_s[:, 8] == ref_s[:, 8]) & ((got_s[:, 3] > 0) == (ref_s[:, 3] > 0))
        ok = same & (ref_s[:, 3] > 0)
        ddir = np.abs(got_s[ok, 4:7] - ref_s[ok, 4:7]).max(axis=1) if ok.any() else np.zeros(1)
        wo2 = G.eval_directions(ref_s, wo)
        ev_all, ep_all = [], []
        for tag, meas in (("sa", 1), ("disc", 4)):
            v, pdf = gpu.ctx.bsdf_eval_pdf(b, surf, wi, wo2, meas)
            rv, rp = gold["bsdf/%s/eval_%s" % (name, tag)], gold["bsdf/%s/pdf_%s" % (name, tag)]
            nz = (v != 0).any(axis=1) & (rv != 0).any(axis=1)
            if nz.any():
                ev_all.append(util.rel_err(v[nz], rv[nz], 1e-7).max(axis=1))
                ep_all.append(util.rel_err(pdf[nz], rp[nz], 1e-7))
        ev = np.concatenate(ev_all) if ev_all else np.zeros(1)
        ep = np.concatenate(ep_all) if ep_all else np.zeros(1)
        pc = lambda a, q: float(np.percentile(a, q))
        print("| %s | %.1e | %.1e | %.1e | %.1e | %.1e | %.1e | %.4f | %.1e |"
              % (name, pc(ev, 99), pc(ev, 99.9), ev.max(), pc(ep, 99), pc(ep, 99.9), ep.max(), same.mean(), pc(ddir, 99)))
    gpu.close()


print("# BSDF parity of libmtscuda against the reference (768 seeded queries per model and measure; relative error)")
table("the nine models of SURVEY 8(a) and their wrappers (probes.npz)", G.zoo(), G.load("probes.npz"), G.bsdf_inputs)
table("twosided / thindielectric / mask / roughdielectric (twosided.npz)", G.twosided(), G.load("twosided.npz"), G.bsdf_inputs)
table("mixturebsdf / roughcoating / normalmap (nextrow.npz)", G.nextrow(), G.load("nextrow.npz"),
      lambda n: G.mip_bsdf_inputs(n) if n in G.NEXTROW_WITH_PARTIALS else G.bsdf_inputs(n))
