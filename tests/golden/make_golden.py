#!/usr/bin/env python
"""Generate the golden vectors of tests/golden/ by running the REFERENCE ITSELF (oracle/_ref: the unchanged
Mitsuba 0.6 sources of /root/reference compiled by oracle/refbuild/Makefile) on seeded inputs.

    python tests/golden/make_golden.py            # needs /root/reference + a built oracle/_ref; ~5 minutes on 8 cores

The inputs are regenerated from fixed seeds by tests/golden_io.py (shared with the tests); only the reference's
outputs are stored.  The files travel with the repository, so the tests need neither /root/reference nor oracle/_ref.
"""
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "refbuild"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import numpy as np  # noqa: E402
import mts_b200 as M  # noqa: E402,F401
import golden_io as G  # noqa: E402
import mtsref  # noqa: E402


def probes():
    sc, z = G.zoo()
    rs = mtsref.RefScene(sc)
    out = {}
    for name, p in z.items():
        surf, wi, wo, u = G.bsdf_inputs(name)
        rb = rs.bsdf(p)
        smp = rs.bsdf_sample(rb, surf, wi, u)
        wo2 = G.eval_directions(smp, wo)
        out["bsdf/%s/sample" % name] = smp
        for tag, meas in (("sa", 1), ("disc", 4)):
            v, pdf = rs.bsdf_eval_pdf(rb, surf, wi, wo2, meas)
            out["bsdf/%s/eval_%s" % (name, tag)] = v
            out["bsdf/%s/pdf_%s" % (name, tag)] = pdf
        out["bsdf/%s/type" % name] = np.array([rs.bsdf_type(rb)], dtype=np.int64)
    rays = G.trace_rays()
    h, ids = rs.trace(rays)
    out["trace/t"], out["trace/ids"] = h[:, 0].copy(), ids
    sr = G.shadow_rays(rays, h)
    out["trace/shadow"] = rs.trace(sr, shadow=True)[0][:, 0].copy()
    rec, ids2 = rs.hit_records(rays[:G.N_RECORDS])
    out["records/rec"], out["records/ids"] = rec, ids2
    ref, refn, u = G.emitter_inputs()
    out["emitter/sample"] = rs.sample_emitter_direct(ref, refn, u)
    out["env/eval"] = rs.eval_environment(G.env_dirs())
    out["sensor/rays"] = rs.sample_rays(G.sensor_positions())
    pos, val = G.splat_inputs()
    out["film/gaussian"] = rs.splat(pos, val)
    # box filter: the Cornell film
    rc = mtsref.RefScene(G.cornell_small(spp=1))
    pos, val = G.splat_inputs(w=G.CORNELL_RES, h=G.CORNELL_RES)
    out["film/box"] = rc.splat(pos, val)
    out["sensor/rays_cornell"] = rc.sample_rays(G.sensor_positions(w=G.CORNELL_RES, h=G.CORNELL_RES))
    return out


def images():
    out = {}
    for name, make in G.IMAGE_SCENES.items():
        sc = make()
        t0 = time.time()
        rs = mtsref.RefScene(sc)
        film = rs.render()
        st = rs.stats
        w, h, spp = G.scene_dims(sc)
        out["image/%s/rgb" % name] = M.develop(film).astype(np.float16)   # converged image; fp16 keeps the fixture small
        out["image/%s/stats" % name] = np.array([st.get("normal_rays", 0) / (w * h * spp), st.get("shadow_rays", 0) / (w * h * spp),
                                                 st.get("avg_path_length", 0), rs.render_s, mtsref.load().mtsref_thread_count()], dtype=np.float64)
        print("%-10s %dx%d %d spp: reference render %.1f s (total %.1f s), rays/sample %.3f + %.3f, avg path length %.2f"
              % (name, w, h, spp, rs.render_s, time.time() - t0, out["image/%s/stats" % name][0], out["image/%s/stats" % name][1],
                 out["image/%s/stats" % name][2]), flush=True)
    return out


def twosided():
    """twosided.npz: BSDF probes of the TwoSidedBRDF configurations + a converged render of the two-sided scene."""
    sc, z = G.twosided()
    rs = mtsref.RefScene(sc)
    out = {}
    for name, p in z.items():
        surf, wi, wo, u = G.bsdf_inputs(name)
        rb = rs.bsdf(p)
        smp = rs.bsdf_sample(rb, surf, wi, u)
        wo2 = G.eval_directions(smp, wo)
        out["bsdf/%s/sample" % name] = smp
        for tag, meas in (("sa", 1), ("disc", 4)):
            v, pdf = rs.bsdf_eval_pdf(rb, surf, wi, wo2, meas)
            out["bsdf/%s/eval_%s" % (name, tag)] = v
            out["bsdf/%s/pdf_%s" % (name, tag)] = pdf
        out["bsdf/%s/type" % name] = np.array([rs.bsdf_type(rb)], dtype=np.int64)
    for name, make in G.TWOSIDED_IMAGE_SCENES.items():
        scn = make()
        rsi = mtsref.RefScene(scn)
        film = rsi.render()
        st = rsi.stats
        w, h, spp = G.scene_dims(scn)
        out["image/%s/rgb" % name] = M.develop(film).astype(np.float16)
        out["image/%s/stats" % name] = np.array([st.get("normal_rays", 0) / (w * h * spp), st.get("shadow_rays", 0) / (w * h * spp),
                                                 st.get("avg_path_length", 0), rsi.render_s, mtsref.load().mtsref_thread_count()], dtype=np.float64)
        print("%-10s %dx%d %d spp: reference render %.1f s" % (name, w, h, spp, rsi.render_s), flush=True)
    return out


def nextrow():
    """nextrow.npz: BSDF probes of the mixturebsdf / roughcoating / normalmap configurations, the albedo / specular texture
    pair looked up with uv partials (shared EWA walk), and a converged render of a scene that uses them all."""
    sc, z = G.nextrow()
    rs = mtsref.RefScene(sc)
    out = {}
    for name, p in z.items():
        surf, wi, wo, u = G.mip_bsdf_inputs(name) if name in G.NEXTROW_WITH_PARTIALS else G.bsdf_inputs(name)
        rb = rs.bsdf(p)
        smp = rs.bsdf_sample(rb, surf, wi, u)
        wo2 = G.eval_directions(smp, wo)
        out["bsdf/%s/sample" % name] = smp
        for tag, meas in (("sa", 1), ("disc", 4)):
            v, pdf = rs.bsdf_eval_pdf(rb, surf, wi, wo2, meas)
            out["bsdf/%s/eval_%s" % (name, tag)] = v
            out["bsdf/%s/pdf_%s" % (name, tag)] = pdf
        out["bsdf/%s/type" % name] = np.array([rs.bsdf_type(rb)], dtype=np.int64)
    for name, make in G.NEXTROW_IMAGE_SCENES.items():
        scn = make()
        rsi = mtsref.RefScene(scn)
        film = rsi.render()
        st = rsi.stats
        w, h, spp = G.scene_dims(scn)
        out["image/%s/rgb" % name] = M.develop(film).astype(np.float16)
        out["image/%s/stats" % name] = np.array([st.get("normal_rays", 0) / (w * h * spp), st.get("shadow_rays", 0) / (w * h * spp),
                                                 st.get("avg_path_length", 0), rsi.render_s, mtsref.load().mtsref_thread_count()], dtype=np.float64)
        print("%-10s %dx%d %d spp: reference render %.1f s" % (name, w, h, spp, rsi.render_s), flush=True)
    return out


def fullsize():
    """fullsize.npz: closest hits of the reference's kd-tree on the full 9,999,392-triangle cfg3 terrain (fallback fixture of
    tests/test_gpu_fullsize.py, which asks oracle/_ref directly where it is present)."""
    rs = mtsref.RefScene(G.fullsize_scene())
    rays = G.fullsize_rays(G.N_FULLSIZE_GOLDEN, rs.sample_rays)
    h, ids = rs.trace(rays)
    prim = np.where(np.isfinite(h[:, 0]), ids[:, 1], -1).astype(np.int32)
    print("fullsize: %d rays, %.1f %% hit, kd-tree build %.1f s" % (len(rays), 100 * np.isfinite(h[:, 0]).mean(), rs.kd_build_s), flush=True)
    return {"trace/t": h[:, 0].copy(), "trace/shape": ids[:, 0].astype(np.int8), "trace/prim": prim}


def boundary():
    """boundary.npz: crop window, sphere transform composition and the alpha channel, rendered by the reference."""
    out = {}
    for name, make in G.BOUNDARY_IMAGE_SCENES.items():
        scn = make()
        rsi = mtsref.RefScene(scn)
        film, alpha = rsi.render(alpha=True)
        out["image/%s/rgb" % name] = M.develop(film).astype(np.float16)
        w = film[..., 3]
        out["image/%s/alpha" % name] = np.divide(alpha, w, out=np.zeros_like(alpha), where=w > 0).astype(np.float16)
        st = rsi.stats
        h, wd = film.shape[:2]
        spp = G.scene_dims(scn)[2]
        out["image/%s/stats" % name] = np.array([st.get("normal_rays", 0) / (wd * h * spp), st.get("shadow_rays", 0) / (wd * h * spp),
                                                 st.get("avg_path_length", 0), rsi.render_s, mtsref.load().mtsref_thread_count()], dtype=np.float64)
        print("%-12s %dx%d %d spp: reference render %.1f s" % (name, wd, h, spp, rsi.render_s), flush=True)
    return out


def mip():
    """mip.npz: TMIPMap lookups with uv partials, Intersection::computePartials of camera rays, the environment map seen
    through ray differentials, BSDF queries on intersections that carry partials, and a converged render of a scene
    whose bitmap textures use the default filterType (ewa)."""
    out = {}
    scn, tex, mats = G.mip()
    rs = mtsref.RefScene(scn)
    for name, t in tex.items():
        uv, duv = G.texture_inputs(name)
        out["tex/%s/filtered" % name] = rs.texture_eval(t, uv, duv)
        out["tex/%s/plain" % name] = rs.texture_eval(t, uv)
    pos = G.partial_positions()
    for cnt in G.MIP_SAMPLE_COUNTS:
        out["partials/%d" % cnt] = rs.hit_partials(pos, cnt)
        out["envcam/%d" % cnt] = rs.eval_environment_camera(pos, cnt)
    for name, p in mats.items():
        surf, wi, wo, u = G.mip_bsdf_inputs(name)
        rb = rs.bsdf(p)
        smp = rs.bsdf_sample(rb, surf, wi, u)
        wo2 = G.eval_directions(smp, wo)
        out["bsdf/%s/sample" % name] = smp
        for tag, meas in (("sa", 1), ("disc", 4)):
            v, pdf = rs.bsdf_eval_pdf(rb, surf, wi, wo2, meas)
            out["bsdf/%s/eval_%s" % (name, tag)] = v
            out["bsdf/%s/pdf_%s" % (name, tag)] = pdf
    film = rs.render()
    st = rs.stats
    w, h, spp = G.scene_dims(scn)
    out["image/mip/rgb"] = M.develop(film).astype(np.float16)
    out["image/mip/stats"] = np.array([st.get("normal_rays", 0) / (w * h * spp), st.get("shadow_rays", 0) / (w * h * spp),
                                       st.get("avg_path_length", 0), rs.render_s, mtsref.load().mtsref_thread_count()], dtype=np.float64)
    print("mip        %dx%d %d spp: reference render %.1f s" % (w, h, spp, rs.render_s), flush=True)
    return out


if __name__ == "__main__":
    if not mtsref.available():
        raise SystemExit("oracle/_ref is not built: make -C oracle/refbuild (needs /root/reference)")
    if "--only" in sys.argv:   # re-render single entries of images.npz (a scene generator changed), keep the rest
        names = sys.argv[sys.argv.index("--only") + 1].split(",")
        old = dict(np.load(os.path.join(HERE, "images.npz")))
        keep = dict(G.IMAGE_SCENES)
        for k in list(G.IMAGE_SCENES):
            if k not in names:
                del G.IMAGE_SCENES[k]
        old.update(images())
        G.IMAGE_SCENES.update(keep)
        np.savez_compressed(os.path.join(HERE, "images.npz"), **old)
        print("images.npz: re-rendered %s" % names)
        raise SystemExit(0)
    if "--mip" in sys.argv:
        t = mip()
        np.savez_compressed(os.path.join(HERE, "mip.npz"), **t)
        print("mip.npz: %d arrays" % len(t))
        raise SystemExit(0)
    if "--boundary" in sys.argv:
        t = boundary()
        np.savez_compressed(os.path.join(HERE, "boundary.npz"), **t)
        print("boundary.npz: %d arrays" % len(t))
        raise SystemExit(0)
    if "--fullsize" in sys.argv:
        t = fullsize()
        np.savez_compressed(os.path.join(HERE, "fullsize.npz"), **t)
        print("fullsize.npz: %d arrays" % len(t))
        raise SystemExit(0)
    if "--nextrow" in sys.argv:
        t = nextrow()
        np.savez_compressed(os.path.join(HERE, "nextrow.npz"), **t)
        print("nextrow.npz: %d arrays" % len(t))
        raise SystemExit(0)
    if "--twosided" in sys.argv:
        t = twosided()
        np.savez_compressed(os.path.join(HERE, "twosided.npz"), **t)
        print("twosided.npz: %d arrays" % len(t))
        raise SystemExit(0)
    p = probes()
    np.savez_compressed(os.path.join(HERE, "probes.npz"), **p)
    print("probes.npz: %d arrays" % len(p))
    if "--no-images" not in sys.argv:
        im = images()
        np.savez_compressed(os.path.join(HERE, "images.npz"), **im)
        print("images.npz: %d arrays" % len(im))
