"""Checks of one backend (the oracle restatement or the CUDA library, both behind the C ABI of include/mtscuda.h)
against the golden vectors the REFERENCE build produced (tests/golden/probes.npz, images.npz).

Tolerances (written here once, used by the CPU and GPU tests alike):
  * BSDF / emitter eval and pdf on identical inputs: 1e-5 relative (north_star) for 99% of the samples of every model,
    3e-5 for 99.9%; the remaining tail is fp32 conditioning (exp(-tan^2/alpha^2), 1 - cos^2 at grazing angles) and the
    reference's own -funsafe-math-optimizations build.  Two models carry a documented looser bar: `plastic_nl` (its
    Gauss-Lobatto-integrated constant fdrInt is itself only defined to the integrator's 1e-5 tolerance) and
    `bump_coating_rc` (three nested frames).
  * intersection: hit/miss, shape id and primitive id identical except for ties / grazing rays.
  * images: relMSE against the converged reference render.
"""
import numpy as np

import golden_io as G
import mts_b200 as M
import util

F = np.float32
REL = 1e-5
LOOSE = {"plastic_nl": 5.0, "bump_coating_rc": 3.0, "coating_bump_blend": 2.0,
         # two EWA lookups per query: a tap whose running quadratic sits on an integer picks the neighbouring weight-table entry
         # (see check_texture), which moves the 99.9th percentile of 768 queries
         "rp_pair_ewa": 3.0}
# models whose sampling runs a numerical inversion (Beckmann visible normals) or an ill-conditioned closed form (GGX)
NEWTON = ("rc_beckmann", "rc_beckmann_aniso", "rp_beckmann", "bump_coating_rc", "blend_tex_rp_diel")


def _pct(a, q):
    return float(np.percentile(a, q)) if len(a) else 0.0


def check_bsdf(ctx, flat, z, name, gold, inputs=None):
    b = flat._bsdf_ids[id(z[name])]
    surf, wi, wo, u = (inputs or G.bsdf_inputs)(name)
    ref_s = gold["bsdf/%s/sample" % name]
    got_s = ctx.bsdf_sample(b, surf, wi, u)
    # sampled lobe type and validity
    same = (got_s[:, 8] == ref_s[:, 8]) & ((got_s[:, 3] > 0) == (ref_s[:, 3] > 0))
    assert same.mean() >= 0.998, (name, same.mean())
    ok = same & (ref_s[:, 3] > 0)
    if ok.any():
        ddir = np.abs(got_s[ok, 4:7] - ref_s[ok, 4:7]).max(axis=1)
        newton = name in NEWTON
        assert _pct(ddir, 95) < (3e-4 if newton else 3e-5), (name, _pct(ddir, 95))
        assert _pct(ddir, 99) < (2e-3 if newton else 3e-4), (name, _pct(ddir, 99))
        e_w = util.rel_err(got_s[ok, 0:3], ref_s[ok, 0:3], 1e-4).max(axis=1)
        e_p = util.rel_err(got_s[ok, 3], ref_s[ok, 3], 1e-4)
        tol = 1e-2 if newton else 2e-3     # conditioned by the direction difference above (sharp lobes amplify it)
        assert _pct(e_w, 95) < tol and _pct(e_p, 95) < tol, (name, _pct(e_w, 95), _pct(e_p, 95))
        assert (got_s[ok, 7] == ref_s[ok, 7]).mean() > 0.999 or np.abs(got_s[ok, 7] - ref_s[ok, 7]).max() < 1e-6   # eta
    # eval / pdf on IDENTICAL directions
    wo2 = G.eval_directions(ref_s, wo)
    k = LOOSE.get(name, 1.0)
    worst = 0.0
    for tag, meas in (("sa", 1), ("disc", 4)):
        v, pdf = ctx.bsdf_eval_pdf(b, surf, wi, wo2, meas)
        rv, rp = gold["bsdf/%s/eval_%s" % (name, tag)], gold["bsdf/%s/pdf_%s" % (name, tag)]
        nzg, nzr = (v != 0).any(axis=1), (rv != 0).any(axis=1)
        assert (nzg == nzr).mean() >= 0.998, (name, tag, (nzg == nzr).mean())
        nz = nzg & nzr
        if not nz.any():
            continue
        ev = util.rel_err(v[nz], rv[nz], 1e-7).max(axis=1)
        ep = util.rel_err(pdf[nz], rp[nz], 1e-7)
        assert _pct(ev, 99) < k * REL and _pct(ep, 99) < k * REL, (name, tag, _pct(ev, 99), _pct(ep, 99))
        assert _pct(ev, 99.9) < 10 * k * REL and _pct(ep, 99.9) < 10 * k * REL, (name, tag, _pct(ev, 99.9), _pct(ep, 99.9))
        assert ev.max() < 5e-3 and ep.max() < 5e-3, (name, tag, ev.max(), ep.max())
        worst = max(worst, _pct(ev, 99), _pct(ep, 99))
    return worst


def compare_trace(rays, got, ref, perturb_trace, max_mismatch=1e-3, jitter=1e-4, dump=None, perturb_got=None):
    """Hit/miss + shape id + primitive id of `got` = (hits, ids) against `ref` = (t, ids): identical except for ties (both hit
    at the same distance) and grazing rays (the checker `perturb_trace(rays) -> ids` changes its own answer under a 1e-4
    perturbation of the direction -- `jitter`, to be chosen well below the angular size of a primitive; with `perturb_got` also
    when the tested implementation's answer turns into the checker's under the same perturbation: an edge-on triangle that
    one of the two finds and the other does not).  Returns (mismatches,
    excused as ties, excused as grazing)."""
    h, ids = got
    rt, rids = ref
    hit_g, hit_r = np.isfinite(h[:, 0]), np.isfinite(rt)
    same = (hit_g == hit_r) & (ids == rids).all(axis=1)
    bad = np.nonzero(~same)[0]
    assert len(bad) / len(rays) < max_mismatch, "hit/prim-id mismatch on %d of %d rays" % (len(bad), len(rays))
    n_tie = n_graze = 0
    if len(bad):
        tie = hit_g[bad] & hit_r[bad] & (np.abs(h[bad, 0] - rt[bad]) <= 1e-4 * np.maximum(1, rt[bad]))
        n_tie = int(tie.sum())
        rest = bad[~tie]
        if len(rest):
            assert perturb_trace is not None, ("non-tie intersection mismatch", rays[rest][:3], ids[rest][:3], rids[rest][:3])
            excused = np.zeros(len(rest), dtype=bool)
            jr = np.random.RandomState(11)
            for _ in range(16):
                jit = rays[rest].copy()
                jit[:, 4:7] += jr.normal(scale=jitter, size=(len(rest), 3)).astype(F)
                jit[:, 4:7] /= np.linalg.norm(jit[:, 4:7], axis=1, keepdims=True)
                excused |= (perturb_trace(jit) == ids[rest]).all(axis=1)
                if perturb_got is not None:
                    excused |= (perturb_got(jit) == rids[rest]).all(axis=1)
            if dump is not None and not excused.all():
                np.savez(dump, rays=rays[rest[~excused]], got_ids=ids[rest[~excused]], ref_ids=rids[rest[~excused]], got_t=h[rest[~excused], 0],
                         ref_t=rt[rest[~excused]], n_bad=len(bad), n_tie=n_tie, n_rest=len(rest))
            assert excused.all(), ("non-grazing intersection mismatch", rays[rest[~excused]][:3], ids[rest[~excused]][:3], rids[rest[~excused]][:3])
            n_graze = int(excused.sum())
    ok = same & hit_g
    assert np.all(np.abs(h[ok, 0] - rt[ok]) <= 1e-4 * np.maximum(1, rt[ok]))
    return len(bad), n_tie, n_graze


def check_trace(ctx, gold, oracle_ctx=None):
    rays = G.trace_rays()
    h, ids = ctx.trace(rays)
    rt, rids = gold["trace/t"], gold["trace/ids"]
    hit_g, hit_r = np.isfinite(h[:, 0]), np.isfinite(rt)
    same = (hit_g == hit_r) & (ids == rids).all(axis=1)
    bad = np.nonzero(~same)[0]
    assert len(bad) / len(rays) < 1e-3, "hit/prim-id mismatch on %d of %d rays" % (len(bad), len(rays))
    if len(bad):
        # excused: ties (both hit at the same distance) or grazing rays (the checker's own answer flips under a 1e-4
        # perturbation of the direction)
        tie = hit_g[bad] & hit_r[bad] & (np.abs(h[bad, 0] - rt[bad]) <= 1e-4 * np.maximum(1, rt[bad]))
        rest = bad[~tie]
        if len(rest):
            assert oracle_ctx is not None, ("non-tie intersection mismatch", rays[rest][:3], ids[rest][:3], rids[rest][:3])
            excused = np.zeros(len(rest), dtype=bool)
            jr = np.random.RandomState(11)
            for _ in range(16):
                jit = rays[rest].copy()
                jit[:, 4:7] += jr.normal(scale=1e-4, size=(len(rest), 3)).astype(F)
                jit[:, 4:7] /= np.linalg.norm(jit[:, 4:7], axis=1, keepdims=True)
                _, ij = oracle_ctx.trace(jit)
                excused |= (ij == ids[rest]).all(axis=1)
            assert excused.all(), ("non-grazing intersection mismatch", rays[rest[~excused]][:3])
    ok = same & hit_g
    assert np.all(np.abs(h[ok, 0] - rt[ok]) <= 1e-4 * np.maximum(1, rt[ok]))
    s, _ = ctx.trace(G.shadow_rays(rays, np.stack([rt] * 4, axis=1)), shadow=True)
    assert np.mean(s[:, 0] != gold["trace/shadow"]) < 1e-3
    return len(bad)


def check_records(ctx, gold):
    rays = G.trace_rays()[:G.N_RECORDS]
    rec, ids = ctx.hit_records(rays)
    rrec, rids = gold["records/rec"], gold["records/ids"]
    ok = (ids == rids).all(axis=1) & np.isfinite(rrec[:, 3]) & np.isfinite(rec[:, 3])
    assert ok.mean() > 0.55 and (ids == rids).all(axis=1).mean() > 0.999
    err = np.abs(rec[ok] - rrec[ok]) / np.maximum(1.0, np.abs(rrec[ok]))
    wrap = np.abs(rec[ok][:, 7] - rrec[ok][:, 7]) > 0.5       # sphere uv wraps at phi = 0
    assert wrap.mean() < 1e-3
    assert _pct(err[~wrap], 99.9) < 2e-4, _pct(err[~wrap], 99.9)
    assert err[~wrap].max() < 5e-2   # grazing hits amplify fp32 noise in dpdu/dpdv (sphere poles, tiny triangles)
    return float(_pct(err[~wrap], 99.9))


def check_emitters(ctx, gold):
    ref, refn, u = G.emitter_inputs()
    out, gref = ctx.sample_emitter_direct(ref, refn, u), gold["emitter/sample"]
    assert (out[:, 15] == gref[:, 15]).mean() > 0.999, (out[:, 15] == gref[:, 15]).mean()
    ok = (out[:, 15] == gref[:, 15]) & (gref[:, 15] >= 0)
    # value(3) pdf d(3) dist p(3) measure n(3): radiance / pdf / direction are the estimator inputs
    e = util.rel_err(out[ok][:, :8], gref[ok][:, :8], 1e-3)
    assert _pct(e.max(axis=1), 99) < 1e-4, _pct(e.max(axis=1), 99)      # cone / grazing-area samples are ill-conditioned
    assert _pct(e.max(axis=1), 90) < REL, _pct(e.max(axis=1), 90)
    assert (out[ok][:, 11] == gref[ok][:, 11]).all()
    e2 = util.rel_err(out[ok][:, 8:15], gref[ok][:, 8:15], 1e-3)
    assert _pct(e2.max(axis=1), 99) < 2e-2
    d = G.env_dirs()
    ev, gev = ctx.eval_environment(d), gold["env/eval"]
    e = util.rel_err(ev, gev, 1e-6)
    # the synthetic map holds 5e3-peaked 1.5-degree "suns": one ulp of atan2/acos in the uv moves the bilinear lookup
    # by up to ~5e-5 relative on their flanks
    assert _pct(e.max(axis=1), 99) < REL and _pct(e.max(axis=1), 99.9) < 5e-5 and e.max() < 5e-4, (e.max(), _pct(e.max(axis=1), 99.9))
    return float(e.max())


def check_sensor_film(ctx, gold, which="zoo"):
    if which == "zoo":
        sp = G.sensor_positions()
        a, b = ctx.sample_rays(sp), gold["sensor/rays"]
        pos, val = G.splat_inputs()
        film, gfilm = ctx.splat(pos, val), gold["film/gaussian"]
    else:
        sp = G.sensor_positions(w=G.CORNELL_RES, h=G.CORNELL_RES)
        a, b = ctx.sample_rays(sp), gold["sensor/rays_cornell"]
        pos, val = G.splat_inputs(w=G.CORNELL_RES, h=G.CORNELL_RES)
        film, gfilm = ctx.splat(pos, val), gold["film/box"]
    # origin, direction, mint/maxt exactly to fp32 noise; the differential directions (cols 8..14) to 3e-5
    assert util.rel_err(a[:, :8], b[:, :8], 1e-3).max() < 3e-6
    assert util.rel_err(a[:, 8:], b[:, 8:], 1e-3).max() < 3e-5
    assert np.abs(film - gfilm).max() < 1e-5 * max(1.0, np.abs(gfilm).max()), np.abs(film - gfilm).max()
    assert (film[..., 3] > 0).sum() == (gfilm[..., 3] > 0).sum()
    return float(np.abs(film - gfilm).max())


def check_image(path_tracer, name, gold_images, spp, bar, lazy_shadow=False):
    """relMSE of a render against the converged reference image + the reference's own ray statistics."""
    ref = gold_images["image/%s/rgb" % name].astype(np.float32)
    st_ref = gold_images["image/%s/stats" % name]
    img = M.develop(path_tracer.render(spp=spp, seed=7))
    assert np.isfinite(img).all()
    err = M.rel_mse(img, ref)
    st = path_tracer.stats
    rays = (st.normal_rays / st.camera_samples, st.shadow_rays / st.camera_samples, st.path_length_sum / st.camera_samples)
    assert abs(img.mean() / ref.mean() - 1) < 0.02, (name, img.mean(), ref.mean())
    assert err < bar, (name, err, bar)
    # same number of rays per camera sample / average path length as the reference's own counters
    assert abs(rays[0] / st_ref[0] - 1) < 0.01, (name, rays, st_ref[:3])
    if lazy_shadow:
        # the CUDA wavefront queues a shadow ray only after the BSDF value of the light sample turned out non-zero;
        # the reference traces it inside Scene::sampleEmitterDirect before evaluating the BSDF (scene.cpp:839-843).
        # Same estimator, fewer rays.
        assert 0.5 * st_ref[1] < rays[1] <= 1.01 * st_ref[1], (name, rays, st_ref[:3])
    else:
        assert abs(rays[1] / st_ref[1] - 1) < 0.01, (name, rays, st_ref[:3])
    assert abs(rays[2] - st_ref[2]) < 0.02, (name, rays, st_ref[:3])
    return err


def check_texture(ctx, flat, tex, name, gold):
    """TMIPMap::eval / evalBilinear / evalBox through Texture2D::eval(its, filter=True) against the reference."""
    tid = flat._tex_ids[id(tex[name])][0]
    uv, duv = G.texture_inputs(name)
    plain = np.abs(ctx.texture_eval(tid, uv) - gold["tex/%s/plain" % name]).max(axis=1)
    # texel coordinates reach 2.5 * uscale * width: the bilinear weights carry that many ulps
    assert _pct(plain, 99) < 2e-5 and plain.max() < 2e-4, (name, _pct(plain, 99), plain.max())
    e = np.abs(ctx.texture_eval(tid, uv, duv) - gold["tex/%s/filtered" % name]).max(axis=1)
    assert _pct(e, 95) < 1e-5, (name, _pct(e, 95))
    assert _pct(e, 99) < 1e-4, (name, _pct(e, 99))
    # the EWA weight table is indexed by (int) q of a running quadratic: one ulp in q across an integer changes one tap's
    # weight, and one ulp in the level selection swaps the pyramid level of a lookup that sits on the boundary
    assert (e > 2e-3).mean() <= 2e-3 and e.max() < 0.1, (name, (e > 2e-3).mean(), e.max())
    return _pct(e, 99)


def check_partials(ctx, gold):
    pos = G.partial_positions()
    worst = 0.0
    for cnt in G.MIP_SAMPLE_COUNTS:
        ref, got = gold["partials/%d" % cnt], ctx.hit_partials(pos, cnt)
        assert (ref[:, 5] == got[:, 5]).mean() > 0.999
        hit = (ref[:, 5] > 0) & (got[:, 5] > 0)
        assert np.abs(ref[hit, 4] - got[hit, 4]).max() < 1e-3       # t
        assert np.abs(ref[hit, 6:8] - got[hit, 6:8]).max() < 1e-3   # uv
        # the partials are differences of nearly equal points (offset-ray hit minus hit): relative to the largest component
        rel = (np.abs(ref[hit, :4] - got[hit, :4]) / (np.abs(ref[hit, :4]).max(axis=1, keepdims=True) + 1e-9)).max(axis=1)
        # ... and the offset shrinks with 1/sqrt(sample count) while the ulp of the hit point does not
        tol = 3e-4 * max(1.0, cnt ** 0.5 / 4)
        assert _pct(rel, 99) < tol and rel.max() < 20 * tol, (cnt, _pct(rel, 99), rel.max())
        worst = max(worst, _pct(rel, 99))
        env_ref, env = gold["envcam/%d" % cnt], ctx.eval_environment_camera(pos, cnt)
        e = util.rel_err(env, env_ref, 1e-3).max(axis=1)
        assert _pct(e, 99) < 1e-3 and (e > 0.05).mean() < 2e-3, (cnt, _pct(e, 99), e.max())
    return worst


def check_boundary_image(path_tracer, name, gold_boundary, spp, bar):
    """Crop window / sphere transform composition / alpha channel against the reference's own `path` render."""
    ref = gold_boundary["image/%s/rgb" % name].astype(np.float32)
    ref_alpha = gold_boundary["image/%s/alpha" % name].astype(np.float32)
    film = path_tracer.render(spp=spp, seed=5)
    img = M.develop(film)
    assert img.shape == ref.shape, (img.shape, ref.shape)
    err = M.rel_mse(img, ref)
    assert abs(img.mean() / ref.mean() - 1) < 0.03, (name, img.mean(), ref.mean())
    assert err < bar, (name, err, bar)
    if path_tracer.flat.film.has_alpha:
        w = film[..., 3]
        alpha = np.divide(path_tracer.ctx.read_film_alpha(), w, out=np.zeros_like(w), where=w > 0)
        # pixel-wise coverage: identical where the pixel footprint is fully inside / outside the geometry
        solid = (ref_alpha == 0) | (ref_alpha == 1)
        assert np.abs(alpha[solid] - ref_alpha[solid]).max() < 1e-3
        assert np.abs(alpha - ref_alpha).mean() < 5e-3 and np.abs(alpha - ref_alpha).max() < 0.15
    return err
