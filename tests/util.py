"""Shared helpers for the parity tests: a 'material zoo' scene that instantiates every in-scope BSDF
(incl. nested wrappers, both microfacet distributions, textured parameters) and every emitter kind."""
import numpy as np

import mts_b200 as M
from oracle_lib import OraclePath
from mitsuba0_6_b200 import hostmath as hm, scenes
from mitsuba0_6_b200.scene import Plugin, Point, Spectrum

F = np.float32


def _tex(seed, channels=3, size=64, **kw):
    rng = np.random.RandomState(seed)
    t = Plugin("texture", "bitmap", filterType="bilinear", gamma=1.0, **kw)
    t.data["bitmap"] = rng.uniform(0.05, 0.95, (size, size, channels)).astype(F)
    return t


def zoo_bsdfs():
    CU, AU, AL = scenes.CU, scenes.AU, scenes.AL
    z = {}
    z["diffuse"] = Plugin("bsdf", "diffuse", reflectance=Spectrum(0.2, 0.5, 0.7))
    d = Plugin("bsdf", "diffuse")
    d.add(_tex(1, uvscale=3.0), "reflectance")
    z["diffuse_tex"] = d
    z["conductor"] = Plugin("bsdf", "conductor", **CU)
    z["conductor_spec"] = Plugin("bsdf", "conductor", specularReflectance=Spectrum(0.9, 0.8, 0.7), extEta=1.33, **AU)
    z["rc_beckmann"] = Plugin("bsdf", "roughconductor", distribution="beckmann", alpha=0.3, **CU)
    z["rc_ggx"] = Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.1, **AU)
    z["rc_beckmann_aniso"] = Plugin("bsdf", "roughconductor", distribution="beckmann", alphaU=0.1, alphaV=0.3, **AL)
    z["rc_ggx_aniso"] = Plugin("bsdf", "roughconductor", distribution="ggx", alphaU=0.2, alphaV=0.3, **CU)
    z["rc_ggx_all"] = Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.25, sampleVisible=False, **CU)
    z["rc_beckmann_all"] = Plugin("bsdf", "roughconductor", distribution="beckmann", alphaU=0.15, alphaV=0.4, sampleVisible=False, **AL)
    rct = Plugin("bsdf", "roughconductor", distribution="ggx", **AU)
    at = Plugin("texture", "scale", scale=0.5)
    at.add(_tex(2, channels=1))
    rct.add(at, "alpha")
    z["rc_ggx_texalpha"] = rct
    z["plastic"] = Plugin("bsdf", "plastic", intIOR=1.49, diffuseReflectance=Spectrum(0.5, 0.3, 0.2))
    z["plastic_nl"] = Plugin("bsdf", "plastic", intIOR=1.9, extIOR="water", nonlinear=True,
                             diffuseReflectance=Spectrum(0.8, 0.7, 0.1), specularReflectance=Spectrum(0.6))
    z["rp_beckmann"] = Plugin("bsdf", "roughplastic", distribution="beckmann", alpha=0.7, intIOR=1.49)
    z["rp_ggx"] = Plugin("bsdf", "roughplastic", distribution="ggx", alpha=0.15, intIOR=1.6, nonlinear=True,
                         diffuseReflectance=Spectrum(0.2, 0.4, 0.8))
    rpt = Plugin("bsdf", "roughplastic", distribution="ggx", intIOR=1.5)
    at2 = Plugin("texture", "scale", scale=0.4)
    at2.add(_tex(3, channels=1))
    rpt.add(at2, "alpha")
    rpt.add(_tex(4), "diffuseReflectance")
    z["rp_ggx_tex"] = rpt
    z["dielectric"] = Plugin("bsdf", "dielectric", intIOR="bk7")
    z["dielectric_water_air"] = Plugin("bsdf", "dielectric", intIOR="water", extIOR="air",
                                       specularTransmittance=Spectrum(0.9, 0.95, 1.0))
    c = Plugin("bsdf", "coating", intIOR=1.5, thickness=1.0, sigmaA=Spectrum(0.1, 0.2, 0.3))
    c.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.6, 0.5, 0.3)))
    z["coating_diffuse"] = c
    c2 = Plugin("bsdf", "coating", intIOR=1.7, thickness=2.0)
    c2.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.2, **CU))
    z["coating_rc"] = c2
    b = Plugin("bsdf", "blendbsdf", weight=0.3)
    b.add(Plugin("bsdf", "conductor", **AU))
    b.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5)))
    z["blend_cond_diff"] = b
    b2 = Plugin("bsdf", "blendbsdf")
    b2.add(_tex(5, channels=1), "weight")
    b2.add(Plugin("bsdf", "roughplastic", distribution="beckmann", alpha=0.2, intIOR=1.49))
    b2.add(Plugin("bsdf", "dielectric", intIOR=1.5))
    z["blend_tex_rp_diel"] = b2
    bm = Plugin("bsdf", "bumpmap")
    bm.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.7, 0.6, 0.5)))
    bt = Plugin("texture", "scale", scale=0.05)
    bt.add(_tex(6, channels=1, uvscale=2.0))
    bm.add(bt)
    z["bump_diffuse"] = bm
    bm2 = Plugin("bsdf", "bumpmap")
    cc = Plugin("bsdf", "coating", intIOR=1.5, sigmaA=Spectrum(0.05, 0.1, 0.2))
    cc.add(Plugin("bsdf", "roughconductor", distribution="beckmann", alpha=0.2, **CU))
    bm2.add(cc)
    bt2 = Plugin("texture", "scale", scale=0.02)
    bt2.add(_tex(7, channels=1))
    bm2.add(bt2)
    z["bump_coating_rc"] = bm2
    cb = Plugin("bsdf", "coating", intIOR=1.5)   # coating over a bump-mapped blend: 3 wrapper levels
    inner_b = Plugin("bsdf", "blendbsdf", weight=0.6)
    inner_b.add(Plugin("bsdf", "plastic", intIOR=1.5))
    inner_b.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.3, **AL))
    ib = Plugin("bsdf", "bumpmap")
    ib.add(inner_b)
    bt3 = Plugin("texture", "scale", scale=0.03)
    bt3.add(_tex(8, channels=1))
    ib.add(bt3)
    cb.add(ib)
    z["coating_bump_blend"] = cb
    return z


def zoo_scene(width=64, height=48, spp=4, emitters=("area_rect", "area_sphere", "area_mesh", "point", "envmap")):
    """Every BSDF on its own sphere + one of each emitter kind + a textured cube and a mesh."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=6))
    s = Plugin("sensor", "perspective", toWorld=hm.look_at([0, 6, 14], [0, 1, 0], [0, 1, 0]), fov=50.0)
    s.add(Plugin("sampler", "independent", sampleCount=spp))
    film = Plugin("film", "hdrfilm", width=width, height=height)
    film.add(Plugin("rfilter", "gaussian"))
    s.add(film)
    sc.add(s)
    z = zoo_bsdfs()
    names = list(z)
    for i, nm in enumerate(names):
        sp = Plugin("shape", "sphere", center=Point((i % 7 - 3) * 2.2, 1.0 + 2.2 * (i // 7), -(i // 7)), radius=1.0)
        sp.add(z[nm])
        sc.add(sp)
    floor = Plugin("shape", "rectangle", toWorld=hm.rotate([1, 0, 0], -90) @ hm.scale([12, 12, 1]))
    floor.add(z["diffuse_tex"])
    sc.add(floor)
    cube = Plugin("shape", "cube", toWorld=hm.translate([5, 0.8, 4]) @ hm.rotate([0, 1, 0], 25) @ hm.scale([0.8, 0.8, 0.8]))
    cube.add(z["bump_diffuse"])
    sc.add(cube)
    pos, idx, nrm, uv = scenes.torus_knot(60, 16)
    knot = Plugin("shape", "mesh", toWorld=hm.translate([-5, 1.5, 4]))
    knot.data.update(positions=pos, indices=idx, normals=nrm, texcoords=uv)
    knot.add(z["rc_ggx"])
    sc.add(knot)
    if "area_rect" in emitters:
        r = Plugin("shape", "rectangle", toWorld=hm.translate([0, 9, 2]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([1.5, 1.0, 1]))
        r.add(Plugin("emitter", "area", radiance=Spectrum(20, 18, 15)))
        sc.add(r)
    if "area_sphere" in emitters:
        sp = Plugin("shape", "sphere", center=Point(-7, 5, 3), radius=0.5)
        sp.add(Plugin("emitter", "area", radiance=Spectrum(30, 10, 10), samplingWeight=2.0))
        sc.add(sp)
    if "area_mesh" in emitters:
        cb = Plugin("shape", "cube", toWorld=hm.translate([7, 5, 3]) @ hm.scale([0.4, 0.3, 0.5]))
        cb.add(Plugin("emitter", "area", radiance=Spectrum(5, 15, 25)))
        sc.add(cb)
    if "point" in emitters:
        sc.add(Plugin("emitter", "point", position=Point(0, 4, 8), intensity=Spectrum(40, 40, 60)))
    if "envmap" in emitters:
        env = Plugin("emitter", "envmap", scale=0.7, toWorld=hm.rotate([1, 0, 0], 40.0))
        env.data["bitmap"] = scenes.synthetic_envmap(128, 64)
        sc.add(env)
    return sc, z


def make_pair(scene):
    """(cuda, oracle) CudaPath objects of the same flattened scene."""
    flat = M.flatten(scene)
    gpu, cpu = M.CudaPath(flat), OraclePath(flat)
    gpu.preprocess()
    cpu.preprocess()
    return gpu, cpu, flat


def random_dirs(rng, n, hemisphere=None):
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    if hemisphere == "+":
        d[:, 2] = np.abs(d[:, 2])
    return d.astype(F)


def rel_err(a, b, floor=1e-6):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)


# ---- twosided (src/bsdfs/twosided.cpp) ------------------------------------------------
def twosided_bsdfs():
    """TwoSidedBRDF wrappers: one nested BRDF (used for both sides), two different ones, a wrapper below the
    adapter, and the adapter below a wrapper."""
    CU, AU = scenes.CU, scenes.AU
    z = {}
    t = Plugin("bsdf", "twosided")
    t.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.7, 0.4, 0.2)))
    z["ts_diffuse"] = t
    t2 = Plugin("bsdf", "twosided")
    t2.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.25, **CU))
    t2.add(Plugin("bsdf", "plastic", intIOR=1.5, diffuseReflectance=Spectrum(0.2, 0.5, 0.3)))
    z["ts_rc_plastic"] = t2
    t3 = Plugin("bsdf", "twosided")
    bm = Plugin("bsdf", "bumpmap")
    bm.add(Plugin("bsdf", "roughplastic", distribution="beckmann", alpha=0.15, intIOR=1.49,
                  diffuseReflectance=Spectrum(0.5, 0.3, 0.6)))
    bt = Plugin("texture", "scale", scale=0.04)
    bt.add(_tex(11, channels=1, uvscale=2.0))
    bm.add(bt)
    t3.add(bm)
    z["ts_bump_rp"] = t3
    b = Plugin("bsdf", "blendbsdf", weight=0.4)
    inner = Plugin("bsdf", "twosided")
    inner.add(Plugin("bsdf", "conductor", **AU))
    b.add(inner)
    b.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5)))
    z["blend_ts_cond"] = b
    # thindielectric (same 8(f) row): delta reflection + null transmission, internal reflections folded into R
    z["thin_glass"] = Plugin("bsdf", "thindielectric", intIOR=1.5)
    # mask (same row): opacity-weighted nested BSDF + null transmission; constant and textured opacity
    mk = Plugin("bsdf", "mask", opacity=Spectrum(0.3, 0.6, 0.8))
    mk.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.8, 0.7, 0.3)))
    z["mask_diffuse"] = mk
    mk2 = Plugin("bsdf", "mask")
    mk2.add(_tex(12, channels=3, uvscale=3.0), "opacity")
    mk2.add(Plugin("bsdf", "roughconductor", distribution="beckmann", alpha=0.2, **AU))
    z["mask_tex_rc"] = mk2
    # roughdielectric (same row): GGX with visible-normal sampling; anisotropic Beckmann with sampleVisible=false
    # (Walter's alpha scaling) and tinted transmission.  Uses the third random number (bRec.sampler->next1D()).
    z["rd_ggx"] = Plugin("bsdf", "roughdielectric", distribution="ggx", alpha=0.2, intIOR=1.5)
    z["rd_beck_all"] = Plugin("bsdf", "roughdielectric", distribution="beckmann", alphaU=0.15, alphaV=0.3, sampleVisible=False,
                              intIOR=1.33, extIOR=1.0, specularReflectance=Spectrum(0.9),
                              specularTransmittance=Spectrum(0.9, 0.8, 0.6))
    z["thin_tinted"] = Plugin("bsdf", "thindielectric", intIOR=1.33, extIOR=1.0, specularReflectance=Spectrum(0.8),
                              specularTransmittance=Spectrum(0.9, 0.7, 0.5))
    return z



# ---- mixturebsdf, roughcoating, normalmap (the rest of the row) --------------------------------
def _normal_tex(seed, size=64, **kw):
    """Tangent-space normal map: unit normals around +z stored as 0.5 * (n + 1) (src/bsdfs/normalmap.cpp:110-112)."""
    rng = np.random.RandomState(seed)
    n = np.concatenate([rng.uniform(-0.35, 0.35, (size, size, 2)), np.ones((size, size, 1))], -1)
    n /= np.linalg.norm(n, axis=2, keepdims=True)
    t = Plugin("texture", "bitmap", filterType="bilinear", gamma=1.0, **kw)
    t.data["bitmap"] = (0.5 * (n + 1)).astype(F)
    return t


def nextrow_bsdfs():
    """MixtureBSDF (2 and 3 nested models, weights above and below one), RoughCoating (constant / textured roughness,
    absorbing layer, back side) and NormalMap (below and above other adapters)."""
    CU, AU = scenes.CU, scenes.AU
    z = {}
    m = Plugin("bsdf", "mixturebsdf", weights="0.7, 0.2")
    m.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.6, 0.3, 0.2)))
    m.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.2, **CU))
    z["mix_diff_rc"] = m
    m3 = Plugin("bsdf", "mixturebsdf", weights="0.8 0.6 0.4")          # sums to 1.8: rescaled (mixturebsdf.cpp:131-143)
    m3.add(Plugin("bsdf", "plastic", intIOR=1.5, diffuseReflectance=Spectrum(0.2, 0.4, 0.6)))
    m3.add(Plugin("bsdf", "conductor", **AU))
    d = Plugin("bsdf", "diffuse")
    d.add(_tex(21, uvscale=2.0), "reflectance")
    m3.add(d)
    z["mix3_rescaled"] = m3
    mg = Plugin("bsdf", "mixturebsdf", weights="0.5;0.5")                # transmission below a mixture
    mg.add(Plugin("bsdf", "dielectric", intIOR=1.5))
    mg.add(Plugin("bsdf", "roughplastic", distribution="beckmann", alpha=0.2, intIOR=1.5, diffuseReflectance=Spectrum(0.7, 0.2, 0.2)))
    z["mix_glass_rp"] = mg
    rc = Plugin("bsdf", "roughcoating", distribution="ggx", alpha=0.15, intIOR=1.5)
    rc.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.7, 0.5, 0.2)))
    z["rcoat_diffuse"] = rc
    rc2 = Plugin("bsdf", "roughcoating", distribution="beckmann", intIOR=1.7, extIOR=1.0, thickness=2.0,
                 sigmaA=Spectrum(0.3, 0.1, 0.6), specularReflectance=Spectrum(0.9, 0.95, 0.8))
    rc2.add(_alpha_tex(22), "alpha")
    rc2.add(Plugin("bsdf", "roughconductor", distribution="beckmann", alpha=0.3, **AU))
    z["rcoat_tex_rcond"] = rc2
    rc3 = Plugin("bsdf", "roughcoating", distribution="ggx", alpha=0.3, sampleVisible=False, intIOR=1.33)
    rc3.add(Plugin("bsdf", "plastic", intIOR=1.6, diffuseReflectance=Spectrum(0.3, 0.6, 0.4)))
    z["rcoat_all_plastic"] = rc3
    # albedo / specular bitmaps of one geometry with the default filterType (ewa): the two lookups share one EWA walk
    rp = Plugin("bsdf", "roughplastic", distribution="ggx", alpha=0.2, intIOR=1.5)
    for k, slot in ((31, "diffuseReflectance"), (32, "specularReflectance")):
        t = _tex(k, uvscale=3.0, size=128)
        t.props["filterType"] = "ewa"
        rp.add(t, slot)
    z["rp_pair_ewa"] = rp
    nm = Plugin("bsdf", "normalmap")
    nm.add(_normal_tex(23, uvscale=2.0))
    nm.add(Plugin("bsdf", "roughplastic", distribution="ggx", alpha=0.1, intIOR=1.49, diffuseReflectance=Spectrum(0.3, 0.5, 0.7)))
    z["nmap_rp"] = nm
    ts = Plugin("bsdf", "twosided")
    nm2 = Plugin("bsdf", "normalmap")
    nm2.add(_normal_tex(24, uvscale=3.0))
    nm2.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.6, 0.6, 0.3)))
    ts.add(nm2)
    z["ts_nmap_diffuse"] = ts
    nm3 = Plugin("bsdf", "normalmap")                                     # normalmap above a wrapper: generic shade class
    nm3.add(_normal_tex(25, uvscale=1.5))
    ct = Plugin("bsdf", "coating", intIOR=1.5, sigmaA=Spectrum(0.2, 0.4, 0.1))
    ct.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5, 0.4, 0.6)))
    nm3.add(ct)
    z["nmap_coating"] = nm3
    return z


def _alpha_tex(seed, size=64):
    """Roughness texture inside the range of the rough-transmittance tables."""
    rng = np.random.RandomState(seed)
    t = Plugin("texture", "bitmap", filterType="bilinear", gamma=1.0, uvscale=2.0)
    t.data["bitmap"] = rng.uniform(0.08, 0.45, (size, size, 1)).astype(F)
    return t


def nextrow_scene(width=96, height=64, spp=4):
    """Spheres and a knot carrying the nextrow materials over a diffuse floor, lit by an area light and an envmap."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=6))
    s = Plugin("sensor", "perspective", toWorld=hm.look_at([0, 4.0, 11], [0, 1.0, 0], [0, 1, 0]), fov=45.0)
    s.add(Plugin("sampler", "independent", sampleCount=spp))
    film = Plugin("film", "hdrfilm", width=width, height=height)
    film.add(Plugin("rfilter", "gaussian"))
    s.add(film)
    sc.add(s)
    z = nextrow_bsdfs()
    floor = Plugin("shape", "rectangle", toWorld=hm.rotate([1, 0, 0], -90) @ hm.scale([12, 12, 1]))
    floor.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5)))
    sc.add(floor)
    names = list(z)
    for i, nm in enumerate(names):
        col, row = i % 5, i // 5
        sp = Plugin("shape", "sphere", center=Point((col - 2) * 2.3, 1.0, -2.6 * row), radius=1.0)
        sp.add(z[nm])
        sc.add(sp)
    pos, idx, nrm, uv = scenes.torus_knot(40, 12)
    knot = Plugin("shape", "mesh", toWorld=hm.translate([0, 0.9, 3.2]) @ hm.scale(0.6))
    knot.data.update(positions=pos, indices=idx, normals=nrm, texcoords=uv)
    knot.add(z["nmap_rp"])
    sc.add(knot)
    lt = Plugin("shape", "rectangle", toWorld=hm.translate([0, 8, 2]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([1.5, 1.5, 1]))
    lt.add(Plugin("emitter", "area", radiance=Spectrum(22, 20, 18)))
    sc.add(lt)
    env = Plugin("emitter", "envmap", scale=0.6)
    env.data["bitmap"] = scenes.synthetic_envmap(128, 64)
    sc.add(env)
    return sc, z


def twosided_scene(width=96, height=64, spp=4, hide_env=False):
    """Open surfaces seen and lit from both sides: free-standing rectangles carrying the twosided materials (front
    faces turned towards, away from and edge-on to the light), one plain one-sided rectangle for contrast."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=6))
    s = Plugin("sensor", "perspective", toWorld=hm.look_at([0, 3.5, 10], [0, 1.6, 0], [0, 1, 0]), fov=45.0)
    s.add(Plugin("sampler", "independent", sampleCount=spp))
    film = Plugin("film", "hdrfilm", width=width, height=height)
    film.add(Plugin("rfilter", "gaussian"))
    s.add(film)
    sc.add(s)
    z = twosided_bsdfs()
    floor = Plugin("shape", "rectangle", toWorld=hm.rotate([1, 0, 0], -90) @ hm.scale([10, 10, 1]))
    floor.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5)))
    sc.add(floor)
    for i, nm in enumerate(z):
        # normals alternate between facing the camera (+z) and facing away (-z), tilted about the vertical axis
        yaw = (180.0 if i % 2 else 0.0) + 25.0 * (i - 1.5)
        r = Plugin("shape", "rectangle", toWorld=hm.translate([(i - 1.5) * 2.6, 1.6, -0.5 * i]) @ hm.rotate([0, 1, 0], yaw)
                   @ hm.scale([1.0, 1.4, 1]))
        r.add(z[nm])
        sc.add(r)
    one = Plugin("shape", "rectangle", toWorld=hm.translate([0, 4.2, -3]) @ hm.rotate([0, 1, 0], 180) @ hm.scale([1.5, 0.7, 1]))
    one.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.2, 0.3, 0.8)))      # back face towards the camera: black
    sc.add(one)
    pos, idx, nrm, uv = scenes.torus_knot(40, 12)
    knot = Plugin("shape", "mesh", toWorld=hm.translate([0, 0.9, 3]) @ hm.scale(0.6))
    knot.data.update(positions=pos, indices=idx, normals=nrm, texcoords=uv)
    knot.add(z["ts_rc_plastic"])
    sc.add(knot)
    lt = Plugin("shape", "rectangle", toWorld=hm.translate([0, 7, 2]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([1.2, 1.2, 1]))
    lt.add(Plugin("emitter", "area", radiance=Spectrum(25, 23, 20)))
    sc.add(lt)
    sc.add(Plugin("emitter", "point", position=Point(0, 2, -6), intensity=Spectrum(60, 60, 80)))   # lights the far sides
    if hide_env:
        # hideEmitters with an environment map seen directly, through the null-transmission panes (still hidden:
        # `scattered` stays false, path.cpp:213,238) and by reflection (visible)
        integ = sc.child("integrator")
        integ.props["hideEmitters"] = True
        env = Plugin("emitter", "envmap", scale=0.5)
        env.data["bitmap"] = scenes.synthetic_envmap(128, 64)
        sc.add(env)
    return sc, z


# ---- boundary cases of the plugin / host flattening -----------------------------------------
def cropped_cornell(spp=4):
    """Cornell box whose film renders a crop window (cropOffsetX/Y > 0): the crop is part of the camera's
    sample-to-camera transform (perspective.cpp:140-160), block offsets are relative to the window."""
    sc = scenes.cornell_box(96, 96, spp)
    film = sc.child("sensor").child("film")
    film.props.update(cropOffsetX=20, cropOffsetY=30, cropWidth=48, cropHeight=40)
    return sc


def sphere_alpha_scene(width=80, height=60, spp=4):
    """Spheres with center != 0 under a scaled + rotated toWorld (sphere.cpp:108-121 composes toWorld * scale(1/r) *
    translate(center)), no environment, an `rgba` film: the alpha channel is 1 where the camera ray hits something."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=5))
    s = Plugin("sensor", "perspective", toWorld=hm.look_at([0, 3, 9], [0, 1, 0], [0, 1, 0]), fov=45.0)
    s.add(Plugin("sampler", "independent", sampleCount=spp))
    film = Plugin("film", "hdrfilm", width=width, height=height, pixelFormat="rgba")
    film.add(Plugin("rfilter", "gaussian"))
    s.add(film)
    sc.add(s)
    floor = Plugin("shape", "rectangle", toWorld=hm.rotate([1, 0, 0], -90) @ hm.scale([4, 4, 1]))
    floor.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5)))
    sc.add(floor)
    tw = hm.translate([-1.0, 0.2, 0.5]) @ hm.rotate([0, 1, 0], 35) @ hm.rotate([0, 0, 1], 20) @ hm.scale(1.6)
    a = Plugin("shape", "sphere", center=Point(1.2, 0.9, -0.4), radius=0.5, toWorld=tw)
    a.add(Plugin("bsdf", "diffuse", reflectance=Spectrum(0.7, 0.3, 0.2)))
    sc.add(a)
    b = Plugin("shape", "sphere", center=Point(-0.8, 0.5, 0.6), radius=0.3, toWorld=hm.translate([0.5, 0, 1.0]) @ hm.scale(2.0))
    b.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.2, **scenes.CU))
    sc.add(b)
    lt = Plugin("shape", "rectangle", toWorld=hm.translate([0, 6, 1]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([1.2, 1.2, 1]))
    lt.add(Plugin("emitter", "area", radiance=Spectrum(25, 23, 20)))
    sc.add(lt)
    return sc


def _pattern(w, h, channels=3, seed=11):
    """High-frequency texture (checker + stripes + noise) whose minified appearance depends on the filter."""
    rng = np.random.RandomState(seed)
    y, x = np.mgrid[0:h, 0:w]
    chk = ((x // 3 + y // 2) % 2).astype(F)
    img = np.stack([0.1 + 0.8 * chk, 0.5 + 0.45 * np.sin(x * 1.3), 0.5 + 0.45 * np.cos(y * 0.9)], axis=2).astype(F)
    img = np.clip(img + rng.uniform(-0.08, 0.08, img.shape).astype(F), 0.0, 1.0)
    return img[:, :, :channels] if channels == 3 else img[:, :, :1]


def mip_textures():
    """Bitmap textures that exercise the MIP map: default filter (ewa, maxAnisotropy 20), low anisotropy, trilinear, odd
    sizes (the Lanczos pyramid rounds sizes up), every wrap mode, uv scale / offset, one channel."""
    def tex(w, h, ch=3, seed=11, **kw):
        t = Plugin("texture", "bitmap", gamma=1.0, **kw)
        t.data["bitmap"] = _pattern(w, h, ch, seed)
        return t
    return {
        "ewa": tex(96, 80, uscale=400.0, vscale=400.0),   # on the floor: hundreds of texels per pixel even at 1/64 footprint
        "ewa_aniso2": tex(61, 47, seed=12, maxAnisotropy=2.0, uscale=5.0, vscale=7.0, wrapMode="mirror"),
        "trilinear": tex(64, 64, seed=13, filterType="trilinear", uscale=9.0, vscale=9.0),
        "ewa_clamp_mono": tex(50, 33, ch=1, seed=14, wrapModeU="clamp", wrapModeV="zero", uoffset=0.1, voffset=-0.2),
        "ewa_one": tex(37, 64, seed=15, wrapModeU="one", wrapModeV="repeat", uscale=3.0, vscale=2.0),
        "bilinear": tex(48, 48, seed=16, filterType="bilinear", uscale=6.0, vscale=6.0),
        "nearest": tex(48, 48, seed=17, filterType="nearest", uscale=6.0, vscale=6.0),
    }


def mip_scene(width=96, height=64, spp=4, envmap=True):
    """A textured ground plane seen at a grazing angle (strong minification + anisotropy), textured spheres and a
    bump-mapped cube, under an environment map that camera rays see through the EWA lookup.  All bitmap textures
    use the reference's default filterType (ewa) unless they say otherwise."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=4))
    s = Plugin("sensor", "perspective", toWorld=hm.look_at([0, 1.2, 9], [0, 0.6, 0], [0, 1, 0]), fov=55.0)
    s.add(Plugin("sampler", "independent", sampleCount=spp))
    film = Plugin("film", "hdrfilm", width=width, height=height)
    film.add(Plugin("rfilter", "gaussian"))
    s.add(film)
    sc.add(s)
    t = mip_textures()
    floor = Plugin("shape", "rectangle", toWorld=hm.rotate([1, 0, 0], -90) @ hm.scale([30, 30, 1]))
    fb = Plugin("bsdf", "diffuse")
    fb.add(t["ewa"], "reflectance")
    floor.add(fb)
    sc.add(floor)
    mats = {}
    b = Plugin("bsdf", "diffuse"); b.add(t["ewa_aniso2"], "reflectance"); mats["diffuse_aniso2"] = b
    b = Plugin("bsdf", "roughplastic", distribution="ggx", alpha=0.15); b.add(t["trilinear"], "diffuseReflectance"); mats["rp_trilinear"] = b
    b = Plugin("bsdf", "roughconductor", distribution="beckmann", **scenes.CU); b.add(t["ewa_clamp_mono"], "alpha"); mats["rc_alpha_tex"] = b
    inner = Plugin("bsdf", "diffuse"); inner.add(t["ewa_one"], "reflectance")
    b = Plugin("bsdf", "twosided"); b.add(inner); mats["twosided_one"] = b
    b = Plugin("bsdf", "diffuse"); b.add(t["bilinear"], "reflectance"); mats["diffuse_bilinear"] = b
    for i, nm in enumerate(mats):
        sp = Plugin("shape", "sphere", center=Point((i - 2) * 2.1, 1.0, 2.0 - 0.8 * i), radius=1.0)
        sp.add(mats[nm])
        sc.add(sp)
    cube = Plugin("shape", "cube", toWorld=hm.translate([0, 0.7, 5.5]) @ hm.rotate([0, 1, 0], 30) @ hm.scale([0.7, 0.7, 0.7]))
    inner = Plugin("bsdf", "diffuse"); inner.add(t["nearest"], "reflectance")
    bump = Plugin("bsdf", "bumpmap")
    disp = Plugin("texture", "scale", scale=0.05); disp.add(_tex(31, channels=1, size=32))
    bump.add(disp); bump.add(inner)
    cube.add(bump)
    sc.add(cube)
    mats["bump_nearest"] = bump
    mats["floor"] = fb
    lt = Plugin("shape", "rectangle", toWorld=hm.translate([0, 8, 3]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([2.0, 2.0, 1]))
    lt.add(Plugin("emitter", "area", radiance=Spectrum(18, 17, 15)))
    sc.add(lt)
    if envmap:
        env = Plugin("emitter", "envmap", scale=0.8, toWorld=hm.rotate([0, 1, 0], 25.0))
        env.data["bitmap"] = scenes.synthetic_envmap(160, 80)
        sc.add(env)
    return sc, t, mats


class HalfTexels:
    """float16 texels for Context.add_texture with texel_format = MTS_TEXEL_F16 (passed through as raw memory)."""

    def __init__(self, a):
        self.a = np.ascontiguousarray(a, dtype=np.float16)
        self.nbytes = self.a.nbytes
