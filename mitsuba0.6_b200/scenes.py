"""Synthetic scenes for BASELINE.json's configs, generated with fixed seeds as
Plugin trees (scene.py) so the same description feeds the CUDA library, the CPU oracle and --
written out with scene.to_xml -- the reference build.
"""
import math

import numpy as np

from . import hostmath as hm
from .scene import Plugin, Point, Spectrum

F = np.float32


def _sensor(origin, target, up, fov, w, h, spp, rfilter, fov_axis="x"):
    s = Plugin("sensor", "perspective", toWorld=hm.look_at(origin, target, up), fov=float(fov), fovAxis=fov_axis)
    s.add(Plugin("sampler", "independent", sampleCount=int(spp)))
    film = Plugin("film", "hdrfilm", width=int(w), height=int(h))
    film.add(Plugin("rfilter", rfilter))
    s.add(film)
    return s


def _diffuse(rgb):
    return Plugin("bsdf", "diffuse", reflectance=Spectrum(rgb))


def _rect(to_world, bsdf, emitter=None):
    r = Plugin("shape", "rectangle", toWorld=to_world)
    r.add(bsdf)
    if emitter is not None:
        r.add(emitter)
    return r


def cornell_box(width=512, height=512, spp=64, max_depth=5, rfilter="box"):
    """cfg1: unit-scale Cornell box from 5 rectangles, 2 cubes, 1 rectangle area light; diffuse only."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=int(max_depth)))
    sc.add(_sensor([0, 1, 3.9], [0, 1, 0], [0, 1, 0], 39.3077, width, height, spp, rfilter))
    white, red, green = (0.725, 0.71, 0.68), (0.63, 0.065, 0.05), (0.14, 0.45, 0.091)
    T, S, R = hm.translate, hm.scale, hm.rotate
    # rectangle object space: [-1,1]^2 in xy, normal +z
    sc.add(_rect(T([0, 0, 0]) @ R([1, 0, 0], -90), _diffuse(white)))                 # floor  (normal +y)
    sc.add(_rect(T([0, 2, 0]) @ R([1, 0, 0], 90), _diffuse(white)))                  # ceiling (normal -y)
    sc.add(_rect(T([0, 1, -1]), _diffuse(white)))                                    # back wall (normal +z)
    sc.add(_rect(T([-1, 1, 0]) @ R([0, 1, 0], 90), _diffuse(red)))                   # left wall (normal +x)
    sc.add(_rect(T([1, 1, 0]) @ R([0, 1, 0], -90), _diffuse(green)))                 # right wall (normal -x)
    light = Plugin("emitter", "area", radiance=Spectrum(17.0, 12.0, 4.0))
    sc.add(_rect(T([0, 1.98, 0]) @ R([1, 0, 0], 90) @ S([0.25, 0.2, 1]), _diffuse((0, 0, 0)), light))
    tall = Plugin("shape", "cube", toWorld=T([-0.33, 0.6, -0.3]) @ R([0, 1, 0], 17) @ S([0.3, 0.6, 0.3]))
    tall.add(_diffuse(white))
    short = Plugin("shape", "cube", toWorld=T([0.33, 0.3, 0.35]) @ R([0, 1, 0], -17) @ S([0.3, 0.3, 0.3]))
    short.add(_diffuse(white))
    sc.add(tall)
    sc.add(short)
    return sc


def synthetic_envmap(width=2048, height=1024, seed=1):
    """Analytic sky gradient (0.2..2.0) + 3 Gaussian 'suns' (peak 5e3, sigma 1.5 deg) so the
    hierarchical CDF is strongly non-uniform."""
    rng = np.random.RandomState(seed)
    v = (np.arange(height, dtype=np.float64) + 0.5) / height
    u = (np.arange(width, dtype=np.float64) + 0.5) / width
    theta, phi = np.pi * v[:, None], 2 * np.pi * u[None, :]
    d = np.stack([np.sin(phi) * np.sin(theta), np.cos(theta) * np.ones_like(phi), -np.cos(phi) * np.sin(theta)], -1)
    sky = 0.2 + 1.8 * np.clip(d[..., 1] * 0.5 + 0.5, 0, 1) ** 2
    img = np.stack([sky * 0.75, sky * 0.85, sky * 1.0], -1)
    sig = math.radians(1.5)
    for col in ((1.0, 0.9, 0.7), (0.9, 0.95, 1.0), (1.0, 0.6, 0.4)):
        el, az = rng.uniform(0.25, 1.2), rng.uniform(0, 2 * np.pi)
        sd = np.array([math.sin(az) * math.cos(el), math.sin(el), -math.cos(az) * math.cos(el)])
        ang = np.arccos(np.clip(d @ sd, -1, 1))
        img += 5e3 * np.exp(-0.5 * (ang / sig) ** 2)[..., None] * np.array(col)
    return img.astype(F)


CU = dict(eta=Spectrum(0.200438, 0.924033, 1.102212), k=Spectrum(3.912949, 2.452848, 2.142188))
AU = dict(eta=Spectrum(0.143119, 0.374957, 1.442479), k=Spectrum(3.983160, 2.385721, 1.603215))
AL = dict(eta=Spectrum(1.657460, 0.880369, 0.521229), k=Spectrum(9.223869, 6.269523, 4.837001))


def material_spheres(width=1024, height=1024, spp=256, max_depth=8, env_size=(2048, 1024), rfilter="gaussian"):
    """cfg2: 5x5 grid of unit spheres (roughconductor / roughplastic / dielectric / coating variants,
    GGX + Beckmann) on a rectangle floor under a synthetic envmap."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=int(max_depth)))
    sc.add(_sensor([0, 9, 16], [0, 0.5, 0], [0, 1, 0], 40, width, height, spp, rfilter))
    env = Plugin("emitter", "envmap", scale=1.0, toWorld=hm.rotate([0, 1, 0], 30.0))
    env.data["bitmap"] = synthetic_envmap(*env_size)
    sc.add(env)
    floor = _rect(hm.rotate([1, 0, 0], -90) @ hm.scale([14, 14, 1]), _diffuse((0.4, 0.4, 0.4)))
    sc.add(floor)
    mats = []
    for metal in (CU, AU, AL):
        for distr, alpha in (("ggx", 0.05), ("beckmann", 0.15), ("ggx", 0.3), ("beckmann", 0.5)):
            mats.append(Plugin("bsdf", "roughconductor", distribution=distr, alpha=alpha, **metal))
    for distr, alpha in (("ggx", 0.05), ("beckmann", 0.2), ("ggx", 0.4), ("beckmann", 0.1)):
        mats.append(Plugin("bsdf", "roughplastic", distribution=distr, alpha=alpha, intIOR=1.49,
                           diffuseReflectance=Spectrum(0.2, 0.35, 0.6)))
    mats.append(Plugin("bsdf", "dielectric", intIOR=1.5046))
    mats.append(Plugin("bsdf", "dielectric", intIOR="water"))
    c1 = Plugin("bsdf", "coating", intIOR=1.5, thickness=1.0, sigmaA=Spectrum(0.1, 0.2, 0.3))
    c1.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.2, **CU))
    c2 = Plugin("bsdf", "coating", intIOR=1.5, thickness=1.0, sigmaA=Spectrum(0.3, 0.1, 0.1))
    c2.add(_diffuse((0.6, 0.5, 0.3)))
    mats += [c1, c2]
    mats.append(Plugin("bsdf", "plastic", intIOR=1.49, diffuseReflectance=Spectrum(0.6, 0.1, 0.1)))
    mats.append(Plugin("bsdf", "conductor", **AU))
    mats.append(Plugin("bsdf", "roughconductor", distribution="beckmann", alphaU=0.1, alphaV=0.3, **AL))
    mats.append(Plugin("bsdf", "roughconductor", distribution="ggx", alphaU=0.3, alphaV=0.08, **CU))
    mats.append(_diffuse((0.5, 0.5, 0.5)))
    assert len(mats) == 25, len(mats)
    for i, m in enumerate(mats):
        x, z = (i % 5 - 2) * 2.5, (i // 5 - 2) * 2.5
        s = Plugin("shape", "sphere", center=Point(x, 1.0, z), radius=1.0)
        s.add(m)
        sc.add(s)
    return sc


def fbm_heightfield(n, octaves=6, seed=7):
    """n x n fBm height grid (value-noise octaves, bilinear-upsampled)."""
    rng = np.random.RandomState(seed)
    h = np.zeros((n, n), dtype=np.float64)
    amp, freq = 1.0, 4
    xs = np.linspace(0, 1, n)
    for _ in range(octaves):
        g = rng.rand(freq + 2, freq + 2)
        fx = xs * freq
        ix = np.minimum(fx.astype(int), freq - 1)
        tx = fx - ix
        tx = tx * tx * (3 - 2 * tx)
        a = g[ix][:, ix] * (1 - tx)[None, :] + g[ix][:, ix + 1] * tx[None, :]
        b = g[ix + 1][:, ix] * (1 - tx)[None, :] + g[ix + 1][:, ix + 1] * tx[None, :]
        h += amp * (a * (1 - tx)[:, None] + b * tx[:, None])
        amp *= 0.5
        freq *= 2
    return h


def terrain_mesh(n=2237, size=20.0, height=2.5, seed=7):
    """(n x n)-vertex terrain -> 2(n-1)^2 triangles; n=2237 gives 9 999 392 ~ 10M triangles."""
    hgt = fbm_heightfield(n, seed=seed)
    hgt = (hgt - hgt.mean()) * height
    xs = np.linspace(-size / 2, size / 2, n)
    X, Z = np.meshgrid(xs, xs, indexing="xy")
    pos = np.stack([X, hgt, Z], -1).reshape(-1, 3).astype(F)
    uv = np.stack(np.meshgrid(np.linspace(0, 1, n), np.linspace(0, 1, n), indexing="xy"), -1).reshape(-1, 2).astype(F)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="xy")
    v00 = (j * n + i).reshape(-1)
    v10, v01, v11 = v00 + 1, v00 + n, v00 + n + 1
    idx = np.concatenate([np.stack([v00, v01, v10], -1), np.stack([v10, v01, v11], -1)], 0).astype(np.uint32)
    # smooth normals from the height gradient (cheap and exact enough for a generated asset)
    step = size / (n - 1)
    gz, gx = np.gradient(hgt, step, step)
    nrm = np.stack([-gx, np.ones_like(gx), -gz], -1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    return pos, idx, nrm.reshape(-1, 3).astype(F), uv


def procedural_texture(size=4096, seed=11, channels=3, base=None):
    """fBm pattern mapped through a seeded colour ramp.  `base`: a normalised heightfield to reuse (the fBm synthesis
    dominates the cost; scenes with several large textures derive them from one field)."""
    rng = np.random.RandomState(seed)
    if base is None:
        base = fbm_heightfield(size, octaves=7, seed=seed)
        base = (base - base.min()) / (base.max() - base.min())
    if channels == 1:
        return base.astype(F)[:, :, None]
    c0, c1 = rng.uniform(0.05, 0.4, 3), rng.uniform(0.5, 0.95, 3)
    return (c0[None, None, :] * (1 - base[..., None]) + c1[None, None, :] * base[..., None]).astype(F)


def terrain_scene(width=1920, height=1080, spp=128, max_depth=8, grid=2237, tex_size=4096, rfilter="gaussian",
                  env_size=(2048, 1024), filter_type="ewa"):
    """cfg3 / cfg5 (synthetic inputs): ~10M-triangle terrain, four tex_size^2 RGB bitmap textures
    and a tex_size^2 luminance bump map -- `bumpmap` over roughplastic (terrain) and over diffuse (slab) -- one area
    light + envmap.  The bitmaps use the reference's default filterType (ewa) unless told otherwise."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=int(max_depth)))
    sc.add(_sensor([0, 6.5, 15], [0, 0.2, 0], [0, 1, 0], 45, width, height, spp, rfilter))
    env = Plugin("emitter", "envmap", scale=0.5)
    env.data["bitmap"] = synthetic_envmap(*env_size)
    sc.add(env)
    pos, idx, nrm, uv = terrain_mesh(grid)
    mesh = Plugin("shape", "mesh")
    mesh.data.update(positions=pos, indices=idx, normals=nrm, texcoords=uv)
    field = fbm_heightfield(tex_size, octaves=7, seed=11)
    field = (field - field.min()) / (field.max() - field.min())

    def rgb(k, **kw):   # four different textures from one fBm field: rotated / mirrored, own colour ramp
        t = Plugin("texture", "bitmap", filterType=filter_type, gamma=1.0, **kw)
        t.data["bitmap"] = procedural_texture(tex_size, seed=11 + 10 * k, base=np.rot90(field, k) if k < 3 else field[::-1])
        return t
    albedo, spec, slab_tex, block_tex = rgb(0, uvscale=4.0), rgb(1, uvscale=4.0), rgb(2), rgb(3)
    bumptex = Plugin("texture", "bitmap", filterType=filter_type, gamma=1.0, uvscale=8.0)
    bumptex.data["bitmap"] = procedural_texture(tex_size, seed=12, channels=1)
    sc_tex = Plugin("texture", "scale", scale=0.02)
    sc_tex.add(bumptex)
    rp = Plugin("bsdf", "roughplastic", distribution="ggx", alpha=0.15, intIOR=1.49)
    rp.add(albedo, "diffuseReflectance")
    rp.add(spec, "specularReflectance")
    bump = Plugin("bsdf", "bumpmap")
    bump.add(rp)
    bump.add(sc_tex)
    mesh.add(bump)
    sc.add(mesh)
    # bumpmap over diffuse: a textured slab ...
    slab = Plugin("shape", "cube", toWorld=hm.translate([3.0, 2.2, 1.0]) @ hm.rotate([0, 1, 0], 30) @ hm.scale([1.2, 0.8, 1.2]))
    db = Plugin("bsdf", "diffuse")
    db.add(slab_tex, "reflectance")
    sbump = Plugin("bsdf", "bumpmap")
    sbump.add(db)
    sbump.add(sc_tex)
    slab.add(sbump)
    sc.add(slab)
    # ... and a plain textured block
    block = Plugin("shape", "cube", toWorld=hm.translate([-3.5, 1.8, 3.0]) @ hm.rotate([0, 1, 0], -20) @ hm.scale([0.9, 0.6, 0.9]))
    bb = Plugin("bsdf", "diffuse")
    bb.add(block_tex, "reflectance")
    block.add(bb)
    sc.add(block)
    light = Plugin("emitter", "area", radiance=Spectrum(60.0, 55.0, 45.0))
    sc.add(_rect(hm.translate([-4, 7, 2]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([1.0, 1.0, 1.0]), _diffuse((0, 0, 0)), light))
    return sc


def torus_knot(nu=500, nv=200, R=1.0, r=0.28, p=2, q=3):
    """(p,q) torus-knot tube mesh: 2*nu*nv triangles (200k at the defaults)."""
    t = np.linspace(0, 2 * np.pi, nu, endpoint=False)
    c = np.stack([(R + 0.4 * np.cos(q * t)) * np.cos(p * t), 0.4 * np.sin(q * t), (R + 0.4 * np.cos(q * t)) * np.sin(p * t)], -1)
    tang = np.roll(c, -1, 0) - np.roll(c, 1, 0)
    tang /= np.linalg.norm(tang, axis=1, keepdims=True)
    up = np.array([0, 1, 0.0])
    n1 = np.cross(tang, up)
    n1 /= np.linalg.norm(n1, axis=1, keepdims=True)
    n2 = np.cross(tang, n1)
    a = np.linspace(0, 2 * np.pi, nv, endpoint=False)
    ring = np.cos(a)[None, :, None] * n1[:, None, :] + np.sin(a)[None, :, None] * n2[:, None, :]
    pos = (c[:, None, :] + r * ring).reshape(-1, 3)
    nrm = ring.reshape(-1, 3)
    uv = np.stack(np.meshgrid(np.arange(nu) / nu * 8, np.arange(nv) / nv, indexing="ij"), -1).reshape(-1, 2)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    v00 = (i * nv + j).reshape(-1)
    v10 = (((i + 1) % nu) * nv + j).reshape(-1)
    v01 = (i * nv + (j + 1) % nv).reshape(-1)
    v11 = (((i + 1) % nu) * nv + (j + 1) % nv).reshape(-1)
    idx = np.concatenate([np.stack([v00, v10, v01], -1), np.stack([v10, v11, v01], -1)], 0).astype(np.uint32)
    return pos.astype(F), idx, nrm.astype(F), uv.astype(F)


def caustic_scene(width=1024, height=1024, spp=512, max_depth=16, rr_depth=5, rfilter="gaussian", knot=(500, 200)):
    """cfg4: glass sphere + torus knot over a blendbsdf(conductor, diffuse) floor, mirror wall, small light."""
    sc = Plugin("scene", "")
    sc.add(Plugin("integrator", "path", maxDepth=int(max_depth), rrDepth=int(rr_depth)))
    sc.add(_sensor([0, 3.2, 7.5], [0, 0.9, 0], [0, 1, 0], 38, width, height, spp, rfilter))
    checker = np.indices((1024, 1024)).sum(0) // 64 % 2
    wtex = Plugin("texture", "bitmap", filterType="bilinear", gamma=1.0, uvscale=2.0)
    wtex.data["bitmap"] = (0.15 + 0.7 * checker).astype(F)[:, :, None]
    blend = Plugin("bsdf", "blendbsdf")
    blend.add(wtex, "weight")
    blend.add(Plugin("bsdf", "conductor", **AU))
    blend.add(_diffuse((0.55, 0.5, 0.45)))
    sc.add(_rect(hm.rotate([1, 0, 0], -90) @ hm.scale([6, 6, 1]), blend))
    sc.add(_rect(hm.translate([0, 3, -4]) @ hm.scale([6, 3, 1]), Plugin("bsdf", "conductor", **AL)))
    sc.add(_rect(hm.translate([-5, 3, 0]) @ hm.rotate([0, 1, 0], 90) @ hm.scale([5, 3, 1]), _diffuse((0.6, 0.2, 0.2))))
    glass = Plugin("shape", "sphere", center=Point(-1.3, 1.0, 0.8), radius=1.0)
    glass.add(Plugin("bsdf", "dielectric", intIOR=1.5))
    sc.add(glass)
    pos, idx, nrm, uv = torus_knot(*knot)
    k = Plugin("shape", "mesh", toWorld=hm.translate([1.4, 1.15, -0.3]) @ hm.scale(0.75))
    k.data.update(positions=pos, indices=idx, normals=nrm, texcoords=uv)
    k.add(Plugin("bsdf", "roughconductor", distribution="ggx", alpha=0.08, **CU))
    sc.add(k)
    light = Plugin("emitter", "area", radiance=Spectrum(400.0, 380.0, 340.0))
    sc.add(_rect(hm.translate([2.5, 5.5, 2.5]) @ hm.rotate([1, 0, 0], 90) @ hm.scale([0.25, 0.25, 1]), _diffuse((0, 0, 0)), light))
    return sc


CONFIGS = {
    "cornell": cornell_box,
    "spheres": material_spheres,
    "terrain": terrain_scene,
    "caustics": caustic_scene,
}
