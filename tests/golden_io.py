"""Seeded inputs shared by tests/golden/make_golden.py (which feeds them to the REFERENCE build) and by the tests
(which feed them to the oracle restatement and to the CUDA library).  Changing anything here invalidates the
fixtures: regenerate them with make_golden.py."""
import os
import zlib

import numpy as np

import mts_b200 as M  # noqa: F401
from mitsuba0_6_b200 import abi, scenes
import util

F = np.float32
HERE = os.path.dirname(os.path.abspath(__file__))
N_BSDF, N_RAYS, N_RECORDS, N_EMIT, N_ENV, N_SENSOR, N_SPLAT = 768, 24000, 6000, 8000, 6000, 2000, 6000
ZOO_W, ZOO_H, CORNELL_RES = 96, 64, 96


def zoo(spp=4):
    return util.zoo_scene(width=ZOO_W, height=ZOO_H, spp=spp)


def twosided(spp=4, hide_env=False):
    return util.twosided_scene(width=ZOO_W, height=ZOO_H, spp=spp, hide_env=hide_env)


def nextrow(spp=4):
    return util.nextrow_scene(width=ZOO_W, height=ZOO_H, spp=spp)


def cornell_small(spp=64):
    return scenes.cornell_box(CORNELL_RES, CORNELL_RES, spp)


def bsdf_inputs(name):
    rng = np.random.RandomState(zlib.crc32(name.encode()))
    n = N_BSDF
    surf = abi.default_surface(n, rng.rand(n, 2))
    wi = util.random_dirs(rng, n)
    wo = util.random_dirs(rng, n)
    u = rng.rand(n, 2).astype(F)
    return surf, wi, wo, u


def eval_directions(reference_sample, wo):
    """eval/pdf are queried at the direction the REFERENCE sampled (so that delta lobes are exercised) where it
    sampled one, at a random direction otherwise."""
    ok = reference_sample[:, 3] > 0
    return np.where(ok[:, None], reference_sample[:, 4:7], wo).astype(F)


def trace_rays():
    rng = np.random.RandomState(3)
    lo, hi = np.array([-9, 0, -4.0]), np.array([9, 9, 6.0])
    a, b = rng.uniform(lo, hi, (N_RAYS, 3)), rng.uniform(lo, hi, (N_RAYS, 3))
    d = b - a
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((N_RAYS, 8), dtype=F)
    rays[:, 0:3], rays[:, 3], rays[:, 4:7], rays[:, 7] = a, 1e-4, d, np.inf
    return rays


def shadow_rays(rays, reference_hits):
    sr = rays.copy()
    t = reference_hits[:, 0]
    sr[:, 7] = np.where(np.isfinite(t), t * 1.5, 10.0)
    return sr


def emitter_inputs():
    rng = np.random.RandomState(5)
    n = N_EMIT
    ref = rng.uniform([-8, 0.1, -3], [8, 6, 5], (n, 3)).astype(F)
    refn = util.random_dirs(rng, n)
    refn[::2] = 0
    return ref, refn, rng.rand(n, 2).astype(F)


def env_dirs():
    return util.random_dirs(np.random.RandomState(6), N_ENV)


def sensor_positions(w=ZOO_W, h=ZOO_H):
    rng = np.random.RandomState(7)
    return (rng.rand(N_SENSOR, 2) * [w, h]).astype(F)


def splat_inputs(w=ZOO_W, h=ZOO_H):
    rng = np.random.RandomState(8)
    pos = (rng.rand(N_SPLAT, 2) * [w + 2, h + 2] - 1).astype(F)   # includes samples outside the film
    val = rng.rand(N_SPLAT, 3).astype(F)
    val[::97] = np.nan                                            # invalid samples are rejected by put()
    val[5::101, 1] = -1.0
    return pos, val


# ---- MIP maps / ray differentials (fixture mip.npz) -----------------------------------------------------------------
MIP_W, MIP_H, MIP_SPP = 96, 64, 4096
N_TEX, N_PARTIALS = 4000, 3000
MIP_SAMPLE_COUNTS = (1, 16, 4096)


def mip(spp=MIP_SPP):
    return util.mip_scene(width=MIP_W, height=MIP_H, spp=spp)


def texture_inputs(name):
    """uv beyond [0,1] (wrap modes) and uv partials from magnification to strong, anisotropic minification."""
    rng = np.random.RandomState(zlib.crc32(("tex/" + name).encode()))
    uv = rng.uniform(-1.5, 2.5, (N_TEX, 2)).astype(F)
    mag = 10 ** rng.uniform(-4, -0.3, (N_TEX, 1))
    duv = (rng.normal(size=(N_TEX, 4)) * mag).astype(F)
    duv[::50, 2:] = 0            # degenerate footprints take the trilinear branch of TMIPMap::eval
    duv[7::50] = 0
    return uv, duv


def mip_bsdf_inputs(name):
    """bsdf_inputs with Intersection::hasUVPartials set (surf slots 2, 3, 7, 11, 15)."""
    surf, wi, wo, u = bsdf_inputs("mip/" + name)
    rng = np.random.RandomState(zlib.crc32(("mipduv/" + name).encode()))
    n = surf.shape[0]
    duv = (rng.normal(size=(n, 4)) * 10 ** rng.uniform(-3.5, -0.5, (n, 1))).astype(F)
    surf[:, 2], surf[:, 3], surf[:, 7], surf[:, 11], surf[:, 15] = duv[:, 0], duv[:, 1], duv[:, 2], duv[:, 3], 1.0
    return surf, wi, wo, u


def partial_positions():
    rng = np.random.RandomState(21)
    return (rng.rand(N_PARTIALS, 2) * [MIP_W, MIP_H]).astype(F)


def scene_dims(sc):
    s = sc.child("sensor")
    f = s.child("film")
    return int(f.get("width")), int(f.get("height")), int(s.child("sampler").get("sampleCount"))


# converged reference renders (small frames, many samples) of the zoo scene and of reduced versions of the
# BASELINE.json configurations
IMAGE_SCENES = {
    "zoo": lambda: zoo(spp=131072)[0],
    "cornell": lambda: scenes.cornell_box(CORNELL_RES, CORNELL_RES, 4096),
    "spheres": lambda: scenes.material_spheres(128, 128, 65536, env_size=(512, 256)),
    "caustics": lambda: scenes.caustic_scene(64, 64, 1048576, knot=(120, 40)),
    "terrain": lambda: scenes.terrain_scene(160, 90, 4096, grid=200, tex_size=256, env_size=(512, 256)),
}


# scenes of the "next" rows; stored in their own fixture so that the files above stay untouched
TWOSIDED_IMAGE_SCENES = {
    "twosided": lambda: twosided(spp=8192)[0],
    "thin_hide": lambda: twosided(spp=4096, hide_env=True)[0],
}


# boundary cases (crop window, sphere transform composition, alpha channel); fixture boundary.npz
NEXTROW_WITH_PARTIALS = ("rp_pair_ewa", "nmap_rp")   # probed on intersections that carry uv partials (mip_bsdf_inputs)
NEXTROW_IMAGE_SCENES = {
    "nextrow": lambda: nextrow(spp=16384)[0],
}
BOUNDARY_IMAGE_SCENES = {
    "crop": lambda: util.cropped_cornell(spp=4096),
    "sphere_alpha": lambda: util.sphere_alpha_scene(spp=4096),
}


# ---- the full-size cfg3 geometry (9,999,392 triangles): closest-hit parity on the BVH that carries the headline number --------
N_FULLSIZE_LIVE, N_FULLSIZE_GOLDEN = 1 << 20, 1 << 17


def fullsize_scene():
    """terrain_scene at BASELINE.json's full triangle count; small textures (the test traces rays, it does not shade)."""
    return scenes.terrain_scene(1920, 1080, 1, grid=2237, tex_size=64, env_size=(64, 32))


def fullsize_rays(n, camera_rays):
    """Half camera rays (`camera_rays(film positions) -> (n, 16)` rows o, d as sample_rays returns them), half random chords
    through the terrain's bounding volume and rays leaving points above the surface (secondary-ray statistics)."""
    rng = np.random.RandomState(77)
    nc = n // 2
    pos = rng.uniform([0, 0], [1920, 1080], (nc, 2)).astype(F)
    cam = camera_rays(pos)
    rays = np.zeros((n, 8), dtype=F)
    rays[:nc, 0:3], rays[:nc, 3], rays[:nc, 4:7], rays[:nc, 7] = cam[:, 0:3], cam[:, 3], cam[:, 4:7], cam[:, 7]
    lo, hi = np.array([-11, -4.0, -11.0]), np.array([11, 6.0, 11.0])
    a, b = rng.uniform(lo, hi, (n - nc, 3)), rng.uniform(lo, hi, (n - nc, 3))
    d = b - a
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays[nc:, 0:3], rays[nc:, 3], rays[nc:, 4:7], rays[nc:, 7] = a, 1e-4, d, np.inf
    return rays


def load(name):
    path = os.path.join(HERE, "golden", name)
    if not os.path.exists(path):
        raise FileNotFoundError("%s is missing: run tests/golden/make_golden.py where /root/reference is mounted" % path)
    return np.load(path)
