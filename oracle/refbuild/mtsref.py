"""ctypes front end of oracle/_ref/libmtsref.so -- the REFERENCE's own code (compiled unchanged from
/root/reference by oracle/refbuild/Makefile), driven from the same Plugin trees the CUDA library renders.

TEST INFRASTRUCTURE ONLY (checker for the oracle restatement, golden-vector generator, timed CPU baseline).
`RefScene(plugin_tree)` performs what src/librender/scenehandler.cpp does for an XML file: Properties ->
PluginManager::createObject -> addChild -> configure, children first.  In-memory meshes / bitmaps are written to a
scratch directory as `.serialized` / `.pfm` files, which the reference's own loaders then read.
"""
import ctypes as C
import os
import re
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(os.path.dirname(_HERE), "_ref")
LIB = os.path.join(REF_DIR, "libmtsref.so")

_lib = None


class RefError(RuntimeError):
    pass


def available():
    return os.path.exists(LIB)


def load(threads=0, log_level=300):   # ELogLevel: 0 trace, 100 debug, 200 info, 300 warn, 400 error
    global _lib
    if _lib is not None:
        return _lib
    if not available():
        raise RefError("oracle/_ref is not built (make -C oracle/refbuild; needs /root/reference)")
    lib = C.CDLL(LIB, mode=C.RTLD_GLOBAL)   # plugins resolve libmitsuba-* symbols through the global scope
    vp, f, i32, i64, cp = C.c_void_p, C.c_float, C.c_int32, C.c_int64, C.c_char_p
    sig = {
        "mtsref_last_error": (cp, []), "mtsref_init": (C.c_int, [cp, C.c_int, C.c_int]), "mtsref_thread_count": (C.c_int, []), "mtsref_shutdown": (None, []),
        "mtsref_props_new": (vp, [cp]), "mtsref_props_free": (None, [vp]),
        "mtsref_props_set_bool": (None, [vp, cp, C.c_int]), "mtsref_props_set_int": (None, [vp, cp, C.c_int]),
        "mtsref_props_set_float": (None, [vp, cp, f]), "mtsref_props_set_string": (None, [vp, cp, cp]),
        "mtsref_props_set_point": (None, [vp, cp, f, f, f]), "mtsref_props_set_vector": (None, [vp, cp, f, f, f]),
        "mtsref_props_set_rgb": (None, [vp, cp, f, f, f]), "mtsref_props_set_transform": (None, [vp, cp, vp]),
        "mtsref_create": (vp, [cp, vp]), "mtsref_add_child": (C.c_int, [vp, cp, vp]), "mtsref_configure": (C.c_int, [vp]),
        "mtsref_release": (None, [vp]), "mtsref_scene_initialize": (C.c_int, [vp, C.POINTER(C.c_double)]),
        "mtsref_render": (C.c_int, [vp, vp, C.POINTER(C.c_double), cp, vp]), "mtsref_stats": (C.c_int, [cp, C.c_int]),
        "mtsref_bsdf_eval_pdf": (C.c_int, [vp, i64, vp, vp, vp, C.c_int, vp, vp]),
        "mtsref_bsdf_sample": (C.c_int, [vp, i64, vp, vp, vp, vp]), "mtsref_bsdf_type": (C.c_int, [vp]),
        "mtsref_trace": (C.c_int, [vp, i64, vp, C.c_int, vp, vp]), "mtsref_hit_records": (C.c_int, [vp, i64, vp, vp, vp]),
        "mtsref_sample_emitter_direct": (C.c_int, [vp, i64, vp, vp, vp, vp]),
        "mtsref_eval_environment": (C.c_int, [vp, i64, vp, vp]), "mtsref_sample_rays": (C.c_int, [vp, i64, vp, vp]),
        "mtsref_splat": (C.c_int, [vp, i64, vp, vp, vp]),
        "mtsref_texture_eval": (C.c_int, [vp, i64, vp, vp, vp]), "mtsref_hit_partials": (C.c_int, [vp, i64, vp, C.c_int, vp]),
        "mtsref_eval_environment_camera": (C.c_int, [vp, i64, vp, C.c_int, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.mtsref_init(REF_DIR.encode(), int(threads), int(log_level)) != 0:
        raise RefError(lib.mtsref_last_error().decode())
    _lib = lib
    import atexit
    atexit.register(lib.mtsref_shutdown)   # joins the scheduler's worker threads; the process would not exit otherwise
    return lib


def _chk(rc):
    if rc != 0:
        raise RefError(_lib.mtsref_last_error().decode())


def _f32(a, shape):
    return np.ascontiguousarray(a, dtype=np.float32).reshape(shape)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class RefObject:
    def __init__(self, handle, plugin):
        self.h, self.plugin = handle, plugin


class RefScene:
    """The reference object graph of one Plugin tree."""

    def __init__(self, scene, threads=0, scratch=None, integrator=None):
        self.integrator = integrator   # e.g. "cudapath": render with the drop-in plugin instead of the scene's `path`
        from mitsuba0_6_b200 import scene as ms   # Plugin / Spectrum / Point, asset writers
        self.ms = ms
        self.lib = load(threads)
        self.tmp = scratch or tempfile.mkdtemp(prefix="mtsref_")
        self.n_assets = 0
        self.objects = {}          # id(plugin) -> RefObject (shared instances are created once, like <ref id=...>)
        self.bsdfs = {}
        self.kd_build_s = 0.0
        self.scene = self._build(scene)
        s = C.c_double(0)
        _chk(self.lib.mtsref_scene_initialize(self.scene.h, C.byref(s)))
        self.kd_build_s = s.value

    # -- object construction --------------------------------------------------------------------
    def _asset(self, ext):
        self.n_assets += 1
        return os.path.join(self.tmp, "asset%d.%s" % (self.n_assets, ext))

    def _build(self, p):
        if id(p) in self.objects:
            return self.objects[id(p)]
        ms, lib = self.ms, self.lib
        ptype, props = p.type, dict(p.props)
        if p.kind == "shape" and p.type == "mesh":
            from mitsuba0_6_b200 import hostmath as hm
            fn = self._asset("serialized")
            d = p.data
            pos, nrm = np.asarray(d["positions"], dtype=np.float64), d.get("normals")
            tw = props.pop("toWorld", None)
            if tw is not None:
                tw = np.asarray(tw, dtype=np.float64)
                pos = (np.c_[pos, np.ones(len(pos))] @ tw.T)[:, :3]
                if nrm is not None:
                    nrm = np.asarray(nrm, dtype=np.float64) @ np.linalg.inv(tw)[:3, :3]
                    nrm = nrm / np.linalg.norm(nrm, axis=1, keepdims=True)
            ms.write_serialized(fn, pos.astype(np.float32), d["indices"], None if nrm is None else np.asarray(nrm, dtype=np.float32),
                                d.get("texcoords"))
            ptype = "serialized"
            props["filename"] = fn
        if p.kind in ("texture", "emitter") and "bitmap" in p.data:
            fn = self._asset("pfm")
            ms.write_pfm(fn, p.data["bitmap"])
            props["filename"] = fn
        if p.kind == "film":
            # OpenEXR is not compiled into the oracle build: films are written as PFM.  PFM forces pixelFormat=rgb
            # (hdrfilm.cpp:326-334), so films that ask for an alpha channel keep the default `openexr`: writing the file
            # then fails inside RenderJob (logged, harmless) after the film storage -- which is what the tests read --
            # has been filled.
            if "a" not in str(props.get("pixelFormat", "rgb")).lower().replace("luminance", "").replace("spectrum", "") \
                    and "alpha" not in str(props.get("pixelFormat", "rgb")).lower():
                props.setdefault("fileFormat", "pfm")
            props.setdefault("banner", False)
            props.setdefault("componentFormat", "float32")
        if p.kind == "integrator":
            ptype = self.integrator or ("path" if ptype == "cudapath" else ptype)
        pr = lib.mtsref_props_new((ptype or "scene").encode())
        for k, v in props.items():
            kb = k.encode()
            if isinstance(v, bool):
                lib.mtsref_props_set_bool(pr, kb, int(v))
            elif isinstance(v, (int, np.integer)):
                lib.mtsref_props_set_int(pr, kb, int(v))
            elif isinstance(v, (float, np.floating)):
                lib.mtsref_props_set_float(pr, kb, float(v))
            elif isinstance(v, str):
                lib.mtsref_props_set_string(pr, kb, v.encode())
            elif isinstance(v, ms.Spectrum):
                lib.mtsref_props_set_rgb(pr, kb, *[float(x) for x in v])
            elif isinstance(v, ms.Point):
                lib.mtsref_props_set_point(pr, kb, *[float(x) for x in v])
            elif isinstance(v, np.ndarray) and v.shape == (4, 4):
                m = _f32(v, (16,))
                lib.mtsref_props_set_transform(pr, kb, _p(m))
            else:
                raise RefError("cannot pass property %s=%r to the reference" % (k, v))
        children = [(n, self._build(c)) for n, c in p.children]
        h = lib.mtsref_create(p.kind.encode(), pr)
        lib.mtsref_props_free(pr)
        if not h:
            raise RefError(lib.mtsref_last_error().decode())
        for n, c in children:
            _chk(lib.mtsref_add_child(h, n.encode(), c.h))
        _chk(lib.mtsref_configure(h))
        o = RefObject(h, p)
        self.objects[id(p)] = o
        if p.kind == "bsdf":
            self.bsdfs[id(p)] = o
        return o

    def bsdf(self, plugin):
        return self.objects[id(plugin)]

    # -- render ---------------------------------------------------------------------------------
    def render(self, alpha=False):
        """The developed film of the crop window as (H, W, 4) = sum w*RGB, sum w; with alpha=True also the alpha plane
        (sum w*alpha, meaningful for films with an alpha pixel format)."""
        film = self.scene.plugin.child("sensor").child("film")
        w, h = int(film.get("width", 768)), int(film.get("height", 576))
        w, h = int(film.get("cropWidth", w)), int(film.get("cropHeight", h))
        out = np.zeros((h, w, 4), dtype=np.float32)
        a = np.zeros((h, w), dtype=np.float32)
        s = C.c_double(0)
        _chk(self.lib.mtsref_render(self.scene.h, _p(out), C.byref(s), os.path.join(self.tmp, "out").encode(), _p(a)))
        self.render_s = s.value
        self.stats = self.read_stats()
        return (out, a) if alpha else out

    def read_stats(self):
        buf = C.create_string_buffer(1 << 16)
        self.lib.mtsref_stats(buf, len(buf))
        txt = buf.value.decode("utf-8", "replace")
        out = {"text": txt}
        units = {"K": 1e3, "M": 1e6, "G": 1e9, "T": 1e12}
        for key, pat in (("normal_rays", r"Normal rays traced\s*:\s*([0-9.]+)\s*([KMGT]?)"),
                         ("shadow_rays", r"Shadow rays traced\s*:\s*([0-9.]+)\s*([KMGT]?)"),
                         ("avg_path_length", r"Average path length\s*:\s*([0-9.]+)")):
            m = re.search(pat, txt)
            if m:
                out[key] = float(m.group(1)) * (units.get(m.group(2), 1.0) if m.lastindex and m.lastindex > 1 else 1.0)
        return out

    # -- probes (same array layouts as mitsuba0_6_b200.abi.Context) ---------------------------------
    def bsdf_eval_pdf(self, bsdf, surf, wi, wo, measure=1):
        s, a, b = _f32(surf, (-1, 16)), _f32(wi, (-1, 3)), _f32(wo, (-1, 3))
        n = a.shape[0]
        val, pdf = np.empty((n, 3), np.float32), np.empty((n,), np.float32)
        _chk(self.lib.mtsref_bsdf_eval_pdf(bsdf.h, n, _p(s), _p(a), _p(b), measure, _p(val), _p(pdf)))
        return val, pdf

    def bsdf_sample(self, bsdf, surf, wi, u):
        s, a, uu = _f32(surf, (-1, 16)), _f32(wi, (-1, 3)), _f32(u, (-1, 2))
        out = np.empty((a.shape[0], 12), np.float32)
        _chk(self.lib.mtsref_bsdf_sample(bsdf.h, a.shape[0], _p(s), _p(a), _p(uu), _p(out)))
        return out

    def bsdf_type(self, bsdf):
        return self.lib.mtsref_bsdf_type(bsdf.h)

    def trace(self, rays, shadow=False):
        r = _f32(rays, (-1, 8))
        hits, ids = np.empty((r.shape[0], 4), np.float32), np.empty((r.shape[0], 2), np.int32)
        _chk(self.lib.mtsref_trace(self.scene.h, r.shape[0], _p(r), int(shadow), _p(hits), _p(ids)))
        return hits, ids

    def hit_records(self, rays):
        r = _f32(rays, (-1, 8))
        out, ids = np.empty((r.shape[0], 24), np.float32), np.empty((r.shape[0], 2), np.int32)
        _chk(self.lib.mtsref_hit_records(self.scene.h, r.shape[0], _p(r), _p(out), _p(ids)))
        return out, ids

    def sample_emitter_direct(self, ref, ref_n, u):
        r, rn, uu = _f32(ref, (-1, 3)), _f32(ref_n, (-1, 3)), _f32(u, (-1, 2))
        out = np.empty((r.shape[0], 16), np.float32)
        _chk(self.lib.mtsref_sample_emitter_direct(self.scene.h, r.shape[0], _p(r), _p(rn), _p(uu), _p(out)))
        return out

    def eval_environment(self, d):
        dd = _f32(d, (-1, 3))
        out = np.empty((dd.shape[0], 4), np.float32)
        _chk(self.lib.mtsref_eval_environment(self.scene.h, dd.shape[0], _p(dd), _p(out)))
        return out

    def texture_eval(self, texture_plugin, uv, duv=None):
        """Texture::eval(its, filter=True) of a texture plugin of this scene; duv = dudx dvdx dudy dvdy or None."""
        t = self.objects[id(texture_plugin)]
        u = _f32(uv, (-1, 2))
        d = None if duv is None else _f32(duv, (-1, 4))
        out = np.empty((u.shape[0], 3), np.float32)
        _chk(self.lib.mtsref_texture_eval(t.h, u.shape[0], _p(u), None if d is None else _p(d), _p(out)))
        return out

    def hit_partials(self, pos, sample_count):
        sp = _f32(pos, (-1, 2))
        out = np.empty((sp.shape[0], 8), np.float32)
        _chk(self.lib.mtsref_hit_partials(self.scene.h, sp.shape[0], _p(sp), int(sample_count), _p(out)))
        return out

    def eval_environment_camera(self, pos, sample_count):
        sp = _f32(pos, (-1, 2))
        out = np.empty((sp.shape[0], 3), np.float32)
        _chk(self.lib.mtsref_eval_environment_camera(self.scene.h, sp.shape[0], _p(sp), int(sample_count), _p(out)))
        return out

    def sample_rays(self, pos):
        sp = _f32(pos, (-1, 2))
        out = np.empty((sp.shape[0], 16), np.float32)
        _chk(self.lib.mtsref_sample_rays(self.scene.h, sp.shape[0], _p(sp), _p(out)))
        return out

    def splat(self, pos, value):
        p, v = _f32(pos, (-1, 2)), _f32(value, (-1, 3))
        film = self.scene.plugin.child("sensor").child("film")
        w, h = int(film.get("width", 768)), int(film.get("height", 576))
        out = np.zeros((h, w, 4), np.float32)
        _chk(self.lib.mtsref_splat(self.scene.h, p.shape[0], _p(p), _p(v), _p(out)))
        return out
