"""ctypes binding of the C ABI declared in include/mtscuda.h.

The package binds exactly one library: ``csrc/libmtscuda.so`` (prefix ``mtscuda_``), the hand-written sm_100a
kernels.  There is no CPU path and no fallback: when the library has not been built, or no GPU is present at call
time, the call raises.  ``bind()`` types the entry points of *a* library exporting the header's surface under a given
prefix; the test infrastructure (oracle/oracle_lib.py) uses it to put the CPU restatement behind the same wrappers.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)

FILTER_RESOLUTION = 31
EPSILON = 1e-4

# ---- enums (mirror include/mtscuda.h) --------------------------------------
(BSDF_DIFFUSE, BSDF_CONDUCTOR, BSDF_ROUGHCONDUCTOR, BSDF_PLASTIC, BSDF_ROUGHPLASTIC,
    BSDF_DIELECTRIC, BSDF_COATING, BSDF_BLEND, BSDF_BUMPMAP, BSDF_TWOSIDED, BSDF_THINDIELECTRIC, BSDF_MASK, BSDF_ROUGHDIELECTRIC,
    BSDF_MIXTURE, BSDF_ROUGHCOATING, BSDF_NORMALMAP) = range(16)
MIXTURE_MAX = 16
NOALLOC_MAX_RAYS = 16384   # MTS_NOALLOC_MAX_RAYS
BF_GGX, BF_SAMPLE_VISIBLE, BF_NONLINEAR, BF_RT_2D, BF_RAY_DIFF = 1, 2, 4, 8, 16
ENULL, EDIFFUSE_REFLECTION, EDIFFUSE_TRANSMISSION, EGLOSSY_REFLECTION, EGLOSSY_TRANSMISSION, \
    EDELTA_REFLECTION, EDELTA_TRANSMISSION = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40
EFRONTSIDE = 0x08000
EBACKSIDE = 0x10000
ESMOOTH = EDIFFUSE_REFLECTION | EDIFFUSE_TRANSMISSION | EGLOSSY_REFLECTION | EGLOSSY_TRANSMISSION
EDELTA = ENULL | EDELTA_REFLECTION | EDELTA_TRANSMISSION
ETRANSMISSION = EDIFFUSE_TRANSMISSION | EGLOSSY_TRANSMISSION | EDELTA_TRANSMISSION | ENULL
MEASURE_SOLID_ANGLE, MEASURE_DISCRETE = 1, 4
WRAP = {"repeat": 0, "mirror": 1, "clamp": 2, "zero": 3, "one": 4}
FILTER = {"nearest": 0, "bilinear": 1, "trilinear": 2, "ewa": 3}
EMITTER_AREA, EMITTER_POINT, EMITTER_ENVMAP = 0, 1, 2


class TexRef(C.Structure):
    _fields_ = [("scale", C.c_float * 3), ("tex", C.c_int32)]


class BsdfNode(C.Structure):
    _fields_ = [("type", C.c_int32), ("flags", C.c_int32), ("child", C.c_int32 * 2),
                ("t0", TexRef), ("t1", TexRef), ("t2", TexRef),
                ("eta", C.c_float * 3), ("k", C.c_float * 3), ("f", C.c_float * 4),
                ("lut", C.c_int32 * 4), ("lobes", C.c_uint32), ("pad0", C.c_uint32)]


class TextureDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32),
                ("wrap_u", C.c_int32), ("wrap_v", C.c_int32), ("filter", C.c_int32),
                ("uv_scale", C.c_float * 2), ("uv_offset", C.c_float * 2), ("max_anisotropy", C.c_float), ("texel_format", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("sample_to_camera", C.c_float * 16), ("camera_to_world", C.c_float * 16),
                ("dx", C.c_float * 3), ("dy", C.c_float * 3),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("inv_resolution", C.c_float * 2)]


class Film(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("filter_radius", C.c_float),
                ("filter_table", C.c_float * (FILTER_RESOLUTION + 1)), ("has_alpha", C.c_int32)]


class MeshDesc(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("normals", C.c_void_p), ("texcoords", C.c_void_p),
                ("indices", C.c_void_p), ("tangents", C.c_void_p),
                ("nverts", C.c_uint32), ("ntris", C.c_uint32), ("bsdf", C.c_int32), ("uv_tangents", C.c_int32)]


class EmitterDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("shape", C.c_int32), ("radiance", C.c_float * 3),
                ("position", C.c_float * 3), ("sampling_weight", C.c_float), ("scale", C.c_float),
                ("to_world", C.c_float * 16), ("env_width", C.c_int32), ("env_height", C.c_int32),
                ("env_texels", C.c_void_p)]


class RenderParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("sample_offset", C.c_int32), ("sample_stride", C.c_int32),
                ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("strict_normals", C.c_int32),
                ("hide_emitters", C.c_int32), ("seed", C.c_uint32), ("clear_film", C.c_int32),
                ("wavefront_size", C.c_int32), ("instrument", C.c_int32), ("sample_count", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("camera_samples", C.c_uint64), ("normal_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("path_length_sum", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("bvh_nodes_visited", C.c_uint64), ("bvh_prims_tested", C.c_uint64),
                ("bvh_nodes_visited_shadow", C.c_uint64), ("bvh_prims_tested_shadow", C.c_uint64),
                ("intersect_launches", C.c_uint64), ("shade_launches", C.c_uint64),
                ("shadow_launches", C.c_uint64), ("splat_launches", C.c_uint64),
                ("render_ms", C.c_float), ("intersect_ms", C.c_float), ("shade_ms", C.c_float),
                ("shadow_ms", C.c_float), ("splat_ms", C.c_float), ("raygen_ms", C.c_float),
                ("build_ms", C.c_float), ("upload_ms", C.c_float),
                ("sort_ms", C.c_float), ("miss_ms", C.c_float),
                ("shaded_paths", C.c_uint64), ("miss_paths", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# every symbol include/mtscuda.h declares (tests check that the libraries export all of them)
ABI_VERSION = 6   # MTS_ABI_VERSION of include/mtscuda.h these mirrors were written against
ABI_SYMBOLS = [
    "abi_version", "last_error", "create", "create_multi", "device_count", "destroy", "set_camera", "set_film", "add_texture", "set_float_pool",
    "set_bsdfs", "add_mesh", "add_rectangle", "add_sphere", "add_emitter", "build", "render",
    "read_film", "read_film_alpha", "bind_film", "set_stream", "get_stats", "sample_rays", "trace", "trace_noalloc", "hit_records", "bsdf_eval_pdf",
    "bsdf_sample", "sample_emitter_direct", "eval_environment", "texture_eval", "hit_partials", "eval_environment_camera", "splat",
]

# MTS_CUDA_LIB selects an experimental build of the same library (kernel tuning sweeps); never a different back end
CUDA_LIB_PATH = os.environ.get("MTS_CUDA_LIB") or os.path.join(_HERE, "csrc", "libmtscuda.so")
CUDA_PREFIX = "mtscuda_"


class MtsError(RuntimeError):
    """Raised when a C-ABI call fails; mirrors Log(EError, ...) throwing std::runtime_error
    in the reference (src/libcore/logger.cpp:147)."""


_libs = {}


def bind(path, prefix):
    """dlopen a library that exports the surface of include/mtscuda.h under `prefix` and type its entry points."""
    key = (path, prefix)
    if key in _libs:
        return _libs[key]
    if not os.path.exists(path):
        raise MtsError("%s is not built: run `python -c 'import __graft_entry__ as g; g.build()'`" % path)
    lib = C.CDLL(path)
    fn = {}
    vp, i32, i64, fp = C.c_void_p, C.c_int32, C.c_int64, C.c_void_p
    sigs = {
        "abi_version": (C.c_int32, []),
        "create_multi": (C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.POINTER(vp)]),
        "device_count": (C.c_int, [vp]),
        "last_error": (C.c_char_p, []),
        "create": (C.c_int, [C.c_int, C.POINTER(vp)]),
        "destroy": (None, [vp]),
        "set_camera": (C.c_int, [vp, C.POINTER(Camera)]),
        "set_film": (C.c_int, [vp, C.POINTER(Film)]),
        "add_texture": (C.c_int, [vp, C.POINTER(TextureDesc), fp]),
        "set_float_pool": (C.c_int, [vp, fp, i64]),
        "set_bsdfs": (C.c_int, [vp, C.POINTER(BsdfNode), i32]),
        "add_mesh": (C.c_int, [vp, C.POINTER(MeshDesc)]),
        "add_rectangle": (C.c_int, [vp, fp, i32]),
        "add_sphere": (C.c_int, [vp, fp, C.c_float, i32, i32]),
        "add_emitter": (C.c_int, [vp, C.POINTER(EmitterDesc)]),
        "build": (C.c_int, [vp]),
        "render": (C.c_int, [vp, C.POINTER(RenderParams), C.POINTER(C.c_int), C.POINTER(Stats)]),
        "read_film": (C.c_int, [vp, fp]),
        "read_film_alpha": (C.c_int, [vp, fp]),
        "bind_film": (C.c_int, [vp, vp]),
        "set_stream": (C.c_int, [vp, vp]),
        "get_stats": (C.c_int, [vp, C.POINTER(Stats)]),
        "sample_rays": (C.c_int, [vp, i64, fp, fp]),
        "trace": (C.c_int, [vp, i64, fp, i32, fp, fp]),
        "trace_noalloc": (C.c_int, [vp, i64, fp, i32, fp, fp]),
        "hit_records": (C.c_int, [vp, i64, fp, fp, fp]),
        "bsdf_eval_pdf": (C.c_int, [vp, i32, i64, fp, fp, fp, i32, fp, fp]),
        "bsdf_sample": (C.c_int, [vp, i32, i64, fp, fp, fp, fp]),
        "sample_emitter_direct": (C.c_int, [vp, i64, fp, fp, fp, fp]),
        "eval_environment": (C.c_int, [vp, i64, fp, fp]),
        "texture_eval": (C.c_int, [vp, i32, i64, fp, fp, fp]),
        "hit_partials": (C.c_int, [vp, i64, fp, i32, fp]),
        "eval_environment_camera": (C.c_int, [vp, i64, fp, i32, fp]),
        "splat": (C.c_int, [vp, i64, fp, fp, i32]),
    }
    for name in ABI_SYMBOLS:
        f = getattr(lib, prefix + name)  # AttributeError if a declared symbol is not exported
        f.restype, f.argtypes = sigs[name]
        fn[name] = f
    lib._fn = fn
    if fn["abi_version"]() != ABI_VERSION:
        raise MtsError("%s implements ABI version %d, this binding was written for version %d -- rebuild the library"
                       % (path, fn["abi_version"](), ABI_VERSION))
    lib._device_tangents = prefix == CUDA_PREFIX   # libmtscuda runs TriMesh::computeUVTangents on the device
    _libs[key] = lib
    return lib


def load():
    """The product library.  Fails loudly when it is missing."""
    return bind(CUDA_LIB_PATH, CUDA_PREFIX)


class DeviceArray:
    """A buffer that already lives in device memory (for instance the result of an NVLink all-gather of a scene that the
    ranks of a job uploaded in slices): the library copies it device to device (cudaMemcpyDefault / UVA) exactly like a
    host buffer.  `owner` keeps the allocation alive."""

    def __init__(self, data_ptr, shape, dtype, owner=None):
        self.data_ptr, self.shape, self.dtype, self.owner = int(data_ptr), tuple(int(x) for x in shape), np.dtype(dtype), owner

    @property
    def size(self):
        return int(np.prod(self.shape))

    @property
    def nbytes(self):
        return self.size * self.dtype.itemsize

    def host_view(self):
        """numpy view of a buffer that lives in host memory after all (a CPU torch tensor as `owner`: the gloo tests)."""
        if getattr(self.owner, "is_cuda", True):
            raise TypeError("DeviceArray.host_view: the buffer is device memory")
        return self.owner.numpy().reshape(-1)[:self.size].view(self.dtype).reshape(self.shape)

    def reshape(self, shape):
        shape = tuple(shape) if isinstance(shape, (tuple, list)) else (shape,)
        if -1 in shape:
            known = int(np.prod([x for x in shape if x != -1]))
            shape = tuple(self.size // known if x == -1 else x for x in shape)
        assert int(np.prod(shape)) == self.size
        return DeviceArray(self.data_ptr, shape, self.dtype, self.owner)


def f32(a, shape=None):
    if isinstance(a, DeviceArray):
        assert a.dtype == np.float32
        return a.reshape(shape) if shape is not None else a
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def ptr(a):
    if isinstance(a, DeviceArray):
        return C.c_void_p(a.data_ptr)
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    """Owns one MtsContext handle; thin, typed wrappers over the C entry points."""

    def __init__(self, device=0, lib=None, devices=None):
        """device: one GPU.  devices: a list of GPUs of this box rendering together (mtscuda_create_multi; devices[0]
        is the primary) -- cudapath's `devices` parameter."""
        self.lib = lib if lib is not None else load()
        self.fn = self.lib._fn
        h = C.c_void_p()
        if devices is not None and len(devices) > 0:
            arr = (C.c_int32 * len(devices))(*[int(d) for d in devices])
            self._check(self.fn["create_multi"](arr, len(devices), C.byref(h)))
        else:
            self._check(self.fn["create"](device, C.byref(h)))
        self.h = h
        self._keep = []

    def _check(self, rc):
        if rc < 0:
            raise MtsError(self.fn["last_error"]().decode("utf-8", "replace") or ("error %d" % rc))
        return rc

    def device_count(self):
        return int(self.fn["device_count"](self.h))

    def close(self):
        if getattr(self, "h", None):
            self.fn["destroy"](self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene description ----
    def set_camera(self, cam):
        self._check(self.fn["set_camera"](self.h, C.byref(cam)))

    def set_film(self, film):
        self.film_size = (film.width, film.height)
        self._check(self.fn["set_film"](self.h, C.byref(film)))

    def add_texture(self, desc, texels):
        if desc.texel_format == 1:      # MTS_TEXEL_F16: an object holding a contiguous float16 array `.a`
            return self._check(self.fn["add_texture"](self.h, C.byref(desc), texels.a.ctypes.data_as(C.c_void_p)))
        t = f32(texels)
        return self._check(self.fn["add_texture"](self.h, C.byref(desc), ptr(t)))

    def set_float_pool(self, data):
        d = f32(data)
        self._check(self.fn["set_float_pool"](self.h, ptr(d), d.size))

    def set_bsdfs(self, nodes):
        arr = (BsdfNode * len(nodes))(*nodes)
        self._check(self.fn["set_bsdfs"](self.h, arr, len(nodes)))

    def add_mesh(self, positions, indices, normals=None, texcoords=None, tangents=None, bsdf=0, uv_tangents=False):
        """uv_tangents=True asks for TriMesh::computeUVTangents: on the device for libmtscuda (flag in the descriptor),
        on the host (hostmath.uv_tangents, the numpy mirror) for a library that takes explicit tangents only."""
        pos = f32(positions, (-1, 3))
        idx = indices.reshape((-1, 3)) if isinstance(indices, DeviceArray) else np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1, 3)
        nrm = None if normals is None else f32(normals, (-1, 3))
        uv = None if texcoords is None else f32(texcoords, (-1, 2))
        flag = 0
        if uv_tangents and uv is not None and tangents is None:
            if self.lib._device_tangents:
                flag = 1
            else:
                from . import hostmath
                hv = lambda a: a.host_view() if isinstance(a, DeviceArray) else a
                tangents = hostmath.uv_tangents(hv(pos), hv(uv), hv(idx))
        tg = None if tangents is None else f32(tangents, (-1, 6))
        d = MeshDesc(ptr(pos), ptr(nrm), ptr(uv), ptr(idx), ptr(tg), pos.shape[0], idx.shape[0], bsdf, flag)
        return self._check(self.fn["add_mesh"](self.h, C.byref(d)))

    def add_rectangle(self, to_world, bsdf):
        m = f32(to_world, (16,))
        return self._check(self.fn["add_rectangle"](self.h, ptr(m), bsdf))

    def add_sphere(self, object_to_world, radius, flip, bsdf):
        m = f32(object_to_world, (16,))
        return self._check(self.fn["add_sphere"](self.h, ptr(m), float(radius), int(bool(flip)), bsdf))

    def add_emitter(self, desc, texels=None):
        if texels is not None:
            t = f32(texels)
            desc.env_texels = t.data_ptr if isinstance(t, DeviceArray) else t.ctypes.data
        return self._check(self.fn["add_emitter"](self.h, C.byref(desc)))

    def build(self):
        self._check(self.fn["build"](self.h))

    # ---- hot path ----
    def render(self, params, cancel=None):
        st = Stats()
        self._check(self.fn["render"](self.h, C.byref(params), cancel, C.byref(st)))
        return st

    def read_film(self):
        w, h = self.film_size
        out = np.empty((h, w, 4), dtype=np.float32)
        self._check(self.fn["read_film"](self.h, ptr(out)))
        return out

    def read_film_alpha(self):
        w, h = self.film_size
        out = np.empty((h, w), dtype=np.float32)
        self._check(self.fn["read_film_alpha"](self.h, ptr(out)))
        return out

    def bind_film(self, device_ptr):
        self._check(self.fn["bind_film"](self.h, C.c_void_p(device_ptr)))

    def set_stream(self, cuda_stream):
        self._check(self.fn["set_stream"](self.h, C.c_void_p(cuda_stream)))

    def get_stats(self):
        st = Stats()
        self._check(self.fn["get_stats"](self.h, C.byref(st)))
        return st

    # ---- per-function entry points ----
    def sample_rays(self, sample_pos):
        sp = f32(sample_pos, (-1, 2))
        out = np.empty((sp.shape[0], 16), dtype=np.float32)
        self._check(self.fn["sample_rays"](self.h, sp.shape[0], ptr(sp), ptr(out)))
        return out

    def trace(self, rays, shadow=False):
        r = f32(rays, (-1, 8))
        hits = np.empty((r.shape[0], 4), dtype=np.float32)
        ids = np.empty((r.shape[0], 2), dtype=np.int32)
        self._check(self.fn["trace"](self.h, r.shape[0], ptr(r), int(shadow), ptr(hits), ptr(ids)))
        return hits, ids

    def trace_noalloc(self, rays, shadow=False):
        """mtscuda_trace_noalloc: trace() in the library's fixed device scratch (no allocation; at most NOALLOC_MAX_RAYS rays)."""
        r = f32(rays, (-1, 8))
        hits = np.empty((r.shape[0], 4), dtype=np.float32)
        ids = np.empty((r.shape[0], 2), dtype=np.int32)
        self._check(self.fn["trace_noalloc"](self.h, r.shape[0], ptr(r), int(shadow), ptr(hits), ptr(ids)))
        return hits, ids

    def hit_records(self, rays):
        r = f32(rays, (-1, 8))
        out = np.empty((r.shape[0], 24), dtype=np.float32)
        ids = np.empty((r.shape[0], 2), dtype=np.int32)
        self._check(self.fn["hit_records"](self.h, r.shape[0], ptr(r), ptr(out), ptr(ids)))
        return out, ids

    def bsdf_eval_pdf(self, bsdf, surf, wi, wo, measure=MEASURE_SOLID_ANGLE):
        s, a, b = f32(surf, (-1, 16)), f32(wi, (-1, 3)), f32(wo, (-1, 3))
        n = a.shape[0]
        val = np.empty((n, 3), dtype=np.float32)
        pdf = np.empty((n,), dtype=np.float32)
        self._check(self.fn["bsdf_eval_pdf"](self.h, bsdf, n, ptr(s), ptr(a), ptr(b), measure, ptr(val), ptr(pdf)))
        return val, pdf

    def bsdf_sample(self, bsdf, surf, wi, u):
        s, a, uu = f32(surf, (-1, 16)), f32(wi, (-1, 3)), f32(u, (-1, 2))
        n = a.shape[0]
        out = np.empty((n, 12), dtype=np.float32)
        self._check(self.fn["bsdf_sample"](self.h, bsdf, n, ptr(s), ptr(a), ptr(uu), ptr(out)))
        return out

    def sample_emitter_direct(self, ref, ref_n, u):
        r, rn, uu = f32(ref, (-1, 3)), f32(ref_n, (-1, 3)), f32(u, (-1, 2))
        out = np.empty((r.shape[0], 16), dtype=np.float32)
        self._check(self.fn["sample_emitter_direct"](self.h, r.shape[0], ptr(r), ptr(rn), ptr(uu), ptr(out)))
        return out

    def eval_environment(self, d):
        dd = f32(d, (-1, 3))
        out = np.empty((dd.shape[0], 4), dtype=np.float32)
        self._check(self.fn["eval_environment"](self.h, dd.shape[0], ptr(dd), ptr(out)))
        return out

    def texture_eval(self, tex, uv, duv=None):
        u = f32(uv, (-1, 2))
        d = None if duv is None else f32(duv, (-1, 4))
        out = np.empty((u.shape[0], 3), dtype=np.float32)
        self._check(self.fn["texture_eval"](self.h, int(tex), u.shape[0], ptr(u), ptr(d), ptr(out)))
        return out

    def hit_partials(self, sample_pos, sample_count):
        sp = f32(sample_pos, (-1, 2))
        out = np.empty((sp.shape[0], 8), dtype=np.float32)
        self._check(self.fn["hit_partials"](self.h, sp.shape[0], ptr(sp), int(sample_count), ptr(out)))
        return out

    def eval_environment_camera(self, sample_pos, sample_count):
        sp = f32(sample_pos, (-1, 2))
        out = np.empty((sp.shape[0], 3), dtype=np.float32)
        self._check(self.fn["eval_environment_camera"](self.h, sp.shape[0], ptr(sp), int(sample_count), ptr(out)))
        return out

    def splat(self, pos, value, clear_first=True):
        p, v = f32(pos, (-1, 2)), f32(value, (-1, 3))
        self._check(self.fn["splat"](self.h, p.shape[0], ptr(p), ptr(v), int(clear_first)))
        return self.read_film()


def default_surface(n, uv=None):
    """n surface contexts for bsdf_* queries: uv, dpdu=(1,0,0), dpdv=(0,1,0), n=(0,0,1)."""
    s = np.zeros((n, 16), dtype=np.float32)
    if uv is not None:
        s[:, 0:2] = uv
    s[:, 4] = 1.0
    s[:, 9] = 1.0
    s[:, 14] = 1.0
    return s
