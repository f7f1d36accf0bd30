"""Scene description mirroring Mitsuba's plugin tree, an XML-subset loader/writer, and the
flattening step of `cudapath::render()` (walk shapes / BSDF trees / emitters /
sensor / film, recompute the derived constants the reference computes in configure(), emit the
plain arrays the C ABI takes).

The same `Plugin` tree is produced by `load_xml` (a Mitsuba scene file) and by the synthetic
scene generators in scenes.py, and can be written back as Mitsuba XML (+ .serialized / .pfm
assets) so that the reference build renders identical inputs.
"""
import math
import os
import struct
import xml.etree.ElementTree as ET
import zlib

import numpy as np

from . import abi, hostmath as hm

F = np.float32


# ---------------------------------------------------------------------------
# plugin tree
# ---------------------------------------------------------------------------
class Spectrum(tuple):
    """An RGB triple (the reference build has SPECTRUM_SAMPLES=3)."""

    def __new__(cls, *v):
        if len(v) == 1 and not isinstance(v[0], (int, float)):
            v = tuple(v[0])
        if len(v) == 1:
            v = (v[0],) * 3
        return tuple.__new__(cls, (float(x) for x in v))


class Point(tuple):
    def __new__(cls, *v):
        if len(v) == 1:
            v = tuple(v[0])
        return tuple.__new__(cls, (float(x) for x in v))


class Plugin:
    """One XML node: <kind type="..."> with typed properties and named children."""

    def __init__(self, kind, type, _children=(), **props):
        self.kind, self.type = kind, type
        self.props = dict(props)
        self.children = list(_children)  # (name, Plugin)
        self.data = {}                   # in-memory payloads (mesh arrays, bitmaps)
        self.id = None

    def add(self, child, name=""):
        self.children.append((name, child))
        return self

    def get(self, name, default=None):
        return self.props.get(name, default)

    def child(self, kind, name=None):
        for n, c in self.children:
            if c.kind == kind and (name is None or n == name):
                return c
        return None

    def children_of(self, kind):
        return [(n, c) for n, c in self.children if c.kind == kind]


KINDS = ("scene", "integrator", "sensor", "film", "sampler", "rfilter", "shape", "bsdf", "emitter", "texture")


# ---------------------------------------------------------------------------
# asset I/O: PFM images and Mitsuba `.serialized` meshes
# ---------------------------------------------------------------------------
def write_pfm(path, img):
    img = np.asarray(img, dtype=F)
    if img.ndim == 2:
        img = img[:, :, None]
    h, w, c = img.shape
    assert c in (1, 3)
    with open(path, "wb") as f:
        f.write(("PF\n" if c == 3 else "Pf\n").encode())
        f.write(("%d %d\n-1.0\n" % (w, h)).encode())
        f.write(np.ascontiguousarray(img[::-1]).astype("<f4").tobytes())  # PFM stores bottom row first


def read_pfm(path):
    with open(path, "rb") as f:
        tag = f.readline().strip()
        c = 3 if tag == b"PF" else 1
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline().strip())
        data = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4", count=w * h * c)
    return np.ascontiguousarray(data.reshape(h, w, c)[::-1]).astype(F)


def read_image(path):
    ext = os.path.splitext(path)[1].lower()
    if ext == ".pfm":
        return read_pfm(path), "float"
    if ext == ".npy":
        a = np.load(path).astype(F)
        return (a[:, :, None] if a.ndim == 2 else a), "float"
    from PIL import Image  # 8-bit formats
    im = np.asarray(Image.open(path).convert("RGB"), dtype=F) / F(255)
    return im, "srgb"


def srgb_to_linear(v):
    v = np.asarray(v, dtype=F)
    return np.where(v <= 0.04045, v / F(12.92), np.power((v + F(0.055)) / F(1.055), F(2.4))).astype(F)


def write_serialized(path, positions, indices, normals=None, texcoords=None, name="mesh"):
    """Mitsuba binary mesh format v4 (src/librender/trimesh.cpp:145-290): u16 0x041C, u16 version,
    zlib stream {u32 flags, utf8 name\\0, u64 nverts, u64 ntris, positions, [normals], [uvs], indices};
    followed by the dictionary (offset of each mesh, count)."""
    pos = np.asarray(positions, dtype="<f4").reshape(-1, 3)
    idx = np.asarray(indices, dtype="<u4").reshape(-1, 3)
    flags = 0x1000  # single precision
    if normals is not None:
        flags |= 0x0001
    if texcoords is not None:
        flags |= 0x0002
    body = struct.pack("<I", flags) + name.encode() + b"\0" + struct.pack("<QQ", pos.shape[0], idx.shape[0])
    body += pos.tobytes()
    if normals is not None:
        body += np.asarray(normals, dtype="<f4").tobytes()
    if texcoords is not None:
        body += np.asarray(texcoords, dtype="<f4").tobytes()
    body += idx.tobytes()
    with open(path, "wb") as f:
        f.write(struct.pack("<HH", 0x041C, 4))
        f.write(zlib.compress(body, 1))
        f.write(struct.pack("<Q", 0))  # offset of mesh 0
        f.write(struct.pack("<I", 1))  # mesh count


def read_serialized(path, shape_index=0):
    raw = open(path, "rb").read()
    count = struct.unpack("<I", raw[-4:])[0]
    fmt, ver = struct.unpack("<HH", raw[0:4])
    if fmt != 0x041C:
        raise ValueError("%s: not a Mitsuba serialized mesh" % path)
    off = 0
    if shape_index:
        if ver == 4:
            off = struct.unpack("<Q", raw[len(raw) - 4 - 8 * (count - shape_index):][:8])[0]
        else:
            off = struct.unpack("<I", raw[len(raw) - 4 - 4 * (count - shape_index):][:4])[0]
    fmt, ver = struct.unpack("<HH", raw[off:off + 4])
    body = zlib.decompressobj().decompress(raw[off + 4:])
    p = 0
    flags = struct.unpack("<I", body[p:p + 4])[0]
    p += 4
    if ver == 4:
        e = body.index(b"\0", p)
        p = e + 1
    nv, nt = struct.unpack("<QQ", body[p:p + 16])
    p += 16
    dbl = bool(flags & 0x2000)
    ft, fs = ("<f8", 8) if dbl else ("<f4", 4)
    pos = np.frombuffer(body, ft, nv * 3, p).reshape(-1, 3).astype(F)
    p += nv * 3 * fs
    nrm = uv = None
    if flags & 0x0001:
        nrm = np.frombuffer(body, ft, nv * 3, p).reshape(-1, 3).astype(F)
        p += nv * 3 * fs
    if flags & 0x0002:
        uv = np.frombuffer(body, ft, nv * 2, p).reshape(-1, 2).astype(F)
        p += nv * 2 * fs
    if flags & 0x0008:
        p += nv * 3 * fs
    idx = np.frombuffer(body, "<u4", nt * 3, p).reshape(-1, 3)
    return pos, idx, nrm, uv, bool(flags & 0x0010)


# ---------------------------------------------------------------------------
# XML subset loader (replaces src/librender/scenehandler.cpp for the in-scope plugins)
# ---------------------------------------------------------------------------
def _floats(s):
    return [float(x) for x in s.replace(",", " ").split()]


def _parse_transform(node, subst):
    m = hm.identity()
    for op in node:
        a = {k: subst(v) for k, v in op.attrib.items()}
        if op.tag == "translate":
            t = hm.translate([float(a.get("x", 0)), float(a.get("y", 0)), float(a.get("z", 0))])
        elif op.tag == "scale":
            if "value" in a:
                t = hm.scale(float(a["value"]))
            else:
                t = hm.scale([float(a.get("x", 1)), float(a.get("y", 1)), float(a.get("z", 1))])
        elif op.tag == "rotate":
            t = hm.rotate([float(a.get("x", 0)), float(a.get("y", 0)), float(a.get("z", 0))], float(a["angle"]))
        elif op.tag == "lookat" or op.tag == "lookAt":
            t = hm.look_at(_floats(a["origin"]), _floats(a["target"]), _floats(a.get("up", "0, 1, 0")))
        elif op.tag == "matrix":
            t = np.array(_floats(a["value"]), dtype=np.float64).reshape(4, 4)
        else:
            raise ValueError("unsupported transform element <%s>" % op.tag)
        m = t @ m  # later operations are applied after earlier ones
    return m


def _parse_spectrum(tag, a):
    v = a["value"]
    if tag == "rgb" or tag == "srgb":
        if v.startswith("#"):
            rgb = [int(v[i:i + 2], 16) / 255.0 for i in (1, 3, 5)]
        else:
            rgb = _floats(v)
            if len(rgb) == 1:
                rgb = rgb * 3
        if tag == "srgb":   # <rgb> is linear whatever its spelling, only <srgb> is converted (scenehandler.cpp:461-528)
            rgb = [float(x) for x in srgb_to_linear(rgb)]
        return Spectrum(rgb)
    vals = _floats(v) if ":" not in v else None
    if vals is None or len(vals) not in (1, 3):
        raise ValueError("only uniform <spectrum value=\"x\"/> values are supported in the RGB build")
    return Spectrum(vals)


def load_xml(source, defines=None, base_dir=None):
    """Parse a Mitsuba 0.6 scene (file path or XML string) into a Plugin tree."""
    defines = defines or {}
    if os.path.exists(source):
        base_dir = base_dir or os.path.dirname(os.path.abspath(source))
        root = ET.parse(source).getroot()
    else:
        root = ET.fromstring(source)
    base_dir = base_dir or os.getcwd()
    ids = {}

    def subst(v):  # -D name=value substitution, src/mitsuba/mitsuba.cpp:168-173
        for k, val in defines.items():
            v = v.replace("$" + k, str(val))
        return v

    def parse(node):
        p = Plugin(node.tag, subst(node.attrib.get("type", "")))
        p.base_dir = base_dir
        if "id" in node.attrib:
            p.id = node.attrib["id"]
            ids[p.id] = p
        for ch in node:
            a = {k: subst(v) for k, v in ch.attrib.items()}
            name = a.get("name", "")
            if ch.tag in ("float",):
                p.props[name] = float(a["value"])
            elif ch.tag == "integer":
                p.props[name] = int(a["value"])
            elif ch.tag == "boolean":
                p.props[name] = a["value"].lower() == "true"
            elif ch.tag == "string":
                p.props[name] = a["value"]
            elif ch.tag in ("rgb", "srgb", "spectrum"):
                p.props[name] = _parse_spectrum(ch.tag, a)
            elif ch.tag in ("point", "vector"):
                p.props[name] = Point(float(a.get("x", 0)), float(a.get("y", 0)), float(a.get("z", 0)))
            elif ch.tag == "transform":
                p.props[name] = _parse_transform(ch, subst)
            elif ch.tag == "ref":
                if a["id"] not in ids:
                    raise ValueError('reference to unknown id "%s"' % a["id"])
                p.children.append((name, ids[a["id"]]))
            elif ch.tag in KINDS:
                p.children.append((name, parse(ch)))
            elif ch.tag in ("default", "include", "alias"):
                raise ValueError("<%s> is not supported by this loader" % ch.tag)
            else:
                raise ValueError("unsupported scene element <%s>" % ch.tag)
        return p

    if root.tag != "scene":
        raise ValueError("root element must be <scene>")
    return parse(root)


def to_xml(scene, path):
    """Write the Plugin tree as a Mitsuba 0.6 scene file; in-memory meshes / bitmaps are written
    next to it as .serialized / .pfm so the reference build reads identical inputs."""
    out_dir = os.path.dirname(os.path.abspath(path))
    stem = os.path.splitext(os.path.basename(path))[0]
    counter = [0]

    def asset(ext):
        counter[0] += 1
        return "%s_asset%d.%s" % (stem, counter[0], ext)

    def emit(p, name=None):
        el = ET.Element(p.kind)
        ptype, props = p.type, dict(p.props)
        if p.kind == "shape" and p.type == "mesh":
            fn = asset("serialized")
            d = p.data
            pos, nrm = d["positions"], d.get("normals")
            tw = props.pop("toWorld", None)
            if tw is not None:  # bake, the arrays are world-space in the flattened scene as well
                pos = np.array([hm.xform_point(tw, q) for q in pos], dtype=F) if len(pos) < 100000 else \
                    (np.c_[pos, np.ones(len(pos))] @ np.asarray(tw).T)[:, :3].astype(F)
                if nrm is not None:
                    nrm = (np.asarray(nrm) @ np.linalg.inv(tw)[:3, :3]).astype(F)
                    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
            write_serialized(os.path.join(out_dir, fn), pos, d["indices"], nrm, d.get("texcoords"))
            ptype = "serialized"
            props["filename"] = fn
        if p.kind in ("texture", "emitter") and "bitmap" in p.data:
            fn = asset("pfm")
            write_pfm(os.path.join(out_dir, fn), p.data["bitmap"])
            props["filename"] = fn
        if p.kind == "scene":
            el.set("version", "0.6.0")
        else:
            el.set("type", ptype)
        if name:
            el.set("name", name)
        if p.id:
            el.set("id", p.id)
        for k, v in props.items():
            if isinstance(v, bool):
                ET.SubElement(el, "boolean", name=k, value="true" if v else "false")
            elif isinstance(v, int):
                ET.SubElement(el, "integer", name=k, value=str(v))
            elif isinstance(v, float):
                ET.SubElement(el, "float", name=k, value=repr(v))
            elif isinstance(v, str):
                ET.SubElement(el, "string", name=k, value=v)
            elif isinstance(v, Spectrum):
                ET.SubElement(el, "rgb", name=k, value=", ".join(repr(x) for x in v))
            elif isinstance(v, Point):
                ET.SubElement(el, "point", name=k, x=repr(v[0]), y=repr(v[1]), z=repr(v[2]))
            elif isinstance(v, np.ndarray) and v.shape == (4, 4):
                t = ET.SubElement(el, "transform", name=k)
                ET.SubElement(t, "matrix", value=" ".join(repr(float(x)) for x in v.reshape(-1)))
            else:
                raise ValueError("cannot serialise property %s=%r" % (k, v))
        for n, c in p.children:
            el.append(emit(c, n))
        return el

    tree = ET.ElementTree(emit(scene))
    ET.indent(tree)
    tree.write(path, xml_declaration=True, encoding="utf-8")
    return path


# ---------------------------------------------------------------------------
# flattening
# ---------------------------------------------------------------------------
# Measured RGB (eta, k) of a few `material` presets of conductor/roughconductor, i.e. the values
# InterpolatedSpectrum("data/ior/<name>.{eta,k}.spd").fromContinuousSpectrum() yields in the RGB
# build (src/bsdfs/conductor.cpp:161-170).  Regenerated by tests/golden/make_ior_rgb.py when the
# reference build is available; other names need explicit `eta`/`k`.
CONDUCTOR_PRESETS = {
    "cu": ((0.200438, 0.924033, 1.102212), (3.912949, 2.452848, 2.142188)),
    "au": ((0.143119, 0.374957, 1.442479), (3.983160, 2.385721, 1.603215)),
    "al": ((1.657460, 0.880369, 0.521229), (9.223869, 6.269523, 4.837001)),
    "ag": ((0.155265, 0.116723, 0.138342), (4.827090, 3.122490, 2.146980)),
    "none": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
}

SUPPORTED_BSDFS = ("diffuse", "conductor", "roughconductor", "plastic", "roughplastic", "dielectric",
                   "coating", "blendbsdf", "bumpmap", "twosided", "thindielectric", "mask", "roughdielectric",
                   "mixturebsdf", "roughcoating", "normalmap")


class Unsupported(ValueError):
    """The plugin rejects scenes that use out-of-scope plugins with a clear error."""


class FlatScene:
    def __init__(self):
        self.textures = []   # (TextureDesc, texels float32 HxWxC)
        self.pool = []       # float LUT pool
        self.nodes = []      # BsdfNode
        self.shapes = []     # ("mesh", dict) / ("rectangle", m, bsdf) / ("sphere", m, radius, flip, bsdf)
        self.emitters = []   # (EmitterDesc, texels or None)
        self.camera = None
        self.film = None
        self.params = {}
        self.film_size = (0, 0)
        self.spp = 4
        self._tex_ids = {}
        self._bsdf_ids = {}
        self._mix_children = {}   # node index -> nested node indices of a mixturebsdf

    # ---- textures ---------------------------------------------------------
    def _bitmap(self, p):
        key = id(p)
        if key in self._tex_ids:
            return self._tex_ids[key]
        if "bitmap" in p.data:
            img, kind = np.asarray(p.data["bitmap"], dtype=F), "float"
            if img.ndim == 2:
                img = img[:, :, None]
        else:
            fn = p.get("filename")
            if fn is None:
                raise ValueError("bitmap texture needs a filename")
            if not os.path.isabs(fn):
                fn = os.path.join(getattr(p, "base_dir", os.getcwd()), fn)
            img, kind = read_image(fn)
        gamma = p.get("gamma", None)
        if gamma is None and kind == "srgb":
            img = srgb_to_linear(img)  # Bitmap gamma = -1 (sRGB) -> linear, bitmap.cpp:246-250
        elif gamma is not None and float(gamma) not in (1.0,) and kind == "srgb":
            img = np.power(img, F(gamma))
        ft = p.get("filterType", "ewa").lower()
        if ft not in abi.FILTER:   # bitmap.cpp:220-231
            raise ValueError("Unknown filter type '%s' -- must be 'ewa', 'trilinear', or 'nearest'!" % ft)
        max_aniso = float(p.get("maxAnisotropy", 20.0)) if ft == "ewa" else 1.0   # bitmap.cpp:230-233
        wm = p.get("wrapMode", "repeat")
        uvscale = p.get("uvscale", 1.0)
        d = abi.TextureDesc(img.shape[1], img.shape[0], img.shape[2],
                            abi.WRAP[p.get("wrapModeU", wm)], abi.WRAP[p.get("wrapModeV", wm)], abi.FILTER[ft],
                            (abi.C.c_float * 2)(p.get("uscale", uvscale), p.get("vscale", uvscale)),
                            (abi.C.c_float * 2)(p.get("uoffset", 0.0), p.get("voffset", 0.0)), max_aniso)
        if p.get("channel") is not None:
            raise Unsupported("bitmap `channel` selection is not supported")
        half = np.maximum(img, 0).astype(np.float16).astype(F)  # TMIPMap stores half and clamps negative texels (mipmap.h:228-236)
        stats = dict(min=half.reshape(-1, img.shape[2]).min(0), max=half.reshape(-1, img.shape[2]).max(0),
                     avg=half.reshape(-1, img.shape[2]).astype(np.float64).mean(0))
        for k in stats:
            if img.shape[2] == 1:
                stats[k] = np.repeat(stats[k], 3)
        self.textures.append((d, np.ascontiguousarray(img)))
        self._tex_ids[key] = (len(self.textures) - 1, stats)
        return self._tex_ids[key]

    def texref(self, owner, name, default, as_float=False):
        """Resolve a Texture slot of a BSDF (addChild(name, Texture) or a plain property)."""
        child = owner.child("texture", name)
        scale = np.ones(3, dtype=np.float64)
        if child is None:
            v = owner.get(name, default)
            v = np.array(Spectrum(v), dtype=np.float64)
            return dict(tex=-1, scale=v, min=v, max=v, avg=v, const=True)
        p = child
        while p.type == "scale":  # ScalingTexture, src/textures/scale.cpp
            s = p.get("scale", 1.0)
            scale = scale * np.array(Spectrum(s), dtype=np.float64)
            if "value" in p.props:
                v = np.array(Spectrum(p.props["value"]), dtype=np.float64) * scale
                return dict(tex=-1, scale=v, min=v, max=v, avg=v, const=True)
            p = p.child("texture")
            if p is None:
                raise ValueError("The scale plugin needs a nested texture!")
        if p.type != "bitmap":
            raise Unsupported("texture type '%s' is out of scope (supported: bitmap, scale, constants)" % p.type)
        tid, st = self._bitmap(p)
        return dict(tex=tid, scale=scale, min=st["min"] * scale, max=st["max"] * scale, avg=st["avg"] * scale,
                    const=False, plugin=p)

    @staticmethod
    def _conserve(t, max_allowed=1.0):
        """BSDF::ensureEnergyConservation (src/librender/bsdf.cpp:85-113)."""
        actual = float(np.max(t["max"]))
        if actual > max_allowed:
            s = F(0.99) * F(max_allowed / actual)
            for k in ("scale", "min", "max", "avg"):
                t[k] = t[k] * float(s)
        return t

    @staticmethod
    def _ref(t):
        return abi.TexRef((abi.C.c_float * 3)(*[float(x) for x in t["scale"]]), int(t["tex"]))

    # ---- BSDFs --------------------------------------------------------------
    def bsdf(self, p):
        key = id(p)
        if key in self._bsdf_ids:
            return self._bsdf_ids[key]
        if p.type not in SUPPORTED_BSDFS:
            raise Unsupported("BSDF plugin '%s' is not supported by cudapath (supported: %s)"
                              % (p.type, ", ".join(SUPPORTED_BSDFS)))
        n = abi.BsdfNode()
        n.child[0] = n.child[1] = -1
        n.t0.tex = n.t1.tex = n.t2.tex = -1  # unused texture slots
        idx = len(self.nodes)
        self.nodes.append(n)
        self._bsdf_ids[key] = idx
        lum = lambda v: 0.212671 * v[0] + 0.715160 * v[1] + 0.072169 * v[2]
        const3 = lambda v: dict(tex=-1, scale=np.array([v] * 3, dtype=np.float64), const=True)
        t = p.type

        def microfacet():
            distr = p.get("distribution", "beckmann").lower()
            if distr not in ("beckmann", "ggx"):
                raise Unsupported("microfacet distribution '%s' is out of scope (beckmann, ggx)" % distr)
            flags = (abi.BF_GGX if distr == "ggx" else 0)
            if p.get("sampleVisible", True):
                flags |= abi.BF_SAMPLE_VISIBLE
            return flags

        def conductor_ior():
            mat = p.get("material", "Cu")
            ext = hm.lookup_ior(p.get("extEta"), "air")
            if "eta" in p.props and "k" in p.props:
                eta, k = Spectrum(p.props["eta"]), Spectrum(p.props["k"])
            elif mat.lower() in CONDUCTOR_PRESETS:
                eta, k = CONDUCTOR_PRESETS[mat.lower()]
                eta = Spectrum(p.props["eta"]) if "eta" in p.props else eta
                k = Spectrum(p.props["k"]) if "k" in p.props else k
            else:
                raise Unsupported("conductor material preset '%s' needs explicit RGB `eta`/`k`" % mat)
            for i in range(3):
                n.eta[i] = F(F(eta[i]) / F(ext))
                n.k[i] = F(F(k[i]) / F(ext))

        if t == "diffuse":
            r = self.texref(p, "reflectance", Spectrum(0.5))
            if p.child("texture", "diffuseReflectance") is not None or "diffuseReflectance" in p.props:
                r = self.texref(p, "diffuseReflectance", Spectrum(0.5))
            r = self._conserve(r)
            n.type, n.t0 = abi.BSDF_DIFFUSE, self._ref(r)
            n.lobes = abi.EDIFFUSE_REFLECTION if float(np.max(r["max"])) > 0 else 0
        elif t == "conductor":
            r = self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0)))
            conductor_ior()
            n.type, n.t0, n.lobes = abi.BSDF_CONDUCTOR, self._ref(r), abi.EDELTA_REFLECTION
        elif t == "roughconductor":
            r = self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0)))
            conductor_ior()
            n.flags = microfacet()
            if p.child("texture", "alpha") is not None or "alpha" in p.props:
                au = av = self.texref(p, "alpha", 0.1)
            else:
                au, av = self.texref(p, "alphaU", 0.1), self.texref(p, "alphaV", 0.1)
            n.type, n.t0, n.t1, n.t2 = abi.BSDF_ROUGHCONDUCTOR, self._ref(r), self._ref(au), self._ref(av)
            n.lobes = abi.EGLOSSY_REFLECTION
        elif t == "roughdielectric":
            # RoughDielectric ctor / addChild / configure (src/bsdfs/roughdielectric.cpp:183-211, 246-270, 618-636)
            int_ior = hm.lookup_ior(p.get("intIOR"), "bk7")
            ext_ior = hm.lookup_ior(p.get("extIOR"), "air")
            if int_ior < 0 or ext_ior < 0 or int_ior == ext_ior:
                raise ValueError("The interior and exterior indices of refraction must be positive and differ!")
            n.type = abi.BSDF_ROUGHDIELECTRIC
            n.eta[0] = F(int_ior) / F(ext_ior)
            n.flags = microfacet()
            if p.child("texture", "alpha") is not None or "alpha" in p.props:
                au = av = self.texref(p, "alpha", 0.1)
            else:
                au, av = self.texref(p, "alphaU", 0.1), self.texref(p, "alphaV", 0.1)
            r = self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0)))
            tr = self._ref(self._conserve(self.texref(p, "specularTransmittance", Spectrum(1.0))))
            n.t0, n.t1, n.t2 = self._ref(r), self._ref(au), self._ref(av)
            for i in range(3):          # fourth texture reference lives in the k / lut slots (include/mtscuda.h)
                n.k[i] = tr.scale[i]
            n.lut[0] = tr.tex
            n.lobes = abi.EGLOSSY_REFLECTION | abi.EGLOSSY_TRANSMISSION | abi.EFRONTSIDE | abi.EBACKSIDE
        elif t in ("plastic", "roughplastic"):
            int_ior = hm.lookup_ior(p.get("intIOR"), "polypropylene")
            ext_ior = hm.lookup_ior(p.get("extIOR"), "air")
            if int_ior < 0 or ext_ior < 0 or (t == "roughplastic" and int_ior == ext_ior):
                raise ValueError("The interior and exterior indices of refraction must be positive and differ!")
            eta = float(F(int_ior) / F(ext_ior))
            s = self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0)))
            d = self._conserve(self.texref(p, "diffuseReflectance", Spectrum(0.5)))
            d_avg, s_avg = lum(d["avg"]), lum(s["avg"])
            n.eta[0] = eta
            n.t0, n.t1 = self._ref(d), self._ref(s)
            n.f[2] = F(s_avg / (d_avg + s_avg))  # m_specularSamplingWeight
            if p.get("nonlinear", False):
                n.flags |= abi.BF_NONLINEAR
            if t == "plastic":
                n.type = abi.BSDF_PLASTIC
                n.f[0] = hm.fresnel_diffuse_reflectance(1.0 / eta)  # m_fdrInt, plastic.cpp:193
                n.f[1] = hm.fresnel_diffuse_reflectance(eta)        # m_fdrExt
                n.lobes = abi.EDELTA_REFLECTION | abi.EDIFFUSE_REFLECTION
            else:
                n.type = abi.BSDF_ROUGHPLASTIC
                n.flags |= microfacet()
                if "alphaU" in p.props or "alphaV" in p.props:
                    raise ValueError("The 'roughplastic' plugin currently does not support anisotropic "
                                     "microfacet distributions!")
                a = self.texref(p, "alpha", 0.1)
                n.t2 = self._ref(a)
                rt = hm.RoughTransmittance("ggx" if n.flags & abi.BF_GGX else "beckmann")
                rt.check_eta(eta)
                rt.check_alpha(float(np.mean(a["min"])))
                rt.check_alpha(float(np.mean(a["max"])))
                ext2d, _ = rt.reduce_eta(eta)                 # m_externalRoughTransmittance->setEta(eta)
                _, int_diff = rt.reduce_eta(1.0 / eta)        # m_internalRoughTransmittance->setEta(1/eta)
                n.f[0], n.f[1] = rt.alpha_min, rt.alpha_max
                n.lut[1], n.lut[2] = rt.n_theta, rt.n_alpha
                if a["const"]:                                # setAlpha(): 1-D slice
                    alpha = max(float(F(np.mean(a["scale"].astype(F)))), 1e-4)  # MicrofacetDistribution clamp
                    alpha_cfg = float(F(np.mean(a["scale"].astype(F))))
                    ext1d, _ = rt.reduce_alpha(ext2d, np.zeros(rt.n_alpha, dtype=F), alpha_cfg)
                    n.lut[0] = len(self.pool)
                    self.pool += [float(x) for x in ext1d]
                    fdr_t = hm.cubic_interp_1d(rt.warp_alpha(alpha), int_diff)
                    n.f[3] = F(1.0) - min(F(1.0), max(F(0.0), fdr_t))
                else:
                    n.flags |= abi.BF_RT_2D
                    n.lut[0] = len(self.pool)
                    self.pool += [float(x) for x in ext2d.reshape(-1)]
                    n.lut[3] = len(self.pool)
                    self.pool += [float(x) for x in int_diff]
                n.lobes = abi.EGLOSSY_REFLECTION | abi.EDIFFUSE_REFLECTION
        elif t == "dielectric":
            int_ior = hm.lookup_ior(p.get("intIOR"), "bk7")
            ext_ior = hm.lookup_ior(p.get("extIOR"), "air")
            if int_ior < 0 or ext_ior < 0:
                raise ValueError("The interior and exterior indices of refraction must be positive!")
            n.type = abi.BSDF_DIELECTRIC
            n.eta[0] = F(int_ior) / F(ext_ior)
            n.t0 = self._ref(self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0))))
            n.t1 = self._ref(self._conserve(self.texref(p, "specularTransmittance", Spectrum(1.0))))
            n.lobes = abi.EDELTA_REFLECTION | abi.EDELTA_TRANSMISSION | abi.EBACKSIDE
        elif t == "coating":
            int_ior = hm.lookup_ior(p.get("intIOR"), "bk7")
            ext_ior = hm.lookup_ior(p.get("extIOR"), "air")
            if int_ior < 0 or ext_ior < 0 or int_ior == ext_ior:
                raise ValueError("The interior and exterior indices of refraction must be positive and differ!")
            nested = p.child("bsdf")
            if nested is None:
                raise ValueError("A child BSDF instance is required")
            n.type = abi.BSDF_COATING
            n.eta[0] = F(int_ior) / F(ext_ior)
            thickness = float(p.get("thickness", 1.0))
            n.f[0] = thickness
            sig = self.texref(p, "sigmaA", Spectrum(0.0))
            # m_specularSamplingWeight = 1/(avgAbsorption+1), coating.cpp:181-183
            avg_abs = float(np.mean([F(math.exp(float(F(x) * F(-2 * thickness)))) for x in sig["avg"]]))
            n.f[2] = F(1.0) / (F(avg_abs) + F(1.0))
            n.t0 = self._ref(sig)
            n.t1 = self._ref(self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0))))
            n.child[0] = self.bsdf(nested)
            n.lobes = self.nodes[n.child[0]].lobes | abi.EDELTA_REFLECTION | abi.EBACKSIDE
        elif t == "blendbsdf":
            kids = [c for _, c in p.children_of("bsdf")]
            if len(kids) != 2:
                raise ValueError("BSDF count mismatch: expected two nested BSDF instances!")
            n.type = abi.BSDF_BLEND
            n.t0 = self._ref(self.texref(p, "weight", 0.5))
            n.child[0], n.child[1] = self.bsdf(kids[0]), self.bsdf(kids[1])
            n.lobes = self.nodes[n.child[0]].lobes | self.nodes[n.child[1]].lobes
        elif t == "bumpmap":
            nested = p.child("bsdf")
            disp = p.child("texture")
            if nested is None:
                raise ValueError("A child BSDF instance is required")
            if disp is None:
                raise ValueError("A displacement texture must be specified")
            n.type = abi.BSDF_BUMPMAP
            name = [nm for nm, c in p.children if c is disp][0]
            n.t0 = self._ref(self.texref(p, name, 0.0))
            n.child[0] = self.bsdf(nested)
            n.lobes = self.nodes[n.child[0]].lobes
        elif t == "thindielectric":
            # ThinDielectric ctor / configure (src/bsdfs/thindielectric.cpp:71-124)
            int_ior = hm.lookup_ior(p.get("intIOR"), "bk7")
            ext_ior = hm.lookup_ior(p.get("extIOR"), "air")
            if int_ior < 0 or ext_ior < 0:
                raise ValueError("The interior and exterior indices of refraction must be positive!")
            n.type = abi.BSDF_THINDIELECTRIC
            n.eta[0] = F(int_ior) / F(ext_ior)
            n.t0 = self._ref(self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0))))
            n.t1 = self._ref(self._conserve(self.texref(p, "specularTransmittance", Spectrum(1.0))))
            n.lobes = abi.EDELTA_REFLECTION | abi.ENULL | abi.EFRONTSIDE | abi.EBACKSIDE
        elif t == "mask":
            # Mask ctor / configure / addChild (src/bsdfs/mask.cpp:62-104, 232-239)
            nested = p.child("bsdf")
            if nested is None:
                raise ValueError("A child BSDF is required")
            n.type = abi.BSDF_MASK
            n.t0 = self._ref(self._conserve(self.texref(p, "opacity", Spectrum(0.5))))
            n.child[0] = self.bsdf(nested)
            n.lobes = self.nodes[n.child[0]].lobes | abi.ENULL | abi.EFRONTSIDE | abi.EBACKSIDE
        elif t == "twosided":
            # TwoSidedBRDF::addChild / configure (src/bsdfs/twosided.cpp:86-105, 182-193)
            kids = [c for _, c in p.children_of("bsdf")]
            if len(kids) == 0:
                raise ValueError("A nested one-sided material is required!")
            if len(kids) > 2:
                raise ValueError("No more than two nested BRDFs can be added!")
            n.type = abi.BSDF_TWOSIDED
            n.child[0] = self.bsdf(kids[0])
            n.child[1] = self.bsdf(kids[1]) if len(kids) == 2 else n.child[0]
            front, back = self.nodes[n.child[0]].lobes, self.nodes[n.child[1]].lobes
            n.lobes = ((front & ~abi.EBACKSIDE) | abi.EFRONTSIDE) | ((back & ~abi.EFRONTSIDE) | abi.EBACKSIDE)
            if n.lobes & abi.ETRANSMISSION:
                raise ValueError("Only materials without a transmission component can be nested!")
        elif t == "mixturebsdf":
            # MixtureBSDF ctor / addChild / configure (src/bsdfs/mixturebsdf.cpp:60-98, 114-172)
            kids = [c for _, c in p.children_of("bsdf")]
            weights = [F(x) for x in str(p.get("weights", "")).replace(",", " ").replace(";", " ").split()]
            if not weights:
                raise ValueError("No weights were supplied!")
            if len(kids) != len(weights):
                raise ValueError("BSDF count mismatch: %d bsdfs, but specified %d weights" % (len(kids), len(weights)))
            if len(kids) > abi.MIXTURE_MAX:
                raise Unsupported("mixturebsdf with more than %d nested BSDFs" % abi.MIXTURE_MAX)
            for w in weights:
                if w < 0:
                    raise ValueError("Invalid BSDF weight: %f" % w)
            total = F(0)
            for w in weights:
                total = F(total + w)
            if total <= 0:
                raise ValueError("The weights must sum to a value greater than zero!")
            if p.get("ensureEnergyConservation", True) and total > 1:
                scale = F(1.0) / total
                weights = [F(w * scale) for w in weights]
            cdf = [F(0)]                                    # DiscreteDistribution::append / normalize (pmf.h:58-66, 101-114)
            for w in weights:
                cdf.append(F(cdf[-1] + w))
            norm = F(1.0) / cdf[-1]
            cdf = [cdf[0]] + [F(c * norm) for c in cdf[1:]]
            cdf[-1] = F(1.0)
            children = [self.bsdf(k) for k in kids]
            n.type = abi.BSDF_MIXTURE
            n.lut[0], n.lut[1] = len(self.pool), len(kids)
            self.pool += [float(c) for c in children] + [float(w) for w in weights] + [float(c) for c in cdf]
            lobes = 0
            for c in children:
                lobes |= self.nodes[c].lobes
            n.lobes = lobes
            self._mix_children[idx] = children
        elif t == "roughcoating":
            # RoughCoating ctor / addChild / configure (src/bsdfs/roughcoating.cpp:113-148, 180-225, 450-470)
            int_ior = hm.lookup_ior(p.get("intIOR"), "bk7")
            ext_ior = hm.lookup_ior(p.get("extIOR"), "air")
            if int_ior < 0 or ext_ior < 0 or int_ior == ext_ior:
                raise ValueError("The interior and exterior indices of refraction must be positive and differ!")
            nested = p.child("bsdf")
            if nested is None:
                raise ValueError("A child BSDF instance is required")
            if "alphaU" in p.props or "alphaV" in p.props:
                raise ValueError("The 'roughplastic' plugin currently does not support anisotropic microfacet distributions!")
            eta = float(F(int_ior) / F(ext_ior))
            n.type = abi.BSDF_ROUGHCOATING
            n.eta[0] = eta
            n.flags = microfacet()
            thickness = float(p.get("thickness", 1.0))
            n.f[3] = thickness
            sig = self.texref(p, "sigmaA", Spectrum(0.0))
            avg_abs = float(np.mean([F(math.exp(float(F(x) * F(-2 * thickness)))) for x in sig["avg"]]))
            n.f[2] = F(1.0) / (F(avg_abs) + F(1.0))        # m_specularSamplingWeight, roughcoating.cpp:200-203
            a = self.texref(p, "alpha", 0.1)
            n.t0 = self._ref(sig)
            n.t1 = self._ref(self._conserve(self.texref(p, "specularReflectance", Spectrum(1.0))))
            n.t2 = self._ref(a)
            rt = hm.RoughTransmittance("ggx" if n.flags & abi.BF_GGX else "beckmann")
            rt.check_eta(eta)
            rt.check_alpha(float(np.mean(a["min"])))
            rt.check_alpha(float(np.mean(a["max"])))
            ext2d, _ = rt.reduce_eta(eta)
            n.f[0], n.f[1] = rt.alpha_min, rt.alpha_max
            n.lut[1], n.lut[2] = rt.n_theta, rt.n_alpha
            n.lut[0] = len(self.pool)
            if a["const"]:
                alpha_cfg = float(F(np.mean(a["scale"].astype(F))))
                ext1d, _ = rt.reduce_alpha(ext2d, np.zeros(rt.n_alpha, dtype=F), alpha_cfg)
                self.pool += [float(x) for x in ext1d]
            else:
                n.flags |= abi.BF_RT_2D
                self.pool += [float(x) for x in ext2d.reshape(-1)]
            n.child[0] = self.bsdf(nested)
            n.lobes = self.nodes[n.child[0]].lobes | abi.EGLOSSY_REFLECTION | abi.EFRONTSIDE | abi.EBACKSIDE
        elif t == "normalmap":
            # NormalMap addChild / configure (src/bsdfs/normalmap.cpp:58-104)
            nested = p.child("bsdf")
            normals = p.child("texture")
            if nested is None:
                raise ValueError("A child BSDF instance is required")
            if normals is None:
                raise ValueError("A normal map texture must be specified")
            if normals.type == "bitmap" and normals.get("gamma") is None:
                raise ValueError("When using a bitmap texture as a normal map, please explicitly specify the 'gamma' "
                                 "parameter of the bitmap plugin.")
            n.type = abi.BSDF_NORMALMAP
            name = [nm for nm, c in p.children if c is normals][0]
            n.t0 = self._ref(self.texref(p, name, 0.0))
            n.child[0] = self.bsdf(nested)
            n.lobes = self.nodes[n.child[0]].lobes
        if self._uses_ray_differentials(n):
            n.flags |= abi.BF_RAY_DIFF
        return idx

    def _uses_ray_differentials(self, n):
        """BSDF::usesRayDifferentials() as every plugin's configure() computes it: a bitmap texture always asks for
        them (src/textures/bitmap.cpp:541-543), constants never (include/mitsuba/hw/basicshader.h:72); bumpmap always
        (bumpmap.cpp:96); blendbsdf / mask / twosided only look at the nested BSDFs (blendbsdf.cpp:130, mask.cpp:108,
        twosided.cpp:91-92), not at their own weight / opacity texture."""
        kids = any(self.nodes[c].flags & abi.BF_RAY_DIFF for c in n.child if c >= 0)
        if n.type in (abi.BSDF_BUMPMAP, abi.BSDF_NORMALMAP):     # bumpmap.cpp:96, normalmap.cpp:69
            return True
        if n.type in (abi.BSDF_BLEND, abi.BSDF_MASK, abi.BSDF_TWOSIDED):
            return kids
        if n.type == abi.BSDF_MIXTURE:                           # mixturebsdf.cpp:166
            return any(self.nodes[c].flags & abi.BF_RAY_DIFF for c in self._mix_children[self.nodes.index(n)])
        own = any(r.tex >= 0 for r in (n.t0, n.t1, n.t2))
        if n.type == abi.BSDF_ROUGHDIELECTRIC:
            own = own or n.lut[0] >= 0      # specularTransmittance lives in the k / lut slots
        return own or kids

    def uses_uv_tangents(self, node):
        """bumpmap forces UV tangents / anisotropic models want them; TriMesh::configure computes them
        whenever texcoords exist (src/librender/trimesh.cpp:376-386)."""
        return True


def _emitter_desc(**kw):
    d = abi.EmitterDesc()
    d.shape = -1
    d.sampling_weight = 1.0
    d.scale = 1.0
    ident = np.eye(4, dtype=F).reshape(-1)
    for i in range(16):
        d.to_world[i] = ident[i]
    for k, v in kw.items():
        if k in ("radiance", "position"):
            for i in range(3):
                getattr(d, k)[i] = float(v[i])
        elif k == "to_world":
            m = np.asarray(v, dtype=F).reshape(-1)
            for i in range(16):
                d.to_world[i] = m[i]
        else:
            setattr(d, k, v)
    return d


def flatten(scene):
    """Scene -> FlatScene (what cudapath::render() hands to libmtscuda)."""
    fs = FlatScene()
    # ---- integrator (MonteCarloIntegrator ctor, src/librender/integrator.cpp:190-225) ----
    integ = scene.child("integrator") or Plugin("integrator", "path")
    if integ.type not in ("path", "cudapath"):
        raise Unsupported("integrator '%s' is out of scope (path / cudapath)" % integ.type)
    rr, md = int(integ.get("rrDepth", 5)), int(integ.get("maxDepth", -1))
    if rr <= 0:
        raise ValueError("'rrDepth' must be set to a value greater than zero!")
    if md <= 0 and md != -1:
        raise ValueError("'maxDepth' must be set to -1 (infinite) or a value greater than zero!")
    fs.params = dict(max_depth=md, rr_depth=rr, strict_normals=bool(integ.get("strictNormals", False)),
                     hide_emitters=bool(integ.get("hideEmitters", False)),
                     device=int(integ.get("device", 0)), devices=int(integ.get("devices", 1)))   # cudapath extras
    if not 1 <= fs.params["devices"] <= 16:
        raise ValueError("'devices' must be between 1 and 16!")
    # ---- sensor / film / sampler ----
    sensor = scene.child("sensor")
    if sensor is None or sensor.type != "perspective":
        raise Unsupported("a `perspective` sensor is required (other sensors are out of scope)")
    film = sensor.child("film") or Plugin("film", "hdrfilm")
    w, h = int(film.get("width", 768)), int(film.get("height", 576))
    cw, ch = int(film.get("cropWidth", w)), int(film.get("cropHeight", h))
    cox, coy = int(film.get("cropOffsetX", 0)), int(film.get("cropOffsetY", 0))
    rf = film.child("rfilter") or Plugin("rfilter", "gaussian")  # Film::configure default, film.cpp:89-96
    radius, table = hm.filter_table(rf.type, rf.get("radius"), rf.get("stddev", 0.5))
    # HDRFilm pixelFormat (src/films/hdrfilm.cpp:240-275): the alpha variants make Film::hasAlpha() true
    pixel_formats = [f.strip().lower() for f in str(film.get("pixelFormat", "rgb")).split(",")]
    has_alpha = any(f in ("luminancealpha", "rgba", "xyza", "spectrumalpha") for f in pixel_formats)
    fs.film = abi.Film(cw, ch, radius, (abi.C.c_float * 32)(*[float(x) for x in table]), int(has_alpha))
    fs.film_size = (cw, ch)
    sampler = sensor.child("sampler") or Plugin("sampler", "independent", sampleCount=4)
    if sampler.type != "independent":
        raise Unsupported("sampler '%s' is out of scope (independent)" % sampler.type)
    fs.spp = int(sampler.get("sampleCount", 4))
    to_world = np.asarray(sensor.get("toWorld", hm.identity()), dtype=np.float64)
    aspect = w / float(h)
    if "focalLength" in sensor.props and "fov" in sensor.props:
        raise ValueError("Please specify either a focal length ('focalLength') or a field of view ('fov')!")
    if "fov" in sensor.props:
        xfov = hm.x_fov(float(sensor.get("fov")), sensor.get("fovAxis", "x"), aspect)
    else:
        fl = str(sensor.get("focalLength", "50mm")).replace("mm", "")
        xfov = hm.x_fov(2 * math.degrees(math.atan(math.sqrt(36 * 36 + 24 * 24) / (2 * float(fl)))), "diagonal", aspect)
    near, far = float(sensor.get("nearClip", 1e-2)), float(sensor.get("farClip", 1e4))
    s2c, dx, dy, inv_res = hm.camera_matrices(to_world, xfov, near, far, (w, h), (cw, ch), (cox, coy))
    cam = abi.Camera()
    for i, v in enumerate(s2c.reshape(-1)):
        cam.sample_to_camera[i] = v
    for i, v in enumerate(to_world.reshape(-1)):
        cam.camera_to_world[i] = v
    for i in range(3):
        cam.dx[i], cam.dy[i] = dx[i], dy[i]
    cam.near_clip, cam.far_clip = near, far
    cam.inv_resolution[0], cam.inv_resolution[1] = inv_res
    fs.camera = cam

    # ---- shapes (+ their area emitters), in scene order like Scene::addChild ----
    default_bsdf = Plugin("bsdf", "diffuse", reflectance=Spectrum(0.5))   # shape.cpp:49-76
    black_bsdf = Plugin("bsdf", "diffuse", reflectance=Spectrum(0.0))

    def add_shape(sp):
        em = sp.child("emitter")
        bp = sp.child("bsdf")
        if bp is None:
            bp = black_bsdf if em is not None else default_bsdf
        b = fs.bsdf(bp)
        tw = np.asarray(sp.get("toWorld", hm.identity()), dtype=np.float64)
        for kind in ("subsurface", "medium"):
            if sp.child(kind) is not None:
                raise Unsupported("%s is out of scope for the path integrator plugin" % kind)
        if sp.type == "rectangle":
            if sp.get("flipNormals", False):
                tw = tw @ hm.scale([1, 1, -1])
            du, dv = hm.xform_vector(tw, [2, 0, 0]), hm.xform_vector(tw, [0, 2, 0])
            if abs(np.dot(du / np.linalg.norm(du), dv / np.linalg.norm(dv))) > 1e-4:
                raise ValueError("Error: 'toWorld' transformation contains shear!")
            fs.shapes.append(("rectangle", tw.astype(F), b))
        elif sp.type == "sphere":
            center = np.array(sp.get("center", Point(0, 0, 0)), dtype=np.float64)
            radius = float(sp.get("radius", 1.0))
            o2w = hm.translate(center)
            if "toWorld" in sp.props:  # sphere.cpp:111-120: strip the scale from the transform
                r = np.linalg.norm(hm.xform_vector(tw, [1, 0, 0]))
                o2w = tw @ hm.scale(1.0 / r) @ o2w
                radius *= r
            if radius <= 0:
                raise ValueError("Cannot create spheres of radius <= 0")
            fs.shapes.append(("sphere", o2w.astype(F), radius, bool(sp.get("flipNormals", False)), b))
        elif sp.type in ("cube", "mesh", "serialized"):
            if sp.type == "cube":
                pos, nrm, uv, idx = hm.cube_mesh(tw)
                if sp.get("flipNormals", False):
                    nrm = -nrm
            else:
                if sp.type == "serialized":
                    fn = sp.get("filename")
                    if not os.path.isabs(fn):
                        fn = os.path.join(getattr(sp, "base_dir", os.getcwd()), fn)
                    pos, idx, nrm, uv, face_n = read_serialized(fn, int(sp.get("shapeIndex", 0)))
                    face_n = face_n or sp.get("faceNormals", False)
                else:
                    d = sp.data
                    pos, idx, nrm, uv = d["positions"], d["indices"], d.get("normals"), d.get("texcoords")
                    face_n = sp.get("faceNormals", False)
                pos = np.asarray(pos, dtype=np.float64)
                if "toWorld" in sp.props:
                    pos = (np.c_[pos, np.ones(len(pos))] @ tw.T)[:, :3]
                    if nrm is not None:
                        nrm = np.asarray(nrm, dtype=np.float64) @ np.linalg.inv(tw)[:3, :3]
                        nrm = nrm / np.linalg.norm(nrm, axis=1, keepdims=True)
                if face_n:
                    nrm = None
                elif nrm is None:
                    nrm = hm.vertex_normals(pos, idx)  # TriMesh::computeNormals
                elif sp.get("flipNormals", False):
                    nrm = -np.asarray(nrm)
            fs.shapes.append(("mesh", dict(positions=np.asarray(pos, dtype=F), indices=np.asarray(idx, dtype=np.uint32),
                                           normals=None if nrm is None else np.asarray(nrm, dtype=F),
                                           texcoords=None if uv is None else np.asarray(uv, dtype=F),
                                           tangents=None, uv_tangents=uv is not None, bsdf=b)))
        else:
            raise Unsupported("shape '%s' is out of scope (rectangle, sphere, cube, serialized/in-memory meshes)" % sp.type)
        sid = len(fs.shapes) - 1
        if em is not None:
            if em.type != "area":
                raise Unsupported("only `area` emitters can be attached to shapes")
            area_emitters.append((_emitter_desc(type=abi.EMITTER_AREA, shape=sid, radiance=Spectrum(em.get("radiance", Spectrum(1.0))),
                                              sampling_weight=float(em.get("samplingWeight", 1.0))), None))

    # Emitter order = Scene::m_emitters order: scene-level emitters as they are added (Scene::addChild,
    # scene.cpp:497-516), then the area emitters of the shapes in shape order (Scene::initialize -> addShape,
    # scene.cpp:322-336, 570-571).  The order fixes which emitter a given sample selects (Scene::m_emitterPDF).
    area_emitters = []
    for _, c in scene.children:
        if c.kind == "shape":
            add_shape(c)
        elif c.kind == "emitter":
            sw = float(c.get("samplingWeight", 1.0))
            if c.type == "point":
                tw = np.asarray(c.get("toWorld", hm.identity()), dtype=np.float64)
                pos = hm.xform_point(tw, c.get("position", Point(0, 0, 0)))
                fs.emitters.append((_emitter_desc(type=abi.EMITTER_POINT, radiance=Spectrum(c.get("intensity", Spectrum(1.0))),
                                                  position=pos, sampling_weight=sw), None))
            elif c.type == "envmap":
                if "bitmap" in c.data:
                    img = np.asarray(c.data["bitmap"], dtype=F)
                else:
                    fn = c.get("filename")
                    if not os.path.isabs(fn):
                        fn = os.path.join(getattr(c, "base_dir", os.getcwd()), fn)
                    img, kind = read_image(fn)
                    if kind == "srgb":
                        img = srgb_to_linear(img)
                if img.shape[2] == 1:
                    img = np.repeat(img, 3, axis=2)
                if max(img.shape[0], img.shape[1]) > 0xFFFF:
                    raise ValueError("Environment maps images must be smaller than 65536 pixels in width and height")
                tw = np.asarray(c.get("toWorld", hm.identity()), dtype=np.float64)
                fs.emitters.append((_emitter_desc(type=abi.EMITTER_ENVMAP, scale=float(c.get("scale", 1.0)), to_world=tw,
                                                  env_width=img.shape[1], env_height=img.shape[0], sampling_weight=sw),
                                    np.ascontiguousarray(img, dtype=F)))
            else:
                raise Unsupported("emitter '%s' is out of scope (area, point, envmap)" % c.type)
        elif c.kind in ("integrator", "sensor"):
            pass
        elif c.kind in ("bsdf", "texture"):
            pass  # top-level definitions that shapes <ref> to
        else:
            raise Unsupported("scene-level <%s> is out of scope" % c.kind)
    fs.emitters += area_emitters
    if not fs.emitters:
        raise Unsupported("the scene has no emitters; the reference would silently add a `sunsky` emitter "
                          "(src/librender/scene.cpp:358-373), which is out of scope")
    return fs


def upload(fs, ctx):
    """Issue the C-ABI calls for a flattened scene.  Geometry goes first: texture uploads run on a stream of their own
    inside the library, so the copy engine is busy with the texels while build() constructs the BVH."""
    ctx.set_camera(fs.camera)
    ctx.set_film(fs.film)
    for s in fs.shapes:
        if s[0] == "rectangle":
            ctx.add_rectangle(s[1], s[2])
        elif s[0] == "sphere":
            ctx.add_sphere(s[1], s[2], s[3], s[4])
        else:
            d = s[1]
            ctx.add_mesh(d["positions"], d["indices"], d["normals"], d["texcoords"], d["tangents"], d["bsdf"],
                         uv_tangents=d.get("uv_tangents", False))
    for d, tex in fs.textures:
        ctx.add_texture(d, tex)
    if fs.pool:
        ctx.set_float_pool(np.asarray(fs.pool, dtype=F))
    ctx.set_bsdfs(fs.nodes)
    for d, tex in fs.emitters:
        ctx.add_emitter(d, tex)
    ctx.build()
    return ctx
