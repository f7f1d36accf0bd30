"""mitsuba0.6_b200 -- host side of the B200-native `cudapath` integrator.

The reference's boundary for this path is the Integrator plugin interface
(include/mitsuba/render/integrator.h:61-118: preprocess / render / cancel / postprocess and the
four MonteCarloIntegrator parameters).  `CudaPath` mirrors it in Python over the C ABI of
include/mtscuda.h; plugin/cudapath.cpp is the C++ plugin a Mitsuba tree builds against the
same library.  The directory name contains a dot, so import it through the `mts_b200` shim at
the repository root.
"""
import ctypes as _C

import numpy as np

from . import abi, hostmath, scene as _scene, sharding
from .abi import Context, MtsError
from .scene import Plugin, Spectrum, Point, load_xml, to_xml, flatten, upload, Unsupported

__all__ = ["CudaPath", "Context", "MtsError", "Plugin", "Spectrum", "Point", "load_xml", "to_xml",
           "flatten", "upload", "Unsupported", "develop", "rel_mse", "abi", "hostmath", "sharding"]


def develop(film_rgbw):
    """Film::develop: divide the accumulated colour by the accumulated filter weight
    (src/libcore/fmtconv.cpp:955-975)."""
    f = np.asarray(film_rgbw, dtype=np.float32)
    w = f[..., 3:4]
    return np.divide(f[..., :3], w, out=np.zeros_like(f[..., :3]), where=w > 0)


def rel_mse(img, ref, eps=1e-2, trim=0.001):
    """Firefly-robust relative MSE used for converged-image parity: per-pixel
    (a-b)^2 / (b^2 + eps) on luminance-free RGB, with the top `trim` fraction of pixels dropped."""
    a, b = np.asarray(img, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    e = ((a - b) ** 2 / (b ** 2 + eps)).mean(axis=-1).reshape(-1)
    k = int(len(e) * (1.0 - trim))
    return float(np.sort(e)[:max(k, 1)].mean())


class CudaPath:
    """Integrator-plugin mirror: CudaPath(scene).render() == `mitsuba scene.xml` with
    <integrator type="cudapath">.  Runs on libmtscuda (sm_100a) only; there is no CPU path.

    `lib` lets the test infrastructure drive another library that exports the same C ABI (the CPU oracle,
    oracle/oracle_lib.py) through these wrappers; the package itself never passes it.
    """

    def __init__(self, scene, device=0, lib=None, devices=None):
        """devices: list of GPUs of this box that share the render (cudapath's `devices` parameter counts them from
        `device`); None / one entry = a single GPU."""
        if isinstance(scene, str):
            scene = load_xml(scene)
        self.scene = scene
        self.flat = scene if isinstance(scene, _scene.FlatScene) else flatten(scene)
        if devices is None and int(self.flat.params.get("devices", 1)) > 1:
            devices = [int(self.flat.params.get("device", device)) + i for i in range(int(self.flat.params["devices"]))]
        self.ctx = Context(device, lib, devices=devices if devices and len(devices) > 1 else None)
        self._cancel = _C.c_int(0)
        self.preprocessed = False
        self.stats = None

    # Integrator::preprocess -- scene upload + acceleration structure build
    def preprocess(self):
        upload(self.flat, self.ctx)
        self.preprocessed = True
        return True

    # Integrator::cancel (may be called from another thread)
    def cancel(self):
        self._cancel.value = 1

    def params(self, spp=None, seed=0, sample_offset=0, sample_stride=1, clear_film=True, wavefront_size=0, instrument=False,
               sample_count=0, **over):
        """sample_count: Sampler::getSampleCount() of the whole job (scales the camera-ray differentials,
        integrator.cpp:144-145); 0 = spp * sample_stride, i.e. this call renders the job alone or as one of `stride` ranks."""
        p = dict(self.flat.params)
        p.update(over)
        return abi.RenderParams(int(self.flat.spp if spp is None else spp), int(sample_offset), int(sample_stride),
                                int(p["max_depth"]), int(p["rr_depth"]), int(p["strict_normals"]),
                                int(p["hide_emitters"]), int(seed), int(clear_film), int(wavefront_size), int(instrument),
                                int(sample_count))

    # Integrator::render -- returns the film as (H, W, 4) = sum w*RGB, sum w
    def render(self, spp=None, read_back=True, **kw):
        if not self.preprocessed:
            self.preprocess()
        self._cancel.value = 0
        self.stats = self.ctx.render(self.params(spp, **kw), _C.byref(self._cancel))
        return self.ctx.read_film() if read_back else None

    def image(self, **kw):
        return develop(self.render(**kw))

    def close(self):
        self.ctx.close()
