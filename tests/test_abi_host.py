"""CPU tests of the boundary and the host logic: the C-ABI library loads and exports every symbol include/mtscuda.h
declares (no compute calls without a GPU), the ctypes mirrors have the C struct sizes, scene flattening reproduces the
reference's validation / defaults, XML round trip."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import mts_b200 as M
from oracle_lib import OraclePath
from mitsuba0_6_b200 import abi, hostmath as hm, scene as ms, scenes
from mitsuba0_6_b200.scene import Plugin, Point, Spectrum

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mtscuda.h")


def declared_symbols():
    txt = open(HEADER).read()
    return sorted(set(re.findall(r"MTS_FN\((\w+)\)\s*\(", txt)) - {"name"})


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(abi.ABI_SYMBOLS)


@pytest.mark.parametrize("backend", ["cuda", "oracle"])
def test_library_exports_every_declared_symbol(backend):
    import oracle_lib
    path, prefix = {"cuda": (abi.CUDA_LIB_PATH, abi.CUDA_PREFIX), "oracle": (oracle_lib.ORACLE_LIB_PATH, oracle_lib.ORACLE_PREFIX)}[backend]
    lib = ctypes.CDLL(path)       # dlopen only: no CUDA call is made
    for s in declared_symbols():
        assert hasattr(lib, prefix + s), s


def test_struct_sizes_match_the_header(tmp_path):
    names = ["MtsTexRef", "MtsBsdfNode", "MtsTextureDesc", "MtsCamera", "MtsFilm", "MtsMeshDesc", "MtsEmitterDesc",
             "MtsRenderParams", "MtsStats"]
    src = '#include <stdio.h>\n#include "mtscuda.h"\nint main(){' + "".join('printf("%%zu\\n", sizeof(%s));' % n for n in names) + "return 0;}"
    c = tmp_path / "s.c"
    c.write_text(src)
    exe = str(tmp_path / "s")
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", exe])
    sizes = [int(x) for x in subprocess.check_output([exe]).split()]
    mirrors = [abi.TexRef, abi.BsdfNode, abi.TextureDesc, abi.Camera, abi.Film, abi.MeshDesc, abi.EmitterDesc,
               abi.RenderParams, abi.Stats]
    assert sizes == [ctypes.sizeof(m) for m in mirrors]


def test_cuda_library_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(M.MtsError, match="no CUDA device|CPU fallback"):
        M.CudaPath(scenes.cornell_box(8, 8, 1))


def test_integrator_parameter_validation():
    """MonteCarloIntegrator ctor, src/librender/integrator.cpp:190-225."""
    sc = scenes.cornell_box(8, 8, 1)
    sc.child("integrator").props.update(rrDepth=0)
    with pytest.raises(ValueError, match="rrDepth"):
        M.flatten(sc)
    sc.child("integrator").props.update(rrDepth=5, maxDepth=0)
    with pytest.raises(ValueError, match="maxDepth"):
        M.flatten(sc)
    sc.child("integrator").props.update(maxDepth=-1)
    assert M.flatten(sc).params["max_depth"] == -1
    cpu = OraclePath(scenes.cornell_box(8, 8, 1))
    with pytest.raises(M.MtsError, match="rrDepth"):
        cpu.render(spp=1, rr_depth=0)


def test_out_of_scope_plugins_are_rejected():
    sc = scenes.cornell_box(8, 8, 1)
    sc.add(Plugin("shape", "cylinder"))
    with pytest.raises(M.Unsupported):
        M.flatten(sc)
    sc = scenes.cornell_box(8, 8, 1)
    sc.children = [c for c in sc.children if not (c[1].kind == "shape" and c[1].child("emitter"))]
    with pytest.raises(M.Unsupported, match="no emitters"):       # the reference would add a sunsky (scene.cpp:358-373)
        M.flatten(sc)
    sc = scenes.cornell_box(8, 8, 1)
    sc.child("sensor").child("sampler").type = "sobol"
    with pytest.raises(M.Unsupported):
        M.flatten(sc)


def test_emitter_order_matches_scene_initialize():
    """Scene::m_emitters: scene-level emitters first, then area emitters in shape order (scene.cpp:322-336, 497-516)."""
    import util
    sc, _ = util.zoo_scene()
    flat = M.flatten(sc)
    types = [d.type for d, _ in flat.emitters]
    assert types == [abi.EMITTER_POINT, abi.EMITTER_ENVMAP, abi.EMITTER_AREA, abi.EMITTER_AREA, abi.EMITTER_AREA]


def test_cube_matches_reference_tables():
    """src/shapes/cube.cpp:22-28: first face is -y with vertices (1,-1,-1) (1,-1,1) (-1,-1,1) (-1,-1,-1)."""
    pos, nrm, uv, idx = hm.cube_mesh(hm.identity())
    assert pos.shape == (24, 3) and idx.shape == (12, 3)
    assert np.array_equal(pos[:4], [[1, -1, -1], [1, -1, 1], [-1, -1, 1], [-1, -1, -1]])
    assert np.array_equal(pos[20:24], [[1, 1, -1], [1, -1, -1], [-1, -1, -1], [-1, 1, -1]])
    assert np.array_equal(uv[:4], [[0, 1], [1, 1], [1, 0], [0, 0]]) and np.array_equal(idx[:2], [[0, 1, 2], [3, 0, 2]])
    assert np.array_equal(nrm[8], [1, 0, 0]) and np.array_equal(nrm[12], [0, 0, 1])


def test_filter_tables():
    """rfilter.cpp:37-56: 31 samples + a trailing zero, normalised so that the Riemann sum is one."""
    for kind in ("box", "gaussian"):
        r, t = hm.filter_table(kind)
        t = np.asarray(t, dtype=np.float64)
        assert len(t) == 32 and t[31] == 0
        assert abs(2 * t[:31].sum() * r / 31 - 1) < 1e-5
    assert abs(hm.filter_table("box")[0] - 0.5) < 1e-4 and hm.filter_table("gaussian")[0] == 2.0


def test_xml_round_trip(tmp_path):
    sc = scenes.cornell_box(16, 16, 2)
    path = ms.to_xml(sc, str(tmp_path / "cbox.xml"))
    sc2 = ms.load_xml(path)
    f1, f2 = M.flatten(sc), M.flatten(sc2)
    assert f1.params == f2.params and f1.spp == f2.spp and len(f1.shapes) == len(f2.shapes)
    assert np.allclose(list(f1.camera.sample_to_camera), list(f2.camera.sample_to_camera))
    a = OraclePath(f1).render(spp=2, seed=1)
    b = OraclePath(f2).render(spp=2, seed=1)
    assert np.allclose(a, b, atol=1e-5)


def test_conductor_presets_and_ior_names():
    assert hm.lookup_ior("bk7", 1.0) == pytest.approx(1.5046)
    assert hm.lookup_ior("water", 1.0) == pytest.approx(1.3330)
    p = Plugin("bsdf", "conductor", material="Au")
    sc = scenes.cornell_box(8, 8, 1)
    sc.children[3][1].children = [("", p)]
    flat = M.flatten(sc)
    node = [n for n in flat.nodes if n.type == abi.BSDF_CONDUCTOR][0]
    # eta, k are divided by the exterior index (air, 1.000277): src/bsdfs/conductor.cpp:161-176
    assert np.allclose(list(node.eta), np.array(ms.CONDUCTOR_PRESETS["au"][0]) / 1.000277, atol=1e-6)
