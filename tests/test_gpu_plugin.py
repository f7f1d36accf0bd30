"""The drop-in boundary for real: plugin/cudapath.cpp, compiled against the reference sources (oracle/_ref), is loaded
by the REFERENCE's own PluginManager as <integrator type="cudapath">, flattens the live reference object graph, renders
through libmtscuda on the GPU and hands the film back through Film::put -- the unchanged reference then develops it.
Compared with the reference's own `path` renders (tests/golden/images.npz)."""
import os
import sys

import numpy as np
import pytest

import mts_b200 as M
import golden_io as G

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle", "refbuild"))
import mtsref  # noqa: E402

needs_ref = pytest.mark.skipif(not os.path.exists(os.path.join(mtsref.REF_DIR, "plugins", "cudapath.so")),
                               reason="oracle/_ref (reference build + cudapath.so) is not present")


@needs_ref
@pytest.mark.parametrize("name,bar", [("cornell", 1e-3), ("zoo", 1.2e-3), ("terrain", 1e-3), ("spheres", 1.5e-3)])   # measured 7.8e-5, 7.9e-4, 2.6e-4, 9.3e-4
def test_cudapath_plugin_in_reference_host(name, bar):
    gold = G.load("images.npz")
    sc = G.IMAGE_SCENES[name]()
    rs = mtsref.RefScene(sc, integrator="cudapath")          # the reference's Scene + RenderJob drive our plugin
    img = M.develop(rs.render())
    ref = gold["image/%s/rgb" % name].astype(np.float32)
    assert np.isfinite(img).all() and img.mean() > 0
    err = M.rel_mse(img, ref)
    print("cudapath.so inside the reference host: relMSE(%s) = %.3g" % (name, err))
    assert abs(img.mean() / ref.mean() - 1) < 0.02
    assert err < bar, (name, err)


@needs_ref
@pytest.mark.parametrize("name,bar", [("twosided", 1e-3), ("thin_hide", 1.5e-3)])   # measured 5.7e-5, 7.7e-4
def test_cudapath_plugin_next_row_bsdfs(name, bar):
    # TwoSidedBRDF / ThinDielectric / Mask parsed from the live reference objects (InstanceManager serialization);
    # thin_hide also passes hideEmitters through the plugin's Properties
    gold = G.load("twosided.npz")
    sc = G.TWOSIDED_IMAGE_SCENES[name]()
    rs = mtsref.RefScene(sc, integrator="cudapath")
    img = M.develop(rs.render())
    ref = gold["image/%s/rgb" % name].astype(np.float32)
    assert np.isfinite(img).all() and img.mean() > 0
    err = M.rel_mse(img, ref)
    print("cudapath.so inside the reference host: relMSE(%s) = %.3g" % (name, err))
    assert abs(img.mean() / ref.mean() - 1) < 0.02
    assert err < bar, (name, err)


@needs_ref
@pytest.mark.parametrize("name,bar", [("crop", 1.5e-3), ("sphere_alpha", 1.5e-3)])
def test_cudapath_plugin_boundary_cases(name, bar):
    # crop window (the block handed to Film::put sits at offset 0 of the window), Sphere `center` under a scaled
    # toWorld, and the alpha channel of an rgba film -- all through the live reference objects
    gold = G.load("boundary.npz")
    sc = G.BOUNDARY_IMAGE_SCENES[name]()
    rs = mtsref.RefScene(sc, integrator="cudapath")
    film, alpha = rs.render(alpha=True)
    img = M.develop(film)
    ref = gold["image/%s/rgb" % name].astype(np.float32)
    assert img.shape == ref.shape
    err = M.rel_mse(img, ref)
    print("cudapath.so inside the reference host: relMSE(%s) = %.3g" % (name, err))
    assert abs(img.mean() / ref.mean() - 1) < 0.02 and err < bar, (name, err)
    w = film[..., 3]
    a = np.divide(alpha, w, out=np.zeros_like(w), where=w > 0)
    ra = gold["image/%s/alpha" % name].astype(np.float32)
    assert np.abs(a - ra).mean() < 5e-3, np.abs(a - ra).mean()


@needs_ref
@pytest.mark.parametrize("exact", [True, False])
def test_cudapath_plugin_default_filter_type_scene(exact, monkeypatch):
    """Stock bitmap textures (filterType ewa, maxAnisotropy 20) and the filtered environment map through the plugin:
    usesRayDifferentials() of the live BSDFs decides which hits get uv partials.  Both texture routes of the plugin:
    the re-decoded image file (`exactTextures`, the default) and level 0 of the live MIP map as half texels."""
    if not exact:
        monkeypatch.setenv("MTS_CUDAPATH_FAST_TEXTURES", "1")
    gold = G.load("mip.npz")
    sc = G.mip()[0]
    rs = mtsref.RefScene(sc, integrator="cudapath")
    img = M.develop(rs.render())
    ref = gold["image/mip/rgb"].astype(np.float32)
    err = M.rel_mse(img, ref)
    print("cudapath.so inside the reference host: relMSE(mip%s) = %.3g" % ("" if exact else ", fast texture route", err))
    assert abs(img.mean() / ref.mean() - 1) < 0.02 and err < 3e-3, err


@needs_ref
def test_cudapath_plugin_parameter_errors():
    sc = G.cornell_small(spp=1)
    sc.child("integrator").props["rrDepth"] = 0
    with pytest.raises(mtsref.RefError, match="rrDepth"):
        mtsref.RefScene(sc, integrator="cudapath")
