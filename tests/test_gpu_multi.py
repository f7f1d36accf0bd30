"""Multi-device groups (mtscuda_create_multi / cudapath's `devices` parameter): the devices of a box split the sample
indices of every pixel; the union is the single-device sample set, so the reduced film equals the single-device film up
to the order of the float additions.  Needs two GPUs (gpurun --gpus 2); skipped on a one-GPU box."""
import os
import sys

import numpy as np
import pytest

import mts_b200 as M
from mitsuba0_6_b200 import scenes
import golden_io as G
import util

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


needs2 = pytest.mark.skipif("_gpus() < 2", reason="needs two GPUs")


@needs2
@pytest.mark.parametrize("make", [lambda: scenes.cornell_box(128, 128, 16), lambda: util.zoo_scene(96, 64, 8)[0],
                                  lambda: util.sphere_alpha_scene(spp=6)], ids=["cornell", "zoo", "alpha"])
def test_two_devices_render_the_single_device_sample_set(make):
    sc = make()
    one = M.CudaPath(sc)
    two = M.CudaPath(sc, devices=[0, 1])
    assert two.ctx.device_count() == 2
    a = one.render(seed=3)
    b = two.render(seed=3)
    sa, sb = one.stats, two.stats
    # identical sample set: identical integer statistics
    assert (sa.camera_samples, sa.normal_rays, sa.shadow_rays, sa.path_length_sum) == (sb.camera_samples, sb.normal_rays, sb.shadow_rays, sb.path_length_sum)
    scale = np.abs(a).max()
    assert np.abs(a - b).max() <= 2e-5 * scale, np.abs(a - b).max() / scale
    if one.flat.film.has_alpha:
        assert np.abs(one.ctx.read_film_alpha() - two.ctx.read_film_alpha()).max() <= 2e-5 * scale
    # a second frame into the same contexts (film cleared on every device, wavefront kept)
    b2 = two.render(seed=4)
    a2 = one.render(seed=4)
    assert np.abs(a2 - b2).max() <= 2e-5 * np.abs(a2).max()
    # odd sample counts: the remainder goes to the low devices; fewer samples than devices leaves a device idle
    for spp in (5, 1):
        x, y = one.render(spp=spp, seed=9), two.render(spp=spp, seed=9)
        assert one.stats.camera_samples == two.stats.camera_samples
        assert np.abs(x - y).max() <= 2e-5 * np.abs(x).max()
    one.close()
    two.close()


@needs2
def test_two_devices_mip_scene_uses_the_job_sample_count():
    """The ray differentials are scaled by the sample count of the whole job, not of one device's share."""
    sc = G.mip(spp=8)[0]
    one, two = M.CudaPath(sc), M.CudaPath(sc, devices=[0, 1])
    a, b = one.render(seed=1), two.render(seed=1)
    assert np.abs(a - b).max() <= 2e-5 * np.abs(a).max()
    one.close()
    two.close()


@needs2
def test_cudapath_plugin_with_two_devices_in_reference_host():
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "refbuild"))
    import mtsref
    if not os.path.exists(os.path.join(mtsref.REF_DIR, "plugins", "cudapath.so")):
        pytest.skip("oracle/_ref (reference build + cudapath.so) is not present")
    gold = G.load("images.npz")
    sc = G.IMAGE_SCENES["cornell"]()
    sc.child("integrator").props["devices"] = 2
    rs = mtsref.RefScene(sc, integrator="cudapath")
    img = M.develop(rs.render())
    ref = gold["image/cornell/rgb"].astype(np.float32)
    err = M.rel_mse(img, ref)
    print("cudapath.so, devices=2, inside the reference host: relMSE(cornell) = %.3g" % err)
    assert abs(img.mean() / ref.mean() - 1) < 0.02 and err < 1e-3


def test_bad_device_lists_are_rejected():
    with pytest.raises(M.MtsError, match="twice"):
        M.CudaPath(scenes.cornell_box(8, 8, 1), devices=[0, 0])
    with pytest.raises(M.MtsError, match="bad device"):
        M.CudaPath(scenes.cornell_box(8, 8, 1), devices=[0, 99])
