"""mtscuda_trace_noalloc (no counterpart in the reference): the trace entry
point that works in the library's fixed device scratch instead of allocating.  Checked through the C ABI against the
existing mtscuda_trace on the same rays -- same kernels, so the results must be identical bit for bit -- on calls in a
row that share the scratch (closest hit, then a shorter shadow batch, then closest hit again).  Collected last on
purpose."""
import numpy as np
import pytest

import mts_b200 as M
from mitsuba0_6_b200 import abi
import golden_io as G
import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def zoo_gpu():
    sc, _ = util.zoo_scene(96, 64, 4)
    gpu = M.CudaPath(sc)
    gpu.preprocess()
    yield gpu
    gpu.close()


def test_trace_noalloc_equals_trace_on_consecutive_calls(zoo_gpu):
    ctx = zoo_gpu.ctx
    rays = G.trace_rays()                       # 24000 seeded chords through the zoo scene
    n_max = abi.NOALLOC_MAX_RAYS
    a, b, c = rays[:n_max], rays[n_max:n_max + 5000], rays[3000:3000 + 7777]
    # call 1: a full batch of closest-hit rays
    h_ref, id_ref = ctx.trace(a)
    h, ids = ctx.trace_noalloc(a)
    assert np.array_equal(ids, id_ref) and np.array_equal(h, h_ref)
    assert np.isfinite(h[:, 0]).mean() > 0.5    # the batch really hits things
    # call 2 (same scratch, different kind and size): shadow rays
    sb = G.shadow_rays(b, ctx.trace(b)[0])
    s_ref, _ = ctx.trace(sb, shadow=True)
    s, _ = ctx.trace_noalloc(sb, shadow=True)
    assert np.array_equal(s[:, 0], s_ref[:, 0])
    assert 0 < (s[:, 0] == 0).mean() < 1        # both verdicts occur
    # call 3: closest hit again, ragged size -- nothing of calls 1 and 2 may leak through the shared scratch
    h_ref, id_ref = ctx.trace(c)
    h, ids = ctx.trace_noalloc(c)
    assert np.array_equal(ids, id_ref) and np.array_equal(h, h_ref)
    # and the existing entry point is unaffected by the calls in between
    h2, id2 = ctx.trace(a)
    h3, id3 = ctx.trace_noalloc(a)
    assert np.array_equal(h2, h3) and np.array_equal(id2, id3)


def test_trace_noalloc_limits(zoo_gpu):
    ctx = zoo_gpu.ctx
    rays = G.trace_rays()
    with pytest.raises(abi.MtsError, match="exceed the fixed scratch"):
        ctx.trace_noalloc(rays[:abi.NOALLOC_MAX_RAYS + 1])
    h, ids = ctx.trace_noalloc(rays[:0])        # empty batch: nothing to do
    assert h.shape == (0, 4) and ids.shape == (0, 2)
    h, ids = ctx.trace_noalloc(rays[:1])
    hr, ir = ctx.trace(rays[:1])
    assert np.array_equal(h, hr) and np.array_equal(ids, ir)
