#!/usr/bin/env python
"""Small fixed-size render used under ncu (launch lists / full captures): one render() of a bench workload at a
reduced spp.  Never a bench number."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="terrain")
    ap.add_argument("--spp", type=int, default=2)
    ap.add_argument("--renders", type=int, default=1)
    ap.add_argument("--wavefront", type=int, default=0)
    ap.add_argument("--sample-count", type=int, default=0, help="sample count of the whole job (scales the ray differentials): 128 = the footprints of the bench configuration at a reduced spp")
    a = ap.parse_args()
    import mts_b200 as M
    sc, flat, kw, desc, gen_s = bench.build_scene(a.workload, a.spp)
    gpu = M.CudaPath(flat)
    gpu.preprocess()
    for i in range(a.renders):
        gpu.render(spp=a.spp, read_back=False, seed=i, wavefront_size=a.wavefront, sample_count=a.sample_count)
        st = gpu.stats
        print("render %d: %.2f ms, %.1f Msamples/s, %.1f Mrays/s | raygen %.2f intersect %.2f shade %.2f shadow %.2f splat %.2f ms | launches %d"
              % (i, st.render_ms, st.camera_samples / st.render_ms / 1e3, (st.normal_rays + st.shadow_rays) / st.render_ms / 1e3,
                 st.raygen_ms, st.intersect_ms, st.shade_ms, st.shadow_ms, st.splat_ms, st.kernel_launches))


if __name__ == "__main__":
    main()
