#!/usr/bin/env python
"""Wall-clock breakdown of the plugin-facing call sequence (upload, build, render, read-back)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import mts_b200 as M
from mitsuba0_6_b200 import abi, scene as ms
import numpy as np

sc, flat, kw, desc, gen = bench.build_scene(sys.argv[1] if len(sys.argv) > 1 else "terrain", int(sys.argv[2]) if len(sys.argv) > 2 else 16)
for it in range(3):
    t0 = time.time(); ctx = abi.Context("cuda"); t1 = time.time()
    ctx.set_camera(flat.camera); ctx.set_film(flat.film)
    tt = time.time()
    for d, tex in flat.textures: ctx.add_texture(d, tex)
    t_tex = time.time() - tt
    if flat.pool: ctx.set_float_pool(np.asarray(flat.pool, dtype=np.float32))
    ctx.set_bsdfs(flat.nodes)
    tt = time.time()
    for s in flat.shapes:
        if s[0] == "rectangle": ctx.add_rectangle(s[1], s[2])
        elif s[0] == "sphere": ctx.add_sphere(s[1], s[2], s[3], s[4])
        else:
            d = s[1]; ctx.add_mesh(d["positions"], d["indices"], d["normals"], d["texcoords"], d["tangents"], d["bsdf"], uv_tangents=d.get("uv_tangents", False))
    t_mesh = time.time() - tt
    tt = time.time()
    for d, tex in flat.emitters: ctx.add_emitter(d, tex)
    t_em = time.time() - tt
    tt = time.time(); ctx.build(); t_build = time.time() - tt
    p = abi.RenderParams(kw["spp"], 0, 1, 8, 5, 0, 0, it, 1, 0, 0)
    tt = time.time(); st = ctx.render(p); t_render = time.time() - tt
    tt = time.time(); film = ctx.read_film(); t_read = time.time() - tt
    tt = time.time(); ctx.close(); t_close = time.time() - tt
    print("create %.0f tex %.0f mesh %.0f emit %.0f build %.0f (dev %.0f) render %.0f (dev %.0f) read %.0f close %.0f | total %.0f ms"
          % ((t1 - t0) * 1e3, t_tex * 1e3, t_mesh * 1e3, t_em * 1e3, t_build * 1e3, st.build_ms, t_render * 1e3, st.render_ms, t_read * 1e3, t_close * 1e3, (time.time() - t0) * 1e3))
