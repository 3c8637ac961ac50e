"""Run by tests/test_checked_build.py with PRT_CHECKED=1: triangles with NaN and +-inf vertices among 3000 ordinary ones.
Both builders, both node forms, trace batches and renders: nothing may hang or trip a bounds check, the non-finite
triangles are never hit and every walk still equals the list loop on the rest."""
import sys

import numpy as np

import util
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context
from test_fuzz_parity import same_records

rng = np.random.default_rng(5)
nt = 3000
c = rng.uniform(-10, 10, (nt, 1, 3))
good = (c + rng.standard_normal((nt, 3, 3)) * 0.5).astype(np.float32)
tv = good.copy()
tv[0:30, rng.integers(0, 3, 30), rng.integers(0, 3, 30)] = np.nan
tv[30:40] = np.nan
tv[40:70, rng.integers(0, 3, 30), rng.integers(0, 3, 30)] = np.inf
tv[70:100, rng.integers(0, 3, 30), rng.integers(0, 3, 30)] = -np.inf
bad = ~np.isfinite(tv).all(axis=(1, 2))
with np.errstate(all="ignore"):
    b = S.SceneBuilder("nonfinite"); b.add_triangles(tv, None, b.lambertian((0.5, 0.5, 0.5))); sd = b.build()
bg = S.SceneBuilder("finite"); bg.add_triangles(good, None, bg.lambertian((0.5, 0.5, 0.5))); sdg = bg.build()
rays = np.concatenate([util.fixed_ray_batch(sdg, 20000, seed=3), util.incoherent_rays(20000, 10.0, seed=4)])
ok = True
with Context(0) as ctx:
    for builder in (1, 0):
        for wide in (1, 0):
            ctx.set_option("bvh_builder", builder); ctx.set_option("bvh_wide", wide)
            ctx.set_scene(sd); ctx.build_bvh()
            brute = ctx.trace_batch(rays, use_bvh=0)
            hit = brute["prim_id"] >= 0
            ok &= int(hit.sum()) > 10000 and not bad[brute["prim_id"][hit]].any()
            for mode in (1, 2):
                ok &= same_records(ctx.trace_batch(rays, use_bvh=mode), brute)
            ctx.set_camera(S.Camera(pos=(0.0, 0.0, 40.0), front=(0.0, 0.0, -1.0), width=64, height=48))
            for fused in (0, 1):
                ctx.set_option("fused", fused); ctx.clear_accum(); ctx.render(2, 6, seed=1)
                ok &= bool(np.isfinite(ctx.read_accum()).all())
            ctx.set_option("fused", -1)
    print("nonfinite ok %d check_flags 0x%x" % (ok, ctx.check_flags()), flush=True)
sys.exit(0 if ok else 1)
