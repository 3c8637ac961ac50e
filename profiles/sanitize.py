"""Small configurations of every kernel family for `compute-sanitizer`:
  compute-sanitizer --tool memcheck  python profiles/sanitize.py
  compute-sanitizer --tool racecheck python profiles/sanitize.py
  compute-sanitizer --tool synccheck python profiles/sanitize.py
  PRT_CHECKED=1 python profiles/sanitize.py         (the repo's own bounds-checked build: compute-sanitizer is closed on this pool)
Covers: BVH build of a 20 K-triangle soup with both hierarchies (PLOC: cooperative kernel, chained scan in shared memory;
Karras + the fence/atomic refit), the wide-node emission, one staged render (persistent extend kernels, per-warp shade queue,
block compaction, tail kernel), one fused render (list loop) on two streams, the deep adversarial scene (local-memory part of
the traversal stacks), ray re-sorting, the film resolve, the scatter entry."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from parallelraytracing_b200 import scene as S  # noqa: E402
from parallelraytracing_b200.capi import Context, material_dt  # noqa: E402

c = Context(0)
cam = S.Camera(width=96, height=64)
soup = S.random_triangle_soup(20000)
for builder in (1, 0):
    c.set_option("bvh_builder", builder)
    c.set_scene(soup); c.build_bvh()
    c.set_camera(S.Camera(pos=(0.0, 0.0, 150.0), front=(0.0, 0.0, -1.0), width=96, height=64))
    c.clear_accum(); c.render(2, 8, seed=1)                       # staged: raygen, extend_wide, shade, tail, resolve
    print("soup20k builder", builder, "rays", c.stats().rays_total, flush=True)
c.set_option("bvh_builder", 1)
c.set_scene_build(soup)                                             # Init in one call: positions-only ingest / gather, k_finish_tris
c.clear_accum(); c.render(1, 4, seed=7)
c.build_bvh()                                                      # a rebuild from the deferred arrays
c.set_option("sort_rays", 1); c.set_option("sort_min_rays", 0)
c.clear_accum(); c.render(1, 4, seed=2)
c.set_option("sort_rays", 0)
c.trace_batch(util.fixed_ray_batch(soup, 4096), use_bvh=1); c.trace_batch(util.fixed_ray_batch(soup, 4096), use_bvh=2)
c.trace_batch(util.fixed_ray_batch(soup, 256), use_bvh=0)
c.set_scene(S.make_scene(S.ScenePreset.RANDOM_BALLS_SMALL)); c.build_bvh(); c.set_camera(cam)
c.clear_accum(); c.render(2, 8, seed=3)                           # staged, two-child persistent kernel
c.set_scene(S.cornell()); c.build_bvh(); c.set_camera(cam)
c.set_option("wave_paths", 4000)
c.clear_accum(); c.render(3, 8, seed=4)                           # fused list loop, several waves on two streams
c.set_option("wave_paths", 0)
c.render_fresh(1, 20, seed=5)                                     # direct write, tail kernel
c.resolve_display(1.0, 2.2)
c.set_option("bvh_wide", 1)
deep = util.deep_scene(512)
c.set_scene(deep); c.build_bvh()
h = c.trace_batch(util.rays_from_point((0.45, 0.45, 0.45), 2048), use_bvh=2)
c.trace_batch(util.rays_from_point((0.45, 0.45, 0.45), 2048), use_bvh=1)
c.set_camera(S.Camera(pos=(0.45, 0.45, 0.45), front=(0.0, 0.0, -1.0), width=48, height=32)); c.clear_accum(); c.render(1, 4, seed=6)
m = np.zeros(64, material_dt); m["type"] = 1 + np.arange(64) % 4; m["r"] = m["g"] = m["b"] = 0.5; m["param"] = 1.5
d = np.tile(np.array([[0.0, -1.0, 0.0]], np.float32), (64, 1)); nrm = np.tile(np.array([[0.0, 1.0, 0.0]], np.float32), (64, 1))
c.debug_scatter(m, d, nrm, np.ones(64, np.int32), np.zeros((64, 3), np.uint32))
flags = c.check_flags()
# the checker itself must fire when it should: pretend the stacks hold 28 entries - the deep scene needs more
c.set_option("check_stack_limit", 28)
c.trace_batch(util.rays_from_point((0.45, 0.45, 0.45), 2048), use_bvh=2)
provoked = c.check_flags()
c.set_option("check_stack_limit", 256)
c.close()
print("sanitize.py done: hits %d check_flags 0x%x provoked 0x%x" % (int((h["prim_id"] >= 0).sum()), flags, provoked))
