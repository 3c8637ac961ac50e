"""Where should the in-thread tail kernel start? 1080p, 1 spp (and 2), depth 20, fused scenes: wall ms per call for
tail_bounce = 0 (off), 1..6, next to the fraction of paths still alive at each bounce.
usage: python profiles/tail_sweep.py [SPP]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
from parallelraytracing_b200 import scene as S  # noqa: E402
from parallelraytracing_b200.capi import Context  # noqa: E402

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 1
pos = np.asarray((5.0, 5.0, 8.0), np.float32)
cam = S.Camera(pos=(5.0, 5.0, 8.0), front=S._normalize(-pos), width=1920, height=1080)
scenes = [S.cornell(), S.closed_cornell(), S.make_scene(S.ScenePreset.LIGHT_TEST), S.make_scene(S.ScenePreset.MATERIAL_TEST),
          S.make_scene(S.ScenePreset.DEFAULT)]
c = Context(0)
c.set_option("tail_max_paths", 64 << 20)
for sd in scenes:
    c.set_scene(sd); c.build_bvh(); c.set_camera(cam)
    row = []
    for tail in range(0, 7):
        c.set_option("tail_bounce", tail); c.set_option("tail_adaptive", 0)
        for i in range(5):
            c.render(spp, 20, seed=1, sample_begin=i)
        t0 = time.perf_counter()
        for i in range(40):
            c.render(spp, 20, seed=1, sample_begin=5 + i)
        row.append((time.perf_counter() - t0) / 40 * 1e3)
    st = c.stats()
    r = list(st.rays_per_bounce)
    print("%-16s prims %3d  alive@b1..6 %s  ms: %s  best tail=%d" % (
        sd.name, sd.n_prims, " ".join("%.2f" % (r[b] / r[0]) for b in range(1, 7)), " ".join("%.3f" % x for x in row), int(np.argmin(row))))
