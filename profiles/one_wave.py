"""Render ONE wave of a bench.py workload twice (first = warm-up) so that ncu can capture the second:
  ncu --set full --clock-control none --import-source on -k regex:'k_' --launch-skip L --launch-count L \
      -o gpurun_out/wave python profiles/one_wave.py cornell_1080p_256spp_d8 16
L = launches of one wave: fused 1 + (depth-1) + 1 resolve; staged 1 + 2*depth + 1.
usage: python profiles/one_wave.py WORKLOAD SPP [key=value ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parallelraytracing_b200.capi import Context  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
sd, cam, w = bench.make_workload(name)
c = Context(0)
for kv in sys.argv[3:]:
    k, v = kv.split("=")
    c.set_option(k, int(v))
c.set_scene(sd); c.build_bvh(); c.set_camera(cam)
for i in range(2):
    c.clear_accum()
    c.render(spp, w["depth"], seed=1 + i)
    st = c.stats()
    print("render %d: %.3f ms, %d rays, %d launches, %.1f Mrays/s" % (i, st.render_ms, st.rays_total, st.kernel_launches,
                                                                     st.rays_total / st.render_ms / 1e3))
