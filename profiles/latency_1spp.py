"""Latency of ONE ProgressiveRender-sized call (1 spp, depth 20 = the reference's literals) through the C ABI:
wall time per prt_render call, device time between its two events, kernel launches.
usage: python profiles/latency_1spp.py [WORKLOAD] [SPP] [DEPTH] [key=value ...]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parallelraytracing_b200.capi import Context  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 20
sd, cam, w = bench.make_workload(name)
c = Context(0)
for kv in sys.argv[4:]:
    k, v = kv.split("=")
    c.set_option(k, int(v))
c.set_scene(sd); c.build_bvh(); c.set_camera(cam)
for i in range(5):
    c.render(spp, depth, seed=1, sample_begin=i)
N = 50
dev = 0.0
t0 = time.perf_counter()
for i in range(N):
    c.render(spp, depth, seed=1, sample_begin=5 + i)
    dev += c.stats().render_ms
wall = (time.perf_counter() - t0) / N * 1e3
st = c.stats()
print("%s %d spp depth %d: wall %.4f ms/call, device %.4f ms/call, %d launches, %d rays/call -> %.0f spp/s, %.1f Mrays/s"
      % (name, spp, depth, wall, dev / N, st.kernel_launches, st.rays_total, 1e3 * spp / wall, st.rays_total / wall / 1e3))
