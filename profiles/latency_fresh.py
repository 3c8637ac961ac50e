"""ONE ProgressiveRender-sized call the way the C++ shim issues it: prt_render_fresh(1 spp), depth 20 (the reference's literals).
usage: python profiles/latency_fresh.py [WORKLOAD] [N] [key=value ...]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parallelraytracing_b200.capi import Context  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
N = int(sys.argv[2]) if len(sys.argv) > 2 else 50
sd, cam, w = bench.make_workload(name)
c = Context(0)
for kv in sys.argv[3:]:
    k, v = kv.split("=")
    c.set_option(k, int(v))
c.set_scene(sd); c.build_bvh(); c.set_camera(cam)
for i in range(5):
    c.render_fresh(1, 20, seed=1, sample_begin=i)
dev = 0.0
t0 = time.perf_counter()
for i in range(N):
    c.render_fresh(1, 20, seed=1, sample_begin=5 + i)
    dev += c.stats().render_ms
wall = (time.perf_counter() - t0) / N * 1e3
st = c.stats()
print("%s fresh 1 spp depth 20 %s: wall %.4f ms/call, device %.4f ms/call, %d launches, %d rays/call, per bounce %s"
      % (name, " ".join(sys.argv[3:]), wall, dev / N, st.kernel_launches, st.rays_total, list(st.rays_per_bounce)[:8]))
