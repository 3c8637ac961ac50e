"""Wall-clock breakdown of B200Renderer.Init's parts on the 10 M-triangle soup: set_scene (H2D of the packed host
arrays from pageable memory + k_expand_tris) and build_bvh, three times each.  usage: python profiles/init_breakdown.py [n]"""
import sys, time; sys.path.insert(0, '/root/repo')
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
sd = S.random_triangle_soup(n)
c = Context(0)
for i in range(4):
    t0 = time.perf_counter(); c.set_scene(sd); t1 = time.perf_counter(); c.build_bvh(); t2 = time.perf_counter()
    print(f"set_scene {1e3 * (t1 - t0):.1f} ms  build_bvh wall {1e3 * (t2 - t1):.1f} ms (device {c.stats().build_ms:.2f} ms)")
