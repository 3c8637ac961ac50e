"""Time the GPU BVH build (device time, CUDA events inside prt_build_bvh) of an n-triangle soup or the 960 K dragon scene.
usage: python profiles/build_only.py [n | dragon960k] [key=value ...]   (e.g. bvh_builder=0 for the Karras hierarchy)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from parallelraytracing_b200 import scene as S  # noqa: E402
from parallelraytracing_b200.capi import Context  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "10000000"
sd = bench.make_workload("dragon960k_1080p_64spp_d8")[0] if what == "dragon960k" else S.random_triangle_soup(int(what))
c = Context(0)
for kv in sys.argv[2:]:
    k, v = kv.split("=")
    c.set_option(k, int(v))
c.set_scene(sd)
for i in range(4):
    c.build_bvh()
    st = c.stats()
    print("%s %s: build_ms %.3f builder %d height %d rounds %d" % (what, " ".join(sys.argv[2:]), st.build_ms, st.bvh_builder, st.bvh_height, st.ploc_rounds), flush=True)
cnt, us = c.ploc_trace()
if len(us):
    print("PLOC rounds (clusters at start : microseconds):")
    print("  " + "  ".join("%d:%.1f" % (int(a), b) for a, b in zip(cnt[:-1], us)))
    print("  rounds with > 1 tile: %.1f us, single-CTA rounds: %.1f us" % (us[cnt[:-1] > 1024].sum(), us[cnt[:-1] <= 1024].sum()))
