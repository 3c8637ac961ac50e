import sys; sys.path.insert(0,'/root/repo')
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context
n=int(sys.argv[1]) if len(sys.argv)>1 else 10_000_000
sd=S.random_triangle_soup(n)
c=Context(0); c.set_scene(sd)
for i in range(3):
    c.build_bvh(); print("build_ms", c.stats().build_ms)
