"""Seeded fuzz of the closest-hit path on small, nasty scenes: mixed spheres / quads / triangles with degenerate members
(zero-area and needle triangles, exact duplicates, coplanar overlaps, boxes of zero thickness, 1e-4 .. 1e3 scale ratios) and
rays that the slab / plane / watertight tests find awkward (exact-zero and negative-zero direction components, axis-parallel
rays inside face planes, origins ON surfaces, rays through shared vertices and edges).
  CPU (-m "not gpu"): the oracle's BVH walks (two-child, 4-wide) against its own brute-force list loop - the reference's
       algorithm (primitive.cpp:21-59), which the pinned tests anchor - for both builders.
  GPU (-m gpu): prt_trace_batch in all three modes against the oracle, byte for byte, counters included."""
import numpy as np
import pytest

import util
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import ray_dt

SEEDS = list(range(16))


def fuzz_scene(seed):
    rng = np.random.default_rng(1000 + seed)
    b = S.SceneBuilder("fuzz%d" % seed)
    mats = [b.lambertian((0.7, 0.6, 0.5)), b.metal((0.8, 0.8, 0.8), 0.2), b.dielectric(1.5), b.emissive((3, 3, 3))]
    ext = float(10.0 ** rng.uniform(-1, 2))                                    # scene size 0.1 .. 100
    for _ in range(int(rng.integers(0, 12))):
        r = ext * 10.0 ** rng.uniform(-3, -0.5)
        b.add_sphere(r, mats[int(rng.integers(0, 4))], trans=tuple(rng.uniform(-ext, ext, 3)))
    if seed % 3 == 0:                                                         # concentric and touching spheres
        c = tuple(rng.uniform(-ext, ext, 3))
        for k in range(3):
            b.add_sphere(ext * 0.1 * (k + 1), mats[k], trans=c)
    for _ in range(int(rng.integers(0, 8))):
        b.add_quad(ext * rng.uniform(0.05, 1.0), ext * rng.uniform(0.05, 1.0), mats[int(rng.integers(0, 4))],
                   euler=tuple(rng.choice([0.0, 90.0, 180.0, 37.0, -63.5], 3)), trans=tuple(rng.uniform(-ext, ext, 3)))
    nt = int(rng.integers(1, 400))
    c = rng.uniform(-ext, ext, (nt, 1, 3))
    tv = c + rng.standard_normal((nt, 3, 3)) * ext * 10.0 ** rng.uniform(-3, -0.5, (nt, 1, 1))
    k = max(nt // 8, 1)
    tv[0:k, 2] = tv[0:k, 1]                                                   # zero area: two equal vertices
    tv[k:2 * k, 2] = 0.5 * (tv[k:2 * k, 0] + tv[k:2 * k, 1])                  # zero area: collinear
    m = min(k, max(nt - 3 * k, 0))
    tv[2 * k:2 * k + m] = tv[3 * k:3 * k + m]                                 # exact duplicates: the lowest id wins
    tv[4 * k:5 * k, :, 1] = np.float32(c[4 * k:5 * k, 0, 1:2])                # axis-aligned, flat boxes (y extent 0)
    tv[5 * k:6 * k, 1] = tv[5 * k:6 * k, 0] + (tv[5 * k:6 * k, 1] - tv[5 * k:6 * k, 0]) * 1e-5   # needles
    if seed % 2:                                                              # a fan sharing one vertex and its edges
        apex = rng.uniform(-ext, ext, 3)
        ring = apex + ext * 0.3 * np.stack([np.cos(np.arange(9) * 0.7), np.sin(np.arange(9) * 0.7), np.full(9, 0.4)], 1)
        fan = np.stack([np.repeat(apex[None], 8, 0), ring[:-1], ring[1:]], 1)
        tv = np.concatenate([tv, fan])
    b.add_triangles(tv.astype(np.float32), None, rng.integers(0, 4, len(tv)).astype(np.int32))
    if seed % 4 == 1:
        b.add_quad(ext, ext, mats[0], trans=(0, -ext, 0))                      # a quad AFTER the triangle block of the list
    return b.build(), ext, tv.astype(np.float32)


def fuzz_rays(sd, ext, tv, seed, n=6000):
    rng = np.random.default_rng(2000 + seed)
    parts = [util.fixed_ray_batch(sd, n, seed=seed + 1), util.incoherent_rays(n, ext, seed=seed + 2)]
    # rays with exact-zero / negative-zero components, some of them inside the planes of flat boxes
    r = util.incoherent_rays(n, ext, seed=seed + 3)
    axis = rng.integers(0, 3, n)
    for a, f in enumerate(("dx", "dy", "dz")):
        r[f][axis == a] = rng.choice(np.array([0.0, -0.0], np.float32), (axis == a).sum())
    l = np.sqrt(r["dx"].astype(np.float64) ** 2 + r["dy"].astype(np.float64) ** 2 + r["dz"].astype(np.float64) ** 2)
    for f in ("dx", "dy", "dz"):
        r[f] = (r[f] / l).astype(np.float32)
    parts.append(r)
    ax = np.zeros(n, ray_dt)                                                  # axis-parallel rays aimed at vertices (two zero components)
    v = tv.reshape(-1, 3)[rng.integers(0, len(tv) * 3, n)]
    a = rng.integers(0, 3, n); sgn = rng.choice([-1.0, 1.0], n).astype(np.float32)
    o = v.copy(); o[np.arange(n), a] -= sgn * ext * 3
    ax["ox"], ax["oy"], ax["oz"] = o.T
    for k, f in enumerate(("dx", "dy", "dz")):
        ax[f] = np.where(a == k, sgn, np.float32(0.0))
    parts.append(ax)
    # rays through vertices and edge midpoints from random origins
    tgt = np.where(rng.random((n, 1)) < 0.5, v, 0.5 * (v + tv.reshape(-1, 3)[rng.integers(0, len(tv) * 3, n)]))
    o = rng.uniform(-2 * ext, 2 * ext, (n, 3)).astype(np.float32)
    d = (tgt - o).astype(np.float64); d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-30)
    e = np.zeros(n, ray_dt)
    e["ox"], e["oy"], e["oz"] = o.T
    e["dx"], e["dy"], e["dz"] = d.astype(np.float32).T
    parts.append(e)
    return np.concatenate(parts)


def same_records(a, b):
    """Bit-equal hit records, except that any NaN equals any NaN (a needle triangle whose float32 cross product underflows has
    a NaN normal on both sides; the GPU's canonical NaN is 0x7fffffff, x86's 0xffc00000)."""
    for f in a.dtype.names:
        x, y = a[f], b[f]
        if x.dtype.kind == "f":
            ok = (x.view(np.uint32) == y.view(np.uint32)) | (np.isnan(x) & np.isnan(y))
        else:
            ok = x == y
        if not ok.all():
            return False
    return True


def with_bounces(trace, rays, seed):
    """The batch plus the rays that start ON the surfaces it hit (origins exactly on primitives: the tMin rules)."""
    hits = trace(rays)
    fin = np.isfinite(hits["nx"]) & np.isfinite(hits["ny"]) & np.isfinite(hits["nz"]) & np.isfinite(hits["px"])
    return np.concatenate([rays, util.bounce_rays(hits[fin], seed=seed + 7)])


@pytest.mark.parametrize("seed", SEEDS)
def test_oracle_bvh_walks_equal_the_list_loop(oracle, seed):
    sd, ext, tv = fuzz_scene(seed)
    for builder in (1, 0):
        osc = oracle.scene(sd); osc.set_builder(builder, 4)
        rays = with_bounces(lambda r: osc.trace_batch(r, use_bvh=0), fuzz_rays(sd, ext, tv, seed, n=1500), seed)
        brute = osc.trace_batch(rays, use_bvh=0)
        assert (brute["prim_id"] >= 0).sum() > 50
        for mode in (1, 2):
            assert same_records(osc.trace_batch(rays, use_bvh=mode), brute), (seed, builder, mode)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", SEEDS)
def test_gpu_closest_hit_fuzz_vs_oracle(oracle, seed):
    from parallelraytracing_b200.capi import Context
    sd, ext, tv = fuzz_scene(seed)
    ctx = Context(0)
    try:
        for builder in (1, 0):
            ctx.set_option("bvh_builder", builder); ctx.set_option("bvh_wide", 1)      # 4-wide nodes also below wide_min_prims
            ctx.set_scene(sd); ctx.build_bvh()
            osc = oracle.scene(sd); osc.set_builder(builder, 4)
            g, o = ctx.debug_lbvh(), osc.get_lbvh()
            for key in ("morton", "sorted_prim", "node_children", "node_bounds"):
                assert g[key].tobytes() == o[key].tobytes(), (seed, builder, key)
            rays = with_bounces(lambda r: ctx.trace_batch(r, use_bvh=0), fuzz_rays(sd, ext, tv, seed), seed)
            ctx.enable_counters(True)
            for mode in (0, 1, 2):
                cnt = np.zeros(2, np.uint64)
                want = osc.trace_batch(rays, use_bvh=mode, counters=cnt)
                got = ctx.trace_batch(rays, use_bvh=mode); st = ctx.stats()
                assert same_records(got, want), (seed, builder, mode)
                if mode:
                    assert (st.nodes_visited, st.prims_tested) == (int(cnt[0]), int(cnt[1])), (seed, builder, mode)
            ctx.enable_counters(False)
    finally:
        ctx.close()


def room_scene(seed):
    """A closed room of six quads (random materials, one wall emissive) with spheres of every material, a triangle cluster with
    degenerate members and duplicated / coplanar primitives inside; the camera sits in the room, so paths bounce until the
    depth limit. Returns (scene, camera)."""
    rng = np.random.default_rng(4000 + seed)
    b = S.SceneBuilder("room%d" % seed)
    L = float(10.0 ** rng.uniform(-1, 2))
    mats = [b.lambertian(tuple(rng.uniform(0.2, 0.9, 3))), b.metal(tuple(rng.uniform(0.5, 1.0, 3)), float(rng.choice([0.0, 0.3, 1.0]))),
            b.dielectric(float(rng.choice([1.0, 1.33, 1.5, 2.4]))), b.emissive((4, 3, 2)), b.lambertian((0.7, 0.7, 0.7))]
    walls = [((0, 0, 0), (0, -L, 0)), ((180, 0, 0), (0, L, 0)), ((0, 0, 90), (L, 0, 0)), ((0, 0, -90), (-L, 0, 0)),
             ((90, 0, 0), (0, 0, -L)), ((-90, 0, 0), (0, 0, L))]
    for k, (eul, tr) in enumerate(walls):
        b.add_quad(2 * L, 2 * L, mats[3] if k == seed % 6 else mats[int(rng.integers(0, 5))], euler=eul, trans=tr)
    if seed % 2:
        b.add_quad(2 * L, 2 * L, mats[0], euler=walls[0][0], trans=walls[0][1])       # an exact duplicate of the floor: lowest id wins
    for _ in range(int(rng.integers(2, 10))):
        b.add_sphere(L * rng.uniform(0.05, 0.35), mats[int(rng.integers(0, 5))], trans=tuple(rng.uniform(-0.7, 0.7, 3) * L))
    nt = int(rng.integers(4, 120))
    c = rng.uniform(-0.6, 0.6, (nt, 1, 3)) * L
    tv = c + rng.standard_normal((nt, 3, 3)) * L * rng.uniform(0.02, 0.3, (nt, 1, 1))
    tv[0, 2] = tv[0, 1]; tv[1, 2] = 0.5 * (tv[1, 0] + tv[1, 1]); tv[2] = tv[3]
    b.add_triangles(tv.astype(np.float32), None, rng.integers(0, 5, nt).astype(np.int32))
    pos = rng.uniform(-0.5, 0.5, 3) * L
    front = rng.standard_normal(3); front /= np.linalg.norm(front)
    if seed % 5 == 4:
        front = np.array([0.0, 0.0, -1.0])                                          # looking exactly down an axis
    cam = S.Camera(pos=tuple(float(x) for x in pos.astype(np.float32)), front=tuple(float(x) for x in front.astype(np.float32)), width=40, height=28)
    return b.build(), cam


def same_image(a, b):
    return bool(((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all())


@pytest.mark.gpu
@pytest.mark.parametrize("seed", SEEDS)
def test_gpu_image_fuzz_vs_oracle(oracle, seed):
    """The whole estimator on the fuzz scenes (all four materials, degenerate triangles with NaN normals included): the
    accumulation buffer and the per-bounce ray counts of every pipeline form against the oracle's render."""
    from parallelraytracing_b200.capi import Context
    sd, cam = room_scene(seed)
    spp, depth = 3, 2 + seed % 11
    want, rays = oracle.scene(sd).render(cam, spp, depth, seed=100 + seed)
    ctx = Context(0)
    try:
        for opts in ({}, {"fused": 0}, {"fused": 0, "bvh_wide": 1}, {"fused": 1, "bvh_builder": 0}, {"fused": 0, "tail_bounce": 0}, {"tail_pool": 0}):
            for k, v in opts.items():
                ctx.set_option(k, v)
            ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum()
            ctx.render(spp, depth, seed=100 + seed)
            got = ctx.read_accum(); st = ctx.stats()
            assert st.rays_total == rays, (seed, opts)
            assert same_image(got, want), (seed, opts, util.image_metrics(np.nan_to_num(want) / spp, np.nan_to_num(got) / spp))
            for k in opts:
                ctx.set_option(k, {"fused": -1, "bvh_wide": -1, "bvh_builder": 1, "tail_bounce": 2, "tail_pool": 1}[k])
    finally:
        ctx.close()


BOUNDARY_SIZES = [2, 3, 4, 5, 7, 8, 9, 31, 32, 33, 255, 256, 257, 1023, 1024, 1025, 2049, 4095, 4096, 4097, 8193, 65537,
                  163839, 163840, 163841, 327681]


@pytest.mark.gpu
def test_builder_at_boundary_sizes_vs_oracle(oracle):
    """Primitive counts around every tile / threshold of the build (32-lane warps, 256-thread CTAs, 1024-cluster PLOC tiles,
    4096-key sort tiles and the 4096-primitive wide-node threshold, the 160 K small-round switch of PLOC), plain and with a
    quarter of the triangles duplicated: integer outputs, node boxes, hit records and traversal counters against the oracle,
    both builders, through the one-call Init."""
    import dataclasses
    from parallelraytracing_b200.capi import Context
    ctx = Context(0)
    try:
        for n in BOUNDARY_SIZES:
            for dups in (False, True):
                sd = S.random_triangle_soup(n, seed=n)
                if dups:
                    if n < 8:
                        continue
                    tv = sd.tri_verts.copy(); k = n // 4
                    tv[:k] = tv[k:2 * k]; tv[2 * k:2 * k + k // 2] = tv[2 * k]
                    sd = dataclasses.replace(sd, tri_verts=tv)
                for builder in (1, 0):
                    ctx.set_option("bvh_builder", builder); ctx.set_option("bvh_wide", 1)
                    ctx.set_scene_build(sd)
                    osc = oracle.scene(sd); osc.set_builder(builder, 4)
                    g, o = ctx.debug_lbvh(), osc.get_lbvh()
                    for key in ("morton", "sorted_prim", "node_children", "node_bounds"):
                        assert g[key].tobytes() == o[key].tobytes(), (n, dups, builder, key)
                    rays = util.incoherent_rays(2000, 50.0, seed=n)
                    cnt = np.zeros(2, np.uint64)
                    want = osc.trace_batch(rays, use_bvh=2, counters=cnt, n_threads=16)
                    ctx.enable_counters(True); got = ctx.trace_batch(rays, use_bvh=2); st = ctx.stats(); ctx.enable_counters(False)
                    assert same_records(got, want), (n, dups, builder)
                    assert (st.nodes_visited, st.prims_tested) == (int(cnt[0]), int(cnt[1])), (n, dups, builder)
    finally:
        ctx.close()


def compact_wide_reference(full):
    """Host restatement of lbvh.cu:k_compact_wide: breadth-first from record 0, frontier in order, a node's internal children
    (slots in order, empty slots = qlo byte 255 and qhi byte 0 on x) get the next free indices; empty slots take slot 0's code."""
    out = []
    frontier = [(0, 0)]                                      # (old index, new index)
    placed = 1
    code = full[:, 4:8].view(np.int32)
    while frontier:
        nxt = []
        for old, new in frontier:
            w = full[old].copy(); c = w[4:8].view(np.int32)
            empty = [((int(w[8]) >> (8 * k)) & 255) == 255 and ((int(w[11]) >> (8 * k)) & 255) == 0 for k in range(4)]
            for k in range(4):
                if not empty[k] and code[old, k] >= 0:
                    nxt.append((int(code[old, k]), placed)); c[k] = placed; placed += 1
            for k in range(1, 4):
                if empty[k]:
                    c[k] = c[0]
            out.append((new, w))
        frontier = nxt
    arr = np.zeros((placed, 16), np.uint32)
    for new, w in out:
        arr[new] = w
    return arr


@pytest.mark.gpu
def test_compact_wide_nodes_are_the_breadth_first_copy():
    """The dense array the walk reads = the reachable records of prt_debug_wide's array in breadth-first order with remapped
    child indices, byte for byte (so: deterministic), for both builders and sizes around the CTA / grid chunking of the pass;
    about half of the records are reachable."""
    from parallelraytracing_b200.capi import Context
    ctx = Context(0)
    try:
        for n in (3, 4, 9, 257, 4097, 30000, 200000):
            sd = S.random_triangle_soup(n, seed=3 * n)
            for builder in (1, 0):
                ctx.set_option("bvh_builder", builder); ctx.set_option("bvh_wide", 1)
                ctx.set_scene_build(sd)
                full, got = ctx.debug_wide(), ctx.debug_wide_compact()
                want = compact_wide_reference(full)
                assert got.shape == want.shape and got.tobytes() == want.tobytes(), (n, builder)
                if n >= 4097:
                    assert 0.3 < len(got) / len(full) < 0.7
                ctx.set_scene_build(sd)                        # again: the same bytes
                assert ctx.debug_wide_compact().tobytes() == got.tobytes()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_scatter_fuzz_vs_oracle(oracle):
    """20 000 random MaterialHandle::Scatter / Emit records (all materials, roughness / ior over their ranges, normals facing the
    ray or not, both face flags, random RNG keys) through the device's shading function against the oracle, bit for bit - the
    oracle itself is pinned to the reference's compiled material code on 4 000 such records (test_fuzz_reference.py)."""
    from parallelraytracing_b200.capi import Context, material_dt
    rng = np.random.default_rng(31337)
    n = 20000
    m = np.zeros(n, material_dt)
    m["type"] = rng.integers(1, 5, n)
    m["r"], m["g"], m["b"] = (rng.uniform(0, 1, n).astype(np.float32) for _ in range(3))
    m["param"] = np.where(m["type"] == 2, rng.choice([0.0, 0.05, 0.3, 1.0, 2.5], n), np.where(m["type"] == 3, rng.choice([0.5, 1.0, 1.33, 1.5, 2.4], n), 0.0)).astype(np.float32)
    d = rng.standard_normal((n, 3)); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    nr = rng.standard_normal((n, 3)); nr /= np.linalg.norm(nr, axis=1, keepdims=True)
    flip = (np.arange(n) % 4 != 0) & ((nr * d).sum(1) > 0)
    nr[flip] *= -1
    nr = nr.astype(np.float32)
    fr = rng.integers(0, 2, n).astype(np.int32)
    keys = np.stack([rng.integers(0, 1 << 21, n), rng.integers(0, 4096, n), rng.integers(0, 32, n)], 1).astype(np.uint32)
    with Context(0) as ctx:
        got = ctx.debug_scatter(m, d, nr, fr, keys, seed=77)
    want = np.stack([oracle.scatter(tuple(m[i].tolist()), d[i], nr[i], int(fr[i]), 77, int(keys[i, 0]), int(keys[i, 1]), int(keys[i, 2])) for i in range(n)])
    assert np.array_equal(got[:, 0], want[:, 0]), "scattered / absorbed decision"
    same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
    assert same.all(), np.argwhere(~same)[:5]
    assert 0.5 < got[:, 0].mean() < 0.98
