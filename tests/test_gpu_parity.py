"""GPU parity tests (-m gpu): the sm_100a CUDA path, called through the C ABI (include/prt.h), against
the CPU oracle on the same seeded inputs — bit-exact on integer work (Morton codes, sorted order,
Karras topology, closest-hit ids) and, because both sides run the same IEEE operation order, bit-exact
on t / positions / radiance too; plus the committed reference vectors (tests/golden) with the
tolerances of SURVEY 8c."""
import numpy as np
import pytest

import util
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context, PrtError

pytestmark = pytest.mark.gpu
PRESETS = {p.name.lower(): int(p) for p in S.ScenePreset}


@pytest.fixture(scope="module")
def gold():
    return np.load(util.GOLDEN + "/reference_vectors.npz")


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def dragon_scene(gold, instances=((0, 0, 0),), subdiv=0):
    return S.mesh_scene((gold["mesh_dragon_v"], gold["mesh_dragon_n"], gold["mesh_dragon_f"]), subdiv=subdiv,
                        instances=instances, name="dragon")


SCENES = {
    "cornell": lambda g: S.cornell(),
    "default": lambda g: S.make_scene(S.ScenePreset.DEFAULT),
    "random_balls_large": lambda g: S.make_scene(S.ScenePreset.RANDOM_BALLS_LARGE),
    "closed_cornell": lambda g: S.closed_cornell(),
    "dragon3": lambda g: dragon_scene(g, ((-1.5, 0, 0), (0, 0, 0), (1.5, 0, 0))),
    "soup20k": lambda g: S.random_triangle_soup(20000),
}


# ---------------------------------------------------------------- LBVH: integer work bit-exact
@pytest.mark.parametrize("builder", [1, 0])
@pytest.mark.parametrize("name", sorted(SCENES))
def test_lbvh_bit_exact(ctx, oracle, gold, name, builder):
    """builder 1 = PLOC over the sorted leaves (default; radius 4 and, on one scene, 1 and 8), 0 = Karras hierarchy + refit."""
    sd = SCENES[name](gold)
    for radius in ((4, 1, 8) if builder == 1 and name == "dragon3" else (4,)):
        ctx.set_option("bvh_builder", builder); ctx.set_option("ploc_radius", radius)
        ctx.set_scene(sd); ctx.build_bvh()
        ctx.set_option("bvh_builder", 1); ctx.set_option("ploc_radius", 4)
        g = ctx.debug_lbvh()
        osc = oracle.scene(sd); osc.set_builder(builder, radius)
        o = osc.get_lbvh()
        assert np.array_equal(g["scene_bounds"], o["scene_bounds"])
        assert np.array_equal(g["morton"], o["morton"]), "Morton codes"
        assert np.array_equal(g["sorted_prim"], o["sorted_prim"]), "sorted primitive order (stable radix sort)"
        assert np.array_equal(g["node_children"], o["node_children"]), "topology (PLOC merges / Karras splits)"
        assert np.array_equal(g["node_parent"], o["node_parent"])
        assert np.array_equal(g["node_bounds"], o["node_bounds"]), "node boxes"
        st, info = ctx.stats(), osc.build_info()
        if sd.n_prims > 2:
            assert (st.bvh_builder, st.bvh_height, st.ploc_rounds) == (info["builder"], info["height"], info["rounds"])


def test_ploc_fallback_and_adversarial_scenes(ctx, oracle):
    """GPU mirror of tests/test_bvh_quality.py: (1) nested boxes - PLOC gives up at its round limit and the Karras hierarchy is
    built instead, same bits as the oracle's fallback; (2) the deep lopsided scene, both builders - hit records through the
    two-child and the 4-wide walk equal the oracle's and brute force (the wide walk needs more stack entries than the 64 the
    kernels had in round 1); (3) corner-grazing rays on the 4-wide quantised nodes equal brute force."""
    sd = util.nested_scene(300)
    ctx.set_scene(sd); ctx.build_bvh()
    st = ctx.stats()
    assert st.bvh_builder == 0 and st.ploc_rounds == 192
    g, o = ctx.debug_lbvh(), oracle.scene(sd).get_lbvh()
    for k in g:
        assert np.array_equal(g[k], o[k]), k
    ctx.set_option("bvh_wide", 1)
    try:
        sd = util.deep_scene(); rays = util.rays_from_point((0.45, 0.45, 0.45), 20000)
        for builder in (0, 1):
            ctx.set_option("bvh_builder", builder)
            ctx.set_scene(sd); ctx.build_bvh()
            osc = oracle.scene(sd); osc.set_builder(builder, 4)
            brute = ctx.trace_batch(rays, use_bvh=0)
            ctx.enable_counters(True)
            for mode in (1, 2):
                cnt = np.zeros(2, np.uint64)
                want = osc.trace_batch(rays, use_bvh=mode, counters=cnt)
                got = ctx.trace_batch(rays, use_bvh=mode); stt = ctx.stats()
                assert np.array_equal(got, want) and np.array_equal(got, brute), (builder, mode)
                assert (stt.nodes_visited, stt.prims_tested) == (int(cnt[0]), int(cnt[1]))
            ctx.enable_counters(False)
            # the same rays as a render: one camera in the middle of all boxes, in-thread tail walk included
            cam = S.Camera(pos=(0.45, 0.45, 0.45), front=(0.0, 0.0, -1.0), width=64, height=48)
            ctx.set_camera(cam); ctx.clear_accum(); ctx.render(2, 6, seed=3)
            want_img, _ = osc.render(cam, 2, 6, seed=3)
            assert np.array_equal(ctx.read_accum(), want_img)
        ctx.set_option("bvh_builder", 1)
        sd, tv = util.grazing_scene(); rays = util.grazing_rays(tv, 400000)
        ctx.set_scene(sd); ctx.build_bvh()
        brute = ctx.trace_batch(rays, use_bvh=0)
        assert np.array_equal(ctx.trace_batch(rays, use_bvh=2), brute) and np.array_equal(ctx.trace_batch(rays, use_bvh=1), brute)
        assert np.array_equal(brute, oracle.scene(sd).trace_batch(rays, use_bvh=0))
    finally:
        ctx.set_option("bvh_wide", -1); ctx.set_option("bvh_builder", 1)


@pytest.mark.parametrize("builder", [1, 0])
def test_lbvh_large_properties(ctx, builder):
    """1M-triangle soup (oracle too slow for a full compare in the GPU suite): size-independent
    properties — sorted keys, permutation, every node's box contains its children, one root. Both hierarchies."""
    sd = S.random_triangle_soup(1_000_000)
    ctx.set_option("bvh_builder", builder)
    ctx.set_scene(sd); ctx.build_bvh()
    ctx.set_option("bvh_builder", 1)
    assert ctx.stats().bvh_builder == builder
    g = ctx.debug_lbvh()
    n = sd.n_prims
    keys = g["morton"][g["sorted_prim"]]
    assert np.all(np.diff(keys.astype(np.int64)) >= 0), "sortedness"
    same = np.diff(keys.astype(np.int64)) == 0
    assert np.all(np.diff(g["sorted_prim"].astype(np.int64))[same] > 0), "stability: ties keep input order"
    assert np.array_equal(np.sort(g["sorted_prim"]), np.arange(n, dtype=np.uint32)), "permutation"
    ch = g["node_children"]
    assert (g["node_parent"] == -1).sum() == 1 and g["node_parent"][0] == -1
    internal = ch[ch >= 0]
    assert np.array_equal(np.sort(internal), np.arange(1, n - 1)), "every internal node but the root is a child once"
    leaves = ~ch[ch < 0]
    assert np.array_equal(np.sort(leaves), np.arange(n)), "every leaf is a child once"
    # child boxes of node i enclose the boxes stored in the child node
    nb = g["node_bounds"]
    for side in (0, 1):
        c = ch[:, side]; m = c >= 0
        box_min, box_max = nb[m, 6 * side:6 * side + 3], nb[m, 6 * side + 3:6 * side + 6]
        cmin = np.minimum(nb[c[m], 0:3], nb[c[m], 6:9]); cmax = np.maximum(nb[c[m], 3:6], nb[c[m], 9:12])
        assert np.array_equal(box_min, cmin) and np.array_equal(box_max, cmax)


# ---------------------------------------------------------------- closest hit on fixed ray batches
@pytest.mark.parametrize("name", sorted(SCENES))
def test_closest_hit_bit_exact_vs_oracle(ctx, oracle, gold, name):
    sd = SCENES[name](gold)
    ctx.set_scene(sd); ctx.build_bvh()
    rays = util.fixed_ray_batch(sd, 65536, seed=1, clamp_extent=60.0)
    ctx.enable_counters(True)
    got = ctx.trace_batch(rays, use_bvh=True)
    st = ctx.stats()
    ctx.enable_counters(False)
    cnt = np.zeros(2, np.uint64)
    want = oracle.scene(sd).trace_batch(rays, use_bvh=True, counters=cnt)
    assert np.array_equal(got["prim_id"], want["prim_id"]), "closest-hit primitive ids"
    assert np.array_equal(got, want), "full hit records (t, barycentrics, position, normal, face, material)"
    assert np.array_equal(np.argsort(got["t"], kind="stable"), np.argsort(want["t"], kind="stable")), "t-ordering"
    assert (st.nodes_visited, st.prims_tested) == (int(cnt[0]), int(cnt[1])), "traversal counters reproduce on the CPU"
    assert (got["prim_id"] >= 0).mean() > 0.01


@pytest.mark.parametrize("name", ["cornell", "random_balls_large", "soup20k"])
def test_gpu_brute_force_equals_bvh(ctx, gold, name):
    """The reference's own algorithm (loop over every primitive, primitive.cpp:26) on the GPU gives the
    same records as the LBVH traversal."""
    sd = SCENES[name](gold)
    ctx.set_scene(sd); ctx.build_bvh()
    rays = util.fixed_ray_batch(sd, 4096, seed=3, clamp_extent=60.0)
    assert np.array_equal(ctx.trace_batch(rays, use_bvh=True), ctx.trace_batch(rays, use_bvh=False))


@pytest.mark.parametrize("name", ["default", "material_test", "cornell", "random_balls_large"])
def test_closest_hit_vs_reference_vectors(ctx, gold, name):
    sd = util.scene_from_npz(gold, "scene_" + name)
    ctx.set_scene(sd); ctx.build_bvh()
    rays, ref_hits, second = gold["hits_%s_rays" % name], gold["hits_%s_ref" % name], gold["hits_%s_second" % name]
    got = ctx.trace_batch(rays)
    with np.errstate(invalid="ignore"):
        near = np.abs(second - ref_hits["t"]) <= 1e-4 * np.maximum(1.0, ref_hits["t"])
    idm = ref_hits["prim_id"] != got["prim_id"]
    assert not (idm & ~near).any()
    ok = (ref_hits["prim_id"] >= 0) & ~idm
    assert np.all(np.abs(ref_hits["t"][ok] - got["t"][ok]) <= 1e-4 * np.maximum(1.0, ref_hits["t"][ok]))
    assert np.array_equal(ref_hits["front"][ok], got["front"][ok])


@pytest.mark.parametrize("name", sorted(SCENES))
def test_wide_nodes_equal_two_child_bvh(ctx, oracle, gold, name):
    """The 4-wide quantised nodes (what the staged render path walks) return the same hit records, bit for bit,
    as the two-child LBVH and as brute force, with fewer node visits; leaf tests may differ (other visit order)."""
    sd = SCENES[name](gold)
    ctx.set_option("bvh_wide", 1)                      # also for the small scenes that the auto policy leaves two-child
    ctx.set_scene(sd); ctx.build_bvh()
    ctx.set_option("bvh_wide", -1)
    rays = util.fixed_ray_batch(sd, 65536, seed=11, clamp_extent=60.0)
    ctx.enable_counters(True)
    two = ctx.trace_batch(rays, use_bvh=1); st2 = ctx.stats(); n2 = st2.nodes_visited
    wide = ctx.trace_batch(rays, use_bvh=2); stw = ctx.stats(); nw = stw.nodes_visited
    ctx.enable_counters(False)
    for k in two.dtype.names:
        assert np.array_equal(two[k], wide[k]), k
    # the CPU restatement of the collapse + quantisation gives the same bytes, and its walk the same counters
    if sd.n_prims > 1:
        osc = oracle.scene(sd)
        ow, valid = osc.get_wide()
        gw = ctx.debug_wide()
        assert valid.all()                                 # greedy collapse: every internal node has a record
        assert np.array_equal(gw[valid], ow[valid]), "wide nodes: origin, scales, child codes, quantised boxes"
        cnt = np.zeros(2, np.uint64)
        owide = osc.trace_batch(rays, use_bvh=2, counters=cnt)
        assert np.array_equal(owide, wide)
        assert (stw.nodes_visited, stw.prims_tested) == (int(cnt[0]), int(cnt[1])), "wide traversal counters reproduce on the CPU"
    if sd.n_prims > 2:
        assert 0 < nw < n2, (nw, n2)
    if sd.n_prims <= 20000:
        brute = ctx.trace_batch(rays, use_bvh=0)
        assert np.array_equal(brute["prim_id"], wide["prim_id"]) and np.array_equal(brute["t"], wide["t"])


def test_wide_nodes_off_renders_same_image(ctx, gold):
    """bvh_wide=0 (two-child nodes only) and the default must give identical images and ray counts."""
    sd = SCENES["dragon3"](gold); cam = S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=96, height=64)
    out = []
    for widev in (1, 0):
        ctx.set_option("bvh_wide", widev)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum()
        ctx.render(3, 8, seed=13)
        out.append((ctx.read_accum(), list(ctx.stats().rays_per_bounce)))
    ctx.set_option("bvh_wide", -1)
    assert np.array_equal(out[0][0], out[1][0]) and out[0][1] == out[1][1]


# ---------------------------------------------------------------- radiance
@pytest.mark.parametrize("name,W,H,spp,depth", [("cornell", 128, 128, 8, 8), ("default", 96, 96, 4, 8),
                                                 ("random_balls_large", 96, 54, 4, 8), ("closed_cornell", 96, 96, 4, 8),
                                                 ("dragon3", 64, 64, 2, 8), ("soup20k", 64, 64, 2, 16)])
def test_image_bit_exact_vs_oracle(ctx, oracle, gold, name, W, H, spp, depth):
    sd = SCENES[name](gold)
    cam = S.Camera(width=W, height=H) if name not in ("dragon3",) else \
        S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=W, height=H)
    if name == "soup20k":
        cam = S.Camera(pos=(0.0, 0.0, 150.0), front=(0.0, 0.0, -1.0), width=W, height=H)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum()
    ctx.render(spp, depth, seed=7)
    got = ctx.read_accum()
    want, rays = oracle.scene(sd).render(cam, spp, depth, seed=7)
    st = ctx.stats()
    assert st.rays_total == rays, "closest-hit query count"
    m = util.image_metrics(want / spp, got / spp)
    assert np.array_equal(got, want), m


@pytest.mark.parametrize("name", ["cornell", "default", "material_test", "random_balls_large", "light_test"])
def test_image_vs_reference_vectors(ctx, gold, name):
    """Against the reference tier-2 render (reference Intersect/Scatter/GetCameraRay + deterministic
    rand): RMSE <= 1e-3 and max rel <= 1e-2 over non-flip pixels; flips (coplanar-tie decisions) counted."""
    W, H, spp, depth, seed, ref_rays = (int(x) for x in gold["img_%s_meta" % name])
    sd = util.scene_from_npz(gold, "scene_" + name)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(S.Camera(width=W, height=H)); ctx.clear_accum()
    ctx.render(spp, depth, seed=seed)
    m = util.image_metrics(gold["img_%s" % name] / spp, ctx.read_accum() / spp)
    assert m["flips"] <= (0.01 if name == "cornell" else 0.002) * m["n"], m      # Cornell: 0.6 % measured (coplanar walls)
    assert m["rmse_nonflip"] <= 1e-3 and m["max_rel_nonflip"] <= 1e-2, m


def test_scatter_known_answers(ctx, oracle, gold):
    """The reference's 600 MaterialHandle::Scatter / Emit known answers (tests/golden/make_golden.py section 5) through the
    device function the render path shades with (prt_debug_scatter -> prt_device.cuh:scatter): equal to the reference's
    compiled material code bit for bit, every branch (metal rejection, total internal reflection, ior < 1, emissive)."""
    mats, ins, nrm, fr, want = (gold["scatter_" + k] for k in ("mats", "in", "n", "front", "out"))
    from parallelraytracing_b200.capi import material_dt
    m = np.zeros(len(mats), material_dt)
    m["type"], m["r"], m["g"], m["b"], m["param"] = mats[:, 0].astype(np.int32), mats[:, 1], mats[:, 2], mats[:, 3], mats[:, 4]
    k = np.arange(len(mats), dtype=np.uint32)
    keys = np.stack([k, k // 7, k % 5], 1).astype(np.uint32)
    got = ctx.debug_scatter(m, ins, nrm, fr, keys, seed=1)
    assert np.array_equal(got[:, 0], want[:, 0]), "scattered / absorbed decision"
    assert np.array_equal(got, want), np.abs(got - want).max()
    orc = np.stack([oracle.scatter(tuple(m[i].tolist()), ins[i], nrm[i], int(fr[i]), 1, int(k[i]), int(k[i]) // 7, int(k[i]) % 5) for i in range(len(m))])
    assert np.array_equal(got, orc)


def test_c1_full_size(ctx, oracle):
    """BASELINE.json configs[0] IN FULL: reference CORNELL, 512 x 512, 16 spp, depth 8, seed 1. GPU == oracle
    bit for bit, and - where oracle/_ref exists - within the SURVEY tolerances of the reference's own tier-2 render."""
    sd = S.cornell(); cam = S.Camera(width=512, height=512)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum()
    ctx.render(16, 8, seed=1)
    got, st = ctx.read_accum(), ctx.stats()
    want, rays = oracle.scene(sd).render(cam, 16, 8, seed=1)
    assert np.array_equal(got, want) and st.rays_total == rays
    import oracle_lib
    ref = oracle_lib.load_ref()
    if ref is not None:
        rimg, ref_rays = ref.preset(int(S.ScenePreset.CORNELL)).render(cam, 16, 8, seed=1)
        m = util.image_metrics(rimg / 16, got / 16)
        assert m["flips"] <= 0.01 * m["n"] and m["rmse_nonflip"] <= 1e-3 and m["max_rel_nonflip"] <= 1e-2, m
        assert abs(st.rays_total - ref_rays) <= 0.01 * ref_rays


@pytest.mark.parametrize("name,W,H,spp,depth", [("cornell", 128, 128, 16, 8), ("default", 96, 96, 8, 8), ("random_balls_large", 96, 54, 8, 8),
                                                 ("closed_cornell", 96, 96, 8, 12), ("dragon3", 64, 64, 4, 8)])
def test_fast_math_mode(ctx, gold, name, W, H, spp, depth):
    """Option fast_math (default OFF): SFU approximations for divide / sqrt / rsqrt on the SHADING side only. Integer work is
    untouched - hit records of a fixed batch stay bit-identical - and radiance stays within the north star's tolerance
    against the bit-exact mode (= the oracle): per-pixel RMSE <= 1e-3, max relative error <= 1e-2 over non-flip pixels,
    flips (a path took another discrete decision: Schlick draw, rejection loop, metal horizon) counted and bounded."""
    sd = SCENES[name](gold)
    cam = S.Camera(width=W, height=H) if name != "dragon3" else S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=W, height=H)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
    rays = util.fixed_ray_batch(sd, 8192, seed=5, clamp_extent=60.0)
    out = {}
    for fast in (0, 1):
        ctx.set_option("fast_math", fast)
        ctx.clear_accum(); ctx.render(spp, depth, seed=11)
        out[fast] = (ctx.read_accum(), list(ctx.stats().rays_per_bounce), ctx.trace_batch(rays))
    ctx.set_option("fast_math", 0)
    assert np.array_equal(out[0][2], out[1][2]), "closest-hit records must not depend on fast_math"
    assert out[0][1][0] == out[1][1][0]                                      # same camera rays
    m = util.image_metrics(out[0][0] / spp, out[1][0] / spp)
    drift = abs(sum(out[1][1]) - sum(out[0][1])) / sum(out[0][1])
    print("fast_math %s: %s, ray-count drift %.2e" % (name, m, drift))
    assert m["rmse_nonflip"] <= 1e-3 and m["max_rel_nonflip"] <= 1e-2, m
    assert m["flips"] <= 0.01 * m["n"] and drift <= 1e-3, (m, drift)
    # (the estimator has no cosine or texture term: a path's radiance is a product of material constants, so most images come
    # out bit-identical and a difference can only be a flipped discrete decision)


def test_sample_ranges_compose(ctx, gold):
    """Multi-GPU sharding rule: samples [0,6) == [0,2) + [2,6), any wave size."""
    sd = S.closed_cornell(); cam = S.Camera(width=80, height=60)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
    ctx.clear_accum(); ctx.render(6, 8, seed=3); a = ctx.read_accum()
    ctx.clear_accum(); ctx.render(2, 8, seed=3, sample_begin=0); ctx.render(4, 8, seed=3, sample_begin=2); b = ctx.read_accum()
    assert np.array_equal(a, b) and ctx.accum_spp() == 6
    ctx.set_option("wave_paths", 7000)                 # forces pixel-range waves with ragged tails
    ctx.clear_accum(); ctx.render(6, 8, seed=3); c = ctx.read_accum()          # many waves, two in flight on two streams
    ctx.set_option("overlap", 0)
    ctx.clear_accum(); ctx.render(6, 8, seed=3); d = ctx.read_accum()          # same waves, one stream
    ctx.set_option("overlap", 1); ctx.set_option("wave_paths", 0)
    assert np.array_equal(a, c) and np.array_equal(a, d)


def test_estimator_options_match_oracle_and_are_off_by_default(ctx, oracle, gold):
    """Russian roulette and the firefly clamp exist only behind options, default OFF. With them on, the
    image still equals the oracle's (same rule, same RNG draw) bit for bit, fewer rays are traced, and the estimate
    stays unbiased (roulette) / bounded (clamp)."""
    sd = S.closed_cornell(); cam = S.Camera(width=96, height=64)           # enclosed: paths run to the depth limit
    osc = oracle.scene(sd)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
    ctx.clear_accum(); ctx.render(64, 12, seed=4); base = ctx.read_accum(); rays0 = ctx.stats().rays_total
    ctx.set_option("rr_start", 1)
    ctx.clear_accum(); ctx.render(64, 12, seed=4); rr = ctx.read_accum(); rays1 = ctx.stats().rays_total
    osc.set_estimator(rr_start=1)
    want, orays = osc.render(cam, 64, 12, seed=4)
    assert np.array_equal(rr, want) and rays1 == orays and rays1 < rays0
    assert abs(rr.mean() - base.mean()) <= 0.03 * base.mean()               # unbiased: same mean within MC noise
    ctx.set_option("firefly_clamp_milli", 2500)
    ctx.clear_accum(); ctx.render(8, 12, seed=4); cl = ctx.read_accum()
    osc.set_estimator(rr_start=1, clamp=2.5)
    want, _ = osc.render(cam, 8, 12, seed=4)
    assert np.array_equal(cl, want) and cl.max() <= 8 * 2.5 + 1e-3 and base.max() > 64 * 2.5
    ctx.set_option("rr_start", -1); ctx.set_option("firefly_clamp_milli", 0)
    ctx.clear_accum(); ctx.render(64, 12, seed=4)
    assert np.array_equal(ctx.read_accum(), base)


def test_pixel_regions_compose(ctx, gold):
    """TILE partition rule: rendering disjoint pixel ranges (bands of rows, ragged last range) leaves the
    other pixels untouched and the union equals the whole-frame render bit for bit, fused and staged."""
    for sd, cam in ((S.closed_cornell(), S.Camera(width=80, height=60)),
                    (SCENES["dragon3"](gold), S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=64, height=48))):
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        P = cam.width * cam.height
        ctx.clear_accum(); ctx.render(3, 8, seed=9, sample_begin=2); whole = ctx.read_accum(); rays = ctx.stats().rays_total
        ctx.clear_accum()
        cuts, tot = [0, 17 * cam.width, 17 * cam.width + 5, 40 * cam.width, P], 0
        for b, e in zip(cuts[:-1], cuts[1:]):
            ctx.render_region(3, 8, seed=9, sample_begin=2, pixel_begin=b, pixel_count=e - b); tot += ctx.stats().rays_total
            part = ctx.read_accum().reshape(P, 3)
            assert not part[e:].any(), "pixels after the rendered ranges stay zero"
        assert np.array_equal(ctx.read_accum(), whole) and tot == rays and ctx.accum_spp() == 0
    with pytest.raises(PrtError):
        ctx.render_region(1, 8, pixel_begin=P - 1, pixel_count=2)


def test_staged_pipeline_equals_fused(ctx, gold):
    """fused=0 runs raygen / extend / shade as separate kernels with a hit queue; fused=1 (default)
    removes the stage boundaries. Same device functions, same order: identical bits and ray counts."""
    for name in ("closed_cornell", "dragon3"):
        sd = SCENES[name](gold); cam = S.Camera(width=96, height=64)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        out = []
        for fused in (1, 0):
            ctx.set_option("fused", fused)
            ctx.clear_accum(); ctx.render(3, 8, seed=11)
            out.append((ctx.read_accum(), ctx.stats().rays_total, list(ctx.stats().rays_per_bounce)))
        ctx.set_option("fused", -1)
        assert np.array_equal(out[0][0], out[1][0]) and out[0][1:] == out[1][1:]


def test_tiny_scene_list_loop_equals_bvh(ctx):
    """Sphere/quad scenes of <= brute_max_prims primitives skip the BVH in the fused kernels (a warp-uniform
    loop over the list, the reference's own algorithm primitive.cpp:26). The (t, prim_id) closest-hit rule is
    order-free, so the image and every per-bounce ray count must equal the BVH walk bit for bit."""
    for sd in (S.cornell(), S.closed_cornell(), S.make_scene(S.ScenePreset.MATERIAL_TEST)):
        cam = S.Camera(width=112, height=80)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        out = []
        for limit in (1 << 20, 0):                     # list loop for any size / always the BVH
            ctx.set_option("brute_max_prims", limit)
            ctx.set_option("fused", 1)
            ctx.clear_accum(); ctx.render(3, 8, seed=23)
            out.append((ctx.read_accum(), list(ctx.stats().rays_per_bounce)))
        ctx.set_option("brute_max_prims", 8); ctx.set_option("fused", -1)
        assert np.array_equal(out[0][0], out[1][0]) and out[0][1] == out[1][1], sd.name


def test_small_wave_tail_kernel_is_invisible(ctx):
    """Small waves finish bounces >= tail_bounce inside ONE kernel (k_tail, in-thread loop) instead of one launch per
    bounce. Same device functions and RNG keys: image, per-bounce ray counts and traversal counters must not change,
    wherever the tail starts, at the reference's depth literal 20 (cpu/renderer.h:34)."""
    ctx.enable_counters(True)
    for sd in (S.cornell(), S.closed_cornell(), S.make_scene(S.ScenePreset.MATERIAL_TEST), S.make_scene(S.ScenePreset.DEFAULT)):
        cam = S.Camera(width=96, height=72)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        out = []
        for tail in (0, 1, 2, 3, 7, 19, 25):
            ctx.set_option("fused", 1); ctx.set_option("tail_bounce", tail); ctx.set_option("tail_adaptive", 0)
            ctx.clear_accum(); ctx.render(2, 20, seed=5)
            st = ctx.stats()
            out.append((ctx.read_accum(), list(st.rays_per_bounce), st.nodes_visited, st.prims_tested, st.kernel_launches))
        ctx.set_option("tail_bounce", 2); ctx.set_option("fused", -1)
        assert out[0][4] == 21 and out[3][4] == 5 and out[1][4] == 3        # 1 + 19 + resolve / primary, b1, b2, tail, resolve
        for o in out[1:]:
            assert np.array_equal(o[0], out[0][0]) and o[1:4] == out[0][1:4], sd.name
        ctx.set_option("tail_max_paths", 100)                                # wave larger than the limit: no tail
        ctx.set_option("fused", 1); ctx.clear_accum(); ctx.render(2, 20, seed=5)
        assert ctx.stats().kernel_launches == 21
        ctx.set_option("tail_max_paths", 1 << 26); ctx.set_option("tail_adaptive", 1)
        ctx.clear_accum(); ctx.render(2, 20, seed=5)                         # 13 824 paths <= tail_rays: tail right after the primary
        st = ctx.stats()
        assert st.kernel_launches == 3 and np.array_equal(ctx.read_accum(), out[0][0]) and list(st.rays_per_bounce) == out[0][1]
        ctx.set_option("tail_rays", 1000)                                    # adaptive: starts where the last call's queue fell to 1000
        ctx.clear_accum(); ctx.render(2, 20, seed=5)
        st = ctx.stats()
        first = next((b for b in range(1, 20) if out[0][1][b] <= 1000), 20)
        assert np.array_equal(ctx.read_accum(), out[0][0]) and list(st.rays_per_bounce) == out[0][1]
        assert st.kernel_launches == (first + 2 if first < 20 else 21)      # primary + (first-1) bounces + tail + resolve
        ctx.set_option("tail_rays", 512 << 10); ctx.set_option("fused", -1)
    ctx.enable_counters(False)


def test_small_wave_tail_kernel_staged_scenes(ctx, gold):
    """The same tail on the staged pipeline (scenes above fused_max_prims): k_tail walks in-thread the kind of node the extend kernel uses
    (two-child below 4096 primitives, 4-wide above): image, ray counts and traversal counters must not move."""
    ctx.enable_counters(True)
    for sd, launches0 in ((S.make_scene(S.ScenePreset.RANDOM_BALLS_SMALL), None), (SCENES["dragon3"](gold), None)):
        cam = S.Camera(width=96, height=72)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        out = []
        for tail in (0, 1, 3, 11):
            ctx.set_option("tail_bounce", tail); ctx.set_option("tail_adaptive", 0)
            ctx.clear_accum(); ctx.render(2, 12, seed=9)
            st = ctx.stats()
            out.append((ctx.read_accum(), list(st.rays_per_bounce), st.kernel_launches, st.nodes_visited, st.prims_tested))
        ctx.set_option("tail_bounce", 2); ctx.set_option("tail_adaptive", 1)
        assert [o[2] for o in out] == [25, 4, 8, 24]          # 2/bounce + resolve; 2*tail + k_tail + resolve (no raygen launch: bounce 0 makes its own camera rays)
        for o in out[1:]:      # the tail walks the same kind of node as the extend kernel (two-child / 4-wide): counters equal too
            assert np.array_equal(o[0], out[0][0]) and o[1] == out[0][1] and o[3:] == out[0][3:], sd.name
    ctx.enable_counters(False)


def test_ray_resorting_is_invisible(ctx, gold):
    """Option sort_rays (default off): the staged pipeline sorts the queue of every bounce >= 1 by the Morton code
    of the ray origin's cell and the direction octant (device-side length, radix sort shared with the LBVH build). Queue
    order has no influence on any path: image, per-bounce ray counts and traversal counters must be bit-identical."""
    ctx.enable_counters(True)
    for sd in (S.make_scene(S.ScenePreset.RANDOM_BALLS_SMALL), SCENES["dragon3"](gold), S.random_triangle_soup(20000)):
        cam = S.Camera(width=128, height=96)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        out = []
        for sort, bits in ((0, 24), (1, 24), (1, 16), (1, 8)):
            ctx.set_option("sort_rays", sort); ctx.set_option("sort_key_bits", bits); ctx.set_option("sort_min_rays", 0)
            ctx.set_option("tail_bounce", 0)
            ctx.clear_accum(); ctx.render(3, 10, seed=13)
            st = ctx.stats()
            out.append((ctx.read_accum(), list(st.rays_per_bounce), st.nodes_visited, st.prims_tested, st.kernel_launches))
        ctx.set_option("sort_rays", 0); ctx.set_option("sort_key_bits", 24); ctx.set_option("sort_min_rays", 1 << 20)
        ctx.set_option("tail_bounce", 2)
        assert out[1][4] == out[0][4] + 9 * 11 and out[3][4] == out[0][4] + 9 * 5     # 9 sorted queues: keys + 3 per pass + gather
        for o in out[1:]:
            assert np.array_equal(o[0], out[0][0]) and o[1:4] == out[0][1:4], sd.name
    ctx.enable_counters(False)


def test_render_fresh_equals_clear_plus_render(ctx, gold):
    """prt_render_fresh = clear + render in one call (what one ProgressiveRender needs). With 1 spp in one wave the paths
    write straight into the sample buffer (no memset, no contribution plane, no resolve launch): same bits."""
    for sd in (S.cornell(), S.make_scene(S.ScenePreset.DEFAULT), S.make_scene(S.ScenePreset.RANDOM_BALLS_SMALL), SCENES["dragon3"](gold)):
        cam = S.Camera(width=160, height=90)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
        for spp in (1, 3):
            ctx.clear_accum(); ctx.render(5, 6, seed=2)                      # stale content that fresh must replace
            ctx.clear_accum(); ctx.render(spp, 20, seed=7, sample_begin=4)
            want, l0 = ctx.read_accum(), ctx.stats().kernel_launches
            ctx.render(2, 6, seed=3)                                         # dirty again
            ctx.render_fresh(spp, 20, seed=7, sample_begin=4)
            assert np.array_equal(ctx.read_accum(), want), (sd.name, spp)
            assert ctx.accum_spp() == spp
            assert ctx.stats().kernel_launches == l0 - (1 if spp == 1 else 0)    # 1 spp: no resolve pass
    ctx.set_option("wave_paths", 4000)                                       # 1 spp but several waves: clear + resolve path
    ctx.set_scene(S.cornell()); ctx.build_bvh(); ctx.set_camera(S.Camera(width=160, height=90))
    ctx.clear_accum(); ctx.render(1, 8, seed=7); want = ctx.read_accum()
    ctx.render(3, 8, seed=1); ctx.render_fresh(1, 8, seed=7)
    ctx.set_option("wave_paths", 0)
    assert np.array_equal(ctx.read_accum(), want) and ctx.accum_spp() == 1
    ctx.render_fresh(2, 0, seed=1)                                           # depth 0: TraceRay returns 0
    assert not ctx.read_accum().any() and ctx.accum_spp() == 2


def test_display_resolve_exact(ctx, oracle):
    sd = S.cornell(); cam = S.Camera(width=64, height=64)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum(); ctx.render(4, 8, seed=1)
    want = oracle.resolve_display(ctx.read_accum(), 4.0, 1.0, 2.2)
    got = ctx.resolve_display(1.0, 2.2)
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1 and (got == want).mean() > 0.999


# ---------------------------------------------------------------- edge cases
def test_edge_cases(ctx):
    cam = S.Camera(width=33, height=17)
    empty = S.SceneData()
    ctx.set_scene(empty); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum(); ctx.render(2, 8, seed=1)
    img = ctx.read_accum()
    assert np.allclose(img, np.array(S.SKY, np.float32) * 2)            # every path misses -> sky
    b = S.SceneBuilder("one"); b.add_sphere(1.0, b.emissive((2, 3, 4)), trans=(0, 0, 0))
    ctx.set_scene(b.build()); ctx.build_bvh(); ctx.clear_accum(); ctx.render(1, 8, seed=1)
    img = ctx.read_accum()
    assert np.isfinite(img).all() and (img == np.array([2, 3, 4], np.float32)).all(-1).any()
    ctx.clear_accum(); ctx.render(3, 0, seed=1)                          # depth 0: TraceRay returns 0
    assert not ctx.read_accum().any() and ctx.accum_spp() == 3
    assert len(ctx.trace_batch(np.zeros(0, util.ray_dt))) == 0
    # ONE triangle: the root is the leaf, no refit - the triangle planes (and the material in v1.w) must still be gathered
    b = S.SceneBuilder("one_tri"); m = b.emissive((5, 6, 7))
    b.add_triangles(np.array([[[-1, -1, 0], [1, -1, 0], [0, 1, 0]]], np.float32), None, m)
    ctx.set_scene(b.build()); ctx.build_bvh()
    rays = np.zeros(2, util.ray_dt)
    rays["oz"] = 3.0; rays["dz"] = -1.0; rays["ox"][1] = 5.0            # ray 0 through the centre, ray 1 past it
    h = ctx.trace_batch(rays)
    assert h["prim_id"][0] == 0 and abs(h["t"][0] - 3.0) < 1e-5 and h["material"][0] == m and h["prim_id"][1] < 0
    with pytest.raises(PrtError):
        ctx.render(1, 99)


def test_renderer_interface(gold):
    """Init / SetCamera / ProgressiveRender call pattern of main.cpp:138-170, 506-532."""
    from parallelraytracing_b200 import B200Renderer, Film
    film = Film(64, 48); cam = S.Camera(width=64, height=48)
    r = B200Renderer(max_depth=20)
    r.Init(film, S.cornell(), cam); film.Clear()
    assert r.ok
    for _ in range(3):
        r.ProgressiveRender()
    assert film.Weight() == 3
    a = film.ReadAccum()
    r.SetCamera(cam); film.Clear()
    assert film.Weight() == 0
    r.SetNextSample(0)
    for _ in range(3):
        r.ProgressiveRender()
    assert np.array_equal(a, film.ReadAccum())
    assert film.UpdateDisplay().shape == (48, 64, 4)
    r.close()


def test_cpp_shim_dropin(ctx, tmp_path):
    """The C++ `B200Renderer : public Renderer` (host/b200_renderer.cpp), compiled against the reference's
    UNMODIFIED headers and linked with the reference's own Film/Scene/Camera objects (oracle/_ref, built by
    `make -C oracle shim`), driven with the viewer's call pattern: the reference Film's device accumulator
    must hold exactly what the C-ABI path produces."""
    import os
    import subprocess
    exe = os.path.join(util.ROOT, "oracle", "_ref", "shim_driver")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/shim_driver not built (needs /root/reference at build time)")
    W, H, calls, depth, seed = 160, 90, 5, 20, 3
    for preset in (int(S.ScenePreset.CORNELL), int(S.ScenePreset.DEFAULT)):
        out = str(tmp_path / ("shim%d.bin" % preset))
        r = subprocess.run([exe, str(preset), str(W), str(H), str(calls), "1", str(depth), str(seed), out],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        raw = np.fromfile(out, np.uint8)
        accum = raw[:W * H * 12].view(np.float32).reshape(H, W, 3)
        weights = raw[W * H * 12:W * H * 16].view(np.float32)
        disp = raw[W * H * 16:].reshape(H, W, 4)
        sd = S.make_scene(preset); cam = S.Camera(width=W, height=H)
        ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam); ctx.clear_accum()
        for i in range(calls):
            ctx.render(1, depth, seed=seed, sample_begin=i)
        assert np.all(weights == calls)
        assert np.array_equal(accum, ctx.read_accum())
        mine = ctx.resolve_display(1.0, 2.2)
        assert np.abs(mine.astype(int) - disp.astype(int)).max() <= 1     # reference GPU tonemap (film.cu:101-121)


def test_cpp_shim_add_mesh(ctx, gold, tmp_path):
    """Mesh -> scene glue on the C++ side: the reference's own `Mesh` class loads a PLY, `B200Renderer::AddMesh`
    appends it to the flattened `Scene(preset)`; the reference Film's accumulator must equal, bit for bit, the C-ABI
    render of the host mirror `scene.add_meshes` (binary PLY written from the golden dragon arrays: exact floats)."""
    import os
    import subprocess
    exe = os.path.join(util.ROOT, "oracle", "_ref", "shim_driver")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/shim_driver not built (needs /root/reference at build time)")
    v = np.ascontiguousarray(gold["mesh_dragon_v"], np.float32); n = np.ascontiguousarray(gold["mesh_dragon_n"], np.float32)
    f = np.ascontiguousarray(gold["mesh_dragon_f"], np.uint32).reshape(-1, 3)
    ply = str(tmp_path / "dragon_bin.ply")
    with open(ply, "wb") as fh:
        fh.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
                  "property float nx\nproperty float ny\nproperty float nz\nelement face %d\n"
                  "property list uchar uint vertex_indices\nend_header\n" % (len(v), len(f))).encode())
        fh.write(np.concatenate([v, n], axis=1).astype("<f4").tobytes())
        rec = np.zeros(len(f), dtype=[("k", "u1"), ("i", "<u4", (3,))]); rec["k"] = 3; rec["i"] = f
        fh.write(rec.tobytes())
    got = S.load_ply(ply)                                        # the host loader reads the same file back exactly
    assert np.array_equal(got[0], v) and np.array_equal(got[1], n) and np.array_equal(np.asarray(got[2]).reshape(-1, 3), f)
    W, H, calls, depth, seed = 160, 90, 3, 12, 5
    preset = int(S.ScenePreset.MATERIAL_TEST)
    out = str(tmp_path / "shim_mesh.bin")
    r = subprocess.run([exe, str(preset), str(W), str(H), str(calls), "1", str(depth), str(seed), out,
                        ply, "1", "0.8", "0.3", "0.3", "0", "0", "1", "2", ply, "2", "0.9", "0.9", "0.9", "0.05", "1.5", "1", "2"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    accum = np.fromfile(out, np.uint8)[:W * H * 12].view(np.float32).reshape(H, W, 3)
    sd = S.add_meshes(S.make_scene(preset), [((v, n, f), (1, (0.8, 0.3, 0.3), 0.0), (0, 1, 2)),
                                             ((v, n, f), (2, (0.9, 0.9, 0.9), 0.05), (1.5, 1, 2))])
    assert sd.n_prims == 4 + 2 * len(f) and sd.tri_prim_base == 4
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(S.Camera(width=W, height=H)); ctx.clear_accum()
    for i in range(calls):
        ctx.render(1, depth, seed=seed, sample_begin=i)
    mine = ctx.read_accum()
    assert np.array_equal(accum, mine)
    assert (mine.reshape(-1, 3) != mine.reshape(-1, 3)[0]).any()


def test_reference_cuda_backends_agree(tmp_path):
    """The reference's OWN CUDA backends (cuda_wavefront, cuda_megakernel: compiled unmodified for sm_100 into oracle/_ref by
    `make -C oracle refgpu`) and B200Renderer, behind the same Renderer interface on this GPU. They use cuRAND, so parity
    is statistical: the mean image of 400 spp at 160x90, depth 20, must agree to 1 % (Monte-Carlo error of the mean ~0.2 %).
    Reference Cornell (coplanar wall quads, scene.cpp:342-349: the two reference backends disagree with EACH OTHER by 2 % over
    the whole image) is compared over the pixels that are NOT distance ties for the camera ray, the mask coming from the
    reference's own second-hit distance."""
    import json
    import os
    import subprocess
    exe = os.path.join(util.ROOT, "oracle", "_ref", "ref_gpu_driver")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_gpu_driver not built (needs /root/reference at build time)")
    W, H, frames = 160, 90, 400
    import oracle_lib
    ref_cpu = oracle_lib.load_ref()
    presets = [int(S.ScenePreset.MATERIAL_TEST), int(S.ScenePreset.DEFAULT), int(S.ScenePreset.RANDOM_BALLS_SMALL)]
    if ref_cpu is not None:
        presets.append(int(S.ScenePreset.CORNELL))
    for preset in presets:
        # Reference CORNELL: pixels whose camera ray sees two of the coplanar wall quads at the same distance (the reference's
        # own second-hit distance, oracle/_ref) are tie pixels - every backend resolves them its own way - and are masked out
        keep = np.ones((H, W), bool)
        if preset == int(S.ScenePreset.CORNELL):
            cam = S.Camera(width=W, height=H)
            prim = util.primary_rays(cam, ref_cpu.camera_dirs(cam))
            hits, second = ref_cpu.preset(preset).trace_batch(prim)
            with np.errstate(invalid="ignore"):
                tie = (hits["prim_id"] >= 0) & (np.abs(second - hits["t"]) <= 1e-4 * np.maximum(1.0, hits["t"]))
            keep = ~tie.reshape(H, W)
            assert 0.0 < (~keep).mean() < 0.5
        imgs = {}
        for backend in ("wavefront", "megakernel", "b200"):
            out = str(tmp_path / ("%s%d.bin" % (backend, preset)))
            r = subprocess.run([exe, backend, str(preset), str(W), str(H), "1", str(frames), "1", "0", out],
                               capture_output=True, text=True, timeout=300)
            assert r.returncode == 0, r.stdout + r.stderr
            j = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
            assert j["cuda_error"] == "no error" and abs(j["weight_per_pixel"] - frames) < 1e-3
            raw = np.fromfile(out, np.float32)
            imgs[backend] = raw[:W * H * 3].reshape(H, W, 3) / frames
        for ref in ("wavefront", "megakernel"):
            m_ref, m_own = float(imgs[ref][keep].mean()), float(imgs["b200"][keep].mean())
            print("preset %d vs %s: mean %.5f / %.5f (%.2f %%), %.1f %% of pixels masked" % (preset, ref, m_ref, m_own, 100 * (m_own / m_ref - 1), 100 * (~keep).mean()))
            if preset == int(S.ScenePreset.CORNELL):
                # measured: -0.85 % vs the reference wavefront, -1.40 % vs its megakernel, which differ by 0.55 % from each other
                # on the same pixels: SECONDARY rays still meet the coplanar walls (bounded here at 2 %, not hidden)
                assert abs(m_own - m_ref) <= 0.02 * m_ref, (preset, ref, m_ref, m_own)
                continue            # block means straddle the masked tie regions
            assert abs(m_own - m_ref) <= 0.01 * m_ref, (preset, ref, m_ref, m_own)
            # coarse structure: 8x8 block means within 10 % + 0.02 absolute on >= 97 % of the blocks
            blk = lambda a: a[:88, :160].reshape(11, 8, 20, 8, 3).mean(axis=(1, 3))
            a, b = blk(imgs[ref]), blk(imgs["b200"])
            assert (np.abs(a - b) <= 0.10 * np.abs(a) + 0.02).mean() >= 0.97, (preset, ref)


# ---------------------------------------------------------------- BASELINE.json's full sizes, by size-independent properties
def test_full_size_1080p_properties(ctx):
    """configs[1] geometry (reference Cornell, 1920x1080, depth 8) at a few spp: determinism, sample ranges and pixel
    regions compose, two-stream waves equal one stream - on 2 M pixels instead of the small parity frames."""
    sd = S.cornell(); cam = S.Camera(width=1920, height=1080)
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
    P = 1920 * 1080
    ctx.clear_accum(); ctx.render(6, 8, seed=77); a = ctx.read_accum(); rays = ctx.stats().rays_total
    ctx.clear_accum(); ctx.render(6, 8, seed=77); assert np.array_equal(ctx.read_accum(), a)          # deterministic
    ctx.clear_accum(); ctx.render(2, 8, seed=77); ctx.render(4, 8, seed=77, sample_begin=2)
    assert np.array_equal(ctx.read_accum(), a)
    ctx.set_option("wave_paths", 3_000_000); ctx.set_option("overlap", 0)                            # 4+ waves, one stream
    ctx.clear_accum(); ctx.render(6, 8, seed=77); b = ctx.read_accum()
    ctx.set_option("overlap", 1)
    ctx.clear_accum(); ctx.render(6, 8, seed=77); c2 = ctx.read_accum()
    ctx.set_option("wave_paths", 0)
    assert np.array_equal(a, b) and np.array_equal(a, c2)
    ctx.clear_accum()
    for b0, n in ((0, P // 3), (P // 3, P // 3 + 11), (2 * (P // 3) + 11, P - 2 * (P // 3) - 11)):
        ctx.render_region(6, 8, seed=77, pixel_begin=b0, pixel_count=n)
    assert np.array_equal(ctx.read_accum(), a)
    assert P * 6 < rays <= P * 6 * 8 and np.isfinite(a).all() and a.min() >= 0.0
    assert a.max() <= 6 * 15.0 + 1e-3                                                                  # no pixel brighter than the light


@pytest.fixture(scope="module")
def big_scenes():
    """BASELINE.json's mesh configs at FULL size, built once: C3 = dragon x 3 x 16 = 960 K triangles + ground quad, C5 = 10 M
    random triangles; each with its bench camera."""
    import bench
    out = {}
    for key, wl in (("dragon960k", "dragon960k_1080p_64spp_d8"), ("soup10m", "soup10m_1080p_16spp_d16")):
        sd, cam, w = bench.make_workload(wl)
        out[key] = (sd, cam, w)
        bench._scene_cache.clear()                       # keep both alive here, not in the cache
    return out


@pytest.mark.parametrize("name", ["dragon960k", "soup10m"])
def test_full_size_mesh_vs_oracle(oracle, big_scenes, name):
    """The big scenes against the ORACLE's own BVH (not the GPU against itself): GPU BVH build == oracle build (integer
    outputs and node boxes, bit for bit, 960 K / 10 M primitives), closest hits of a 65 536-ray fixed batch + 16 384
    incoherent rays through the two-child AND the 4-wide walk == the oracle's records, traversal counters equal, and a
    64-row band of the bench frame (depth 8 / 16) == the oracle's image with the same number of rays."""
    sd, cam, w = big_scenes[name]
    osc = oracle.scene(sd)
    o = osc.get_lbvh()
    lo, hi = util.scene_bounds(sd)
    rays = np.concatenate([util.fixed_ray_batch(sd, 65536, seed=1, clamp_extent=60.0),
                           util.incoherent_rays(16384, float(min(50.0, np.abs(np.concatenate([lo, hi])).max())))])
    with Context(0) as c:
        c.set_scene(sd); c.build_bvh()
        g = c.debug_lbvh()
        for k in ("scene_bounds", "morton", "sorted_prim", "node_children", "node_parent", "node_bounds"):
            assert np.array_equal(g[k], o[k]), k
        st, info = c.stats(), osc.build_info()
        assert (st.bvh_builder, st.bvh_height, st.ploc_rounds) == (info["builder"], info["height"], info["rounds"])
        c.enable_counters(True)
        for mode in (1, 2):
            cnt = np.zeros(2, np.uint64)
            want = osc.trace_batch(rays, use_bvh=mode, n_threads=16, counters=cnt)
            got = c.trace_batch(rays, use_bvh=mode); stt = c.stats()
            assert np.array_equal(got["prim_id"], want["prim_id"]) and np.array_equal(got["t"], want["t"]), mode
            assert np.array_equal(got, want), mode
            assert (stt.nodes_visited, stt.prims_tested) == (int(cnt[0]), int(cnt[1])), mode
            assert (got["prim_id"] >= 0).mean() > 0.05
        c.enable_counters(False)
        band = S.Camera(pos=tuple(cam.pos), front=tuple(cam.front_in), width=cam.width, height=cam.height)
        y0 = cam.height // 2 - 32
        c.set_camera(band); c.clear_accum()
        c.render_region(1, w["depth"], seed=3, pixel_begin=y0 * cam.width, pixel_count=64 * cam.width)
        got = c.read_accum()[y0:y0 + 64]
        want, orays = osc.render(band, 1, w["depth"], seed=3, n_threads=16, y_begin=y0, y_end=y0 + 64)
        assert np.array_equal(got, want[y0:y0 + 64]) and c.stats().rays_total == orays
        assert got.max() > 0


def test_full_size_10m_triangles_wide_equals_two_child(big_scenes):
    """configs[4] geometry (10 M random triangles, depth 16) as an IMAGE: the render through the 4-wide quantised nodes
    equals the render through the two-child tree bit for bit, with about half the node visits."""
    sd = big_scenes["soup10m"][0]
    cam = S.Camera(pos=(0.0, 0.0, 150.0), front=(0.0, 0.0, -1.0), width=640, height=360)
    out = []
    with Context(0) as c:
        c.enable_counters(True)
        for wide in (1, 0):
            c.set_option("bvh_wide", wide)
            c.set_scene(sd); c.build_bvh(); c.set_camera(cam); c.clear_accum()
            c.render(1, 16, seed=5)
            st = c.stats()
            out.append((c.read_accum(), st.rays_total, st.nodes_visited))
    assert np.array_equal(out[0][0], out[1][0]) and out[0][1] == out[1][1]
    assert out[0][2] < 0.7 * out[1][2]
    assert out[0][0].max() > 0
