"""CPU tests of the traversal structures the reference does not have (oracle only, no GPU): the PLOC tree-quality pass
(SURVEY 8f-4), its Karras fallback, the conservativeness of the 4-wide quantised nodes and the traversal-stack bound.
The GPU mirrors of these tests are in test_gpu_parity.py; both sides must produce the same bits."""
import numpy as np
import pytest

import util
from parallelraytracing_b200 import scene as S

STACK_TOTAL = 256          # parallelraytracing_b200/csrc/wavefront.cu


@pytest.fixture(scope="module")
def gold():
    return np.load(util.GOLDEN + "/reference_vectors.npz")


def _tree_ok(g, n):
    ch, par = g["node_children"], g["node_parent"]
    assert par[0] == -1 and (par == -1).sum() == 1
    assert np.array_equal(np.sort(ch[ch >= 0]), np.arange(1, n - 1)), "every internal node but the root is a child once"
    assert np.array_equal(np.sort(~ch[ch < 0]), np.arange(n)), "every leaf is a child once"
    nb = g["node_bounds"]
    for side in (0, 1):                                   # the box stored for an internal child encloses that child's own two boxes
        c = ch[:, side]; m = c >= 0
        cmin = np.minimum(nb[c[m], 0:3], nb[c[m], 6:9]); cmax = np.maximum(nb[c[m], 3:6], nb[c[m], 9:12])
        assert np.array_equal(nb[m, 6 * side:6 * side + 3], cmin) and np.array_equal(nb[m, 6 * side + 3:6 * side + 6], cmax)


@pytest.mark.parametrize("name", ["soup", "dragon", "balls"])
def test_ploc_tree_is_valid_and_hits_equal_brute_force(oracle, gold, name):
    if name == "soup":
        sd = S.random_triangle_soup(3000); rays = util.incoherent_rays(4096, 50.0)
    elif name == "dragon":
        f = gold["mesh_dragon_f"].reshape(-1, 3)[:4000]
        sd = S.mesh_scene((gold["mesh_dragon_v"], gold["mesh_dragon_n"], f), instances=((0, 0, 0),), name="dragon4k")
        rays = util.fixed_ray_batch(sd, 4096, seed=2, clamp_extent=3.0)
    else:
        sd = S.make_scene(S.ScenePreset.RANDOM_BALLS_SMALL); rays = util.fixed_ray_batch(sd, 4096, seed=2, clamp_extent=60.0)
    brute = oracle.scene(sd).trace_batch(rays, use_bvh=0)
    for radius in (1, 4, 8):
        osc = oracle.scene(sd); osc.set_builder(1, radius)
        info = osc.build_info()
        assert info["builder"] == 1 and 0 < info["height"] <= 84 and info["rounds"] < 192
        _tree_ok(osc.get_lbvh(), sd.n_prims)
        for mode in (1, 2):
            assert np.array_equal(osc.trace_batch(rays, use_bvh=mode), brute), (radius, mode)


def test_ploc_visits_fewer_nodes_than_the_karras_tree(oracle, gold):
    """The quality claim, on the 60 K-triangle dragon scene with its 20 x 20 ground quad: camera rays and diffuse bounce
    rays fetch clearly fewer nodes through the PLOC tree, two-child and 4-wide, with identical hit records."""
    sd = S.mesh_scene((gold["mesh_dragon_v"], gold["mesh_dragon_n"], gold["mesh_dragon_f"]), name="dragon3")
    cam = S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=128, height=72)
    prim = util.primary_rays(cam, oracle.camera_dirs(cam))
    out = {}
    for builder in (0, 1):
        osc = oracle.scene(sd); osc.set_builder(builder, 4)
        hits = osc.trace_batch(prim, use_bvh=1)
        rays = np.concatenate([prim, util.bounce_rays(hits)])
        for mode in (1, 2):
            cnt = np.zeros(2, np.uint64)
            h = osc.trace_batch(rays, use_bvh=mode, counters=cnt)
            out[builder, mode] = (h, int(cnt[0]))
    for mode in (1, 2):
        assert np.array_equal(out[0, mode][0], out[1, mode][0])
        assert out[1, mode][1] < 0.75 * out[0, mode][1], (mode, out[0, mode][1], out[1, mode][1])


def test_ploc_falls_back_to_karras_when_it_does_not_converge(oracle):
    sd = util.nested_scene(300)
    osc = oracle.scene(sd); osc.set_builder(1, 4)
    info = osc.build_info()
    assert info["builder"] == 0 and info["rounds"] == 192          # gave up at the round limit: Karras hierarchy instead
    k = oracle.scene(sd); k.set_builder(0)
    a, b = osc.get_lbvh(), k.get_lbvh()
    for key in a:
        assert np.array_equal(a[key], b[key]), key
    rays = util.rays_from_point((1.0, 2.0, 3.05), 2000)
    assert np.array_equal(osc.trace_batch(rays, use_bvh=2), osc.trace_batch(rays, use_bvh=0))


def test_identical_boxes_pair_up(oracle):
    """4096 identical boxes: with 'lowest position' as the only tie rule PLOC chains (one merge per round); the i^1 rule
    pairs them, so the build converges in ~log2(n) rounds."""
    osc = oracle.scene(util.deep_scene()); osc.set_builder(1, 4)
    info = osc.build_info()
    assert info["builder"] == 1 and info["rounds"] <= 40 and info["height"] <= 40


def test_wide_nodes_are_conservative_at_box_corners(oracle):
    """Advisor finding (round 1): a child whose low face coincides with the node's own minimum was quantised with no margin
    there, and the fused decode of the walk rounds by up to 2^-9 grid step, so rays clipping that corner skipped the child
    (1.1 % of this batch). The grid now starts one step below the boxes: the 4-wide walk must equal brute force."""
    sd, tv = util.grazing_scene()
    rays = util.grazing_rays(tv, 400000)
    for builder in (0, 1):
        osc = oracle.scene(sd); osc.set_builder(builder, 4)
        brute = osc.trace_batch(rays, use_bvh=0)
        assert (brute["prim_id"] >= 0).mean() > 0.9
        assert np.array_equal(osc.trace_batch(rays, use_bvh=2)["prim_id"], brute["prim_id"])
        assert np.array_equal(osc.trace_batch(rays, use_bvh=1), brute)


def test_traversal_stack_bound(oracle):
    """The device walks hold STACK_TOTAL = 256 entries. The 4-wide walk pushes up to three siblings per binary level, so it can
    need more than the 64 entries the kernels had in round 1 (advisor finding); the oracle reports the deepest stack any ray
    needed, in device entries. Both builders, every ray from inside all boxes."""
    sd = util.deep_scene()
    rays = util.rays_from_point((0.45, 0.45, 0.45), 20000)
    for builder in (0, 1):
        osc = oracle.scene(sd); osc.set_builder(builder, 4)
        brute = osc.trace_batch(rays, use_bvh=0)
        for mode in (1, 2):
            oracle.lib.po_reset_max_stack()
            assert np.array_equal(osc.trace_batch(rays, use_bvh=mode), brute)
            need = oracle.lib.po_get_max_stack()
            assert 20 < need <= STACK_TOTAL, (builder, mode, need)
