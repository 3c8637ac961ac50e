"""Tier-1 golden images: the reference CPU backend EXACTLY AS SHIPPED - CPURenderer::ProgressiveRender
(src/backend/cpu/renderer.cpp:18-103: depth literal 20, std::rand, one std::async task per 16x16 tile) compiled from
/root/reference by oracle/Makefile - rendered at 64x64, 4096 spp from the viewer camera. The run is not
reproducible sample by sample (unseeded, racy rand()), so these are used statistically: an independent-seed
render of ours must agree within Monte-Carlo error. Run here (needs /root/reference); output is committed.

  python tests/golden/make_tier1.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_lib  # noqa: E402
from parallelraytracing_b200 import scene as S  # noqa: E402

W = H = 64
SPP = 4096
ref = oracle_lib.load_ref()
assert ref is not None, "oracle/_ref is not built (needs /root/reference)"
out = {"meta": np.array([W, H, SPP, 20], np.int32)}
for p in (S.ScenePreset.CORNELL, S.ScenePreset.MATERIAL_TEST, S.ScenePreset.DEFAULT, S.ScenePreset.LIGHT_TEST):
    sec, acc = ref.render_shipped(int(p), S.Camera(width=W, height=H), SPP)
    out["mean_" + p.name.lower()] = (acc / SPP).astype(np.float32)
    print(p.name, "%.1f s" % sec, "mean radiance", float(acc.mean() / SPP))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "tier1_shipped.npz"), **out)
