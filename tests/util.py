"""Shared helpers of the test-suite: deterministic ray batches, scene bounds, comparison metrics."""
import os

import numpy as np

from parallelraytracing_b200.capi import ray_dt
from parallelraytracing_b200 import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
ASSETS = os.path.join(ROOT, "tests", "golden")          # small mesh fixtures live beside the vectors


def scene_bounds(sd):
    lo, hi = [], []
    if len(sd.spheres):
        c = np.stack([sd.spheres["cx"], sd.spheres["cy"], sd.spheres["cz"]], 1)
        r = sd.spheres["radius"][:, None]
        lo.append((c - r).min(0)); hi.append((c + r).max(0))
    if len(sd.quads):
        q = sd.quads
        c = np.stack([q["cx"], q["cy"], q["cz"]], 1)
        u = np.stack([q["ux"], q["uy"], q["uz"]], 1) * q["half_w"][:, None]
        v = np.stack([q["vx"], q["vy"], q["vz"]], 1) * q["half_h"][:, None]
        e = np.abs(u) + np.abs(v)
        lo.append((c - e).min(0)); hi.append((c + e).max(0))
    if sd.n_tris:
        v = sd.tri_verts.reshape(-1, 3)
        lo.append(v.min(0)); hi.append(v.max(0))
    return np.min(lo, 0).astype(np.float32), np.max(hi, 0).astype(np.float32)


def fixed_ray_batch(sd, n=65536, seed=1, clamp_extent=None):
    """SURVEY 8(d) fixed batch: origin uniform on the sphere of radius 2R about the scene centre,
    direction towards a uniform point in the scene AABB; unit directions, float32."""
    lo, hi = scene_bounds(sd)
    if clamp_extent is not None:                        # huge ground quads would hide everything else
        c0 = (lo + hi) / 2
        lo, hi = np.maximum(lo, c0 - clamp_extent), np.minimum(hi, c0 + clamp_extent)
    rng = np.random.default_rng(seed)
    c = (lo + hi) / 2
    R = float(np.linalg.norm(hi - lo) / 2)
    g = rng.standard_normal((n, 3))
    g /= np.linalg.norm(g, axis=1, keepdims=True)
    o = (c + 2 * R * g).astype(np.float32)
    tgt = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = (tgt - o).astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    # re-normalise in float32 the way glm would, so |d| is as close to 1 as float32 allows
    rays = np.zeros(n, ray_dt)
    rays["ox"], rays["oy"], rays["oz"] = o[:, 0], o[:, 1], o[:, 2]
    rays["dx"], rays["dy"], rays["dz"] = d[:, 0], d[:, 1], d[:, 2]
    return rays


def primary_rays(cam, dirs):
    n = cam.width * cam.height
    rays = np.zeros(n, ray_dt)
    d = dirs.reshape(-1, 3)
    rays["ox"], rays["oy"], rays["oz"] = cam.pos
    rays["dx"], rays["dy"], rays["dz"] = d[:, 0], d[:, 1], d[:, 2]
    return rays


def image_metrics(a, b, floor=1e-3, rel_flip=1e-2):
    """a = reference mean image, b = candidate. Returns dict(rmse, max_abs, flips, n, rmse_nonflip,
    max_rel_nonflip): a pixel is a 'flip' when its max relative error exceeds rel_flip (a discrete
    decision went the other way somewhere along one of its paths)."""
    d = np.abs(a.astype(np.float64) - b.astype(np.float64))
    rel = d / np.maximum(np.abs(a), floor)
    flip = rel.max(-1) > rel_flip
    nf = ~flip
    return dict(rmse=float(np.sqrt((d ** 2).mean())), max_abs=float(d.max()), flips=int(flip.sum()), n=int(flip.size),
                rmse_nonflip=float(np.sqrt((d[nf] ** 2).mean())) if nf.any() else 0.0,
                max_rel_nonflip=float(rel[nf].max()) if nf.any() else 0.0,
                biteq=float((d.max(-1) == 0).mean()))


def scene_from_npz(z, prefix):
    sd = S.SceneData(name=prefix)
    sd.spheres, sd.quads, sd.materials = z[prefix + "_spheres"], z[prefix + "_quads"], z[prefix + "_materials"]
    if prefix + "_tri_verts" in z.files:
        sd.tri_verts, sd.tri_normals = z[prefix + "_tri_verts"], z[prefix + "_tri_normals"]
        sd.tri_material = z[prefix + "_tri_material"]
    sd.tri_prim_base = int(z[prefix + "_tri_prim_base"])
    return sd


def scene_to_npz(d, prefix, sd):
    d[prefix + "_spheres"], d[prefix + "_quads"], d[prefix + "_materials"] = sd.spheres, sd.quads, sd.materials
    if sd.n_tris:
        d[prefix + "_tri_verts"], d[prefix + "_tri_normals"], d[prefix + "_tri_material"] = sd.tri_verts, sd.tri_normals, sd.tri_material
    d[prefix + "_tri_prim_base"] = np.int64(sd.tri_prim_base)
