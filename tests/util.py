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


# ---------------------------------------------------------------- adversarial scenes for the traversal structures
def _tri_with_box(c, r):
    """A triangle whose AABB is exactly [c - r, c + r] (its centre, which the Morton code is taken from, is c)."""
    c = np.asarray(c, np.float64); r = np.asarray(r, np.float64)
    return [c + (-r[0], -r[1], -r[2]), c + (r[0], -r[1], r[2]), c + (0.0, r[1], 0.0)]


def _tri_scene(name, tris):
    b = S.SceneBuilder(name)
    b.add_triangles(np.array(tris, np.float32), None, b.lambertian((0.5, 0.5, 0.5)))
    return b.build()


def grazing_scene():
    """Three long triangles whose x-min vertex is also the y-max corner of their box, node extent ~100 in x (grid step 0.5
    of the 4-wide quantised nodes): rays that clip that corner found the low face of the quantised box without margin
    (advisor finding, round 1). Returns (scene, triangle vertices)."""
    tv = np.array([[[0.0, 1.0, 0.0], [100.0, 0.0, 0.3], [60.0, 0.2, -0.4]],
                   [[3.0, 5.0, 1.0], [90.0, 4.0, 1.5], [50.0, 4.5, 0.7]],
                   [[-7.0, -3.0, 2.0], [80.0, -4.0, 2.5], [40.0, -3.5, 1.7]]], np.float32)
    return _tri_scene("graze", tv), tv


def grazing_rays(tv, n, seed=1):
    """Steep rays that hit within ~3e-4 of the corner vertex of a triangle of grazing_scene()."""
    rng = np.random.default_rng(seed)
    k = rng.integers(0, len(tv), n)
    vert = tv[k, 0].astype(np.float64)
    w = rng.random((n, 2)) * 3e-6
    tgt = vert + w[:, :1] * (tv[k, 1] - vert) + w[:, 1:] * (tv[k, 2] - vert)
    d = np.stack([-(0.01 + 0.05 * rng.random(n)), -(0.2 + rng.random(n)) * 4, 0.3 * rng.standard_normal(n)], 1)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = tgt - d * (5.0 + 10 * rng.random((n, 1)))
    r = np.zeros(n, ray_dt)
    r["ox"], r["oy"], r["oz"] = o.astype(np.float32).T
    r["dx"], r["dy"], r["dz"] = d.astype(np.float32).T
    return r


def deep_scene(n_center=4096):
    """Deep, lopsided hierarchy: n_center co-centred triangles in ONE Morton cell (identical codes: Karras splits them by
    index) + for every code bit 3..29 three triangles in neighbouring cells whose boxes all contain the point (0.45,
    0.45, 0.45) and grow with the bit. From that point every box is entered at t = 0, so the walks push every sibling."""
    tris = [_tri_with_box((0.5, 0.5, 0.5), (0.3, 0.3, 0.3)) for _ in range(n_center)]
    for b in range(3, 30):
        for low in (0, 1, 2):
            cell = [0, 0, 0]
            cell[2 - b % 3] = 1 << (b // 3)
            if low == 1: cell[2] |= 1
            if low == 2: cell[1] |= 1
            tris.append(_tri_with_box([k + 0.5 for k in cell], [k + 0.5 - 0.45 + 0.05 * (1 + low) for k in cell]))
    tris.append(_tri_with_box((1024.0, 1024.0, 1024.0), (0.2, 0.2, 0.2)))
    return _tri_scene("deep", tris)


def nested_scene(n=300):
    """n concentric triangles with geometrically growing boxes (identical Morton code): every cluster's best partner is a
    smaller one, so PLOC merges ONE pair per round and gives up after its round limit - the Karras fallback must take over."""
    return _tri_scene("nested", [_tri_with_box((1.0, 2.0, 3.0), (0.1 * 1.05 ** i,) * 3) for i in range(n)])


def rays_from_point(p, n, seed=1):
    rng = np.random.default_rng(seed)
    g = rng.standard_normal((n, 3)); g /= np.linalg.norm(g, axis=1, keepdims=True)
    r = np.zeros(n, ray_dt)
    r["ox"], r["oy"], r["oz"] = p
    r["dx"], r["dy"], r["dz"] = g.astype(np.float32).T
    return r


def incoherent_rays(n, extent, seed=3):
    """Origins uniform in [-extent, extent]^3, directions uniform on the sphere (what a diffuse bounce inside a soup looks like)."""
    rng = np.random.default_rng(seed)
    o = (rng.random((n, 3)) * 2 - 1) * extent
    g = rng.standard_normal((n, 3)); g /= np.linalg.norm(g, axis=1, keepdims=True)
    r = np.zeros(n, ray_dt)
    r["ox"], r["oy"], r["oz"] = o.astype(np.float32).T
    r["dx"], r["dy"], r["dz"] = g.astype(np.float32).T
    return r


def bounce_rays(hits, seed=5):
    """Diffuse-like secondary rays from the hit records of a traced batch (origin = hit point, direction = n + unit vector)."""
    m = hits["prim_id"] >= 0
    rng = np.random.default_rng(seed)
    nrm = np.stack([hits["nx"], hits["ny"], hits["nz"]], 1)[m].astype(np.float64)
    p = np.stack([hits["px"], hits["py"], hits["pz"]], 1)[m]
    g = rng.standard_normal(p.shape); g /= np.linalg.norm(g, axis=1, keepdims=True)
    d = nrm + 0.999 * g; d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros(len(p), ray_dt)
    r["ox"], r["oy"], r["oz"] = p.astype(np.float32).T
    r["dx"], r["dy"], r["dz"] = d.astype(np.float32).T
    return r
