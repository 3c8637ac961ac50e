"""Generate tests/golden/*.npz from the REFERENCE's own compiled CPU code (oracle/_ref/libprt_ref.so,
built by `make -C oracle ref` from the sources under /root/reference). Run in the build container:

    python tests/golden/make_golden.py

The reference ships no tests or golden vectors, so these fixtures are the pin:
every value below is an output of reference code (Scene presets, Camera::GetCameraRay,
PrimitiveList::Intersect, MaterialHandle::Scatter, CPURenderer::TraceRay semantics with the
interposed deterministic rand(), Film::UpdateDisplay, Mesh PLY loading). They travel to the GPU
box, where /root/reference does not exist.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O            # noqa: E402
import util                       # noqa: E402
from parallelraytracing_b200 import scene as S   # noqa: E402
from parallelraytracing_b200.capi import MAT_DIELECTRIC, MAT_EMISSIVE, MAT_LAMBERTIAN, MAT_METAL  # noqa: E402

REF_MODELS = "/root/reference/assets/models"
PRESETS = {p.name.lower(): int(p) for p in S.ScenePreset}


def main():
    ref = O.load_ref()
    assert ref is not None, "build oracle/_ref first: make -C oracle ref"
    out = {}

    # 1. every Scene preset through host/flatten.h (scene.cpp:62-350)
    for name, p in PRESETS.items():
        util.scene_to_npz(out, "scene_" + name, ref.preset(p).flatten(name))

    # 2. Camera::GetCameraRay at pixel centres (camera.h:104-132), viewer camera (main.cpp:142-150)
    cam = S.Camera(width=64, height=36)
    out["cam64x36_dirs"] = ref.camera_dirs(cam)
    out["cam64x36_struct"] = np.frombuffer(bytes(ref.camera(cam)), np.uint8).copy()
    cam2 = S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=48, height=48)
    out["cam48_dirs"] = ref.camera_dirs(cam2)

    # 3. PrimitiveList::Intersect on fixed ray batches (primitive.cpp:21-59)
    for name in ("default", "material_test", "cornell", "random_balls_large"):
        rs = ref.preset(PRESETS[name])
        sd = rs.flatten(name)
        rays = np.concatenate([util.fixed_ray_batch(sd, 3072, seed=1, clamp_extent=30.0),
                               util.primary_rays(cam, out["cam64x36_dirs"])])
        hits, second = rs.trace_batch(rays)
        out["hits_%s_rays" % name], out["hits_%s_ref" % name], out["hits_%s_second" % name] = rays, hits, second

    # 4. tier-2 deterministic images: reference Intersect/Scatter/Emit/GetCameraRay + TraceRay recursion
    #    (cpu/renderer.cpp:59-103), rand() interposed with the (seed,pixel,sample,bounce) stream
    for name, (W, H, spp, depth) in {"cornell": (64, 64, 16, 8), "default": (64, 64, 8, 8), "material_test": (64, 64, 8, 8),
                                     "random_balls_large": (64, 36, 4, 8), "light_test": (48, 48, 4, 20)}.items():
        c = S.Camera(width=W, height=H)
        acc, rays = ref.preset(PRESETS[name]).render(c, spp, depth, seed=1)
        out["img_%s" % name] = acc
        out["img_%s_meta" % name] = np.array([W, H, spp, depth, 1, rays], np.int64)

    # 5. MaterialHandle::Scatter known answers (material.h:11-161)
    rng = np.random.default_rng(7)
    mats, ins, nrm, fronts, res = [], [], [], [], []
    table = [(MAT_LAMBERTIAN, 0.8, 0.3, 0.2, 0.0), (MAT_METAL, 0.9, 0.8, 0.7, 0.0), (MAT_METAL, 0.9, 0.8, 0.7, 0.3),
             (MAT_DIELECTRIC, 1, 1, 1, 1.5), (MAT_DIELECTRIC, 1, 1, 1, 0.9), (MAT_EMISSIVE, 15, 15, 15, 0.0)]
    for k in range(600):
        m = table[k % len(table)]
        n = rng.standard_normal(3); n /= np.linalg.norm(n)
        d = rng.standard_normal(3); d /= np.linalg.norm(d)
        n = n.astype(np.float32); d = d.astype(np.float32)
        if np.dot(d, n) > 0:
            n = -n                                            # hit normals face the ray (shape.h)
        front = int(rng.integers(0, 2))
        mat = (m[0], m[1], m[2], m[3], m[4], (0, 0, 0))
        ok, a, o, e = ref.scatter(mat, d, np.zeros(3, np.float32), n, front, 1, k, k // 7, k % 5)
        mats.append(m); ins.append(d); nrm.append(n); fronts.append(front)
        res.append(np.concatenate([[float(ok)], a, o if ok else np.zeros(3), e]))
    out["scatter_mats"], out["scatter_in"], out["scatter_n"] = np.array(mats, np.float32), np.array(ins), np.array(nrm)
    out["scatter_front"], out["scatter_out"] = np.array(fronts, np.int32), np.array(res, np.float32)

    # 6. Film::UpdateDisplay (film.cu:134-194)
    acc = (rng.random((16, 16, 3)) ** 3 * 40).astype(np.float32)
    acc[0, 0] = 0; acc[0, 1] = 1e6
    out["film_accum"], out["film_display"] = acc, ref.film_display(acc, 8, 1.0, 2.2)

    # 7. Mesh (tinyply) loading of the bundled PLY models (mesh.cpp:23-184) + brute-force hits on dragon
    for m in ("icosahedron", "dragon", "bunny"):
        v, n, f = ref.load_mesh(os.path.join(REF_MODELS, m + ".ply"))
        out["mesh_%s_v" % m], out["mesh_%s_f" % m] = v, f
        if n is not None:
            out["mesh_%s_n" % m] = n
    v, n, f = out["mesh_dragon_v"], out["mesh_dragon_n"], out["mesh_dragon_f"]
    # the mesh the bench workloads are built from ships with the package (bench.py must not depend on a test fixture)
    S.write_ply(os.path.join(S.ASSETS, "dragon.ply"), v, n, f,
                comment="arrays the reference's own Mesh loader returns for its assets/models/dragon.ply (10000 vertices, 20000 faces)")
    rs = ref.new_scene()
    lam = (MAT_LAMBERTIAN, 0.8, 0.3, 0.3, 0.0, (0, 0, 0))
    rs.add_quad(20, 20, (MAT_LAMBERTIAN, 0.5, 0.5, 0.5, 0.0, (0, 0, 0)), trans=(0, -1, 0))
    rs.add_triangles(v[f.astype(np.int64)], n[f.astype(np.int64)], lam)
    sd = rs.flatten("dragon1")
    c3 = S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=48, height=48)
    rays = np.concatenate([util.fixed_ray_batch(sd, 2048, seed=2, clamp_extent=2.0), util.primary_rays(c3, out["cam48_dirs"])])
    hits, second = rs.trace_batch(rays)
    out["hits_dragon_rays"], out["hits_dragon_ref"], out["hits_dragon_second"] = rays, hits, second
    acc, nr = rs.render(c3, 2, 8, seed=1)
    out["img_dragon"] = acc
    out["img_dragon_meta"] = np.array([48, 48, 2, 8, 1, nr], np.int64)

    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote", os.path.join(HERE, "reference_vectors.npz"), "%d arrays" % len(out))


if __name__ == "__main__":
    main()
