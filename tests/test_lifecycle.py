"""Life-cycle hygiene of the C-ABI context on the GPU: device memory comes back after prt_destroy, repeated scene / camera
changes do not grow it, and the entry points can be called in any sensible order (the reference's backends leak their scene
buffers, cuda_megakernel/renderer.cu:178-186; this library must not)."""
import numpy as np
import pytest

from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context, PrtError

pytestmark = pytest.mark.gpu


def free_bytes():
    import torch
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def exercise(ctx, n_tris, W, H, spp=2):
    sd = S.random_triangle_soup(n_tris, seed=n_tris)
    ctx.set_scene(sd); ctx.build_bvh()
    ctx.set_camera(S.Camera(pos=(0.0, 0.0, 150.0), front=(0.0, 0.0, -1.0), width=W, height=H))
    ctx.clear_accum(); ctx.render(spp, 4, seed=1)
    img = ctx.read_accum()
    assert np.isfinite(img).all() and img.sum() > 0
    ctx.resolve_display(1.0, 2.2)
    return img


def test_destroy_returns_device_memory():
    c = Context(0); exercise(c, 2000, 64, 48); c.close()          # warm: CUDA modules, torch context
    base = free_bytes()
    for k in range(12):
        c = Context(0)
        exercise(c, 1000 + 3000 * (k % 4), 96 + 32 * (k % 3), 64)
        c.close()
    assert base - free_bytes() <= (2 << 20), "device memory lost across create/destroy cycles"


def test_repeated_scene_and_camera_changes_do_not_grow():
    c = Context(0)
    try:
        exercise(c, 40000, 256, 192)                               # the largest configuration first: buffers only ever grow to the maximum
        c.trace_batch(np.zeros(70000, dtype=[("ox", "f4"), ("oy", "f4"), ("oz", "f4"), ("dx", "f4"), ("dy", "f4"), ("dz", "f4")]))
        high = free_bytes()
        ref = None
        for k in range(20):
            img = exercise(c, [500, 40000, 7000, 1][k % 4], [64, 256, 128][k % 3], [48, 192, 96][k % 3])
            if k % 12 == 1:                                         # the same configuration gives the same image every time round
                ref = img if ref is None else ref
                assert np.array_equal(img, ref)
        assert high - free_bytes() <= (2 << 20), "device memory grows with repeated prt_set_scene / prt_set_camera"
    finally:
        c.close()


def test_calls_out_of_order_fail_cleanly():
    c = Context(0)
    try:
        with pytest.raises(PrtError):
            c.build_bvh()                                           # no scene
        with pytest.raises(PrtError):
            c.render(1, 4, seed=1)                                  # no scene, no camera
        sd = S.cornell()
        c.set_scene(sd)
        with pytest.raises(PrtError):
            c.render(1, 4, seed=1)                                  # no BVH, no camera
        c.build_bvh()
        with pytest.raises(PrtError):
            c.render(1, 4, seed=1)                                  # no camera
        c.set_camera(S.Camera(width=32, height=32))
        with pytest.raises(PrtError):
            c.render(-1, 4, seed=1)
        with pytest.raises(PrtError):
            c.render(1, -1, seed=1)
        c.clear_accum(); c.render(2, 0, seed=1)                     # TraceRay(depth <= 0) = 0 (cpu/renderer.cpp:61): weight only
        assert c.accum_spp() == 2 and not c.read_accum().any()
        with pytest.raises(PrtError):
            c.render(1, 33, seed=1)                                 # PRT_MAX_DEPTH = 32
        c.clear_accum(); c.render(1, 4, seed=1)
        a = c.read_accum()
        c.set_scene(sd)                                             # a new scene invalidates the BVH, not the film
        with pytest.raises(PrtError):
            c.render(1, 4, seed=1)
        c.build_bvh(); c.clear_accum(); c.render(1, 4, seed=1)
        assert np.array_equal(a, c.read_accum())
    finally:
        c.close()


def test_set_scene_build_equals_the_two_calls():
    """prt_set_scene_build (positions first, BVH build under the upload of normals and materials) leaves the device in the state
    prt_set_scene + prt_build_bvh leave it in: integer outputs, node bytes, hit records with their normals and materials, and
    images - with vertex normals, with flat normals (NULL), from pageable and from page-locked arrays, and without triangles."""
    import dataclasses
    import util
    from parallelraytracing_b200.capi import pinned_copy
    mesh = S.mesh_scene(S.bundled_mesh("dragon"), subdiv=1, name="dragon240k")        # 240 K triangles with vertex normals, 3 materials + ground
    soup = S.random_triangle_soup(30000, seed=9)
    flat = dataclasses.replace(soup, tri_normals=None)
    pinned = dataclasses.replace(mesh, tri_verts=pinned_copy(mesh.tri_verts), tri_normals=pinned_copy(mesh.tri_normals),
                                 tri_material=pinned_copy(mesh.tri_material))
    cam = S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=96, height=64)
    a, b = Context(0), Context(0)
    try:
        for sd in (mesh, flat, pinned, S.make_scene(S.ScenePreset.RANDOM_BALLS_LARGE), S.cornell(), soup):
            for builder in (1, 0):
                for c in (a, b):
                    c.set_option("bvh_builder", builder)
                a.set_scene(sd); a.build_bvh()
                b.set_scene_build(sd)
                ga, gb = a.debug_lbvh(), b.debug_lbvh()
                for k in ga:
                    assert ga[k].tobytes() == gb[k].tobytes(), (sd.name, builder, k)
                if sd.n_prims >= 4096:
                    assert a.debug_wide().tobytes() == b.debug_wide().tobytes()
                rays = util.fixed_ray_batch(sd, 20000, seed=2, clamp_extent=60.0)
                ha, hb = a.trace_batch(rays, use_bvh=2), b.trace_batch(rays, use_bvh=2)
                assert ha.tobytes() == hb.tobytes(), (sd.name, builder)
                assert (ha["prim_id"] >= 0).sum() > 100
                for c in (a, b):
                    c.set_camera(cam); c.clear_accum(); c.render(2, 6, seed=3)
                assert np.array_equal(a.read_accum(), b.read_accum()), (sd.name, builder)
                assert a.stats().rays_total == b.stats().rays_total
            # a later prt_build_bvh with other options rebuilds from what prt_set_scene_build left on the device
            for c in (a, b):
                c.set_option("bvh_builder", 1); c.set_option("ploc_radius", 2); c.build_bvh(); c.set_option("ploc_radius", 4)
                c.clear_accum(); c.render(2, 6, seed=4)
            assert a.debug_lbvh()["node_children"].tobytes() == b.debug_lbvh()["node_children"].tobytes()
            assert np.array_equal(a.read_accum(), b.read_accum()), (sd.name, "rebuild")
    finally:
        a.close(); b.close()


def test_two_contexts_in_two_host_threads():
    """A context is not thread-safe, but contexts are independent of each other: two host threads, one context each, building
    and rendering at the same time on the same GPU, get the images they get one after the other (ctypes drops the GIL
    around every call, so the calls really overlap)."""
    import threading
    jobs = [(S.mesh_scene(S.bundled_mesh("dragon"), subdiv=1, name="dragon240k"), S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=160, height=96), 6),
            (S.cornell(), S.Camera(width=192, height=128), 8)]

    def run(job, out, reps):
        sd, cam, depth = job
        with Context(0) as c:
            for k in range(reps):
                c.set_scene_build(sd); c.set_camera(cam); c.clear_accum()
                c.render(4, depth, seed=11)
                out.append((c.read_accum(), c.stats().rays_total))

    want = []
    for job in jobs:
        o = []; run(job, o, 1); want.append(o[0])
    outs = [[], []]
    th = [threading.Thread(target=run, args=(jobs[k], outs[k], 12)) for k in range(2)]
    for t in th: t.start()
    for t in th: t.join()
    for k in range(2):
        assert len(outs[k]) == 12
        for img, rays in outs[k]:
            assert rays == want[k][1] and np.array_equal(img, want[k][0]), "context %d disturbed by its neighbour" % k
