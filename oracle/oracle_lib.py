"""ctypes bindings of the CPU oracle (oracle/libprt_oracle.so) and of the reference harness
(oracle/_ref/libprt_ref.so). TEST INFRASTRUCTURE: imported only from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs. The product never loads these."""
import ctypes as C
import os
import subprocess

import numpy as np

from parallelraytracing_b200.capi import CCamera, hit_dt, material_dt, quad_dt, ray_dt, sphere_dt
from parallelraytracing_b200.scene import SceneData

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libprt_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libprt_ref.so")
vp, i32, u32, u64, f32 = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_float


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(C.c_void_p)


def build_oracle():
    src = os.path.join(ORACLE_DIR, "prt_oracle.cpp")
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, capture_output=True)
    return ORACLE_SO


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        lib.po_scene_create.restype = vp
        lib.po_scene_create.argtypes = [vp, i32, vp, i32, vp, vp, vp, i32, i32, i32, vp, i32, vp]
        lib.po_scene_destroy.argtypes = [vp]
        lib.po_num_prims.argtypes = [vp]
        lib.po_build_lbvh.argtypes = [vp]
        lib.po_get_lbvh.argtypes = [vp] * 7
        lib.po_get_wide.argtypes = [vp, vp, vp]
        lib.po_set_builder.argtypes = [vp, i32, i32]
        lib.po_ploc_rounds.argtypes = [vp]; lib.po_bvh_height.argtypes = [vp]; lib.po_builder_used.argtypes = [vp]
        lib.po_set_estimator.argtypes = [vp, i32, C.c_float]
        lib.po_trace_batch.argtypes = [vp, vp, i32, vp, i32, vp, i32]
        lib.po_camera_dirs.argtypes = [C.POINTER(CCamera), vp]
        lib.po_render.restype = u64
        lib.po_render.argtypes = [vp, C.POINTER(CCamera), i32, i32, u64, i32, vp, i32, i32, i32, i32]
        lib.po_trace_pixel.argtypes = [vp, C.POINTER(CCamera), i32, i32, i32, i32, u64, i32, vp]
        lib.po_rng_stream.argtypes = [u64, u32, u32, u32, i32, vp]
        lib.po_resolve_display.argtypes = [vp, i32, f32, f32, f32, vp]
        lib.po_scatter.argtypes = [vp, vp, vp, i32, u64, u32, u32, u32, vp]

    def scene(self, sd):
        return OracleScene(self, sd)

    def camera_dirs(self, cam):
        c = cam.to_c()
        out = np.empty((cam.height, cam.width, 3), np.float32)
        self.lib.po_camera_dirs(C.byref(c), _ptr(out))
        return out

    def scatter(self, mat, in_dir, normal, front, seed, pixel, sample, bounce):
        """One MaterialHandle::Scatter/Emit record -> 10 floats {ok, atten.rgb, out_dir.xyz, emit.rgb}."""
        m = np.zeros(1, material_dt); m[0] = mat
        i_, n_ = (np.ascontiguousarray(x, np.float32) for x in (in_dir, normal))
        out = np.zeros(10, np.float32)
        self.lib.po_scatter(_ptr(m), _ptr(i_), _ptr(n_), int(front), seed, pixel, sample, bounce, _ptr(out))
        return out

    def rng_stream(self, seed, pixel, sample, bounce, n):
        out = np.empty(n, np.uint32)
        self.lib.po_rng_stream(seed, pixel, sample, bounce, n, _ptr(out))
        return out

    def resolve_display(self, accum, spp, exposure=1.0, gamma=2.2):
        accum = np.ascontiguousarray(accum, np.float32)
        npix = accum.size // 3
        out = np.empty(accum.shape[:-1] + (4,), np.uint8)
        self.lib.po_resolve_display(_ptr(accum), npix, spp, exposure, gamma, _ptr(out))
        return out


class OracleScene:
    def __init__(self, oracle, sd):
        self.o, self.lib, self.sd = oracle, oracle.lib, sd
        nt = sd.n_tris
        sky = np.asarray(sd.sky, np.float32)
        tn = sd.tri_normals
        self.h = self.lib.po_scene_create(_ptr(sd.spheres), len(sd.spheres), _ptr(sd.quads), len(sd.quads),
                                          _ptr(sd.tri_verts) if nt else None, _ptr(tn) if nt else None,
                                          _ptr(sd.tri_material) if nt else None, -1, nt, int(sd.tri_prim_base),
                                          _ptr(sd.materials), len(sd.materials), _ptr(sky))
        self.built = False

    def __del__(self):
        try:
            self.lib.po_scene_destroy(self.h)
        except Exception:
            pass

    def set_builder(self, builder=1, radius=4):
        """Hierarchy over the sorted leaves: 1 = PLOC (device default; falls back to Karras like lbvh.cu), 0 = Karras (plain LBVH)."""
        self.lib.po_set_builder(self.h, int(builder), int(radius))
        self.built = False

    def build_info(self):
        self.build_lbvh()
        return dict(builder=self.lib.po_builder_used(self.h), height=self.lib.po_bvh_height(self.h), rounds=self.lib.po_ploc_rounds(self.h))

    def build_lbvh(self):
        if not self.built:
            self.lib.po_build_lbvh(self.h)
            self.built = True

    def get_lbvh(self):
        self.build_lbvh()
        n = self.lib.po_num_prims(self.h)
        ni = max(n - 1, 0)
        d = dict(morton=np.zeros(n, np.uint32), sorted_prim=np.zeros(n, np.uint32),
                 node_children=np.zeros((ni, 2), np.int32), node_parent=np.zeros(ni, np.int32),
                 node_bounds=np.zeros((ni, 12), np.float32), scene_bounds=np.zeros(6, np.float32))
        self.lib.po_get_lbvh(self.h, _ptr(d["morton"]), _ptr(d["sorted_prim"]), _ptr(d["node_children"]),
                             _ptr(d["node_parent"]), _ptr(d["node_bounds"]), _ptr(d["scene_bounds"]))
        return d

    def set_estimator(self, rr_start=-1, clamp=0.0):
        """Estimator options of the device path (default off): Russian roulette from bounce rr_start, radiance clamp."""
        self.lib.po_set_estimator(self.h, int(rr_start), float(clamp))

    def get_wide(self):
        """4-wide quantised nodes in the device layout: ((n-1, 16) uint32 words, (n-1,) valid = even depth)."""
        self.build_lbvh()
        ni = max(self.lib.po_num_prims(self.h) - 1, 0)
        w, v = np.zeros((ni, 16), np.uint32), np.zeros(ni, np.uint8)
        self.lib.po_get_wide(self.h, _ptr(w), _ptr(v))
        return w, v.astype(bool)

    def trace_batch(self, rays, use_bvh=True, n_threads=8, counters=None):
        if use_bvh:
            self.build_lbvh()
        rays = np.ascontiguousarray(rays, ray_dt)
        out = np.zeros(len(rays), hit_dt)
        cnt = np.zeros(2, np.uint64)
        self.lib.po_trace_batch(self.h, _ptr(rays), len(rays), _ptr(out), int(use_bvh), _ptr(cnt), n_threads)
        if counters is not None:
            counters[:] = cnt
        return out

    def render(self, cam, spp, max_depth, seed=1, sample_begin=0, use_bvh=True, n_threads=8, accum=None,
               y_begin=0, y_end=0):
        """Returns (accum (H,W,3) float32 sum of samples, number of closest-hit queries)."""
        if use_bvh:
            self.build_lbvh()
        if accum is None:
            accum = np.zeros((cam.height, cam.width, 3), np.float32)
        c = cam.to_c()
        rays = self.lib.po_render(self.h, C.byref(c), spp, max_depth, seed, sample_begin, _ptr(accum),
                                  1 if use_bvh else 0, n_threads, y_begin, y_end)
        return accum, int(rays)

    def trace_pixel(self, cam, x, y, sample, max_depth, seed=1, use_bvh=True):
        if use_bvh:
            self.build_lbvh()
        out = np.zeros(3, np.float32)
        c = cam.to_c()
        self.lib.po_trace_pixel(self.h, C.byref(c), x, y, sample, max_depth, seed, 1 if use_bvh else 0, _ptr(out))
        return out


def load_oracle():
    return Oracle(C.CDLL(build_oracle()))


# ---------------------------------------------------------------- reference harness (oracle/_ref)
class Ref:
    """The reference's own compiled CPU code (oracle/ref_harness.cpp)."""

    def __init__(self, lib):
        self.lib = lib
        fp = C.POINTER(C.c_float)
        lib.ref_scene_preset.restype = vp; lib.ref_scene_preset.argtypes = [i32]
        lib.ref_scene_new.restype = vp
        lib.ref_scene_destroy.argtypes = [vp]
        lib.ref_num_prims.argtypes = [vp]
        lib.ref_scene_add_sphere.argtypes = [vp, f32, vp, vp, vp, vp]
        lib.ref_scene_add_quad.argtypes = [vp, f32, f32, vp, vp, vp, vp]
        lib.ref_scene_add_triangles.argtypes = [vp, vp, vp, i32, vp]
        lib.ref_mesh_load.restype = vp; lib.ref_mesh_load.argtypes = [C.c_char_p]
        lib.ref_mesh_destroy.argtypes = [vp]
        lib.ref_mesh_counts.argtypes = [vp, vp, vp, vp]
        lib.ref_mesh_data.argtypes = [vp, vp, vp, vp]
        lib.ref_flatten_counts.argtypes = [vp] * 6
        lib.ref_flatten.argtypes = [vp] * 7
        lib.ref_camera.argtypes = [vp, vp, i32, i32, C.POINTER(CCamera)]
        lib.ref_camera_dirs.argtypes = [vp, vp, i32, i32, vp]
        lib.ref_trace_batch.argtypes = [vp, vp, i32, vp, vp, i32]
        lib.ref_render.restype = u64
        lib.ref_render.argtypes = [vp, vp, vp, i32, i32, i32, i32, u64, i32, vp, i32, i32, i32]
        lib.ref_trace_pixel.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, u64, vp]
        lib.ref_render_shipped.restype = C.c_double
        lib.ref_render_shipped.argtypes = [i32, vp, vp, i32, i32, i32, vp]
        lib.ref_film_display.argtypes = [vp, i32, i32, i32, f32, f32, vp]
        lib.ref_scatter.argtypes = [vp, vp, vp, vp, i32, u64, u32, u32, u32, vp, vp, vp]

    def preset(self, preset):
        return RefScene(self, self.lib.ref_scene_preset(int(preset)))

    def new_scene(self):
        return RefScene(self, self.lib.ref_scene_new())

    def load_mesh(self, path):
        m = self.lib.ref_mesh_load(path.encode())
        nv, nn, ni = i32(), i32(), i32()
        self.lib.ref_mesh_counts(m, C.byref(nv), C.byref(nn), C.byref(ni))
        v, n, idx = np.zeros((nv.value, 3), np.float32), np.zeros((nn.value, 3), np.float32), np.zeros(ni.value, np.uint32)
        self.lib.ref_mesh_data(m, _ptr(v), _ptr(n), _ptr(idx))
        self.lib.ref_mesh_destroy(m)
        return v, (n if nn.value else None), idx.reshape(-1, 3)

    def camera(self, cam):
        out = CCamera()
        self.lib.ref_camera(_ptr(cam.pos), _ptr(cam.front_in), cam.width, cam.height, C.byref(out))
        return out

    def camera_dirs(self, cam):
        out = np.empty((cam.height, cam.width, 3), np.float32)
        self.lib.ref_camera_dirs(_ptr(cam.pos), _ptr(cam.front_in), cam.width, cam.height, _ptr(out))
        return out

    def render_shipped(self, preset, cam, spp, want_accum=True):
        acc = np.zeros((cam.height, cam.width, 3), np.float32) if want_accum else None
        sec = self.lib.ref_render_shipped(int(preset), _ptr(cam.pos), _ptr(cam.front_in), cam.width, cam.height, spp, _ptr(acc))
        return sec, acc

    def film_display(self, accum_sum, spp, exposure=1.0, gamma=2.2):
        a = np.ascontiguousarray(accum_sum, np.float32)
        H, W = a.shape[:2]
        out = np.empty((H, W, 4), np.uint8)
        self.lib.ref_film_display(_ptr(a), W, H, spp, exposure, gamma, _ptr(out))
        return out

    def scatter(self, mat, in_dir, pos, normal, front, seed, pixel, sample, bounce):
        m = np.zeros(1, material_dt); m[0] = mat
        a, o, e = np.zeros(3, np.float32), np.zeros(3, np.float32), np.zeros(3, np.float32)
        i_, p_, n_ = (np.ascontiguousarray(x, np.float32) for x in (in_dir, pos, normal))
        ok = self.lib.ref_scatter(_ptr(m), _ptr(i_), _ptr(p_), _ptr(n_), int(front), seed, pixel, sample, bounce,
                                  _ptr(a), _ptr(o), _ptr(e))
        return bool(ok), a, o, e


class RefScene:
    def __init__(self, ref, h):
        self.ref, self.lib, self.h = ref, ref.lib, h

    def __del__(self):
        try:
            self.lib.ref_scene_destroy(self.h)
        except Exception:
            pass

    @staticmethod
    def _m(mat):
        m = np.zeros(1, material_dt); m[0] = mat
        return m

    def add_sphere(self, radius, mat, scale=(1, 1, 1), euler=(0, 0, 0), trans=(0, 0, 0)):
        s, e, t = (np.asarray(x, np.float32) for x in (scale, euler, trans))
        m = self._m(mat)
        self.lib.ref_scene_add_sphere(self.h, radius, _ptr(s), _ptr(e), _ptr(t), _ptr(m))

    def add_quad(self, w, h, mat, scale=(1, 1, 1), euler=(0, 0, 0), trans=(0, 0, 0)):
        s, e, t = (np.asarray(x, np.float32) for x in (scale, euler, trans))
        m = self._m(mat)
        self.lib.ref_scene_add_quad(self.h, w, h, _ptr(s), _ptr(e), _ptr(t), _ptr(m))

    def add_triangles(self, verts, normals, mat):
        v = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = np.ascontiguousarray(normals, np.float32).reshape(-1, 9)
        m = self._m(mat)
        self.lib.ref_scene_add_triangles(self.h, _ptr(v), _ptr(n), len(v), _ptr(m))

    def num_prims(self):
        return self.lib.ref_num_prims(self.h)

    def flatten(self, name="ref"):
        """host/flatten.h applied to the reference's own Scene object -> SceneData."""
        ns, nq, nt, nm, base = i32(), i32(), i32(), i32(), i32()
        rc = self.lib.ref_flatten_counts(self.h, C.byref(ns), C.byref(nq), C.byref(nt), C.byref(nm), C.byref(base))
        if rc != 0:
            raise RuntimeError("flatten failed")
        sd = SceneData(name=name)
        sd.spheres, sd.quads = np.zeros(ns.value, sphere_dt), np.zeros(nq.value, quad_dt)
        sd.materials = np.zeros(nm.value, material_dt)
        tv = tn = tm = None
        if nt.value:
            tv, tn, tm = np.zeros((nt.value, 9), np.float32), np.zeros((nt.value, 9), np.float32), np.zeros(nt.value, np.int32)
        self.lib.ref_flatten(self.h, _ptr(sd.spheres), _ptr(sd.quads), _ptr(tv), _ptr(tn), _ptr(tm), _ptr(sd.materials))
        sd.tri_verts, sd.tri_normals, sd.tri_material, sd.tri_prim_base = tv, tn, tm, base.value
        return sd

    def trace_batch(self, rays, n_threads=8, want_second=True):
        rays = np.ascontiguousarray(rays, ray_dt)
        out = np.zeros(len(rays), hit_dt)
        second = np.zeros(len(rays), np.float32) if want_second else None
        self.lib.ref_trace_batch(self.h, _ptr(rays), len(rays), _ptr(out), _ptr(second), n_threads)
        return out, second

    def render(self, cam, spp, depth, seed=1, sample_begin=0, n_threads=8, accum=None, y_begin=0, y_end=0):
        if accum is None:
            accum = np.zeros((cam.height, cam.width, 3), np.float32)
        rays = self.lib.ref_render(self.h, _ptr(cam.pos), _ptr(cam.front_in), cam.width, cam.height, spp, depth, seed,
                                   sample_begin, _ptr(accum), n_threads, y_begin, y_end)
        return accum, int(rays)

    def trace_pixel(self, cam, x, y, sample, depth, seed=1):
        out = np.zeros(3, np.float32)
        self.lib.ref_trace_pixel(self.h, _ptr(cam.pos), _ptr(cam.front_in), cam.width, cam.height, x, y, sample, depth, seed, _ptr(out))
        return out


def build_ref():
    """Build oracle/_ref/libprt_ref.so from the sources under /root/reference (only where they exist)."""
    if os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-C", ORACLE_DIR, "ref", "shim", "refgpu"], check=True, capture_output=True)
    return REF_SO if os.path.exists(REF_SO) else None


def load_ref():
    if not os.path.exists(REF_SO):
        return None
    return Ref(C.CDLL(REF_SO))
