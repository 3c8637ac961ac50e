"""ctypes binding of include/prt.h (libprt_b200.so). There is NO CPU fallback: if the CUDA
library is missing or no sm_100 device is present, calls raise PrtError."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PRT_MAX_DEPTH = 32

sphere_dt = np.dtype([("cx", "f4"), ("cy", "f4"), ("cz", "f4"), ("radius", "f4"), ("tmin", "f4"),
                      ("material", "i4"), ("prim_id", "i4"), ("_pad", "i4")])
quad_dt = np.dtype([("cx", "f4"), ("cy", "f4"), ("cz", "f4"), ("half_w", "f4"),
                    ("ux", "f4"), ("uy", "f4"), ("uz", "f4"), ("half_h", "f4"),
                    ("vx", "f4"), ("vy", "f4"), ("vz", "f4"), ("tmin", "f4"),
                    ("nx", "f4"), ("ny", "f4"), ("nz", "f4"),
                    ("material", "i4"), ("prim_id", "i4"), ("_pad", "i4", (3,))])
material_dt = np.dtype([("type", "i4"), ("r", "f4"), ("g", "f4"), ("b", "f4"), ("param", "f4"), ("_pad", "i4", (3,))])
ray_dt = np.dtype([("ox", "f4"), ("oy", "f4"), ("oz", "f4"), ("dx", "f4"), ("dy", "f4"), ("dz", "f4")])
hit_dt = np.dtype([("t", "f4"), ("prim_id", "i4"), ("b1", "f4"), ("b2", "f4"),
                   ("px", "f4"), ("py", "f4"), ("pz", "f4"), ("nx", "f4"), ("ny", "f4"), ("nz", "f4"),
                   ("front", "i4"), ("material", "i4")])
assert sphere_dt.itemsize == 32 and quad_dt.itemsize == 80 and material_dt.itemsize == 32
assert ray_dt.itemsize == 24 and hit_dt.itemsize == 48

MAT_NONE, MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_EMISSIVE = 0, 1, 2, 3, 4


class CCamera(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("front", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("tan_half_fov", C.c_float), ("width", C.c_int32), ("height", C.c_int32)]


class CStats(C.Structure):
    _fields_ = [("rays_total", C.c_uint64), ("rays_per_bounce", C.c_uint64 * PRT_MAX_DEPTH), ("paths", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("prims_tested", C.c_uint64),
                ("render_ms", C.c_float), ("extend_ms", C.c_float), ("shade_ms", C.c_float), ("build_ms", C.c_float),
                ("primary_ms", C.c_float), ("reduce_ms", C.c_float),
                ("kernel_launches", C.c_int32), ("n_prims", C.c_int32), ("n_nodes", C.c_int32), ("primary_launches", C.c_int32),
                ("bvh_builder", C.c_int32), ("bvh_height", C.c_int32), ("ploc_rounds", C.c_int32), ("n_devices", C.c_int32)]


class PrtError(RuntimeError):
    pass


EXPORTS = [
    "prt_create", "prt_create_multi", "prt_destroy", "prt_last_error", "prt_version", "prt_set_scene", "prt_set_scene_build", "prt_build_bvh",
    "prt_set_camera", "prt_render", "prt_render_fresh", "prt_render_region", "prt_wait_event", "prt_clear_accum", "prt_scale_accum", "prt_accum_device_ptr",
    "prt_set_accum_buffer", "prt_read_accum", "prt_accum_spp", "prt_set_accum_spp", "prt_resolve_display", "prt_trace_batch", "prt_trace_batch_static",
    "prt_debug_lbvh", "prt_debug_wide", "prt_debug_wide_compact", "prt_debug_scatter", "prt_debug_ploc_trace", "prt_debug_check_flags", "prt_host_alloc", "prt_host_free", "prt_num_prims", "prt_enable_counters", "prt_get_stats", "prt_set_option",
]

_lib = None


def lib_path():
    """libprt_b200.so, or - PRT_CHECKED=1 in the environment - the bounds-checked build of the same sources."""
    if os.environ.get("PRT_LIB"):                # an experimental build of the same sources (profiling A/B runs only)
        return os.environ["PRT_LIB"]
    return os.path.join(HERE, "libprt_b200_checked.so" if os.environ.get("PRT_CHECKED") == "1" else "libprt_b200.so")


def load_library():
    """Load libprt_b200.so (built in-tree by build.py). Raises PrtError if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    from . import build as _build
    checked = os.environ.get("PRT_CHECKED") == "1"
    if not os.environ.get("PRT_LIB") and _build.needs_build(checked=checked):        # missing, or built from other sources than the tree holds: never run a stale .so
        try:
            _build.build(checked=checked)
        except Exception as e:  # noqa: BLE001
            raise PrtError(f"{p} is missing or older than its sources and could not be rebuilt ({e}); run "
                           "`python -m parallelraytracing_b200.build` (needs nvcc). There is no CPU fallback.")
    lib = C.CDLL(p)
    vp, i32, u64, f32 = C.c_void_p, C.c_int, C.c_uint64, C.c_float
    lib.prt_create.argtypes = [C.POINTER(vp), i32]
    lib.prt_create_multi.argtypes = [C.POINTER(vp), C.POINTER(C.c_int), i32]
    lib.prt_destroy.argtypes = [vp]; lib.prt_destroy.restype = None
    lib.prt_last_error.argtypes = [vp]; lib.prt_last_error.restype = C.c_char_p
    lib.prt_version.restype = C.c_char_p
    lib.prt_set_scene.argtypes = [vp, vp, i32, vp, i32, vp, vp, vp, i32, i32, i32, vp, i32, vp]
    lib.prt_set_scene_build.argtypes = lib.prt_set_scene.argtypes
    lib.prt_build_bvh.argtypes = [vp]
    lib.prt_set_camera.argtypes = [vp, C.POINTER(CCamera)]
    lib.prt_render.argtypes = [vp, i32, i32, u64, i32]
    lib.prt_render_fresh.argtypes = [vp, i32, i32, u64, i32]
    lib.prt_render_region.argtypes = [vp, i32, i32, u64, i32, i32, i32]
    lib.prt_wait_event.argtypes = [vp, vp]
    lib.prt_clear_accum.argtypes = [vp]
    lib.prt_scale_accum.argtypes = [vp, f32]
    lib.prt_accum_device_ptr.argtypes = [vp]; lib.prt_accum_device_ptr.restype = vp
    lib.prt_set_accum_buffer.argtypes = [vp, vp]
    lib.prt_read_accum.argtypes = [vp, vp]
    lib.prt_accum_spp.argtypes = [vp]
    lib.prt_set_accum_spp.argtypes = [vp, i32]
    lib.prt_resolve_display.argtypes = [vp, f32, f32, vp]
    lib.prt_trace_batch.argtypes = [vp, vp, i32, vp, i32]
    lib.prt_trace_batch_static.argtypes = [vp, vp, i32, vp, i32, vp]
    lib.prt_debug_lbvh.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.prt_debug_wide.argtypes = [vp, vp]
    lib.prt_debug_wide_compact.argtypes = [vp, vp, i32, vp]
    lib.prt_debug_scatter.argtypes = [vp, vp, vp, vp, vp, vp, u64, i32, vp]
    lib.prt_debug_ploc_trace.argtypes = [vp, vp, vp, i32]
    lib.prt_debug_check_flags.argtypes = [vp]
    lib.prt_host_alloc.argtypes = [C.c_size_t]; lib.prt_host_alloc.restype = vp
    lib.prt_host_free.argtypes = [vp]; lib.prt_host_free.restype = None
    lib.prt_num_prims.argtypes = [vp]
    lib.prt_enable_counters.argtypes = [vp, i32]
    lib.prt_get_stats.argtypes = [vp, C.POINTER(CStats)]
    lib.prt_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    _lib = lib
    return lib


class _PinnedBlock:
    """Owner of one prt_host_alloc block (freed when the last numpy view of it goes away)."""

    def __init__(self, lib, nbytes):
        self.lib, self.p = lib, lib.prt_host_alloc(nbytes)
        if not self.p:
            raise PrtError("prt_host_alloc(%d) failed" % nbytes)

    def __del__(self):
        try:
            self.lib.prt_host_free(self.p)
        except Exception:
            pass


def pinned_copy(a):
    """A copy of numpy array `a` in page-locked host memory (prt_host_alloc): prt_set_scene then uploads it by plain DMA."""
    a = np.ascontiguousarray(a)
    blk = _PinnedBlock(load_library(), a.nbytes)
    buf = (C.c_char * max(1, a.nbytes)).from_address(blk.p)
    buf._owner = blk                            # numpy keeps `buf` alive through the buffer protocol, `buf` keeps the block
    out = np.frombuffer(buf, dtype=a.dtype, count=a.size).reshape(a.shape)
    out[...] = a
    return out


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(C.c_void_p)


class Context:
    """Thin RAII wrapper of prt_ctx; every method maps 1:1 to a C-ABI call."""

    def __init__(self, device=0):
        """device: a CUDA ordinal, or a list of ordinals for ONE context over several GPUs of this process
        (prt_create_multi: the first one is the root that holds the frame)."""
        self.lib = load_library()
        h = C.c_void_p()
        if isinstance(device, (list, tuple)):
            ids = (C.c_int * len(device))(*[int(d) for d in device])
            rc = self.lib.prt_create_multi(C.byref(h), ids, len(device))
        else:
            rc = self.lib.prt_create(C.byref(h), device)
        if rc != 0:
            raise PrtError(f"prt_create failed ({rc}): {self.lib.prt_last_error(None).decode()}")
        self.h = h
        self.width = self.height = 0

    def _check(self, rc, what):
        if rc != 0:
            raise PrtError(f"{what} failed ({rc}): {self.lib.prt_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.prt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_scene(self, scene, build=False):
        """prt_set_scene; build=True: prt_set_scene_build (scene + BVH in one call, the build hidden behind the normals' upload)."""
        s = scene
        nt = 0 if s.tri_verts is None else int(s.tri_verts.shape[0])
        sky = np.asarray(s.sky, dtype=np.float32)
        self._keep = (s, sky)
        fn, what = (self.lib.prt_set_scene_build, "prt_set_scene_build") if build else (self.lib.prt_set_scene, "prt_set_scene")
        self._check(fn(
            self.h, _ptr(s.spheres), len(s.spheres), _ptr(s.quads), len(s.quads),
            _ptr(s.tri_verts) if nt else None, _ptr(s.tri_normals) if nt else None,
            _ptr(s.tri_material) if nt else None, -1, nt, int(s.tri_prim_base),
            _ptr(s.materials), len(s.materials), _ptr(sky)), what)

    def set_scene_build(self, scene):
        self.set_scene(scene, build=True)

    def build_bvh(self):
        self._check(self.lib.prt_build_bvh(self.h), "prt_build_bvh")

    def set_camera(self, cam):
        c = cam.to_c()
        self._check(self.lib.prt_set_camera(self.h, C.byref(c)), "prt_set_camera")
        self.width, self.height = cam.width, cam.height

    def render(self, spp, max_depth, seed=1, sample_begin=0):
        self._check(self.lib.prt_render(self.h, spp, max_depth, seed, sample_begin), "prt_render")

    def render_fresh(self, spp, max_depth, seed=1, sample_begin=0):
        """clear + render in one call (prt_render_fresh); 1 spp in one wave goes straight into the sample buffer."""
        self._check(self.lib.prt_render_fresh(self.h, spp, max_depth, seed, sample_begin), "prt_render_fresh")

    def render_region(self, spp, max_depth, seed=1, sample_begin=0, pixel_begin=0, pixel_count=0):
        self._check(self.lib.prt_render_region(self.h, spp, max_depth, seed, sample_begin, pixel_begin, pixel_count), "prt_render_region")

    def clear_accum(self):
        self._check(self.lib.prt_clear_accum(self.h), "prt_clear_accum")

    def scale_accum(self, s):
        self._check(self.lib.prt_scale_accum(self.h, s), "prt_scale_accum")

    def accum_device_ptr(self):
        return self.lib.prt_accum_device_ptr(self.h)

    def set_accum_buffer(self, device_ptr):
        self._check(self.lib.prt_set_accum_buffer(self.h, device_ptr), "prt_set_accum_buffer")

    def read_accum(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 3), dtype=np.float32)
        self._check(self.lib.prt_read_accum(self.h, _ptr(out)), "prt_read_accum")
        return out

    def accum_spp(self):
        return self.lib.prt_accum_spp(self.h)

    def set_accum_spp(self, spp):
        self._check(self.lib.prt_set_accum_spp(self.h, int(spp)), "prt_set_accum_spp")

    def resolve_display(self, exposure=1.0, gamma=2.2):
        out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        self._check(self.lib.prt_resolve_display(self.h, exposure, gamma, _ptr(out)), "prt_resolve_display")
        return out

    def trace_batch(self, rays, use_bvh=True):
        rays = np.ascontiguousarray(rays, dtype=ray_dt)
        out = np.zeros(len(rays), dtype=hit_dt)
        self._check(self.lib.prt_trace_batch(self.h, _ptr(rays), len(rays), _ptr(out), int(use_bvh)), "prt_trace_batch")
        return out

    def trace_batch_static(self, rays_dev, n, out_dev, use_bvh=2, stream=None):
        """prt_trace_batch_static: device pointers (ints) of n prt_ray records in and n prt_hit records out; allocates nothing."""
        self._check(self.lib.prt_trace_batch_static(self.h, C.c_void_p(rays_dev), int(n), C.c_void_p(out_dev), int(use_bvh),
                                                    C.c_void_p(stream)), "prt_trace_batch_static")

    def num_prims(self):
        return self.lib.prt_num_prims(self.h)

    def debug_lbvh(self):
        n = self.num_prims()
        ni = max(n - 1, 0)
        d = dict(morton=np.zeros(n, np.uint32), sorted_prim=np.zeros(n, np.uint32),
                 node_children=np.zeros((ni, 2), np.int32), node_parent=np.zeros(ni, np.int32),
                 node_bounds=np.zeros((ni, 12), np.float32), scene_bounds=np.zeros(6, np.float32))
        self._check(self.lib.prt_debug_lbvh(self.h, _ptr(d["morton"]), _ptr(d["sorted_prim"]), _ptr(d["node_children"]),
                                            _ptr(d["node_parent"]), _ptr(d["node_bounds"]), _ptr(d["scene_bounds"])),
                    "prt_debug_lbvh")
        return d

    def debug_wide_compact(self):
        """(n_live, 16) uint32: the dense breadth-first array of reachable WideNode records the walk reads."""
        n = C.c_int(0)
        self._check(self.lib.prt_debug_wide_compact(self.h, None, 0, C.byref(n)), "prt_debug_wide_compact")
        w = np.zeros((n.value, 16), np.uint32)
        self._check(self.lib.prt_debug_wide_compact(self.h, _ptr(w), n.value, C.byref(n)), "prt_debug_wide_compact")
        return w

    def debug_wide(self):
        """(n-1, 16) uint32: raw WideNode records; rows of odd-depth nodes are undefined."""
        w = np.zeros((max(self.num_prims() - 1, 0), 16), np.uint32)
        self._check(self.lib.prt_debug_wide(self.h, _ptr(w)), "prt_debug_wide")
        return w

    def debug_scatter(self, mats, in_dirs, normals, fronts, keys, seed=1):
        """n MaterialHandle::Scatter/Emit records on the device -> (n, 10) float32 {ok, atten.rgb, out_dir.xyz, emit.rgb}."""
        mats = np.ascontiguousarray(mats, material_dt)
        i_, n_ = np.ascontiguousarray(in_dirs, np.float32), np.ascontiguousarray(normals, np.float32)
        f_, k_ = np.ascontiguousarray(fronts, np.int32), np.ascontiguousarray(keys, np.uint32)
        out = np.zeros((len(mats), 10), np.float32)
        self._check(self.lib.prt_debug_scatter(self.h, _ptr(mats), _ptr(i_), _ptr(n_), _ptr(f_), _ptr(k_), seed, len(mats), _ptr(out)),
                    "prt_debug_scatter")
        return out

    def check_flags(self):
        """Checked library only (PRT_CHECKED=1): PRT_CHK_* bits raised since the last call | 1<<30; 0 from the regular library."""
        return self.lib.prt_debug_check_flags(self.h)

    def ploc_trace(self):
        """(clusters at the start of each PLOC round, microseconds each round took) of the last build."""
        cnt, ns = np.zeros(200, np.int32), np.zeros(200, np.uint32)
        n = self.lib.prt_debug_ploc_trace(self.h, _ptr(cnt), _ptr(ns), 200)
        if n <= 0:
            return cnt[:0], np.zeros(0)
        return cnt[:n], np.diff(ns[:n]).astype(np.uint32).astype(np.float64) * 1e-3

    def enable_counters(self, on=True):
        self._check(self.lib.prt_enable_counters(self.h, 1 if on else 0), "prt_enable_counters")

    def stats(self):
        st = CStats()
        self._check(self.lib.prt_get_stats(self.h, C.byref(st)), "prt_get_stats")
        return st

    def set_option(self, key, value):
        self._check(self.lib.prt_set_option(self.h, key.encode(), int(value)), "prt_set_option")
