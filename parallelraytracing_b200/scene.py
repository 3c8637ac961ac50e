"""Host-side scene, camera and mesh types of the hot path, mirroring the reference's
src/core types (Rickyeeeeee/ParallelRayTracing) as flat POD arrays for the C ABI (include/prt.h).

  ScenePreset / make_scene   Scene presets                  src/core/scene.h:6-15, scene.cpp:62-350
  Transform.Set math         T * eulerAngleXYZ * S          src/core/geometry.h:92-99
  flatten rules              same as host/flatten.h (the C++ shim a maintainer links)
  Camera                     Camera ctor + GetCameraRay     src/core/camera.h:10-16, 104-132
  load_ply                   Mesh (tinyply wrapper)         src/core/mesh.cpp:23-184

All arithmetic that decides bits downstream (camera basis, transforms) is done in float32 in the
same operation order as glm so that the arrays equal what host/flatten.h produces from the
reference's own objects (tests/test_oracle_pinned.py::test_presets_equal_reference_flatten and test_camera_equals_reference check
this against the vectors generated through oracle/_ref).
"""
import enum
import math
import os
import struct
from dataclasses import dataclass, field

import numpy as np

from .capi import (CCamera, MAT_DIELECTRIC, MAT_EMISSIVE, MAT_LAMBERTIAN, MAT_METAL, material_dt, quad_dt,
                   sphere_dt)

f32 = np.float32
K_SHAPE_RAY_TMIN = f32(0.001)          # kShapeRayTMin, shape.h:128
SKY = (0.4, 0.3, 0.6)                  # m_SkyLight, cpu/renderer.h:31


class ScenePreset(enum.IntEnum):       # scene.h:6-15
    DEFAULT = 0
    LIGHT_TEST = 1
    MATERIAL_TEST = 2
    CORNELL = 3
    RANDOM_BALLS_SMALL = 4
    RANDOM_BALLS_MEDIUM = 5
    RANDOM_BALLS_LARGE = 6


@dataclass
class SceneData:
    """Flat world-space scene = the arguments of prt_set_scene."""
    spheres: np.ndarray = field(default_factory=lambda: np.zeros(0, sphere_dt))
    quads: np.ndarray = field(default_factory=lambda: np.zeros(0, quad_dt))
    tri_verts: np.ndarray = None       # (nt, 9) f32 or None
    tri_normals: np.ndarray = None     # (nt, 9) f32
    tri_material: np.ndarray = None    # (nt,) i32
    tri_prim_base: int = 0
    materials: np.ndarray = field(default_factory=lambda: np.zeros(0, material_dt))
    sky: tuple = SKY
    name: str = "scene"

    @property
    def n_tris(self):
        return 0 if self.tri_verts is None else int(self.tri_verts.shape[0])

    @property
    def n_prims(self):
        return len(self.spheres) + len(self.quads) + self.n_tris

    def pin(self):
        """Move the triangle arrays into page-locked host memory (prt_host_alloc), in place: B200Renderer.Init / prt_set_scene
        then upload them by plain asynchronous DMA instead of through the staging ring. Needs the CUDA library. Returns self."""
        if self.n_tris and not getattr(self, "_pinned", False):
            from .capi import pinned_copy
            self.tri_verts, self.tri_normals = pinned_copy(self.tri_verts), pinned_copy(self.tri_normals)
            self.tri_material = pinned_copy(self.tri_material)
            self._pinned = True
        return self


# ---------------------------------------------------------------- float32 glm helpers
def _v3(x, y, z):
    return np.array([x, y, z], dtype=f32)


def _dot(a, b):                         # glm::dot(vec3): (x*x' + y*y') + z*z'
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))


def _normalize(v):                      # glm::normalize: v * inversesqrt(dot(v, v))
    s = f32(f32(1.0) / f32(np.sqrt(_dot(v, v))))
    return (v * s).astype(f32)


def _cross(x, y):                       # glm::cross
    return _v3(f32(x[1] * y[2]) - f32(y[1] * x[2]), f32(x[2] * y[0]) - f32(y[2] * x[0]),
               f32(x[0] * y[1]) - f32(y[0] * x[1]))


def _length(v):
    return f32(np.sqrt(_dot(v, v)))


import ctypes as _C

_libm = _C.CDLL("libm.so.6")
_libm.cosf.restype = _C.c_float; _libm.cosf.argtypes = [_C.c_float]
_libm.sinf.restype = _C.c_float; _libm.sinf.argtypes = [_C.c_float]
_libm.tanf.restype = _C.c_float; _libm.tanf.argtypes = [_C.c_float]


def _transform(scale, euler_deg, trans):
    """Transform::Set (geometry.h:92-99): M = translate * eulerAngleXYZ(radians(euler)) * scale.
    Returns the 3 basis columns (already scaled) and the translation, float32. With T a pure
    translation and S diagonal the 4x4 products reduce exactly to column scaling."""
    rad = [f32(f32(a) * f32(0.01745329251994329576923690768489)) for a in euler_deg]
    c1, c2, c3 = (f32(_libm.cosf(float(-r))) for r in rad)
    s1, s2, s3 = (f32(_libm.sinf(float(-r))) for r in rad)
    r0 = _v3(f32(c2 * c3), f32(f32(-c1 * s3) + f32(f32(s1 * s2) * c3)), f32(f32(s1 * s3) + f32(f32(c1 * s2) * c3)))
    r1 = _v3(f32(c2 * s3), f32(f32(c1 * c3) + f32(f32(s1 * s2) * s3)), f32(f32(-s1 * c3) + f32(f32(c1 * s2) * s3)))
    r2 = _v3(f32(-s2), f32(s1 * c2), f32(c1 * c2))
    sx, sy, sz = (f32(s) for s in scale)
    return (r0 * sx).astype(f32), (r1 * sy).astype(f32), (r2 * sz).astype(f32), _v3(*trans)


def _inverse_row1(c0, c1, c2, t):
    """(Mi[0][1], Mi[1][1], Mi[2][1]) of Mi = glm::inverse(M), M = [c0 c1 c2 t; 0 0 0 1] (glm is column-major: m[col][row]),
    with the operations and their order of glm/detail/func_matrix.inl compute_inverse<4,4> in float32 - what
    TransformNormal(Mi, (0,1,0)) reads (geometry.h:139-148, primitive.cpp:39)."""
    m = [[f32(c0[0]), f32(c0[1]), f32(c0[2]), f32(0)], [f32(c1[0]), f32(c1[1]), f32(c1[2]), f32(0)],
         [f32(c2[0]), f32(c2[1]), f32(c2[2]), f32(0)], [f32(t[0]), f32(t[1]), f32(t[2]), f32(1)]]

    def coef(a, b, c, d):                       # m[a0][a1]*m[b0][b1] - m[c0][c1]*m[d0][d1]
        return f32(f32(m[a[0]][a[1]] * m[b[0]][b[1]]) - f32(m[c[0]][c[1]] * m[d[0]][d[1]]))
    C00 = coef((2, 2), (3, 3), (3, 2), (2, 3)); C04 = coef((2, 1), (3, 3), (3, 1), (2, 3)); C08 = coef((2, 1), (3, 2), (3, 1), (2, 2))
    C12 = coef((2, 0), (3, 3), (3, 0), (2, 3)); C16 = coef((2, 0), (3, 2), (3, 0), (2, 2)); C20 = coef((2, 0), (3, 1), (3, 0), (2, 1))

    def comb(a, fa, b, fb, c, fc):              # a*fa - b*fb + c*fc, left to right
        return f32(f32(f32(a * fa) - f32(b * fb)) + f32(c * fc))
    inv0x = comb(m[1][1], C00, m[1][2], C04, m[1][3], C08); inv0y = comb(m[0][1], C00, m[0][2], C04, m[0][3], C08)
    inv1x = comb(m[1][0], C00, m[1][2], C12, m[1][3], C16); inv1y = comb(m[0][0], C00, m[0][2], C12, m[0][3], C16)
    inv2x = comb(m[1][0], C04, m[1][1], C12, m[1][3], C20); inv2y = comb(m[0][0], C04, m[0][1], C12, m[0][3], C20)
    inv3x = comb(m[1][0], C08, m[1][1], C16, m[1][2], C20)
    row0 = [inv0x, f32(-inv1x), inv2x, f32(-inv3x)]                                   # Inverse[k][0]: SignA / SignB
    d = [f32(m[0][k] * row0[k]) for k in range(4)]
    det = f32(f32(d[0] + d[1]) + f32(d[2] + d[3]))
    ood = f32(f32(1) / det)
    return _v3(f32(f32(-inv0y) * ood), f32(inv1y * ood), f32(f32(-inv2y) * ood))


class SceneBuilder:
    """Appends primitives in list order (list index = prim_id, the tie-break of
    PrimitiveList::Intersect, primitive.cpp:44) and flattens them like host/flatten.h."""

    def __init__(self, name="scene", sky=SKY):
        self.name, self.sky = name, sky
        self._spheres, self._quads, self._mats = [], [], []
        self._tv, self._tn, self._tm = [], [], []
        self._n = 0
        self._tri_base = None

    # materials: one table entry per Add* call, like MaterialPool (material.h:163-200)
    def _mat(self, type_, rgb=(0, 0, 0), param=0.0):
        self._mats.append((type_, f32(rgb[0]), f32(rgb[1]), f32(rgb[2]), f32(param), (0, 0, 0)))
        return len(self._mats) - 1

    def lambertian(self, rgb): return self._mat(MAT_LAMBERTIAN, rgb)
    def metal(self, rgb, rough): return self._mat(MAT_METAL, rgb, rough)
    def dielectric(self, ior): return self._mat(MAT_DIELECTRIC, (1, 1, 1), ior)
    def emissive(self, rgb): return self._mat(MAT_EMISSIVE, rgb)

    def add_sphere(self, radius, mat, scale=(1, 1, 1), euler=(0, 0, 0), trans=(0, 0, 0)):
        c0, _, _, t = _transform(scale, euler, trans)
        s0 = _length(c0)
        self._spheres.append((t[0], t[1], t[2], f32(f32(radius) * s0), f32(K_SHAPE_RAY_TMIN * s0), mat, self._n, 0))
        self._n += 1

    def add_quad(self, width, height, mat, scale=(1, 1, 1), euler=(0, 0, 0), trans=(0, 0, 0)):
        c0, c1, c2, t = _transform(scale, euler, trans)
        s0, s1, s2 = _length(c0), _length(c1), _length(c2)
        u, v, n = (c0 / s0).astype(f32), (c2 / s2).astype(f32), _normalize(_inverse_row1(c0, c1, c2, t))   # the normal goes through glm::inverse, like flatten.h
        self._quads.append((t[0], t[1], t[2], f32(f32(f32(width) / f32(2)) * s0), u[0], u[1], u[2],
                            f32(f32(f32(height) / f32(2)) * s2), v[0], v[1], v[2], f32(K_SHAPE_RAY_TMIN * s1),
                            n[0], n[1], n[2], mat, self._n, (0, 0, 0)))
        self._n += 1

    def add_triangles(self, verts, normals, mat):
        """verts, normals: (nt, 3, 3) world space; mat: index or (nt,) array. Triangles must be
        contiguous in the list (one prim_id base)."""
        verts = np.ascontiguousarray(verts, dtype=f32).reshape(-1, 9)
        nt = verts.shape[0]
        if normals is None:
            v = verts.reshape(-1, 3, 3)
            fn = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]).astype(f32)
            normals = np.repeat(fn[:, None, :], 3, axis=1)
        normals = np.ascontiguousarray(normals, dtype=f32).reshape(-1, 9)
        if self._tri_base is None:
            self._tri_base = self._n
        elif self._tri_base + sum(len(x) for x in self._tv) != self._n:
            raise ValueError("triangles must be contiguous in the primitive list")
        self._tv.append(verts); self._tn.append(normals)
        self._tm.append(np.full(nt, mat, np.int32) if np.isscalar(mat) else np.asarray(mat, np.int32))
        self._n += nt

    def build(self):
        sd = SceneData(name=self.name, sky=self.sky)
        sd.spheres = np.array(self._spheres, dtype=sphere_dt) if self._spheres else np.zeros(0, sphere_dt)
        sd.quads = np.array(self._quads, dtype=quad_dt) if self._quads else np.zeros(0, quad_dt)
        sd.materials = np.array(self._mats, dtype=material_dt) if self._mats else np.zeros(0, material_dt)
        if self._tv:
            sd.tri_verts = np.concatenate(self._tv); sd.tri_normals = np.concatenate(self._tn)
            sd.tri_material = np.concatenate(self._tm)
            sd.tri_prim_base = self._tri_base
        else:
            sd.tri_prim_base = self._n
        return sd


# ---------------------------------------------------------------- presets (scene.cpp:62-350)
class _StdMt19937:
    """std::mt19937(seed) + libstdc++ std::uniform_real_distribution<float> (one 32-bit draw per
    value: generate_canonical<float,24>), so RANDOM_BALLS_* equal the reference built with g++."""

    def __init__(self, seed):
        self._bg = np.random.RandomState(seed)._bit_generator   # init_genrand seeding = std::mt19937

    def uniform(self, a, b):
        r = f32(f32(self._bg.random_raw()) / f32(4294967296.0))
        if r >= f32(1.0):
            r = np.nextafter(f32(1.0), f32(0.0))
        return f32(f32(f32(b) - f32(a)) * r + f32(a))


def _random_balls(n_balls, name):               # Scene::InitRandomBalls, scene.cpp:62-170
    b = SceneBuilder(name)
    b.add_quad(200.0, 200.0, b.lambertian((0.5, 0.5, 0.5)))
    rng = _StdMt19937(1337)
    d01 = lambda: rng.uniform(0.0, 1.0)
    dpos = lambda: rng.uniform(-40.0, 40.0)
    for _ in range(n_balls):
        radius = rng.uniform(0.2, 1.0)
        pz = dpos(); px = dpos()                   # g++ evaluates ctor arguments right to left
        m = d01()
        if m < f32(0.65):
            bl = d01(); g = d01(); r = d01()
            mat = b.lambertian((r, g, bl))
        elif m < f32(0.9):
            rough = f32(f32(0.05) * d01())         # AddMetal(albedo, rough): right-to-left again
            a = f32(f32(0.7) + f32(f32(0.3) * d01()))
            mat = b.metal((a, a, a), rough)
        else:
            mat = b.dielectric(f32(f32(1.3) + f32(f32(0.4) * d01())))
        b.add_sphere(radius, mat, trans=(px, radius, pz))
    for _ in range(8):
        pz = dpos(); px = dpos()
        e = f32(f32(10.0) + f32(f32(10.0) * d01()))
        b.add_sphere(1.5, b.emissive((e, e, e)), trans=(px, 8.0, pz))
    return b.build()


def cornell():                                   # Scene::InitCornell, scene.cpp:332-350
    b = SceneBuilder("cornell")
    red, green, white = b.lambertian((0.75, 0.1, 0.1)), b.lambertian((0.1, 0.75, 0.1)), b.lambertian((0.8, 0.8, 0.8))
    b.add_quad(10, 10, white)
    b.add_quad(10, 10, red, euler=(90, 0, 0), trans=(-5, 5, 0))
    b.add_quad(10, 10, green, euler=(90, 0, 0), trans=(5, 5, 0))
    b.add_quad(10, 10, b.emissive((15, 15, 15)), euler=(90, 0, 0), trans=(0, 9, 0))
    return b.build()


def make_scene(preset=ScenePreset.RANDOM_BALLS_LARGE):
    """Scene::Scene(ScenePreset) (scene.cpp:42-56), flattened."""
    preset = ScenePreset(preset)
    if preset == ScenePreset.CORNELL:
        return cornell()
    if preset == ScenePreset.RANDOM_BALLS_SMALL:
        return _random_balls(100, "random_balls_small")
    if preset == ScenePreset.RANDOM_BALLS_MEDIUM:
        return _random_balls(400, "random_balls_medium")
    if preset == ScenePreset.RANDOM_BALLS_LARGE:
        return _random_balls(800, "random_balls_large")
    if preset == ScenePreset.LIGHT_TEST:         # scene.cpp:285-310
        b = SceneBuilder("light_test")
        b.add_quad(30, 30, b.lambertian((0.6, 0.6, 0.6)))
        for i in range(-5, 6):
            b.add_sphere(0.5, b.emissive((4, 4, 4)), trans=(float(i * 2), 6, 0))
        return b.build()
    if preset == ScenePreset.MATERIAL_TEST:      # scene.cpp:312-330
        b = SceneBuilder("material_test")
        b.add_quad(25, 25, b.lambertian((0.8, 0.8, 0.8)))
        b.add_sphere(1.0, b.lambertian((1, 0, 0)), trans=(-4, 1, 0))
        b.add_sphere(1.0, b.metal((0.9, 0.9, 0.9), 0.0), trans=(0, 1, 0))
        b.add_sphere(1.0, b.dielectric(1.5), trans=(4, 1, 0))
        return b.build()
    b = SceneBuilder("default")                  # Scene::InitDefault, scene.cpp:187-283
    b.add_sphere(1.0, b.emissive((10, 5, 5)), scale=(2, 2, 2), trans=(5, 6, 0))
    b.add_quad(8, 8, b.emissive((3, 4, 2)), euler=(50, 0, 0), trans=(-4, 7, 7))
    b.add_quad(8, 8, b.emissive((3, 2, 1)), euler=(50, 0, 0), trans=(4, 7, 7))
    b.add_sphere(1.0, b.lambertian((0.2, 1.0, 0.2)), trans=(4, 1, 0))
    b.add_sphere(1.0, b.lambertian((1.0, 0.2, 0.2)), trans=(-4, 1, 0))
    b.add_sphere(1.0, b.dielectric(0.9), trans=(0, 1, 4))
    b.add_sphere(1.0, b.metal((1, 0.7, 0.8), 0.01), trans=(0, 1, -4))
    b.add_quad(20, 20, b.lambertian((0.7, 0.7, 0.4)))
    return b.build()


def closed_cornell():
    """A CLOSED box of our own (NOT a reference scene; the reference 'Cornell' is open and
    coplanar): floor, ceiling, back, left, right walls, a ceiling light, a metal and
    a glass sphere. Used as an enclosed-bounce workload."""
    b = SceneBuilder("closed_cornell")
    white, red, green = b.lambertian((0.73, 0.73, 0.73)), b.lambertian((0.65, 0.05, 0.05)), b.lambertian((0.12, 0.45, 0.15))
    b.add_quad(10, 10, white, trans=(0, 0, 0))                                  # floor  y=0
    b.add_quad(10, 10, white, euler=(180, 0, 0), trans=(0, 10, 0))              # ceiling y=10
    b.add_quad(10, 10, white, euler=(90, 0, 0), trans=(0, 5, -5))               # back   z=-5
    b.add_quad(10, 10, red, euler=(0, 0, -90), trans=(-5, 5, 0))                # left   x=-5
    b.add_quad(10, 10, green, euler=(0, 0, 90), trans=(5, 5, 0))                # right  x=+5
    b.add_quad(3, 3, b.emissive((15, 15, 15)), euler=(180, 0, 0), trans=(0, 9.99, 0))
    b.add_sphere(1.5, b.metal((0.9, 0.9, 0.9), 0.02), trans=(-2, 1.5, -1.5))
    b.add_sphere(1.5, b.dielectric(1.5), trans=(2, 1.5, 1))
    return b.build()


# ---------------------------------------------------------------- camera (camera.h:10-16, 104-143)
class Camera:
    """Position + view direction + resolution; right/up recomputed as the reference ctor does."""

    def __init__(self, pos=(5.0, 5.0, 8.0), front=None, width=1920, height=1080):
        self.pos = _v3(*pos)
        if front is None:                          # main.cpp:142-150: normalize(0 - pos)
            front = _normalize((_v3(0, 0, 0) - self.pos).astype(f32))
        self.front_in = _v3(*front)                # what the caller hands to Camera(position, front, ...): pass THIS on when copying a camera
        self.front = _normalize(self.front_in)     # the reference ctor normalises it (camera.h:13), main.cpp's already normalised vector included
        self.width, self.height = int(width), int(height)

    @property
    def right(self):
        return _normalize(_cross(self.front, _v3(0, 1, 0)))

    @property
    def up(self):
        return _normalize(_cross(self.right, self.front))

    def to_c(self):
        c = CCamera()
        for i in range(3):
            c.pos[i], c.front[i], c.right[i], c.up[i] = self.pos[i], self.front[i], self.right[i], self.up[i]
        c.tan_half_fov = _libm.tanf(0.5)           # tan(0.5f), camera.h:111
        c.width, c.height = self.width, self.height
        return c


# ---------------------------------------------------------------- meshes (mesh.cpp:23-184)
_PLY_TYPES = {"char": "b", "int8": "b", "uchar": "B", "uint8": "B", "short": "h", "int16": "h", "ushort": "H",
              "uint16": "H", "int": "i", "int32": "i", "uint": "I", "uint32": "I", "float": "f", "float32": "f",
              "double": "d", "float64": "d"}


def load_ply(path):
    """PLY -> (vertices (nv,3) f32, normals (nv,3) f32 or None, indices (nf,3) u32): the three
    arrays the reference's Mesh exposes (mesh.h:12-14). ASCII and binary_little_endian; faces
    must be triangles (the reference passes list-size hint 3, mesh.cpp:96)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header")
    end = data.index(b"\n", end) + 1
    header = data[:end].decode("ascii", "replace").split("\n")
    fmt, elements = None, []
    for line in header:
        t = line.split()
        if not t:
            continue
        if t[0] == "format":
            fmt = t[1]
        elif t[0] == "element":
            elements.append([t[1], int(t[2]), []])
        elif t[0] == "property":
            elements[-1][2].append(t[1:])
    body = data[end:]
    verts = normals = faces = None
    if fmt == "ascii":
        tok = body.split()
        p = 0
        for name, count, props in elements:
            if name == "vertex":
                k = len(props)
                arr = np.array(tok[p:p + count * k], dtype=np.float64).reshape(count, k)
                p += count * k
                names = [q[-1] for q in props]
                verts = arr[:, [names.index(c) for c in "xyz"]].astype(f32)
                if "nx" in names:
                    normals = arr[:, [names.index(c) for c in ("nx", "ny", "nz")]].astype(f32)
            elif name == "face":
                out = np.empty((count, 3), np.uint32)
                for i in range(count):
                    n = int(tok[p])
                    if n != 3:
                        raise ValueError("non-triangle face (the reference fails on these too, mesh.cpp:96)")
                    out[i] = [int(tok[p + 1]), int(tok[p + 2]), int(tok[p + 3])]
                    p += 1 + n + (len(props) - 1)
                faces = out
            else:
                for _ in range(count):
                    p += len(props)
    elif fmt == "binary_little_endian":
        p = 0
        for name, count, props in elements:
            if name == "vertex":
                dt = np.dtype([(q[-1], "<" + _PLY_TYPES[q[0]]) for q in props])
                arr = np.frombuffer(body, dt, count, p); p += count * dt.itemsize
                verts = np.stack([arr["x"], arr["y"], arr["z"]], 1).astype(f32)
                if "nx" in arr.dtype.names:
                    normals = np.stack([arr["nx"], arr["ny"], arr["nz"]], 1).astype(f32)
            elif name == "face" and len(props) == 1:
                ct, it = _PLY_TYPES[props[0][1]], _PLY_TYPES[props[0][2]]
                dt = np.dtype([("n", "<" + ct), ("i", "<" + it, (3,))])
                arr = np.frombuffer(body, dt, count, p); p += count * dt.itemsize
                if count and not np.all(arr["n"] == 3):
                    raise ValueError("non-triangle face")
                faces = arr["i"].astype(np.uint32)
            else:
                raise ValueError("unsupported PLY element layout: " + name)
    else:
        raise ValueError("unsupported PLY format: %s" % fmt)
    return verts, normals, faces


def write_ply(path, verts, normals, faces, comment=None):
    """Binary little-endian PLY (x y z [nx ny nz] + uchar/uint face lists): the layout load_ply and the reference's tinyply
    loader (mesh.cpp:23-184) both read back exactly (float32 in, float32 out)."""
    v = np.ascontiguousarray(verts, f32)
    f = np.ascontiguousarray(faces, np.uint32).reshape(-1, 3)
    props = "property float x\nproperty float y\nproperty float z\n"
    body = v
    if normals is not None:
        props += "property float nx\nproperty float ny\nproperty float nz\n"
        body = np.concatenate([v, np.ascontiguousarray(normals, f32)], axis=1)
    rec = np.zeros(len(f), dtype=[("k", "u1"), ("i", "<u4", (3,))])
    rec["k"] = 3; rec["i"] = f
    with open(path, "wb") as fh:
        fh.write(("ply\nformat binary_little_endian 1.0\n" + ("comment %s\n" % comment if comment else "") +
                  "element vertex %d\n%selement face %d\nproperty list uchar uint vertex_indices\nend_header\n"
                  % (len(v), props, len(f))).encode())
        fh.write(body.astype("<f4").tobytes())
        fh.write(rec.tobytes())


ASSETS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")


def bundled_mesh(name="dragon"):
    """(verts, normals, faces) of a mesh shipped with the package (assets/<name>.ply)."""
    return load_ply(os.path.join(ASSETS, name + ".ply"))


def subdivide(verts, normals, faces, levels):
    """Midpoint subdivision (x4 triangles per level); normals are averaged, not re-normalised
    (the hit normal is normalised at shading time like primitive.cpp:39)."""
    for _ in range(levels):
        e = np.sort(np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]]), axis=1).astype(np.uint64)
        key = e[:, 0] * np.uint64(1 << 32) + e[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        a, b = (uniq >> np.uint64(32)).astype(np.int64), (uniq & np.uint64(0xffffffff)).astype(np.int64)
        nv = len(verts)
        verts = np.concatenate([verts, ((verts[a] + verts[b]) * f32(0.5)).astype(f32)])
        if normals is not None:
            normals = np.concatenate([normals, ((normals[a] + normals[b]) * f32(0.5)).astype(f32)])
        nf = len(faces)
        m01, m12, m20 = (nv + inv[:nf]), (nv + inv[nf:2 * nf]), (nv + inv[2 * nf:])
        v0, v1, v2 = faces[:, 0], faces[:, 1], faces[:, 2]
        faces = np.concatenate([np.stack([v0, m01, m20], 1), np.stack([v1, m12, m01], 1),
                                np.stack([v2, m20, m12], 1), np.stack([m01, m12, m20], 1)]).astype(np.uint32)
    return verts, normals, faces


def _face_arrays(verts, normals, faces):
    tv = verts[faces.astype(np.int64)]                      # (nf, 3, 3)
    tn = None if normals is None else normals[faces.astype(np.int64)]
    return tv, tn


def mesh_scene(mesh, subdiv=0, instances=((-1.5, 0, 0), (0, 0, 0), (1.5, 0, 0)), name=None):
    """ground quad 20x20 at y=-1 + `instances` copies of the mesh (Lambertian /
    Metal / Dielectric cycling), each `subdiv` x midpoint-subdivided. mesh = PLY path or
    (verts, normals, faces) arrays as Mesh exposes them (mesh.h:12-14)."""
    if isinstance(mesh, (str, os.PathLike)):
        name = name or os.path.splitext(os.path.basename(mesh))[0]
        mesh = load_ply(mesh)
    verts, normals, faces = mesh
    verts, normals, faces = subdivide(verts, normals, np.asarray(faces).reshape(-1, 3), subdiv)
    b = SceneBuilder(name or "mesh")
    b.add_quad(20, 20, b.lambertian((0.5, 0.5, 0.5)), trans=(0, -1, 0))
    mats = [b.lambertian((0.8, 0.3, 0.3)), b.metal((0.9, 0.9, 0.9), 0.05), b.dielectric(1.5)]
    tv, tn = _face_arrays(verts, normals, faces)
    vs, ns, ms = [], [], []
    for i, off in enumerate(instances):
        vs.append((tv + _v3(*off)).astype(f32)); ms.append(np.full(len(tv), mats[i % 3], np.int32))
        if tn is not None:
            ns.append(tn)
    b.add_triangles(np.concatenate(vs), np.concatenate(ns) if ns else None, np.concatenate(ms))
    return b.build()


def add_meshes(sd, meshes):
    """Mesh -> scene glue, the host mirror of `B200Renderer::AddMesh` / `prt_host::AppendMesh`
    (host/flatten.h): every (mesh, material, offset) is appended to a triangle-free scene as world-space triangles after
    its own primitives; material = (type, (r, g, b), param) gets the next index in the table. mesh = PLY path or
    (verts, normals, faces)."""
    if sd.n_tris:
        raise ValueError("add_meshes: the scene already has triangles")
    out = SceneData(spheres=sd.spheres, quads=sd.quads, materials=sd.materials, sky=sd.sky, name=sd.name + "+mesh")
    mats, tv, tn, tm = [tuple(m) for m in sd.materials.tolist()], [], [], []
    for mesh, (mtype, rgb, param), off in meshes:
        verts, normals, faces = load_ply(mesh) if isinstance(mesh, (str, os.PathLike)) else mesh
        v, n = _face_arrays(np.asarray(verts, f32), None if normals is None else np.asarray(normals, f32),
                            np.asarray(faces).reshape(-1, 3))
        v = (v + _v3(*off)).astype(f32)
        if n is None:
            fn = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]).astype(f32)
            fn = (fn / np.sqrt((fn * fn).sum(-1, keepdims=True), dtype=f32)).astype(f32)
            n = np.repeat(fn[:, None, :], 3, axis=1)
        rec = np.zeros(1, material_dt)
        rec["type"] = mtype; rec["r"], rec["g"], rec["b"] = rgb; rec["param"] = param
        mats.append(tuple(rec.tolist()[0]))
        tv.append(v.reshape(-1, 9)); tn.append(n.reshape(-1, 9)); tm.append(np.full(len(v), len(mats) - 1, np.int32))
    out.materials = np.array(mats, dtype=material_dt)
    out.tri_verts, out.tri_normals, out.tri_material = np.concatenate(tv), np.concatenate(tn), np.concatenate(tm)
    out.tri_prim_base = len(sd.spheres) + len(sd.quads)
    return out


def random_triangle_soup(n_tris, seed=0xB200, extent=50.0, edge=0.25, albedo=0.7):
    """n random triangles, centre ~U([-extent,extent]^3), edges ~U([-edge,edge]^3),
    flat normals, all Lambertian(albedo). Counter-based integer generator -> identical bits anywhere."""
    i = np.arange(n_tris * 9, dtype=np.uint64)
    with np.errstate(over="ignore"):               # splitmix64 finaliser, wrap-around intended
        x = i + np.uint64((seed * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)
        x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
    u = ((x >> np.uint64(40)).astype(np.float32) * f32(2.0 ** -24)).reshape(n_tris, 9)
    c = (u[:, 0:3] * f32(2 * extent) - f32(extent)).astype(f32)
    e1 = (u[:, 3:6] * f32(2 * edge) - f32(edge)).astype(f32)
    e2 = (u[:, 6:9] * f32(2 * edge) - f32(edge)).astype(f32)
    tv = np.stack([c, c + e1, c + e2], 1).astype(f32)
    b = SceneBuilder("soup_%d" % n_tris)
    b.add_triangles(tv, None, b.lambertian((albedo, albedo, albedo)))
    return b.build()
