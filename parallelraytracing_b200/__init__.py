"""parallelraytracing_b200 — B200-native path-tracing backend behind the reference's
`Renderer` interface (Rickyeeeeee/ParallelRayTracing, src/core/renderer.h:8-16).

Only the hot path lives here: csrc/ (hand-written sm_100a CUDA + the C ABI of include/prt.h),
host/ (the C++ `B200Renderer : public Renderer` shim for the reference tree) and the Python
mirror of the same interface used by tests and bench.py (capi.py, scene.py, renderer.py).
"""
from .capi import PrtError, Context, lib_path, load_library  # noqa: F401
from .scene import (SceneData, SceneBuilder, Camera, ScenePreset, make_scene, cornell, closed_cornell,  # noqa: F401
                    random_triangle_soup, mesh_scene, load_ply, subdivide)
from .renderer import B200Renderer, Film  # noqa: F401

__version__ = "0.1.0"
