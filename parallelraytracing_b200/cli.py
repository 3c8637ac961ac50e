"""Headless driver in place of the reference's GLFW/ImGui viewer.

Same object sequence and frame loop as main.cpp:138-170 and 506-532 -
    Film film(W, H); Scene scene(preset); Camera camera((5,5,8) -> origin);
    renderer.Init(film, scene, camera); film.Clear();
    per frame: renderer.ProgressiveRender(); film.UpdateDisplay(); <present>
- with the same three timers (render / film update / upload-draw, main.cpp:512-532) logged as JSON,
and "present" = write the RGBA8 display buffer to a PNG/PPM file (no GL on this box).

  python -m parallelraytracing_b200.cli --scene cornell --width 1920 --height 1080 --frames 16 \
         --spp-per-frame 16 --depth 8 --seed 1 --out cornell.png
  torchrun --nproc-per-node 8 ... -m parallelraytracing_b200.cli --scene dragon.ply ...   (sample-range sharding)

There is no CPU path: without a CUDA device / libprt_b200.so the command fails.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

from . import scene as S
from .imageio import write_image
from .renderer import B200Renderer, Film

PRESETS = {p.name.lower(): p for p in S.ScenePreset}


def build_scene(name, subdiv=0):
    """--scene: a ScenePreset name (scene.h:6-15), 'closed_cornell', 'soup:<n_tris>' or a .ply path
    (3 instances + ground quad)."""
    key = name.lower()
    if key in PRESETS:
        return S.make_scene(PRESETS[key]), None
    if key == "closed_cornell":
        return S.closed_cornell(), None
    if key.startswith("soup:"):
        sd = S.random_triangle_soup(int(key.split(":")[1]))
        return sd, S.Camera(pos=(0.0, 0.0, 150.0), front=(0.0, 0.0, -1.0))
    if key.endswith(".ply"):
        sd = S.mesh_scene(S.load_ply(name), subdiv=subdiv, name=os.path.basename(name))
        return sd, S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)))
    raise SystemExit("unknown scene %r (presets: %s, closed_cornell, soup:<n>, <file>.ply)" % (name, ", ".join(PRESETS)))


def main(argv=None):
    ap = argparse.ArgumentParser(prog="parallelraytracing_b200.cli", description=__doc__.split("\n")[0])
    ap.add_argument("--scene", default="random_balls_large")          # the viewer's default preset, scene.cpp:182-185
    ap.add_argument("--subdiv", type=int, default=0, help="midpoint subdivisions of a PLY mesh (x4 triangles each)")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--frames", type=int, default=16, help="ProgressiveRender() calls")
    ap.add_argument("--spp-per-frame", type=int, default=1, help="samples per ProgressiveRender (reference: exactly 1)")
    ap.add_argument("--depth", type=int, default=20, help="max closest-hit queries per path (reference literal: 20)")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--camera-pos", type=float, nargs=3, default=None)
    ap.add_argument("--camera-target", type=float, nargs=3, default=(0.0, 0.0, 0.0))
    ap.add_argument("--exposure", type=float, default=1.0)
    ap.add_argument("--gamma", type=float, default=2.2)
    ap.add_argument("--out", default=None, help="image file (.png or .ppm) written by rank 0")
    ap.add_argument("--raw", default=None, help="also save the float accumulation buffer (sum of samples) as .npy")
    ap.add_argument("--partition", default="samples", choices=["samples", "tiles"],
                    help="multi-GPU split of a frame: sample ranges, or bands of rows (SURVEY 8e)")
    ap.add_argument("--opt", action="append", default=[], help="prt_set_option key=value (applied before Init, so build-time options count)")
    ap.add_argument("--devices", default=None, help="comma-separated CUDA ordinals: ONE process drives them all (prt_create_multi); "
                                                   "alternative to one process per GPU under torchrun")
    args = ap.parse_args(argv)

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    sd, cam = build_scene(args.scene, args.subdiv)
    if args.camera_pos is not None or cam is None:
        pos = np.asarray(args.camera_pos if args.camera_pos is not None else (5.0, 5.0, 8.0), dtype=np.float32)
        front = S._normalize(np.asarray(args.camera_target, dtype=np.float32) - pos)
        cam = S.Camera(pos=tuple(float(x) for x in pos), front=front)
    cam = S.Camera(pos=tuple(float(x) for x in cam.pos), front=cam.front_in, width=args.width, height=args.height)

    film = Film(args.width, args.height)
    devices = [int(d) for d in args.devices.split(",")] if args.devices else None
    r = B200Renderer(device=devices if devices and len(devices) > 1 else (devices[0] if devices else local),
                     max_depth=args.depth, samples_per_call=args.spp_per_frame, seed=args.seed)
    t0 = time.perf_counter()
    # options BEFORE Init: build-time ones (bvh_builder, ploc_radius, bvh_wide, wide_min_prims, morton_centroid) take effect in
    # prt_build_bvh, which Init runs - applied afterwards they silently measured the default tree (advisor finding, round 1)
    from .capi import Context, PrtError
    try:
        r.ctx = Context(r.m_Device)
        for kv in args.opt:
            k, v = kv.split("=")
            r.ctx.set_option(k, int(v))
        if devices and len(devices) > 1 and args.partition == "tiles":
            r.ctx.set_option("multi_partition", 1)
    except PrtError as e:
        raise SystemExit("B200Renderer: %s (there is no CPU path)" % e)
    r.Init(film, sd, cam)
    if not r.ok:
        raise SystemExit("B200Renderer.Init failed: no CUDA device or libprt_b200.so (there is no CPU path)")
    init_ms = 1e3 * (time.perf_counter() - t0)
    film.Clear()

    sharded = None
    if world > 1:
        import torch
        import torch.distributed as dist
        from .dist import ShardedRenderer
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        sharded = ShardedRenderer(r, film, torch.device("cuda", local), partition=args.partition)

    render_ms, film_ms, rays = [], [], 0
    display = None
    for _ in range(args.frames):
        t0 = time.perf_counter()
        if sharded is None:
            r.ProgressiveRender()
        else:
            sharded.ProgressiveRender(args.spp_per_frame * world)
        t1 = time.perf_counter()
        rays += r.Stats().rays_total
        if sharded is None:
            display = film.UpdateDisplay(args.exposure, args.gamma)     # film.UpdateDisplayGPU(), main.cpp:516
        t2 = time.perf_counter()
        render_ms.append(1e3 * (t1 - t0)); film_ms.append(1e3 * (t2 - t1))

    accum, weight = None, 0
    if sharded is None:
        accum, weight = film.ReadAccum(), film.Weight()
    elif rank == 0:
        accum, weight = sharded.total.cpu().numpy().reshape(args.height, args.width, 3), sharded.weight
        # rank 0 resolves the reduced image with the same Film kernel: hand it the reduced sum as its accumulator
        r.ctx.set_accum_buffer(sharded.total.data_ptr())
        r.ctx.set_accum_spp(weight)
        display = film.UpdateDisplay(args.exposure, args.gamma)
    t3 = time.perf_counter()
    if rank == 0:
        if args.out and display is not None:
            write_image(args.out, display)
        if args.raw:
            np.save(args.raw, accum)
        upload_ms = 1e3 * (time.perf_counter() - t3)
        st = r.Stats()
        print(json.dumps({
            "scene": sd.name, "prims": sd.n_prims, "width": args.width, "height": args.height, "frames": args.frames,
            "spp_per_frame": args.spp_per_frame * world, "spp_total": int(weight), "depth": args.depth, "seed": args.seed,
            "n_gpus": world * st.n_devices, "init_ms": init_ms, "bvh_build_ms": st.build_ms,
            "bvh_builder": "ploc" if st.bvh_builder == 1 else "karras", "options": args.opt,
            "render_ms_avg": float(np.mean(render_ms)), "film_update_ms_avg": float(np.mean(film_ms)), "present_ms": upload_ms,
            "rays_rank0": int(rays), "mrays_per_s_rank0_wall": rays / max(1e-9, sum(render_ms)) / 1e3,
            "mean_radiance": float(accum.mean() / max(1, weight)), "out": args.out}))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    r.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
