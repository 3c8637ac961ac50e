#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 path-tracing backend (contract: see README / DESIGN.md).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one pass of the hot path over one batch: render `spp` samples/pixel of the workload into the float
accumulation buffer; with N>1 ranks (torchrun, one per GPU) every rank renders its own sample range of the same frame
and the buffers are sum-reduced with NCCL. Prints ONE JSON line on rank 0.

The default run measures, in one process and with the same timing rules (L2 flush between steps, CUDA events, max over
ranks), BASELINE.json's configs:
  value / e2e / roofline ...   configs[1]  cornell_1080p_256spp_d8      (weak scaling: 256 spp per GPU)
  workloads[...]               configs[2]  dragon960k_1080p_64spp_d8    (960 K-triangle mesh: GPU BVH build + wide-node traversal)
                               configs[4]  soup10m_1080p_16spp_d16      (10 M random triangles, fully incoherent, depth 16)
                               configs[3]  dragon960k_4k_1024spp_d8     (N > 1 only: 4K, 1024 spp IN TOTAL, strong scaling)
`--workload NAME` measures that one workload alone (what the profiling scripts use).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: scene, W, H, spp per step, depth
    "cornell_1080p_256spp_d8": dict(scene="cornell", W=1920, H=1080, spp=256, depth=8),
    "random_balls_large_1080p_64spp_d8": dict(scene="random_balls_large", W=1920, H=1080, spp=64, depth=8),
    "closed_cornell_1080p_64spp_d8": dict(scene="closed_cornell", W=1920, H=1080, spp=64, depth=8),
    # BASELINE.json configs[2] / SURVEY 8d C3: dragon.ply x 3 instances x 16 (two midpoint subdivisions) = 960 K triangles, 64 spp
    "dragon960k_1080p_64spp_d8": dict(scene="dragon960k", W=1920, H=1080, spp=64, depth=8),
    "dragon960k_1080p_16spp_d8": dict(scene="dragon960k", W=1920, H=1080, spp=16, depth=8),
    # BASELINE.json configs[4] / SURVEY 8d C5: 10 M random triangles, all Lambertian, depth 16, 16 spp
    "soup10m_1080p_16spp_d16": dict(scene="soup10m", W=1920, H=1080, spp=16, depth=16),
    "soup10m_1080p_4spp_d16": dict(scene="soup10m", W=1920, H=1080, spp=4, depth=16),
    "soup1m_1080p_8spp_d16": dict(scene="soup1m", W=1920, H=1080, spp=8, depth=16),
    # BASELINE.json configs[3] / SURVEY 8d C4: the C3 mesh scene at 4K, 1024 spp IN TOTAL, sample range split over
    # the ranks (strong scaling: rank r renders samples [r*1024/N, (r+1)*1024/N)), one 99.5 MB NCCL sum-reduce
    "dragon960k_4k_1024spp_d8": dict(scene="dragon960k", W=3840, H=2160, spp=1024, depth=8, strong=True),
    "dragon960k_4k_64spp_d8": dict(scene="dragon960k", W=3840, H=2160, spp=64, depth=8, strong=True),
}
DEFAULT_WORKLOAD = "cornell_1080p_256spp_d8"
EXTRA_WORKLOADS = ["dragon960k_1080p_64spp_d8", "soup10m_1080p_16spp_d16"]        # every default run
EXTRA_WORKLOADS_MULTI = ["dragon960k_4k_1024spp_d8"]                              # default runs with N > 1
_scene_cache = {}


def make_workload(name):
    from parallelraytracing_b200 import scene as S
    w = WORKLOADS[name]
    s = w["scene"]
    cam = S.Camera(width=w["W"], height=w["H"])                       # viewer camera (5,5,8)->0, main.cpp:142-150
    if s == "dragon960k":
        cam = S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=w["W"], height=w["H"])
    elif s in ("soup10m", "soup1m"):
        cam = S.Camera(pos=(0.0, 0.0, 150.0), front=(0.0, 0.0, -1.0), width=w["W"], height=w["H"])
    if s not in _scene_cache:
        if s == "cornell":
            sd = S.cornell()
        elif s == "random_balls_large":
            sd = S.make_scene(S.ScenePreset.RANDOM_BALLS_LARGE)
        elif s == "closed_cornell":
            sd = S.closed_cornell()
        elif s == "dragon960k":                                      # the mesh ships with the package (assets/dragon.ply)
            sd = S.mesh_scene(S.bundled_mesh("dragon"), subdiv=2, name="dragon960k")
        elif s in ("soup10m", "soup1m"):
            sd = S.random_triangle_soup(10_000_000 if s == "soup10m" else 1_000_000)
        else:
            raise ValueError(s)
        _scene_cache.clear()                                         # one big scene at a time (10 M triangles = 760 MB of host arrays)
        _scene_cache[s] = sd
    return _scene_cache[s], cam, w


def scene_bytes(sd):
    n = sd.spheres.nbytes + sd.quads.nbytes + sd.materials.nbytes
    if sd.n_tris:
        n += sd.tri_verts.nbytes + sd.tri_normals.nbytes + sd.tri_material.nbytes
    return int(n)


def config_of(name, sd, w, spp_gpu, args):
    """The `config` object of the JSON line: identical for both arms (the reference arm renders a bounded SAMPLE of this
    workload on the host cores; what the sample was is said under cpu_baseline.sample)."""
    part = "bands of rows" if getattr(args, "partition", "samples") == "tiles" else "sample-range"
    return {"workload": name, "scene": sd.name, "prims": int(sd.n_prims), "width": w["W"], "height": w["H"],
            "spp_per_step_per_gpu": int(spp_gpu), "depth": w["depth"],
            "partition": part + ", NCCL sum-reduce to rank 0",
            "l2": "256 MiB L2 flush between timed steps; ray queues (GBs) also exceed the 126 MB L2",
            "options": list(getattr(args, "opt", []))}


def spp_per_gpu(w, world, tiles=False):
    if w.get("strong") and not tiles:          # fixed total work: every rank takes an equal share of the samples
        if w["spp"] % world:
            raise SystemExit("bench.py: %d spp does not split over %d ranks" % (w["spp"], world))
        return w["spp"] // world
    return w["spp"]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


# ---------------------------------------------------------------- CPU arm (oracle/ is used ONLY here)
PRESET_OF = {"cornell": 3, "random_balls_large": 6}


class CpuArm:
    """The reference's own CPU implementation of the path, timed on this host's cores:
    oracle/_ref (reference Scene::Intersect, MaterialHandle::Scatter/Emit, Camera::GetCameraRay compiled
    from the reference sources + the TraceRay recursion of cpu/renderer.cpp:59-103 with a depth
    parameter and the deterministic rand()) when it was built, else the oracle port."""

    def __init__(self, workload):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_lib
        self.sd, self.cam, self.w = make_workload(workload)
        self.cores = os.cpu_count() or 1
        ref = oracle_lib.load_ref()
        if ref is not None and self.w["scene"] in PRESET_OF:
            self.kind, self.scene = "reference", ref.preset(PRESET_OF[self.w["scene"]])
            self._render = lambda spp, y0, y1, s0: self.scene.render(self.cam, spp, self.w["depth"], seed=1, sample_begin=s0,
                                                                     n_threads=self.cores, y_begin=y0, y_end=y1)
        else:
            self.kind, self.scene = "port", oracle_lib.load_oracle().scene(self.sd)
            self.scene.build_lbvh()
            self._render = lambda spp, y0, y1, s0: self.scene.render(self.cam, spp, self.w["depth"], seed=1, sample_begin=s0,
                                                                     use_bvh=True, n_threads=self.cores, y_begin=y0, y_end=y1)

    def run(self, spp, rows=None):
        """Render `spp` samples of a centred band of `rows` rows (None = whole frame). Returns (seconds, rays)."""
        H = self.cam.height
        y0 = 0 if rows is None else (H - rows) // 2
        t = time.perf_counter()
        _, rays = self._render(spp, y0, H if rows is None else y0 + rows, 0)
        return time.perf_counter() - t, rays

    def calibrate(self, target_s):
        """Pick (spp, rows) so one sample run takes about target_s."""
        H = self.cam.height
        rows = max(8, H // 16)
        dt, rays = self.run(1, rows)
        per_row_spp = max(dt, 1e-4) / rows
        full_spp = per_row_spp * H
        if full_spp <= target_s:
            return max(1, int(target_s / full_spp)), None
        return 1, max(8, min(H, int(target_s / per_row_spp)))

    def as_shipped(self):
        """the reference CPURenderer EXACTLY as shipped (depth literal 20, std::rand behind its global lock,
        one std::async task per 16x16 tile), one ProgressiveRender() = 1 spp of the whole frame, wall clock. Preset scenes
        with oracle/_ref only. It counts no rays, so the figure is spp/s."""
        import oracle_lib
        ref = oracle_lib.load_ref()
        if ref is None or self.w["scene"] not in PRESET_OF:
            return None
        sec, _ = ref.render_shipped(PRESET_OF[self.w["scene"]], self.cam, 1, want_accum=False)
        return {"spp_per_s": 1.0 / sec, "seconds_per_ProgressiveRender": sec, "depth": 20,
                "what": "CPURenderer::ProgressiveRender unmodified (oracle/_ref), %dx%d, %d host threads available" %
                        (self.cam.width, self.cam.height, self.cores)}

    def reference_gpu(self):
        """the reference's own CUDA backends (compiled unmodified for sm_100 into oracle/_ref by `make -C oracle
        refgpu`) and B200Renderer behind the same C++ Renderer interface, on THIS GPU, the viewer's call pattern: one
        ProgressiveRender = 1 spp at the reference's depth literal 20, wall clock per call. Preset scenes only."""
        exe = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_driver")
        if not os.path.exists(exe) or self.w["scene"] not in PRESET_OF:
            return None
        res = {"depth": 20, "spp_per_call": 1, "what": "ms per Renderer::ProgressiveRender(), %dx%d, 30 calls after 3" %
               (self.cam.width, self.cam.height)}
        for backend in ("wavefront", "megakernel", "b200"):
            try:
                r = subprocess.run([exe, backend, str(int(PRESET_OF[self.w["scene"]])), str(self.cam.width), str(self.cam.height),
                                    "3", "30", "1", "0"], capture_output=True, text=True, timeout=300)
                j = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
                res["ref_" + backend if backend != "b200" else "B200Renderer"] = {
                    "ms_per_call": j["ms_per_call"], "spp_per_s": j["spp_per_s"], "mean_radiance": j["mean_radiance"]}
            except Exception as e:  # noqa: BLE001
                res[backend] = {"error": str(e)[:200]}
        return res

    def describe(self, spp, rows):
        H, W = self.cam.height, self.cam.width
        what = ("reference Intersect/Scatter/GetCameraRay (oracle/_ref) + TraceRay recursion, deterministic rand()"
                if self.kind == "reference" else "oracle port (CPU BVH traversal, not the reference's brute force)")
        return "%s: %d spp x %d of %d rows x %d px, depth %d, %d threads" % (what, spp, H if rows is None else rows, H, W,
                                                                            self.w["depth"], self.cores)


def run_reference_arm(args, out_fd=1):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    arm = CpuArm(args.workload)
    target = max(1.0, min(10.0, 150.0 / max(1, args.steps + args.warmup)))
    spp, rows = arm.calibrate(target)
    for _ in range(args.warmup):
        arm.run(spp, rows)
    t_tot, r_tot = 0.0, 0
    for _ in range(args.steps):
        dt, rays = arm.run(spp, rows)
        t_tot += dt; r_tot += rays
    val = r_tot / t_tot / 1e6
    H = arm.cam.height
    frac = spp * (1.0 if rows is None else rows / H)
    out = {"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps, "higher_is_better": True,
           "scaling": "strong" if arm.w.get("strong") else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": config_of(args.workload, arm.sd, arm.w, spp_per_gpu(arm.w, max(1, world)), args),
           "spp_per_s": frac * args.steps / t_tot, "spp_per_s_1080p": (frac * args.steps / t_tot) if H == 1080 else None,
           "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": arm.cores, "kind": arm.kind, "sample": arm.describe(spp, rows),
                            "sample_spp_of_frame_per_step": frac},
           "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    sys.stdout.flush()
    emit(json.dumps(out), out_fd)
    return 0


# ---------------------------------------------------------------- GPU arm
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measure(args, name, steps, warmup, with_clocks=False, devices=None):
    """Device-timed and end-to-end measurement of ONE workload on this rank's GPU (+ the NCCL reduce with N > 1).
    devices = [0..N-1]: ONE process drives N GPUs through a multi-device context (prt_create_multi: the split and the NCCL
    reduce happen inside libprt_b200.so); no torch.distributed on that path."""
    import torch
    import torch.distributed as dist
    from parallelraytracing_b200.capi import Context
    from parallelraytracing_b200.dist import pixel_range

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    n_dev = 1
    if devices:
        rank, world, n_dev = 0, 1, len(devices)
    sd, cam, w = make_workload(name)
    W, H, depth = w["W"], w["H"], w["depth"]
    tiles = args.partition == "tiles"
    if tiles and not w.get("strong"):
        raise SystemExit("bench.py: --partition tiles needs a fixed-total-work workload (dragon960k_4k_*)")
    spp = spp_per_gpu(w, world, tiles) if (w.get("strong") or not devices) else w["spp"] * n_dev   # weak scaling in one process: N x the samples in one call
    pix_begin, pix_count = pixel_range(rank, world, W * H, W) if tiles else (0, W * H)

    def render(seed):
        if tiles:
            ctx.render_region(spp, depth, seed=seed, sample_begin=0, pixel_begin=pix_begin, pixel_count=pix_count)
        else:
            ctx.render(spp, depth, seed=seed, sample_begin=rank * spp)
    ctx = Context(list(devices) if devices else local)
    if devices and tiles:
        ctx.set_option("multi_partition", 1)
    for kv in args.opt:
        k, v = kv.split("=")
        ctx.set_option(k, int(v))
    ctx.set_scene(sd); ctx.build_bvh(); ctx.set_camera(cam)
    ctx.build_bvh()                                                       # second build: allocations are warm, the kernels are timed
    st0 = ctx.stats()
    build_ms, builder, height = st0.build_ms, st0.bvh_builder, st0.bvh_height
    # the accumulation buffer is a torch tensor so that NCCL can reduce it in place
    accum = torch.zeros(H * W * 3, dtype=torch.float32, device="cuda")
    ctx.set_accum_buffer(accum.data_ptr())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2

    def step(i):
        """rank r renders samples [r*spp, (r+1)*spp) of step i's frame; sum-reduce to rank 0."""
        accum.zero_()
        ctx.clear_accum()
        render(1 + i)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for i in range(warmup):
        step(i)
    sync_all()
    clocks = ClockSampler(local)
    if rank == 0 and with_clocks:
        clocks.start()
    # ---- timed region: device time of prt_render (CUDA events on the context's own stream, read back
    # through prt_get_stats) + the reduce (torch events on torch's stream), max over ranks
    rays = 0; launches = 0; render_ms = 0.0; ext_ms = 0.0; shd_ms = 0.0; red_ms = 0.0; prim_ms = 0.0; red_in_ms = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    t_wall = time.perf_counter()
    for i in range(steps):
        flush.zero_()                                                     # L2 flush between timed iterations
        accum.zero_()
        ctx.clear_accum()
        torch.cuda.synchronize()
        render(100 + i)                                                   # synchronous, like ProgressiveRender
        st = ctx.stats()
        rays += st.rays_total; launches += st.kernel_launches
        render_ms += st.render_ms; ext_ms += st.extend_ms; shd_ms += st.shade_ms; prim_ms += st.primary_ms; red_in_ms += st.reduce_ms
        if world > 1:
            e0.record(); dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM); e1.record(); torch.cuda.synchronize()
            red_ms += e0.elapsed_time(e1)
    sync_all()
    wall_ms = 1e3 * (time.perf_counter() - t_wall)
    clk = clocks.stop() if (rank == 0 and with_clocks) else None
    dev_ms = render_ms + red_ms
    if world > 1 and not devices:
        t = torch.tensor([dev_ms, float(rays), float(launches)], dtype=torch.float64, device="cuda")
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms_max, rays_all, launches_all = float(tmax[0]), float(tsum[1]), int(tsum[2])
    else:
        dev_ms_max, rays_all, launches_all = dev_ms, float(rays), launches

    # ---- e2e: the call a user makes (Renderer interface, HOST scene in, HOST image out), every step; the step's inputs
    # (the host scene arrays) sit in page-locked host memory, as the contract asks
    from parallelraytracing_b200 import B200Renderer, Film
    sd.pin()
    film, r = Film(W, H), B200Renderer(device=list(devices) if devices else local, max_depth=depth, samples_per_call=spp, seed=1)
    r.ctx = ctx                                                          # same device context (allocations stay warm)
    ctx.set_accum_buffer(None)
    e2e_rays = 0; init_s = 0.0
    from parallelraytracing_b200.capi import pinned_copy
    host_img = pinned_copy(np.zeros((H, W, 3), np.float32))             # D2H target in page-locked memory
    sync_all()
    t0 = time.perf_counter()
    for i in range(steps):
        r.SetSeed(200 + i); r.SetNextSample(rank * spp)
        ti = time.perf_counter()
        r.Init(film, sd, cam)                 # host scene -> H2D upload + GPU BVH build + camera
        init_s += time.perf_counter() - ti
        film.Clear()
        if tiles:                             # this rank's band of rows, all samples
            ctx.render_region(spp, depth, seed=200 + i, sample_begin=0, pixel_begin=pix_begin, pixel_count=pix_count)
        else:
            r.ProgressiveRender()
        img = film.ReadAccum(host_img)        # D2H of the float image
        e2e_rays += r.Stats().rays_total
    e2e_s = time.perf_counter() - t0
    if world > 1 and not devices:
        t = torch.tensor([e2e_s, float(e2e_rays)], dtype=torch.float64, device="cuda")
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        e2e_s, e2e_rays = float(tmax[0]), float(tsum[1])
    paths_step = pix_count * spp
    if devices:                                  # one process, N devices: report as an N-GPU line; the reduce ran inside prt_render
        world, red_ms = n_dev, red_in_ms         # (render_ms already contains it: root ev0 -> end of the reduce)
        spp = spp // n_dev if not tiles else spp
    R = dict(rank=rank, world=world, inproc=bool(devices), ctx=ctx, sd=sd, cam=cam, w=w, name=name, steps=steps, warmup=warmup, rays_all=rays_all,
             dev_ms_max=dev_ms_max, launches_all=launches_all, ext_ms=ext_ms, shd_ms=shd_ms, prim_ms=prim_ms, render_ms=render_ms,
             red_ms=red_ms, rays=rays, spp=spp, tiles=tiles, paths_step=paths_step, clk=clk, e2e_s=e2e_s, e2e_rays=e2e_rays,
             init_ms=1e3 * init_s / steps, build_ms=build_ms, builder=builder, height=height, wall_ms=wall_ms, img_mean=float(img.mean()))
    out = result_json(args, R) if rank == 0 else None
    ctx.close()
    del accum, flush
    torch.cuda.empty_cache()
    return out


def result_json(args, R):
    """The JSON object of one workload (rank 0): contract keys + roofline."""
    ctx, w, sd, name = R["ctx"], R["w"], R["sd"], R["name"]
    W, H, spp, depth, steps = w["W"], w["H"], R["spp"], w["depth"], R["steps"]
    # per-ray traversal work from the device counters (untimed pass, counters cost speed)
    ctx.enable_counters(True)
    ctx.clear_accum(); ctx.render(min(spp, 4), depth, seed=999, sample_begin=0)
    st = ctx.stats()
    ctx.enable_counters(False)
    n_node = st.nodes_visited / max(1, st.rays_total)
    n_leaf = st.prims_tested / max(1, st.rays_total)
    per = R["world"] if R.get("inproc") else 1         # one process driving N devices: per-device averages for the roofline
    rays_step = R["rays"] / steps / per                # this rank (device)
    paths_step = R["paths_step"] / per                 # this rank (device)
    cont = rays_step - paths_step                      # segments that continued into another bounce
    fused = R["shd_ms"] == 0.0
    # ALGORITHMIC bytes: B_ray = B_queue + B_shade + 64*n_node + 48*n_leaf.
    #   staged: extend 32 in (bounce >= 1 only) + 16 hit out; shade 48 in (bounce >= 1 only) + 16 hit in, 48 out per continuing ray,
    #           16 out per finished path
    #   fused : no hit queue and no queue traffic for camera rays: 48 in per bounce>=1 ray, 48 out per continuing
    #           ray, 16 out per finished path
    b_shade = 64 if sd.n_tris > (len(sd.spheres) + len(sd.quads)) else 16
    b_trav = 64 * n_node + 48 * n_leaf
    if fused:
        queue_bytes = cont * 48 + cont * 48 + paths_step * 16
        kname = "k_primary + k_bounce (fused raygen/extend/shade/compaction, warp-uniform list loop)"
    else:
        # camera rays (bounce 0) never touch a queue: the bounce-0 extend / shade kernels compute them from the path index
        queue_bytes = cont * 32 + rays_step * 16 + cont * 48 + rays_step * 16 + cont * 48 + paths_step * 16
        wide = sd.n_prims >= 4096 and not any(o.startswith("bvh_wide=0") for o in args.opt)
        kname = ("k_extend_wide + k_shade (persistent traversal of the 4-wide quantised nodes, shade/compaction)" if wide else
                 "k_extend_pt + k_shade (persistent two-child BVH traversal, shade/compaction)")
    alg_bytes = queue_bytes + rays_step * (b_shade + b_trav)
    # kernel time: summed per-launch CUDA-event spans; with two waves in flight (two streams) spans of different
    # waves overlap in time, so the sum is capped by the device time of the whole render
    ker_ms = min(R["ext_ms"] + R["shd_ms"], R["render_ms"]) / steps
    peak, peak_src = peaks()
    alg = alg_bytes / (ker_ms * 1e-3) / 1e9
    hbm = queue_bytes / (ker_ms * 1e-3) / 1e9
    prof = {}
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name, {})
    except Exception:
        pass
    traffic = prof.get("dram_bytes_per_step")
    # Which resource binds: every hot kernel of this path is limited by INSTRUCTION ISSUE at partial warp occupancy (ncu:
    # profiles/), as SURVEY 8(d) predicted for the small scenes and as the counters show for the meshes too (DRAM 5-23 %).
    # So `frac` is the honest HBM number: bytes that MUST cross HBM (ray queues, hit queue, contributions - the BVH working
    # set is served by L1/L2) over the measured copy bandwidth; the SURVEY-8(d) algorithmic figure, which also counts 64 B per
    # node visit and 48 B per leaf test wherever they are served from, is kept as `algorithmic_l2_level`.
    roof = {"bound": "issue", "kernel": kname, "achieved": hbm, "peak": peak, "unit": "GB/s", "frac": hbm / peak,
            "frac_what": "HBM-only: queue + hit + contribution bytes that must cross HBM / kernel time / measured copy peak",
            "traffic": traffic, "traffic_frac": (traffic / (ker_ms * 1e-3) / 1e9 / peak) if traffic else None,
            "l2_traffic": prof.get("l2_bytes_per_step"), "peak_source": peak_src,
            "issue_slot_pct": prof.get("issue_slot_pct"), "lanes_per_inst": prof.get("lanes_per_inst"),
            "ncu_source": prof.get("source"),
            "algorithmic_l2_level": {"achieved": alg, "frac": alg / peak, "bytes_per_ray": alg_bytes / max(1.0, rays_step),
                                     "what": "SURVEY 8(d): B_queue + B_shade + 64 B x nodes + 48 B x leaf tests per ray; > HBM-only "
                                             "because node/leaf bytes are served by L1/L2"},
            "kernel_ms_per_step": ker_ms, "extend_ms_per_step": R["ext_ms"] / steps, "shade_ms_per_step": R["shd_ms"] / steps,
            "primary_ms_per_step": R["prim_ms"] / steps, "render_ms_per_step": R["render_ms"] / steps,
            "nodes_per_ray": n_node, "leaf_tests_per_ray": n_leaf,
            "algorithmic_bytes_per_step": alg_bytes, "queue_bytes_per_step": queue_bytes,
            "note": "achieved/frac = bytes that must cross HBM in all launches of these kernels in one step / their summed "
                    "CUDA-event time (events on the library's own stream, read back via prt_get_stats). bound = issue: ncu has "
                    "these kernels at 64-81 % issue-slot utilisation with 16-27 of 32 lanes active (profiles/)."}
    dev_s = R["dev_ms_max"] * 1e-3
    value = R["rays_all"] / dev_s / 1e6
    # samples per pixel of the FRAME completed per step: weak scaling adds every rank's spp, strong scaling splits the total
    spp_frame = spp if R["tiles"] else spp * R["world"]
    out = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": R["world"], "steps": steps, "warmup": R["warmup"],
           "ms_per_step": R["dev_ms_max"] / steps, "higher_is_better": True, "scaling": "strong" if w.get("strong") else "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": config_of(name, sd, w, spp, args),
           "spp_per_s": spp_frame * steps / dev_s, "spp_per_s_1080p": (spp_frame * steps / dev_s) if H == 1080 else None,
           "mpaths_per_s": spp_frame * W * H * steps / dev_s / 1e6,
           "rays_per_path": rays_step / paths_step,
           "e2e": {"value": R["e2e_rays"] / R["e2e_s"] / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": scene_bytes(sd) + 64,
                   "d2h_bytes_per_step": W * H * 12, "init_ms_per_step": R["init_ms"],
                   "what": "B200Renderer.Init(host scene in page-locked memory: H2D + GPU BVH build)+ProgressiveRender+Film.ReadAccum per step, wall clock"},
           "gpu_launches": R["launches_all"], "bvh_build_ms": R["build_ms"],
           "bvh_builder": "ploc" if R["builder"] == 1 else "karras", "bvh_height": R["height"],
           "bvh_build_mprims_per_s": (sd.n_prims / (R["build_ms"] * 1e-3) / 1e6) if R["build_ms"] > 0 else None,
           "multi_gpu": ("one process, prt_create_multi: split + NCCL reduce inside libprt_b200.so" if R.get("inproc") else
                         "one process per GPU (torchrun), torch.distributed NCCL reduce") if R["world"] > 1 else None,
           "reduce_ms_per_step": R["red_ms"] / steps, "wall_ms_per_step": R["wall_ms"] / steps, "clocks": R["clk"], "roofline": roof}
    if name == "cornell_1080p_256spp_d8":
        out["parity_note"] = ("reference CORNELL is the one preset with a documented deviation: its three wall quads are coplanar "
                              "(scene.cpp:342-349), ties go to (t, prim_id) here and to rounded |o-P|^2 there: 0.6 % of pixels take "
                              "another discrete decision, mean radiance 0.2947 vs 0.3010/0.3070 for the reference's two CUDA backends")
    return out


def fast_math_parity(name, spp=8):
    """Per-pixel metrics of fast_math=1 against the bit-exact mode on the workload's own frame (same seed, `spp` samples)."""
    from parallelraytracing_b200.capi import Context
    sd, cam, w = make_workload(name)
    imgs = []
    with Context(int(os.environ.get("LOCAL_RANK", "0"))) as c:
        c.set_scene(sd); c.build_bvh(); c.set_camera(cam)
        for fast in (0, 1):
            c.set_option("fast_math", fast)
            c.clear_accum(); c.render(spp, w["depth"], seed=4242)
            imgs.append(c.read_accum().astype(np.float64) / spp)
    d = np.abs(imgs[0] - imgs[1])
    rel = d / np.maximum(np.abs(imgs[0]), 1e-3)
    flip = rel.max(-1) > 1e-2
    return {"spp": spp, "rmse": float(np.sqrt((d ** 2).mean())), "rmse_nonflip": float(np.sqrt((d[~flip] ** 2).mean())),
            "max_rel_nonflip": float(rel[~flip].max()), "flip_pixels": int(flip.sum()), "pixels": int(flip.size),
            "tolerance": "north star: per-pixel RMSE <= 1e-3, max relative error <= 1e-2 (flips = pixels where a path took another discrete decision)"}


def emit(line, fd):
    os.write(fd, (line + "\n").encode())


def main():
    # stdout carries exactly ONE JSON line: anything libraries print there (e.g. "NCCL version ..." from rank 0)
    # is sent to stderr by pointing fd 1 at fd 2 for the duration of the run
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="measure this workload alone (default: %s + the mesh workloads of BASELINE.json)" % DEFAULT_WORKLOAD)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="default run without the extra mesh workloads")
    ap.add_argument("--extra-steps", type=int, default=5, help="timed steps of each extra workload (<= --steps)")
    ap.add_argument("--inproc-gpus", type=int, default=0, help="ONE process drives this many GPUs through prt_create_multi (no torchrun)")
    ap.add_argument("--partition", default="samples", choices=["samples", "tiles"],
                    help="multi-GPU split: sample ranges (default) or bands of rows (strong-scaling workloads only)")
    ap.add_argument("--opt", action="append", default=[], help="prt_set_option key=value (e.g. fused=0, wave_paths=8388608)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    default_run = args.workload is None
    if default_run:
        args.workload = DEFAULT_WORKLOAD
    if args.impl == "reference":
        return run_reference_arm(args, out_fd)

    import torch
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gloo = None
    if world > 1:
        gloo = dist.new_group(backend="gloo")                            # CPU-side barrier for the in-process leg below
    if args.inproc_gpus > 1:                                             # manual: one process drives N GPUs, nothing else
        out = measure(args, args.workload, args.steps, args.warmup, with_clocks=True, devices=list(range(args.inproc_gpus)))
        emit(json.dumps(out), out_fd)
        return 0
    out = measure(args, args.workload, args.steps, args.warmup, with_clocks=True)
    if default_run and not args.no_extra and args.partition == "samples":
        extra = {}
        names = EXTRA_WORKLOADS + (EXTRA_WORKLOADS_MULTI if world > 1 else [])
        for name in names:
            k = max(1, min(args.steps, args.extra_steps))
            if WORKLOADS[name]["spp"] >= 1024:
                k = min(k, 3)                                            # seconds per step at 2 GPUs
            o = measure(args, name, k, 3)
            if rank == 0:
                extra[name] = o
        if world > 1:
            # the same strong-scaling workload with ONE process driving all N GPUs through the C ABI (prt_create_multi): what a
            # maintainer who adds B200Renderer to main.cpp gets. The other ranks have released their contexts and wait on the CPU.
            name = EXTRA_WORKLOADS_MULTI[0]
            if rank == 0:
                try:
                    extra[name]["inproc"] = measure(args, name, 3, 3, devices=list(range(world)))
                except Exception as e:  # noqa: BLE001
                    extra[name]["inproc"] = {"error": str(e)[:300]}
            dist.barrier(group=gloo)
        if rank == 0:
            out["workloads"] = extra
    if default_run and world == 1 and not args.no_extra:
        # tolerance-mode float math (option fast_math, default off): the same workload, and its parity against the bit-exact mode
        import copy
        a2 = copy.copy(args); a2.opt = list(args.opt) + ["fast_math=1"]
        fm = measure(a2, args.workload, max(1, min(args.steps, args.extra_steps)), 3)
        out["fast_math"] = {"value": fm["value"], "unit": "Mrays/s", "ms_per_step": fm["ms_per_step"], "e2e": fm["e2e"]["value"],
                            "speedup_vs_exact": fm["value"] / out["value"], "parity_vs_exact": fast_math_parity(args.workload),
                            "what": "SFU approximations for divide / sqrt / rsqrt on the shading side only; hit ids unchanged; default OFF"}
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            arm = CpuArm(args.workload)
            spp, rows = arm.calibrate(12.0)
            dt, rays = arm.run(spp, rows)
            out["cpu_baseline"] = {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": arm.cores, "kind": arm.kind,
                                   "sample": arm.describe(spp, rows), "seconds": dt}
            shipped = arm.as_shipped()
            if shipped:
                out["cpu_baseline"]["as_shipped"] = shipped
            refgpu = arm.reference_gpu()
            if refgpu:
                out["reference_cuda_backends"] = refgpu
        else:
            out["cpu_baseline"] = None
        sys.stdout.flush()
        emit(json.dumps(out), out_fd)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
