"""prt_trace_batch_static (include/prt.h): the closest-hit batch with its scratch in a static device array - no cudaMalloc,
no cudaFree. Called through the C ABI with device pointers and compared, bit for bit (the tolerance every closest-hit test
here uses), with prt_trace_batch on the same rays, on calls in a row that share the array. Collected last on purpose
(`zz`). Works alone in a fresh process and behind the other tests."""
import numpy as np
import pytest

import util
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context, PrtError, hit_dt

pytestmark = pytest.mark.gpu
MAX_RAYS = 65536          # PRT_TRACE_STATIC_MAX_RAYS


def header_max_rays():
    import os
    import re
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "prt.h")).read()
    return int(re.search(r"#define\s+PRT_TRACE_STATIC_MAX_RAYS\s+(\d+)", src).group(1))


def to_device(rays):
    import torch
    return torch.from_numpy(np.ascontiguousarray(rays).view(np.uint8).copy()).to("cuda:0")


def static_trace(ctx, rays, out_dev, use_bvh, stream=None):
    """Rays to the device with torch (the caller's memory), the entry point, the records back as hit_dt."""
    import torch
    d_rays = to_device(rays)
    torch.cuda.synchronize()
    out_dev.fill_(0xAB)
    torch.cuda.synchronize()
    ctx.trace_batch_static(d_rays.data_ptr(), len(rays), out_dev.data_ptr(), use_bvh, stream)
    return out_dev[: len(rays) * hit_dt.itemsize].cpu().numpy().view(hit_dt).copy()


def test_static_scratch_matches_trace_batch_two_calls_in_a_row():
    import torch
    assert header_max_rays() == MAX_RAYS
    ctx = Context(0)
    try:
        out_dev = torch.empty(MAX_RAYS * hit_dt.itemsize, dtype=torch.uint8, device="cuda:0")
        sd = S.random_triangle_soup(20000)
        ctx.set_scene(sd); ctx.build_bvh()
        ctx.trace_batch_static(0, 0, 0, 2)                                    # n = 0: the once-per-device set-up alone
        first = util.fixed_ray_batch(sd, n=MAX_RAYS, seed=11)                 # fills the whole array
        second = util.fixed_ray_batch(sd, n=20001, seed=12)                   # a shorter, ragged batch over the first one's leftovers
        want_first = {m: ctx.trace_batch(first, use_bvh=m) for m in (0, 1, 2)}
        want_second = {m: ctx.trace_batch(second, use_bvh=m) for m in (0, 1, 2)}
        assert (want_first[2]["prim_id"] >= 0).sum() > 1000
        for mode in (2, 1, 0):
            got1 = static_trace(ctx, first, out_dev, mode)
            got2 = static_trace(ctx, second, out_dev, mode)                   # the second call in a row: same array, last call's contents
            assert got1.tobytes() == want_first[mode].tobytes(), "first call, use_bvh=%d" % mode
            assert got2.tobytes() == want_second[mode].tobytes(), "second call, use_bvh=%d" % mode

        # another scene (spheres + quads: the two-child and the list walks), the caller's own stream, and three calls back to back
        sd2 = S.make_scene(S.ScenePreset.RANDOM_BALLS_LARGE)
        ctx.set_scene(sd2); ctx.build_bvh()
        rays = util.fixed_ray_batch(sd2, n=4097, seed=13)
        want = ctx.trace_batch(rays, use_bvh=2)
        st = torch.cuda.Stream(device="cuda:0")
        for _ in range(3):
            assert static_trace(ctx, rays, out_dev, 2, stream=st.cuda_stream).tobytes() == want.tobytes()
        assert np.array_equal(want, ctx.trace_batch(rays, use_bvh=0))

        # nothing is allocated: the driver's free-memory figure does not move across a call (warm: modules are loaded)
        d_rays = to_device(rays)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        for _ in range(4):
            ctx.trace_batch_static(d_rays.data_ptr(), len(rays), out_dev.data_ptr(), 2)
        assert torch.cuda.mem_get_info(0)[0] == free0
    finally:
        ctx.close()


def test_static_scratch_limits_and_errors():
    import torch
    ctx = Context(0)
    try:
        buf = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
        with pytest.raises(PrtError, match="scene/BVH not ready"):
            ctx.trace_batch_static(buf.data_ptr(), 1, buf.data_ptr(), 2)
        sd = S.cornell()
        ctx.set_scene(sd); ctx.build_bvh()
        with pytest.raises(PrtError, match="exceed the static scratch"):      # rejected before anything is launched
            ctx.trace_batch_static(buf.data_ptr(), MAX_RAYS + 1, buf.data_ptr(), 2)
        with pytest.raises(PrtError, match="negative"):
            ctx.trace_batch_static(buf.data_ptr(), -1, buf.data_ptr(), 2)
        with pytest.raises(PrtError, match="NULL"):
            ctx.trace_batch_static(0, 4, buf.data_ptr(), 2)
        with pytest.raises(PrtError, match="use_bvh"):
            ctx.trace_batch_static(buf.data_ptr(), 1, buf.data_ptr(), 7)
        # the context still works after the refusals, and one ray is a valid batch
        rays = util.fixed_ray_batch(sd, n=1, seed=3)
        out = torch.zeros(hit_dt.itemsize, dtype=torch.uint8, device="cuda:0")
        d = to_device(rays)
        ctx.trace_batch_static(d.data_ptr(), 1, out.data_ptr(), 2)
        assert out.cpu().numpy().view(hit_dt).tobytes() == ctx.trace_batch(rays, use_bvh=2).tobytes()
    finally:
        ctx.close()
