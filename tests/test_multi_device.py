"""Multi-GPU inside libprt_b200.so (-m gpu, needs >= 2 devices: `gpurun --gpus 2`): ONE context over several GPUs of the
process (prt_create_multi, SURVEY 8b export list / 8e) - scene + BVH replicated, prt_render splits the call over the devices
and one sum-reduce (ncclReduce, or the peer-to-peer kernel) lands the frame on device 0 with no host sync in between.
The N-device image is the same multiset of samples as the one-device image: equal up to fp32 summation order (sample split)
or bit for bit (row bands: disjoint supports)."""
import os
import subprocess

import numpy as np
import pytest

import util
from parallelraytracing_b200 import scene as S
from parallelraytracing_b200.capi import Context, PrtError

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs")


def _scenes():
    z = np.load(util.GOLDEN + "/reference_vectors.npz")
    dragon = S.mesh_scene((z["mesh_dragon_v"], z["mesh_dragon_n"], z["mesh_dragon_f"]), name="dragon3")
    return [(S.closed_cornell(), S.Camera(width=160, height=96)),
            (dragon, S.Camera(pos=(0.0, 1.0, 5.0), front=S._normalize(S._v3(0, -1, -5)), width=128, height=72))]


def test_multi_context_with_one_device_is_the_single_context():
    """prt_create_multi with ONE device id must behave exactly like prt_create (runs on any box)."""
    sd, cam = S.closed_cornell(), S.Camera(width=96, height=64)
    out = []
    for dev in (0, [0]):
        with Context(dev) as c:
            c.set_scene(sd); c.build_bvh(); c.set_camera(cam); c.clear_accum(); c.render(3, 8, seed=5)
            out.append((c.read_accum(), c.stats().rays_total, c.stats().n_devices))
    assert np.array_equal(out[0][0], out[1][0]) and out[0][1] == out[1][1] and out[0][2] == out[1][2] == 1
    with pytest.raises(PrtError):
        Context([0, 0])


@needs2
@pytest.mark.parametrize("reduce_mode", [1, 2])
def test_two_devices_equal_one_device(reduce_mode):
    for sd, cam in _scenes():
        with Context(0) as one:
            one.set_scene(sd); one.build_bvh(); one.set_camera(cam); one.clear_accum()
            one.render(8, 8, seed=3, sample_begin=2)
            want, st1 = one.read_accum(), one.stats()
            one.render(3, 8, seed=3, sample_begin=10); want11 = one.read_accum()
            g1 = one.debug_lbvh()
        with Context([0, 1]) as two:
            try:
                two.set_option("multi_reduce", reduce_mode)
            except PrtError as e:
                pytest.skip(str(e))
            two.set_scene(sd); two.build_bvh(); two.set_camera(cam); two.clear_accum()
            g2 = two.debug_lbvh()
            for k in g1:
                assert np.array_equal(g1[k], g2[k]), k
            # sample split: device 0 renders samples [2,6), device 1 [6,10); the reduce adds device 1's buffer
            two.render(8, 8, seed=3, sample_begin=2)
            got, st2 = two.read_accum(), two.stats()
            assert st2.n_devices == 2 and st2.rays_total == st1.rays_total
            assert list(st2.rays_per_bounce) == list(st1.rays_per_bounce) and two.accum_spp() == 8
            np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-6)              # fp32 summation order only
            assert st2.reduce_ms > 0.0
            # a second call ADDS (uneven split: 2 + 1 samples)
            two.render(3, 8, seed=3, sample_begin=10)
            np.testing.assert_allclose(two.read_accum(), want11, rtol=2e-6, atol=1e-6)
            assert two.accum_spp() == 11
            # clear + render in one call, one sample: device 1 has nothing to render and contributes zeros
            two.render_fresh(1, 8, seed=3, sample_begin=2)
            one1 = Context(0); one1.set_scene(sd); one1.build_bvh(); one1.set_camera(cam); one1.clear_accum(); one1.render(1, 8, seed=3, sample_begin=2)
            assert np.array_equal(two.read_accum(), one1.read_accum()) and two.accum_spp() == 1
            one1.close()
            # row bands: disjoint supports, so the sum is exact
            two.set_option("multi_partition", 1)
            two.clear_accum(); two.render(8, 8, seed=3, sample_begin=2)
            assert np.array_equal(two.read_accum(), want) and two.stats().rays_total == st1.rays_total
            two.set_option("multi_partition", 0)
            # a pixel region through the multi context
            P = cam.width * cam.height
            two.clear_accum(); two.render_region(8, 8, seed=3, sample_begin=2, pixel_begin=0, pixel_count=P)
            np.testing.assert_allclose(two.read_accum(), want, rtol=2e-6, atol=1e-6)
            img = two.resolve_display(1.0, 2.2)
            assert img.shape == (cam.height, cam.width, 4)


@needs2
def test_cpp_shim_two_devices(tmp_path):
    """B200Renderer::SetDevices({0, 1}) behind the reference's own Renderer / Film objects (oracle/_ref/shim_driver): the
    reference Film's accumulator equals the one-device run up to fp32 summation order."""
    exe = os.path.join(util.ROOT, "oracle", "_ref", "shim_driver")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/shim_driver not built (needs /root/reference at build time)")
    W, H, calls, spp, depth, seed = 160, 90, 3, 4, 12, 3
    imgs = []
    for devs in (None, "0,1"):
        out = str(tmp_path / ("shim_%s.bin" % (devs or "one").replace(",", "_")))
        env = dict(os.environ)
        if devs:
            env["PRT_SHIM_DEVICES"] = devs
        r = subprocess.run([exe, str(int(S.ScenePreset.DEFAULT)), str(W), str(H), str(calls), str(spp), str(depth), str(seed), out],
                           capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0, r.stdout + r.stderr
        raw = np.fromfile(out, np.uint8)
        imgs.append((raw[:W * H * 12].view(np.float32).reshape(H, W, 3), raw[W * H * 12:W * H * 16].view(np.float32)))
    assert np.all(imgs[1][1] == calls * spp)
    np.testing.assert_allclose(imgs[1][0], imgs[0][0], rtol=3e-6, atol=1e-6)
    assert imgs[0][0].max() > 0
