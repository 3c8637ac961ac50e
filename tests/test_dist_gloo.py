"""world_size-2 (and 3) gloo tests of the multi-GPU host logic on CPU: the sample-range partition and
the single sum-reduce (parallelraytracing_b200/dist.py). The per-rank renderer is the CPU oracle here
(test infrastructure standing in for the CUDA context, which needs a GPU): the reduced image must
equal the single-rank render of the same samples up to fp32 summation order."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sample_range_partition():
    from parallelraytracing_b200.dist import sample_range
    for world in (1, 2, 3, 4, 8):
        for spp in (0, 1, 5, 8, 256, 1023):
            got = [sample_range(r, world, spp, 10) for r in range(world)]
            assert sum(c for _, c in got) == spp
            pos = 10
            for b, c in got:
                assert b == pos and c >= 0
                pos += c
            assert max(c for _, c in got) - min(c for _, c in got) <= 1


def _worker(rank, world, port, spp, out_path):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_lib
    from parallelraytracing_b200 import scene as S
    from parallelraytracing_b200.dist import render_sharded
    sd, cam = S.closed_cornell(), S.Camera(width=48, height=32)
    sc = oracle_lib.load_oracle().scene(sd)
    accum = torch.zeros(cam.height * cam.width * 3, dtype=torch.float32)

    def render_fn(begin, count):
        a = accum.numpy().reshape(cam.height, cam.width, 3)
        sc.render(cam, count, 8, seed=3, sample_begin=begin, accum=a, n_threads=2)
    begin, count = render_sharded(render_fn, accum, spp)
    if rank == 0:
        np.save(out_path, accum.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,spp", [(2, 6), (3, 7)])
def test_sharded_render_equals_single(tmp_path, oracle, world, spp):
    from parallelraytracing_b200 import scene as S
    out = str(tmp_path / "img.npy")
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, spp, out), nprocs=world, join=True)
    got = np.load(out).reshape(32, 48, 3)
    sd, cam = S.closed_cornell(), S.Camera(width=48, height=32)
    want, _ = oracle.scene(sd).render(cam, spp, 8, seed=3)
    # same multiset of samples; only the fp32 summation order differs (<= 1 ulp per add)
    assert np.allclose(got, want, rtol=1e-6, atol=1e-6)
    assert np.abs(got - want).max() <= 4 * np.finfo(np.float32).eps * np.abs(want).max()


def test_pixel_range_partition():
    from parallelraytracing_b200.dist import pixel_range
    for world in (1, 2, 3, 8):
        for (w, h) in ((48, 32), (1920, 1080), (7, 5)):
            got = [pixel_range(r, world, w * h, w) for r in range(world)]
            assert sum(c for _, c in got) == w * h
            pos = 0
            for b, c in got:
                assert b == pos and c % w == 0          # whole rows
                pos += c
            assert max(c for _, c in got) - min(c for _, c in got) <= w


def _tile_worker(rank, world, port, spp, out_path):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_lib
    from parallelraytracing_b200 import scene as S
    from parallelraytracing_b200.dist import render_tiled
    sd, cam = S.closed_cornell(), S.Camera(width=48, height=32)
    sc = oracle_lib.load_oracle().scene(sd)
    accum = torch.zeros(cam.height * cam.width * 3, dtype=torch.float32)

    def render_fn(pixel_begin, pixel_count):               # a band of whole rows
        a = accum.numpy().reshape(cam.height, cam.width, 3)
        sc.render(cam, spp, 8, seed=3, accum=a, n_threads=2, y_begin=pixel_begin // cam.width,
                  y_end=(pixel_begin + pixel_count) // cam.width)
    render_tiled(render_fn, accum, cam.width * cam.height, cam.width)
    if rank == 0:
        np.save(out_path, accum.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,spp", [(2, 4), (3, 3)])
def test_tiled_render_equals_single(tmp_path, oracle, world, spp):
    """TILE partition: disjoint pixel bands + the same sum-reduce give EXACTLY the single-rank image."""
    from parallelraytracing_b200 import scene as S
    out = str(tmp_path / "img.npy")
    port = 31500 + (os.getpid() % 2000) + world
    mp.spawn(_tile_worker, args=(world, port, spp, out), nprocs=world, join=True)
    got = np.load(out).reshape(32, 48, 3)
    want, _ = oracle.scene(S.closed_cornell()).render(S.Camera(width=48, height=32), spp, 8, seed=3)
    assert np.array_equal(got, want)
