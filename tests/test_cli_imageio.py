"""Headless driver + image writers. CPU: file formats and scene selection;
GPU (-m gpu): the CLI end to end, its PNG equals Film.UpdateDisplay of the same render."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from parallelraytracing_b200 import cli, imageio
from parallelraytracing_b200 import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_png_and_ppm_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, size=(37, 53, 4), dtype=np.uint8)
    imageio.write_png(tmp_path / "a.png", img)
    assert np.array_equal(imageio.read_png_rgb8(tmp_path / "a.png"), img[:, :, :3])
    imageio.write_image(tmp_path / "a.ppm", img)
    raw = open(tmp_path / "a.ppm", "rb").read()
    assert raw.startswith(b"P6\n53 37\n255\n") and raw[len(b"P6\n53 37\n255\n"):] == img[:, :, :3].tobytes()
    with pytest.raises(ValueError):
        imageio.write_png(tmp_path / "b.png", img.astype(np.float32))


def test_cli_scene_selection():
    sd, cam = cli.build_scene("cornell")
    assert sd.n_prims == S.cornell().n_prims and cam is None
    sd, cam = cli.build_scene("closed_cornell")
    assert sd.n_prims == 8
    sd, cam = cli.build_scene("soup:1000")
    assert sd.n_tris == 1000 and cam is not None
    with pytest.raises(SystemExit):
        cli.build_scene("no_such_scene")


def test_cli_fails_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(SystemExit):
        cli.main(["--scene", "cornell", "--width", "32", "--height", "32", "--frames", "1"])


@pytest.mark.gpu
def test_cli_end_to_end(tmp_path):
    out, raw = str(tmp_path / "c.png"), str(tmp_path / "c.npy")
    r = subprocess.run([sys.executable, "-m", "parallelraytracing_b200.cli", "--scene", "closed_cornell", "--width", "96",
                        "--height", "64", "--frames", "3", "--spp-per-frame", "2", "--depth", "8", "--seed", "5",
                        "--out", out, "--raw", raw], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    j = json.loads(r.stdout.strip().splitlines()[-1])
    assert j["spp_total"] == 6 and j["prims"] == 8 and j["rays_rank0"] > 96 * 64 * 6
    # the same render through the Renderer mirror
    from parallelraytracing_b200 import B200Renderer, Film
    film, rr = Film(96, 64), B200Renderer(max_depth=8, samples_per_call=2, seed=5)
    rr.Init(film, S.closed_cornell(), S.Camera(width=96, height=64)); film.Clear()
    for _ in range(3):
        rr.ProgressiveRender()
    assert np.array_equal(np.load(raw), film.ReadAccum())
    assert np.array_equal(imageio.read_png_rgb8(out), film.UpdateDisplay()[:, :, :3])
    rr.close()


@pytest.mark.gpu
def test_cli_build_time_options_take_effect():
    """--opt is applied BEFORE Init builds the BVH (advisor finding, round 1: build-time options were silently ignored)."""
    out = {}
    for opt in ("bvh_builder=1", "bvh_builder=0"):
        r = subprocess.run([sys.executable, "-m", "parallelraytracing_b200.cli", "--scene", "random_balls_small", "--width", "64",
                            "--height", "48", "--frames", "2", "--depth", "6", "--opt", opt], cwd=ROOT, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        out[opt] = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["bvh_builder=1"]["bvh_builder"] == "ploc" and out["bvh_builder=0"]["bvh_builder"] == "karras"
    assert out["bvh_builder=1"]["mean_radiance"] == out["bvh_builder=0"]["mean_radiance"]      # the tree never changes the image


def test_add_meshes_host_mirror():
    """scene.add_meshes = the host mirror of B200Renderer::AddMesh: triangles after the scene's own
    primitives, one new material per mesh, flat unit normals when the mesh has none, offsets applied in fp32."""
    from parallelraytracing_b200 import scene as S
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)
    f = np.array([[0, 1, 2], [0, 1, 3], [0, 2, 3], [1, 2, 3]], np.uint32)
    base = S.make_scene(S.ScenePreset.MATERIAL_TEST)
    sd = S.add_meshes(base, [((v, None, f), (S.MAT_LAMBERTIAN, (0.8, 0.3, 0.3), 0.0), (0, 1, 2)),
                             ((v, None, f), (S.MAT_METAL, (0.9, 0.9, 0.9), 0.05), (1.5, 1, 2))])
    assert sd.n_prims == base.n_prims + 8 and sd.tri_prim_base == base.n_prims and sd.n_tris == 8
    assert len(sd.materials) == len(base.materials) + 2 and list(sd.tri_material) == [4] * 4 + [5] * 4
    assert np.array_equal(sd.tri_verts[0].reshape(3, 3), v[f[0]] + np.array([0, 1, 2], np.float32))
    nrm = sd.tri_normals.reshape(-1, 3, 3)
    assert np.allclose(np.linalg.norm(nrm, axis=-1), 1.0, atol=1e-6) and np.allclose(nrm[0], [[0, 0, 1]] * 3)
    assert float(sd.materials[5]["param"]) == float(np.float32(0.05)) and int(sd.materials[5]["type"]) == S.MAT_METAL
    with pytest.raises(ValueError):
        S.add_meshes(sd, [((v, None, f), (S.MAT_LAMBERTIAN, (1, 1, 1), 0.0), (0, 0, 0))])
