"""The b200pt command keeps the reference's CLI surface (src/main.zig:29-84)."""
import os
import subprocess

import numpy as np
import pytest

import b200pt
from tools import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "vulkan-pathtracer_b200", "b200pt")


def run(*args, cwd=None):
    return subprocess.run([EXE, *args], capture_output=True, text=True, cwd=cwd, timeout=300)


def test_help_exits_zero():
    r = run("--help")
    assert r.returncode == 0 and "--scene-path" in r.stdout and "--num-bounces" in r.stdout     # error.Help -> exit(0), main.zig:75


def test_parse_error_exits_two():
    assert run("--no-such-flag", "1").returncode == 2                                            # error.Parser -> exit(2), main.zig:76
    assert run("--num-samples", "many", "-s", "x.glb").returncode == 2
    assert run("--num-samples").returncode == 2


def test_missing_scene_path_prints_reference_message_and_returns():
    r = run("--num-samples", "4")
    assert r.returncode == 0 and r.stderr == "Missing path to scene from arguments\n"            # main.zig:81-84


def test_loader_errors_surface_with_zig_names(tmp_path):
    r = run("-s", str(tmp_path / "scene.obj"))
    assert r.returncode == 1 and "IoError" in r.stderr
    (tmp_path / "scene.obj").write_text("o")
    r = run("-s", str(tmp_path / "scene.obj"))
    assert r.returncode == 1 and "InvalidFileExtension" in r.stderr


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(v) for v in f.readline().split())
        assert float(f.readline()) < 0
        a = np.frombuffer(f.read(), "<f4").reshape(h, w, 3)
    return a[::-1]


def write_scene(sb, tmp_path, glb):
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        return sb.write_gltf(str(tmp_path), "cornell", glb=glb)
    finally:
        os.chdir(cwd)


@pytest.mark.gpu
@pytest.mark.parametrize("glb", [False, True])
def test_cli_renders_the_same_image_as_the_library(tmp_path, glb):
    sb = S.cornell()
    path = write_scene(sb, tmp_path, glb)
    out = str(tmp_path / "out.pfm")
    r = run("-s", path, "-x", "96", "-y", "64", "-c", "2", "-b", "3", "-f", "2", "-o", out, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert "Mrays/s" in r.stdout
    img = np.ascontiguousarray(read_pfm(out))
    ren = b200pt.Renderer(sb.pack(), 96, 64, 2, 3)
    ren.record(S.camera_default(96, 64, 1)); ren.sync()          # the CLI's last frame has frame_count 1
    want = np.ascontiguousarray(ren.image()[..., :3])
    ren.deinit()
    assert (img.view(np.uint32) == want.view(np.uint32)).all()


@pytest.mark.gpu
def test_cli_divider_exr_and_bgra(tmp_path):
    path = write_scene(S.cornell(), tmp_path, True)
    r = run("-s", path, "-x", "64", "-y", "48", "-d", "2", "-o", str(tmp_path / "o.exr"), "--quiet", cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert os.path.getsize(tmp_path / "o.exr") > 64 * 48 * 16          # full-resolution (up-scaled) float image
    r = run("-s", path, "-x", "64", "-y", "48", "-o", str(tmp_path / "o.bgra"), "--accumulate", "-f", "3", cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert os.path.getsize(tmp_path / "o.bgra") == 64 * 48 * 4
