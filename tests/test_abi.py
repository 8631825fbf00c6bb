"""The C-ABI shared library: loads on a CPU-only box, exports every symbol include/b200pt.h
declares, fails loudly without a GPU (no CPU fallback), and its pure-host entry points work."""
import ctypes as C
import math
import os
import re

import numpy as np
import pytest

import b200pt
from tools import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("pt_create", "pt_render", "pt_sync", "pt_resize", "pt_destroy", "pt_scene_load", "pt_camera_new",
                 "pt_read_rgba32f", "pt_read_ids", "pt_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    L = b200pt.lib()
    missing = [n for n in declared_functions() if not hasattr(L, n)]
    assert not missing, f"declared in include/b200pt.h but not exported: {missing}"


def test_struct_layouts_match_header():
    assert C.sizeof(S.PtPrimitive) == 56 and C.sizeof(S.PtMesh) == 8 and C.sizeof(S.PtMaterial) == 28
    assert C.sizeof(S.PtInstance) == 72 and C.sizeof(S.PtCamera) == 68 and C.sizeof(S.PtRenderConfig) == 40
    assert C.sizeof(S.PtTexture) == 16 and C.sizeof(S.PtSceneDesc) == 96
    assert b200pt.PtHit_dtype.itemsize == 24 and b200pt.PtRay_dtype.itemsize == 32


def _has_gpu():
    return b200pt.lib().pt_device_count() > 0


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU error path")
def test_create_fails_loudly_without_gpu():
    ps = S.cornell().pack()
    with pytest.raises(b200pt.PtError) as e:
        b200pt.Renderer(ps, 64, 64)
    assert e.value.status == -2 and "no CPU fallback" in e.value.message


def test_argument_errors_do_not_crash():
    L = b200pt.lib()
    assert L.pt_create(None, None, None) == -1
    assert b"null" in L.pt_last_error()
    assert L.pt_render(None, None, None) == -1
    assert L.pt_sync(None) == -1
    L.pt_destroy(None)
    assert L.pt_status_name(-16) == b"GltfNoTangents"
    assert L.pt_status_name(-22) == b"TooManyScenes"


def test_camera_matches_restated_camera_zig():
    for (w, h) in ((1920, 1080), (512, 512), (3840, 2160)):
        a = b200pt.camera_new(np.float32(math.pi * 0.25), np.float32(w) / np.float32(h), frame_count=7)
        b = S.camera_default(w, h, 7)
        for f in ("position", "horizontal", "vertical", "forward"):
            assert np.allclose(list(getattr(a, f)), list(getattr(b, f)), rtol=0, atol=1e-7), f
        assert a.frame_count == 7
    # start-up state (main.zig:220-224): forward +z, left -x, up +y; doubled field of view (quirk Q2)
    c = b200pt.camera_new(math.pi / 4, 1.0)
    assert list(c.forward)[:3] == [0, 0, 1]
    assert abs(c.vertical[1] - 2 * math.tan(math.pi / 8)) < 1e-6 and c.horizontal[0] < 0
    # yaw rotates about +y
    y = b200pt.camera_new(math.pi / 4, 1.0, yaw=90.0)
    assert np.allclose(list(y.forward)[:3], [1, 0, 0], atol=1e-6)
    y2 = S.camera_default(100, 100, 0, yaw=90.0)
    assert np.allclose(list(y.forward), list(y2.forward), atol=1e-6) and np.allclose(list(y.horizontal), list(y2.horizontal), atol=1e-6)


def test_pfm_and_exr_writers(tmp_path):
    rng = np.random.default_rng(0)
    img = rng.random((5, 7, 4)).astype(np.float32)
    p = str(tmp_path / "a.pfm")
    b200pt.write_pfm(p, img)
    raw = open(p, "rb").read()
    header, rest = raw.split(b"-1.0\n", 1)
    assert header == b"PF\n7 5\n"
    data = np.frombuffer(rest, "<f4").reshape(5, 7, 3)
    assert np.array_equal(data[::-1], img[..., :3])      # bottom row first
    e = str(tmp_path / "a.exr")
    b200pt.write_exr(e, img)
    raw = open(e, "rb").read()
    assert raw[:4] == bytes([0x76, 0x2f, 0x31, 0x01]) and b"channels" in raw and b"dataWindow" in raw
    # last scanline block: y, size, then A,B,G,R planes
    row = np.frombuffer(raw[-7 * 16:], "<f4").reshape(4, 7)
    assert np.array_equal(row[3], img[4, :, 0]) and np.array_equal(row[0], img[4, :, 3])
    cv2 = pytest.importorskip("cv2")
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    try:
        back = cv2.imread(e, cv2.IMREAD_UNCHANGED)
    except Exception:
        back = None
    if back is not None and back.ndim == 3:
        assert np.allclose(back[..., :3][..., ::-1], img[..., :3])


def test_new_round_one_entry_points_reject_null_arguments():
    """Every entry point added for SURVEY 8(f) returns a status (no crash) on null arguments."""
    L = b200pt.lib()
    cam = b200pt.PtCameraState()
    assert L.pt_camera_state_new(0.5, 1.0, None, C.byref(cam)) == -1
    assert L.pt_camera_update(None, None) == -1
    assert L.pt_camera_update(C.byref(cam), None) == -1
    assert L.pt_camera_update_aspect_ratio(None, 1.0) == -1
    assert L.pt_camera_pack(None, 0, None) == -1
    assert L.pt_set_instance_transforms(None, None, 0, None) == -1
    assert L.pt_debug_bvh_digest(None, None) == -1
    assert L.pt_read_rgba32f_overlapped(None, None) == -1
    assert L.pt_create_sibling(None, None) == -1
    assert L.pt_comm_init(None, 0, 1, None) == -1 and L.pt_comm_init_all(None, 0) == -1
    assert L.pt_reduce_image(None, 0, None) == -1 and L.pt_reduce_accum(None, 0, None) == -1
    assert L.pt_add_accum(None, None) == -1 and L.pt_comm_unique_id(None) == -1
    assert L.pt_debug_read_bandwidth(0, C.c_size_t(1 << 20), 1, None, None) == -1
    assert C.sizeof(b200pt.PtCameraState) == 26 * 4 and C.sizeof(b200pt.PtInput) == 20


def test_create_from_gltf_surfaces_loader_errors_with_zig_names(tmp_path):
    """pt_create_from_gltf = Scene.load + init: the loader's error wins, no renderer is returned."""
    with pytest.raises(b200pt.PtError) as e:
        b200pt.Renderer.from_gltf(str(tmp_path / "missing.glb"), 64, 64)
    assert e.value.name == "IoError"
    (tmp_path / "scene.obj").write_text("o")
    with pytest.raises(b200pt.PtError) as e:
        b200pt.Renderer.from_gltf(str(tmp_path / "scene.obj"), 64, 64)
    assert e.value.name == "InvalidFileExtension"
    assert b200pt.lib().pt_create_from_gltf(None, None, None) == -1
