"""Multi-GPU combine inside the library (include/b200pt.h "multi-GPU combine"): NCCL communicators owned by
the renderers, tile shares reduced to one image bit-identical to the one-GPU image, frame shares reduced to the mean.
The single-rank cases run on any GPU box; the two-device cases are skipped when the box has one GPU."""
import os
import subprocess

import numpy as np
import pytest

import b200pt
from parity_util import bits_equal
from tools import scene as S

pytestmark = pytest.mark.gpu
T = b200pt.PT_PARTITION_TILES


def two_gpus():
    return b200pt.device_count() >= 2


def test_nccl_is_opened_at_run_time():
    assert b200pt.comm_available() >= 20000          # version code of libnccl.so.2 (2.x)


def test_single_rank_communicator_reduce_is_the_identity():
    ps = S.cornell().pack()
    W, H = 160, 96
    r = b200pt.Renderer(ps, W, H, 2, 3)
    with pytest.raises(b200pt.PtError):
        r.reduce_image(0)                            # no communicator yet
    r.comm_init(0, 1, b200pt.comm_unique_id())
    cam = S.camera_default(W, H, 1)
    r.record(cam); r.sync()
    want = r.image()
    r.record(cam)
    r.reduce_image(0)                                # on the communication stream, behind the frame
    assert bits_equal(r.image(), want).all()         # read-backs follow the reduce
    r.record(S.camera_default(W, H, 2))              # the next frame waits for the reduce before it writes the image
    r.sync()
    assert not bits_equal(r.image(), want).all()
    r.set_accumulate(True)
    for f in range(3):
        r.record(S.camera_default(W, H, f))
    r.reduce_accum(0)
    mean = r.accum_mean()
    assert np.isfinite(mean[..., :3]).all() and (mean[..., 3] == 1.0).all()
    with pytest.raises(b200pt.PtError):
        r.reduce_image(1)                            # root outside the communicator
    s = r.sibling()
    s.record(cam); s.reduce_image(0); s.sync()       # a sibling uses its parent's communicator
    assert bits_equal(s.image(), want).all()
    s.deinit(); r.deinit()


@pytest.mark.skipif(not two_gpus(), reason="needs two GPUs")
def test_two_gpu_tile_shares_reduce_to_the_single_gpu_image():
    ps = S.atrium(1234, 20000, 6, 64).pack()
    W, H, spp, seg = 500, 300, 2, 4                  # ragged edge tiles
    cam = S.camera_default(W, H, 3)
    one = b200pt.Renderer(ps, W, H, spp, seg)
    one.record(cam); one.sync(); want = one.image(); one.deinit()
    rs = [b200pt.Renderer(ps, W, H, spp, seg, device=g, rank=g, world=2, partition=T) for g in range(2)]
    b200pt.comm_init_all(rs)
    for rep in range(2):
        for r in rs:
            r.record(cam)
        with b200pt.comm_group():
            for r in rs:
                r.reduce_image(0)
        assert bits_equal(rs[0].image(), want).all(), f"repetition {rep}"
    for r in rs:
        r.deinit()


@pytest.mark.skipif(not two_gpus(), reason="needs two GPUs")
def test_two_gpu_frame_shares_reduce_to_the_one_gpu_accumulation_mean():
    """rank r renders frames r, r + N, ...; the reduced accumulation buffer divided by the total frame count
    equals the N-frame accumulate mean on one GPU (two ranks: a + b is commutative, so bit for bit)."""
    ps = S.cornell().pack()
    W, H = 128, 128
    one = b200pt.Renderer(ps, W, H, 1, 3)
    one.set_accumulate(True)
    for f in range(2):
        one.record(S.camera_default(W, H, f))
    want = one.accum_mean(); one.deinit()
    rs = [b200pt.Renderer(ps, W, H, 1, 3, device=g) for g in range(2)]
    b200pt.comm_init_all(rs)
    for g, r in enumerate(rs):
        r.set_accumulate(True)
        r.record(S.camera_default(W, H, g))
    with b200pt.comm_group():
        for r in rs:
            r.reduce_accum(0)
    assert bits_equal(rs[0].accum_mean(total_frames=2), want).all()
    for r in rs:
        r.deinit()


def _cli(tmp_path, name, *extra):
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vulkan-pathtracer_b200", "b200pt")
    out = tmp_path / name
    r = subprocess.run([exe, "-s", str(tmp_path / "cornell.glb"), "-x", "200", "-y", "120", "-c", "2", "-b", "3", "--frames", "3", "-o", str(out),
                        "--quiet", *extra], capture_output=True, text=True, cwd=tmp_path, timeout=600)
    assert r.returncode == 0, r.stderr
    return out.read_bytes()


def test_cli_gpus_flag(tmp_path):
    cwd = os.getcwd(); os.chdir(tmp_path)
    try:
        S.cornell(1234).write_gltf(str(tmp_path), "cornell", glb=True)
    finally:
        os.chdir(cwd)
    a = _cli(tmp_path, "a.pfm")
    assert _cli(tmp_path, "b.pfm", "--gpus", "1") == a
    acc = _cli(tmp_path, "c.pfm", "--accumulate")
    if two_gpus():
        assert _cli(tmp_path, "d.pfm", "--gpus", "2") == a                       # tile shares + NCCL reduce: the same bytes
        assert _cli(tmp_path, "e.pfm", "--gpus", "2", "--accumulate") == acc
