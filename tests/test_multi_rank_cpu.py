"""Multi-rank partitioning on CPU (gloo, world_size 2): interleaved tile ownership is a partition of
the image and the sum-reduce of the per-rank images equals the single-rank image bit for bit;
frame (sample-range) partitioning sums to the mean over frames."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    from oracle import orc
    from tools import scene as S
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ps = S.cornell().pack()
    osc = orc.OracleScene(ps, True)
    W, H = 96, 40
    cam = S.camera_default(W, H, 2)
    # tiles: rank renders only the tiles it owns (owner = tile % world), everything else stays 0
    cfg = S.render_config(W, H, 2, 3, rank=rank, world=world, partition=1)
    img, _, rays = osc.render(cfg, cam, True, nthreads=1)
    t = torch.from_numpy(img.copy())
    owned = torch.from_numpy((img[..., 3] != 0).astype(np.int32))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    dist.all_reduce(owned, op=dist.ReduceOp.SUM)
    # frames: rank renders frame_count = rank; reduce(sum)/world is the mean over frames
    cfg1 = S.render_config(W, H, 1, 3)
    f, _, _ = osc.render(cfg1, S.camera_default(W, H, rank), True, nthreads=1)
    ft = torch.from_numpy(f.copy())
    dist.reduce(ft, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        full, _, _ = osc.render(S.render_config(W, H, 2, 3), cam, True, nthreads=1)
        frames = [osc.render(cfg1, S.camera_default(W, H, r), True, nthreads=1)[0] for r in range(world)]
        np.savez(out, reduced=t.numpy(), owned=owned.numpy(), full=full, fsum=ft.numpy(), fref=sum(frames[1:], frames[0]))
    dist.destroy_process_group()


def test_tile_and_frame_partition_world2(tmp_path):
    out = str(tmp_path / "r.npz")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    g = np.load(out)
    assert (g["owned"] == 1).all()                                              # every pixel owned exactly once
    assert np.array_equal(g["reduced"].view(np.uint32), g["full"].view(np.uint32))   # bit-exact vs 1 rank
    assert np.array_equal(g["fsum"].view(np.uint32), g["fref"].view(np.uint32))
