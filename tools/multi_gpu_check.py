"""Multi-GPU tile partition check (run under torchrun on N GPUs of one box):
the N-rank image — NCCL sum-reduce of the per-rank tile images, and peer stores into rank 0's buffer over
CUDA IPC / NVLink — must equal the single-GPU image bit for bit.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/multi_gpu_check.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt  # noqa: E402
from tools import scene as S  # noqa: E402


class DevArray:
    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (ptr, False), "version": 2}


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    for name, sb, W, H, spp, seg in (("cornell", S.cornell(), 520, 300, 2, 3), ("atrium-20k", S.atrium(1234, 20000, 6, 64), 640, 360, 2, 4)):
        ps = sb.pack()
        cam = S.camera_default(W, H, 3)
        full = None
        if rank == 0:
            r1 = b200pt.Renderer(ps, W, H, spp, seg, device=local)
            r1.record(cam); r1.sync(); full = r1.image(); r1.deinit()
        stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
        sp = C.c_void_p(stream.cuda_stream)
        # (a) sum-reduce of zero-initialised tile images
        r = b200pt.Renderer(ps, W, H, spp, seg, device=local, rank=rank, world=world, partition=b200pt.PT_PARTITION_TILES)
        img = torch.as_tensor(DevArray(r.device_image_ptr(), (H, W, 4)), device=f"cuda:{local}")
        r.record(cam, sp)
        dist.reduce(img, dst=0, op=dist.ReduceOp.SUM)
        torch.cuda.synchronize()
        if rank == 0:
            same = np.array_equal(img.cpu().numpy().view(np.uint32), full.view(np.uint32))
            print(f"{name}: {world}-rank tiles + NCCL reduce == 1-rank image: {same}"); ok &= same
        r.deinit()
        # (b) peer stores into rank 0's image
        r = b200pt.Renderer(ps, W, H, spp, seg, device=local, rank=rank, world=world, partition=b200pt.PT_PARTITION_TILES)
        handle = torch.zeros(64, dtype=torch.uint8, device="cuda")
        if rank == 0:
            hb = (C.c_uint8 * 64)()
            b200pt.check(b200pt.lib().pt_ipc_export_image(r.handle, hb))
            handle.copy_(torch.tensor(list(hb), dtype=torch.uint8))
        dist.broadcast(handle, src=0)
        if rank != 0:
            hb = (C.c_uint8 * 64)(*handle.cpu().tolist())
            b200pt.check(b200pt.lib().pt_ipc_open_root_image(r.handle, hb))
        fence = torch.zeros(1, dtype=torch.int32, device="cuda")
        for frame in (7, 3):      # two frames: the second must fully overwrite the first
            r.record(S.camera_default(W, H, frame), sp)
            dist.all_reduce(fence)
        torch.cuda.synchronize()
        if rank == 0:
            same = np.array_equal(r.image().view(np.uint32), full.view(np.uint32))
            print(f"{name}: {world}-rank tiles, peer stores over CUDA IPC == 1-rank image: {same}"); ok &= same
        dist.barrier()
        r.deinit()
    if rank == 0:
        print("MULTI-GPU CHECK", "PASS" if ok else "FAIL")
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
