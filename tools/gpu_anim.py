"""Animated instances on C4 (run under gpurun): every frame all 1 000 instance transforms change, the top level
is rebuilt on the device (pt_set_instance_transforms) and one frame is rendered.  Prints the rebuild and
frame times and checks the last frame against a renderer created from scratch with the same transforms."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt
from tools import scene as S


def spin(T0, angle):
    """Every instance rotated about its own y axis by `angle` * (1 + i / n), translation kept."""
    n = len(T0)
    out = T0.copy().reshape(n, 4, 4)
    a = angle * (1.0 + np.arange(n) / n)
    c, s = np.cos(a).astype(np.float32), np.sin(a).astype(np.float32)
    R = np.zeros((n, 3, 3), np.float32)
    R[:, 0, 0] = c; R[:, 0, 2] = s; R[:, 1, 1] = 1; R[:, 2, 0] = -s; R[:, 2, 2] = c
    out[:, :3, :3] = np.einsum("nij,njk->nik", T0.reshape(n, 4, 4)[:, :3, :3], R).astype(np.float32)
    return out.reshape(n, 16)


def main():
    frames = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    builder, W, H, spp, seg = S.CONFIGS["c4"]
    ps = builder(1234).pack()
    T0 = np.array([list(i.transform) for i in ps.instances], np.float32)
    r = b200pt.Renderer(ps, W, H, spp, seg)
    r.set_profiling(True)
    reb, frm, wall = [], [], []
    for f in range(frames):
        T = spin(T0, 0.05 * (f + 1))
        t0 = time.perf_counter()
        reb.append(r.set_instance_transforms(T))
        wall.append((time.perf_counter() - t0) * 1e3)
        r.record(S.camera_default(W, H, f)); st = r.stats()
        frm.append(st.ms_total)
    img = r.image().copy()
    r.deinit()
    import copy, ctypes as C
    ps1 = copy.copy(ps)
    ps1.instances = (S.PtInstance * len(T))()
    for i in range(len(T)):
        ps1.instances[i].mesh_index = ps.instances[i].mesh_index
        ps1.instances[i].transform = (C.c_float * 16)(*T[i].tolist())
    r2 = b200pt.Renderer(ps1, W, H, spp, seg)
    r2.record(S.camera_default(W, H, frames - 1)); r2.sync()
    same = np.array_equal(img.view(np.uint32), r2.image().view(np.uint32))
    r2.deinit()
    print(f"C4 animated: {len(T)} instances, {frames} frames: TLAS rebuild {np.median(reb):.3f} ms device "
          f"({np.median(wall):.3f} ms host wall incl. upload), frame {np.median(frm):.3f} ms; "
          f"last frame == fresh renderer: {same}")


if __name__ == "__main__":
    main()
