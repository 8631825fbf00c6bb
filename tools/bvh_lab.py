"""Driver of tools/bvh_lab.cpp: dumps the triangles of a synthetic configuration (single identity instance: C1, C2, C5)
and runs the CPU model of the builder variants on the reference camera's rays.

    python tools/bvh_lab.py [c1|c2|c5] [pixel step]

Design tool only (no GPU, no oracle): see the header of bvh_lab.cpp.
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from tools import scene as S  # noqa: E402


def world_triangles(packed):
    """(n, 9) float32 triangle soup of a scene whose instances are all identity, (geometry, triangle) order."""
    blob = packed.triangle_data
    out = []
    for p in packed.primitives:
        n = p.triangle_count
        if p.info & 0x80000000:
            idx = np.frombuffer(blob, np.uint32, 3 * n, p.indices_offset)
        else:
            idx = np.frombuffer(blob, np.uint16, 3 * n, p.indices_offset).astype(np.uint32)
        pos = np.frombuffer(blob, np.float32, 3 * (p.max_vertex + 1), p.positions_offset).reshape(-1, 3)
        out.append(pos[idx].reshape(n, 9))
    return np.ascontiguousarray(np.concatenate(out), np.float32)


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c2"
    step = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    if which == "c1":
        sb, (W, H) = S.cornell(), (512, 512)
    elif which == "c2":
        sb, (W, H) = S.sponza_scale(textures=False), (1920, 1080)
    elif which == "c5":
        sb, (W, H) = S.atrium(1234, 1_000_000, 16, 64, (0.05, 0.5), 0.6, False), (1920, 1080)
    elif which == "c4":
        sb, (W, H) = S.instanced(), (1920, 1080)
    else:
        raise SystemExit("c1 | c2 | c4 | c5")
    packed = sb.pack()
    tris = world_triangles(packed)
    build = os.path.join(HERE, "_build")
    os.makedirs(build, exist_ok=True)
    exe = os.path.join(build, "bvh_lab")
    src = os.path.join(HERE, "bvh_lab.cpp")
    deps = [src, os.path.join(ROOT, "vulkan-pathtracer_b200", "csrc", "bvh_sah.cuh")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-Wall", "-Wno-unused-function", "-o", exe, src])
    if which == "c4":
        # two-level model: per-mesh triangles + instance transforms
        path = os.path.join(build, "c4.scene")
        with open(path, "wb") as f:
            f.write(np.uint32(len(packed.meshes)).tobytes())
            for m in packed.meshes:
                sub = S.PackedScene()
                sub.triangle_data = packed.triangle_data
                sub.primitives = [packed.primitives[p] for p in range(m.start, m.end)]
                t = world_triangles(sub)
                f.write(np.uint32(len(t)).tobytes()); f.write(t.tobytes())
            f.write(np.uint32(len(packed.instances)).tobytes())
            for i in packed.instances:
                f.write(np.uint32(i.mesh_index).tobytes()); f.write(np.array(list(i.transform)[:12], np.float32).tobytes())
        subprocess.check_call([exe, path, str(W), str(H), str(step)])
        return
    path = os.path.join(build, f"{which}_tris.bin")
    with open(path, "wb") as f:
        f.write(np.uint32(len(tris)).tobytes())
        f.write(tris.tobytes())
    subprocess.check_call([exe, path, str(W), str(H), str(step)])


if __name__ == "__main__":
    main()
