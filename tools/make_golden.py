"""Generates tests/golden/*.npz (run in the authoring container, where /root/reference exists).

* bsdf_ref_vectors.npz — outputs of the REFERENCE's own brdf.glsl / rand.glsl compiled as C++
  (oracle/_ref/libref_glsl.so) on seeded random inputs: the oracle and the CUDA kernels must
  reproduce them bit for bit.
* cornell_64.npz — the oracle's image / primary-hit ids for the Cornell scene (regression + GPU parity
  without needing the oracle's BVH to agree with anything else).
* raygen_ref_frames.npz — whole frames rendered by the REFERENCE's own ray_gen.rgen main() (+ closest_hit.rchit,
  miss.rmiss, triangle.glsl, material.glsl, brdf.glsl, rand.glsl) compiled as C++, one invocation per pixel, with the
  oracle's closest-hit query behind traceRayEXT (the Vulkan driver's part has no source).  The oracle and the CUDA
  path must reproduce them bit for bit.  REF_FRAMES below is the list of (scene, extent, samples, segments, frame).
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402
from tools import scene as S  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def bsdf_inputs(seed, n):
    rng = np.random.default_rng(seed)
    unit = lambda v: v / np.linalg.norm(v, axis=1, keepdims=True)
    N = unit(rng.normal(size=(n, 3))).astype(np.float32)
    V = unit(N + rng.normal(size=(n, 3)) * 0.8).astype(np.float32)
    kd = rng.random((n, 3)).astype(np.float32)
    f0 = rng.random((n, 3)).astype(np.float32)
    rough = np.maximum(0.006, rng.random(n) ** 2).astype(np.float32)
    xi = rng.random((n, 2)).astype(np.float32)
    # edge cases: grazing view, xi on the lobe boundary, xi == 1.0 (pcg2d can return exactly 1)
    V[:16] = unit(np.cross(N[:16], rng.normal(size=(16, 3))) + 1e-3 * N[:16]).astype(np.float32)
    xi[16:32, 0] = 0.0
    xi[32:48] = 1.0
    kd[48:64] = 0.0
    return N, V, kd, f0, rough, xi


# name -> (scene builder, W, H, samples, segments, frame_count); every generator is seeded and deterministic
REF_FRAMES = {
    "cornell": (lambda: S.cornell(), 64, 64, 2, 3, 5),                              # normal map + albedo map + emissive quad
    "atrium": (lambda: S.atrium(1234, 20000, 6, 64), 96, 54, 2, 4, 1),            # 6 textured PBR materials, sky, u16 + u32 indices
    "instanced": (lambda: S.instanced(1234, 4, 64), 80, 45, 2, 4, 2),             # rotated instances: the literal A^-1 normal matrix
    "soup": (lambda: S.random_soup(3, 3000, 3, 7), 80, 45, 2, 4, 0),              # rotation + non-uniform scale
}


def ref_frames():
    out = {}
    for name, (builder, W, H, spp, seg, frame) in REF_FRAMES.items():
        ps = builder().pack()
        img, rays = orc.OracleScene(ps, True).ref_render(S.render_config(W, H, spp, seg), S.camera_default(W, H, frame), True)
        out[name] = img
        out[name + "_rays"] = np.uint64(rays)
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    orc.build()
    R = orc.ref_lib()
    assert R is not None, "needs /root/reference (oracle/_ref/libref_glsl.so)"
    n = 4096
    N, V, kd, f0, rough, xi = bsdf_inputs(20261104, n)
    L, pdf, f = orc.bsdf_batch(R, N, V, kd, f0, rough, xi, "ref")
    px = np.array([(0, 0, 0), (3, 7, 0), (1919, 1079, 5), (511, 1, 77), (65535, 65535, 0xFFFFFFFF), (12, 900, 123456)], np.uint32)
    seeds = np.zeros((len(px), 2), np.uint32); states = np.zeros((len(px), 3, 2), np.uint32); xis = np.zeros((len(px), 3, 2), np.float32)
    for i, (x, y, fr) in enumerate(px):
        s = (C.c_uint32 * 2)(); R.ref_prng_seed(int(x), int(y), int(fr), s)
        seeds[i] = list(s)
        for k in range(3):
            o = (C.c_float * 2)(); R.ref_pcg2d(s, o)
            states[i, k] = list(s); xis[i, k] = list(o)
    np.savez_compressed(os.path.join(OUT, "bsdf_ref_vectors.npz"), N=N, V=V, kd=kd, f0=f0, rough=rough, xi=xi, L=L, pdf=pdf, f=f,
                        px=px, seeds=seeds, states=states, xis=xis)
    ps = S.cornell().pack()
    osc = orc.OracleScene(ps, True)
    cfg = S.render_config(64, 64, 2, 3); cam = S.camera_default(64, 64, 5)
    img, ids, rays = osc.render(cfg, cam, False)
    np.savez_compressed(os.path.join(OUT, "cornell_64.npz"), img=img, inst=ids[0], prim=ids[1], tri=ids[2], rays=np.uint64(rays))
    np.savez_compressed(os.path.join(OUT, "raygen_ref_frames.npz"), **ref_frames())
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
