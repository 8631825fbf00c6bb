"""GPU diagnostic battery (run under gpurun): builder + traversal + render parity against the CPU
oracle on several scenes, printing mismatch statistics without stopping at the first failure."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt  # noqa: E402
from oracle import orc  # noqa: E402
from tools import scene as S  # noqa: E402


def bits_equal(a, b):
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def random_rays(rng, n, ps_extent=8.0):
    o = (rng.random((n, 3)) - 0.5) * ps_extent * 2
    o[:, 2] -= ps_extent * 0.2
    t = (rng.random((n, 3)) - 0.5) * ps_extent + np.array([0, 0, ps_extent * 0.75])
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros(n, b200pt.PtRay_dtype)
    rays["origin"] = o; rays["dir"] = d; rays["tmin"] = 1e-3; rays["tmax"] = 1e4
    return rays


def cmp_hits(name, a, b):
    bad = np.zeros(len(a), bool)
    for f in ("instance", "primitive_index", "triangle_index"):
        bad |= a[f] != b[f]
    for f in ("t", "b1", "b2"):
        bad |= a[f].view(np.uint32) != b[f].view(np.uint32)
    nh = int((b["t"] < 1e31).sum())
    print(f"  {name}: rays={len(a)} oracle_hits={nh} mismatches={int(bad.sum())}")
    if bad.any():
        i = np.flatnonzero(bad)[:5]
        for k in i:
            print("    gpu", a[k], "orc", b[k])
    return int(bad.sum())


def check_scene(name, sb, W, H, S_, B_, n_rays=200000, brute=False):
    print(f"== {name}")
    ps = sb.pack()
    osc = orc.OracleScene(ps, True)
    total_bad = 0
    for flags, label in ((0, "bvh8"), (b200pt.PT_FLAG_TRACE_BVH2, "bvh2")):
        t0 = time.time()
        r = b200pt.Renderer(ps, W, H, S_, B_, flags=flags)
        st = r.stats()
        print(f"  [{label}] create {time.time()-t0:.2f}s build {st.ms_build:.2f} ms tris {st.num_triangles} nodes8 {st.bvh_nodes} bytes {st.bvh_bytes}")
        rng = np.random.default_rng(7)
        rays = random_rays(rng, n_rays)
        hg = r.trace(rays)
        ho = osc.trace(rays, use_bvh=not brute)
        total_bad += cmp_hits(f"[{label}] trace vs oracle({'brute' if brute else 'bvh'})", hg, ho)
        for frame in (0, 3):
            cam = S.camera_default(W, H, frame)
            r.record(cam); r.sync()
            img = r.image(); ids = r.ids(); st = r.stats()
            cfg = S.render_config(W, H, S_, B_)
            oimg, oids, orays = osc.render(cfg, cam, True)
            idbad = sum(int((a != b).sum()) for a, b in zip(ids, oids))
            eq = bits_equal(img, oimg)
            px_bad = int((~eq.all(axis=2)).sum())
            fin = np.isfinite(img) & np.isfinite(oimg)
            rel = np.abs(img - oimg)[fin] / np.maximum(np.abs(oimg)[fin], 1e-3)
            print(f"  [{label}] frame {frame}: rays gpu {st.rays_traced} orc {orays}  id mismatches {idbad}  pixel bit-mismatches {px_bad}/{W*H}"
                  f"  max rel {rel.max() if rel.size else 0:.3g}  mean gpu {np.nanmean(img[...,:3]):.6f} orc {np.nanmean(oimg[...,:3]):.6f} iters {st.wavefront_iterations}")
            total_bad += idbad + px_bad
        r.deinit()
    return total_bad


def main():
    print(b200pt.lib().pt_version().decode(), "devices", b200pt.lib().pt_device_count())
    bad = 0
    bad += check_scene("cornell 128x128 1spp 2seg", S.cornell(), 128, 128, 1, 2, brute=True)
    bad += check_scene("cornell 96x64 4spp 5seg", S.cornell(), 96, 64, 4, 5, brute=True)
    bad += check_scene("random soup", S.random_soup(3, 3000, 3, 7), 100, 60, 2, 3, brute=True)
    bad += check_scene("random soup, 1-triangle meshes", S.random_soup(5, 3, 3, 4), 64, 64, 1, 2, brute=True)
    bad += check_scene("atrium 20k tris 256px textures", S.atrium(1234, 20000, 6, 64), 160, 90, 2, 3)
    bad += check_scene("instanced 3^3", S.instanced(1234, 3, 32), 160, 90, 2, 3)
    print("TOTAL MISMATCHES", bad)
    return 0


if __name__ == "__main__":
    sys.exit(main())
