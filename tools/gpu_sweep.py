"""Kernel-time sweeps on the C2 workload (run under gpurun): per-kernel CUDA-event times from pt_get_stats."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt
from tools import scene as S

def measure(r, W, H, frames=int(os.environ.get("SWEEP_FRAMES", "8")), warm=2):
    r.set_profiling(True)
    tr = sh = ot = 0.0; rays = 0
    for f in range(warm + frames):
        r.record(S.camera_default(W, H, f)); st = r.stats()
        if f >= warm:
            tr += st.ms_trace; sh += st.ms_shade; ot += st.ms_other; rays += st.rays_traced
    return tr / frames, sh / frames, ot / frames, rays / frames

def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "one"
    name = sys.argv[2] if len(sys.argv) > 2 else "c2"
    builder, W, H, spp, seg = S.CONFIGS[name]
    ps = builder(1234).pack()
    if what == "chunk":
        for v in (8, 16, 32, 64, 128):       # the knob applies to renderers created afterwards
            b200pt.set_option("chunk", v)
            r = b200pt.Renderer(ps, W, H, spp, seg)
            tr, sh, ot, rays = measure(r, W, H)
            print(f"chunk {v:4d}: trace {tr:.3f} ms  shade {sh:.3f}  other {ot:.3f}  total {tr+sh+ot:.3f}")
            r.deinit()
        b200pt.set_option("chunk", 32)
    elif what == "one":
        r = b200pt.Renderer(ps, W, H, spp, seg, flags=int(os.environ.get("SWEEP_FLAGS", "0")))
        tr, sh, ot, rays = measure(r, W, H)
        st = r.stats()
        nn, nt, nr = r.count_traversal()
        print(f"trace {tr:.3f} ms  shade {sh:.3f}  other {ot:.3f}  total {tr+sh+ot:.3f}  {rays/1e6:.2f} Mrays   "
              f"bvh8 nodes {st.bvh_nodes}  per ray: {nn/max(nr,1):.2f} nodes {nt/max(nr,1):.2f} tris  build {st.ms_build:.2f} ms")
        r.deinit()
    elif what == "builders":
        # A/B of the binary-hierarchy stages and collapses: per-ray counts (pt_count_traversal), kernel times, and the image
        # must be bit-identical to the default build (the closest-hit rule does not depend on the hierarchy)
        ref = None
        for flags, label in ((0, "SAH + optimal collapse (default)"), (512, "LBVH + greedy"), (512 | 128, "LBVH + optimal collapse"), (64, "SAH + greedy"),
                             (16, "PLOC + greedy")):
            try:
                r = b200pt.Renderer(ps, W, H, spp, seg, flags=flags)
            except Exception as e:   # an opt-in stage that fails must not hide the others
                print(f"{label:34s}: build failed: {e}")
                continue
            tr, sh, ot, rays = measure(r, W, H)
            st = r.stats()
            nn, nt, nr = r.count_traversal()
            r.record(S.camera_default(W, H, 0)); r.sync()
            img = r.image()
            if ref is None: ref = img
            same = np.array_equal(ref.view(np.uint32), img.view(np.uint32))
            print(f"{label:34s}: trace {tr:.3f} ms  shade {sh:.3f}  other {ot:.3f}  total {tr+sh+ot:.3f}  bvh8 nodes {st.bvh_nodes}  "
                  f"per ray: {nn/max(nr,1):.2f} nodes {nt/max(nr,1):.2f} tris  build {st.ms_build:.2f} ms  same_image={same}")
            r.deinit()
    elif what == "modes":
        for flags, label in ((0, "persistent"), (8, "walker"), (2, "bvh2")):
            r = b200pt.Renderer(ps, W, H, spp, seg, flags=flags)
            tr, sh, ot, rays = measure(r, W, H)
            print(f"{label:10s}: trace {tr:.3f} ms  shade {sh:.3f}  other {ot:.3f}  total {tr+sh+ot:.3f}")
            r.deinit()

if __name__ == "__main__":
    main()
