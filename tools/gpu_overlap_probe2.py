"""Within-frame overlap probe (run under gpurun): one frame split into two interleaved tile sets rendered by two
renderers of the same GPU on their own streams (K-shade of one half can overlap K-trace of the other), against one
renderer doing the whole frame."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt
from tools import scene as S

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 200
builder, W, H, spp, seg = S.CONFIGS[name]
ps = builder(1234).pack()
full = b200pt.Renderer(ps, W, H, spp, seg)
halves = [b200pt.Renderer(ps, W, H, spp, seg, rank=k, world=2, partition=b200pt.PT_PARTITION_TILES) for k in range(2)]
cams = [S.camera_default(W, H, f) for f in range(frames)]
for f in range(5):
    full.record(cams[f]); [h.record(cams[f]) for h in halves]
full.sync(); [h.sync() for h in halves]
t0 = time.perf_counter()
for f in range(frames):
    full.record(cams[f])
full.sync()
t_full = time.perf_counter() - t0
t0 = time.perf_counter()
for f in range(frames):
    halves[0].record(cams[f]); halves[1].record(cams[f])
[h.sync() for h in halves]
t_split = time.perf_counter() - t0
img = halves[0].image() + halves[1].image()
same = np.array_equal(img.view(np.uint32), full.image().view(np.uint32))
print(f"{name}: {frames} frames, whole frame on one stream {1e3*t_full/frames:.3f} ms, two tile halves on two streams "
      f"{1e3*t_split/frames:.3f} ms ({t_full/t_split:.3f}x), images identical: {same}")
