"""Upper bound for overlapping K-shade with K-trace (run under gpurun): two renderers of the same scene on their
own streams, frames submitted alternately, against one renderer doing all frames."""
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
r1 = b200pt.Renderer(ps, W, H, spp, seg)
r2 = b200pt.Renderer(ps, W, H, spp, seg)
cams = [S.camera_default(W, H, f) for f in range(frames)]
for f in range(10):
    r1.record(cams[f]); r2.record(cams[f])
r1.sync(); r2.sync()
t0 = time.perf_counter()
for f in range(frames):
    r1.record(cams[f])
r1.sync()
t_one = time.perf_counter() - t0
t0 = time.perf_counter()
for f in range(0, frames, 2):
    r1.record(cams[f]); r2.record(cams[f + 1])
r1.sync(); r2.sync()
t_two = time.perf_counter() - t0
print(f"{name}: {frames} frames, one stream {1e3*t_one/frames:.3f} ms/frame, two streams {1e3*t_two/frames:.3f} ms/frame "
      f"({t_one/t_two:.3f}x)")
