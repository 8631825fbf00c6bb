import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/vulkan-pathtracer_b200")
import numpy as np, b200pt
from tools import scene as S
ps = S.sponza_scale(1234).pack()
for (W, H) in ((3840, 2160), (1360, 760)):
    r = b200pt.Renderer(ps, W, H, 64, 8)
    r.set_profiling(True)
    for f in range(2): r.record(S.camera_default(W, H, f)); st = r.stats()
    t0 = time.perf_counter(); r.record(S.camera_default(W, H, 2)); r.sync(); wall = (time.perf_counter() - t0) * 1e3
    st = r.stats()
    print(f"{W}x{H}: wall {wall:.1f} ms  trace {st.ms_trace:.1f} shade {st.ms_shade:.1f} other {st.ms_other:.2f} total {st.ms_total:.1f}  rays {st.rays_traced/1e6:.1f}M  iters {st.wavefront_iterations} launches {st.kernel_launches}  -> {st.rays_traced/st.ms_total/1e3:.0f} Mrays/s")
    r.set_profiling(False)
    t0 = time.perf_counter(); r.record(S.camera_default(W, H, 2)); r.sync(); wall = (time.perf_counter() - t0) * 1e3
    print(f"   without per-kernel events: wall {wall:.1f} ms")
    r.deinit()
