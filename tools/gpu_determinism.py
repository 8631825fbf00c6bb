"""Builder determinism probe (run under gpurun): same scene built repeatedly between other work."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt
from oracle import orc
from tools import scene as S
from tools.gpu_check import random_rays, cmp_hits

ps = S.random_soup(3, 3000, 3, 7).pack()
pc = S.cornell().pack()
rays = random_rays(np.random.default_rng(7), 200000)
ho = orc.OracleScene(ps, True).trace(rays, use_bvh=False)
for k in range(8):
    # disturb the allocator between builds
    rc = b200pt.Renderer(pc, 128 - 8 * k, 128, 1 + k % 3, 2 + k % 4); rc.trace(rays[: 1000 * (k + 1)]); rc.record(S.camera_default(128 - 8 * k, 128, k)); rc.sync(); rc.deinit()
    flags = (0, 2)[k & 1]
    r = b200pt.Renderer(ps, 100, 60, 2, 3, flags=flags)
    st = r.stats()
    hg = r.trace(rays)
    cmp_hits(f"flags {flags} build {k} nodes8 {st.bvh_nodes}", hg, ho)
    r.deinit()
