"""Small build + trace + render sequence for compute-sanitizer (initcheck / memcheck) under gpurun."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt
from tools import scene as S
from tools.gpu_check import random_rays

for sb, flags in ((S.cornell(), 0), (S.random_soup(3, 3000, 3, 7), 0), (S.random_soup(3, 3000, 3, 7), 2), (S.random_soup(3, 3000, 3, 7), 16), (S.atrium(1234, 5000, 4, 32), 1)):
    ps = sb.pack()
    r = b200pt.Renderer(ps, 64, 40, 2, 3, flags=flags)
    r.trace(random_rays(np.random.default_rng(7), 2000))
    r.record(S.camera_default(64, 40, 1)); r.sync()
    print("nodes8", r.stats().bvh_nodes, float(np.nanmean(r.image())))
    r.deinit()
