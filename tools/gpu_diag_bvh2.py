"""Diagnose BVH2-walker mismatches on axis-parallel rays."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import b200pt
from oracle import orc
from parity_util import hits_mismatch, random_rays
from tools import scene as S
ps = S.random_soup(3, 3000, 3, 7).pack()
rays = random_rays(3, 100000, dtype=b200pt.PtRay_dtype)
rays["dir"][:300] = np.tile(np.eye(3, dtype=np.float32), (100, 1))
rays["dir"][300:600] = -np.tile(np.eye(3, dtype=np.float32), (100, 1))
ho = orc.OracleScene(ps, False).trace(rays, use_bvh=False)
for flags in (0, 8, 2):
    r = b200pt.Renderer(ps, 32, 32, 1, 1, flags=flags)
    hg = r.trace(rays)
    bad = hits_mismatch(hg, ho)
    print("flags", flags, "mismatches", int(bad.sum()), "indices", np.flatnonzero(bad)[:10])
    for k in np.flatnonzero(bad)[:6]:
        print("  ray", rays[k], "\n   gpu", hg[k], "\n   orc", ho[k])
    r.deinit()
