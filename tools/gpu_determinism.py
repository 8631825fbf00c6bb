"""Builder determinism probe (run under gpurun): the same scenes built repeatedly between other work; every
build must give the same digests (pt_debug_bvh_digest) and trace results equal to the oracle's."""
import collections, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt
from oracle import orc
from tools import scene as S
from tools.gpu_check import random_rays, cmp_hits

N = int(sys.argv[1]) if len(sys.argv) > 1 else 60
ps = S.random_soup(3, 3000, 3, 7).pack()
pc = S.cornell().pack()
pa = S.atrium(1234, 20000, 6, 64).pack()
rays = random_rays(np.random.default_rng(7), 200000)
ho = orc.OracleScene(ps, True).trace(rays, use_bvh=False)
seen = collections.Counter()
bad_total = 0
for k in range(N):
    # disturb the allocator and the caches between builds
    if k % 3 == 0:
        rc = b200pt.Renderer(pc, 128 - 8 * (k % 8), 128, 1 + k % 3, 2 + k % 4); rc.trace(rays[: 1000 * (k % 8 + 1)]); rc.record(S.camera_default(128, 128, k)); rc.sync(); rc.deinit()
    if k % 5 == 0:
        ra = b200pt.Renderer(pa, 160, 90, 1, 2); ra.record(S.camera_default(160, 90, k)); ra.sync(); ra.deinit()
    if k % 7 == 0:
        orc.OracleScene(ps, True)       # CPU work between builds, as tools/gpu_check.py does
    flags = (0, 2)[k & 1]
    r = b200pt.Renderer(ps, 100, 60, 2, 3, flags=flags)
    d = r.bvh_digest()
    seen[d] += 1
    hg = r.trace(rays)
    bad = int(np.any([hg[f] != ho[f] for f in ("instance", "primitive_index", "triangle_index")], axis=0).sum())
    bad_total += bad
    if len(seen) > 1 and seen[d] == 1 or bad:
        print(f"build {k} flags {flags}: nodes8 {d[3]} mismatches {bad} digest {[hex(x)[-8:] for x in d]}")
    r.deinit()
print(f"{N} builds: {len(seen)} distinct digests, counts {sorted(seen.values(), reverse=True)}, trace mismatches {bad_total}")
for d, c in seen.most_common():
    print(c, [hex(x)[-8:] for x in d])
