"""The oracle's closest-hit rule, texture filter and trace loop: internal consistency (BVH == brute
force), facing / interval / tie-break semantics, hand-computed texture taps, golden regression."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import orc
from parity_util import bits_equal, hits_mismatch, random_rays
from tools import scene as S

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _hit(lib, o, d, v, tmin=1e-3, tmax=1e4):
    out = (C.c_float * 3)()
    r = lib.orc_intersect_tri((C.c_float * 3)(*o), (C.c_float * 3)(*d), (C.c_float * 9)(*v), C.c_float(tmin), C.c_float(tmax), out)
    return r, tuple(out)


def test_front_face_is_normal_towards_origin(oracle_lib):
    # gl_RayFlagsCullBackFacingTrianglesEXT (ray_gen.rgen:53): cross(p1-p0, p2-p0) . d < 0 is front
    rng = np.random.default_rng(0)
    n_hit = 0
    for _ in range(5000):
        tri = rng.normal(size=9) * 2
        o = rng.normal(size=3) * 3
        c = (tri[0:3] + tri[3:6] + tri[6:9]) / 3 + rng.normal(size=3) * 0.3
        d = c - o; d /= np.linalg.norm(d)
        r, _ = _hit(oracle_lib, o, d, tri)
        nrm = np.cross(tri[3:6] - tri[0:3], tri[6:9] - tri[0:3])
        if abs(np.dot(nrm, d)) < 1e-4:
            continue
        if r:
            n_hit += 1
            assert np.dot(nrm, d) < 0
        # the flipped triangle is never hit when the original is
        r2, _ = _hit(oracle_lib, o, d, np.concatenate([tri[0:3], tri[6:9], tri[3:6]]))
        assert not (r and r2)
    assert n_hit > 500


def test_hit_interval_is_open(oracle_lib):
    v = [-1, -1, 2, -1, 1, 2, 1, -1, 2]   # normal -z, plane z = 2
    o, d = [-0.5, -0.5, 0], [0, 0, 1]
    assert _hit(oracle_lib, o, d, v)[0] == 1
    assert _hit(oracle_lib, o, d, v, tmin=2.0)[0] == 0      # t == tmin rejected
    assert _hit(oracle_lib, o, d, v, tmax=2.0)[0] == 0      # t == tmax rejected
    r, (t, b1, b2) = _hit(oracle_lib, o, d, v)
    assert t == 2.0 and abs(b1 - 0.25) < 1e-6 and abs(b2 - 0.25) < 1e-6


def test_watertight_shared_edges(oracle_lib):
    """Rays through shared edges/vertices of a fan hit exactly once or more, never zero times."""
    rng = np.random.default_rng(3)
    c = np.array([0.0, 0.0, 5.0])
    ring = [np.array([np.cos(a), np.sin(a), 5.0 + 0.3 * np.sin(3 * a)]) for a in np.linspace(0, 2 * np.pi, 9)[:-1]]
    tris = []
    for i in range(8):
        a, b = ring[i], ring[(i + 1) % 8]
        t = np.concatenate([c, a, b])
        if np.dot(np.cross(a - c, b - c), [0, 0, 1]) > 0:
            t = np.concatenate([c, b, a])
        tris.append(t)
    o = np.array([0.0, 0.0, 0.0])
    for k in range(2000):
        i = rng.integers(0, 8)
        s = rng.random() * 0.95
        target = c * (1 - s) + ring[i] * s if k % 2 else c            # interior shared edge / the vertex shared by all
        d = (target - o).astype(np.float32)
        hits = sum(_hit(oracle_lib, o, d, t)[0] for t in tris)
        assert hits >= 1


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_bvh_equals_brute_force(seed):
    ps = S.random_soup(seed, 1500, 3, 6).pack()
    osc = orc.OracleScene(ps, True)
    rays = random_rays(seed, 30000, dtype=orc.PtRay_dtype)
    a, b = osc.trace(rays, True), osc.trace(rays, False)
    assert (b["t"] < 1e31).sum() > 5000
    assert not hits_mismatch(a, b).any()


def test_tie_break_duplicate_geometry():
    """Two coincident instances of the same mesh: the smaller (instance, primitive, triangle) wins."""
    sb = S.SceneBuilder()
    m = sb.add_material((0.5, 0.5, 0.5))
    s = S.Soup(); S.grid_patch(s, (-1, -1, 3), (2, 0, 0), (0, 2, 0), 2, 2, facing=(0, 0, -1))
    mesh = sb.add_mesh([s.prim(m), s.prim(m)])      # two identical primitives (geometries)
    sb.add_node(mesh); sb.add_node(mesh)            # two identical instances
    ps = sb.pack()
    osc = orc.OracleScene(ps, True)
    rays = np.zeros(1000, orc.PtRay_dtype)
    rng = np.random.default_rng(0)
    rays["origin"] = 0; rays["dir"] = np.column_stack([(rng.random(1000) - 0.5) * 0.5, (rng.random(1000) - 0.5) * 0.5, np.ones(1000)])
    rays["tmin"] = 1e-3; rays["tmax"] = 1e4
    for use_bvh in (True, False):
        h = osc.trace(rays, use_bvh)
        assert (h["t"] < 1e31).all()
        assert (h["instance"] == 0).all() and (h["primitive_index"] == 0).all()


def test_texture_filter_hand_values(oracle_lib):
    tex = np.zeros((2, 2, 4), np.uint8)
    tex[0, 0] = (255, 0, 0, 255); tex[0, 1] = (0, 255, 0, 255); tex[1, 0] = (0, 0, 255, 255); tex[1, 1] = (255, 255, 255, 0)
    out = (C.c_float * 4)()

    def tap(u, v):
        oracle_lib.orc_sample_texture(tex.ctypes.data_as(C.c_void_p), 2, 2, C.c_float(u), C.c_float(v), out)
        return np.array(list(out))
    assert np.allclose(tap(0.25, 0.25), [1, 0, 0, 1])                 # texel centre (0,0)
    assert np.allclose(tap(0.75, 0.25), [0, 1, 0, 1])                 # texel centre (1,0)
    assert np.allclose(tap(0.5, 0.25), [0.5, 0.5, 0, 1])              # halfway in u
    assert np.allclose(tap(0.5, 0.5), [0.5, 0.5, 0.5, 0.75])          # all four
    assert np.allclose(tap(0.0, 0.25), [0.5, 0.5, 0, 1])              # REPEAT: wraps to texel 1 on the left
    assert np.allclose(tap(1.25, 0.25), tap(0.25, 0.25))              # period 1
    assert np.allclose(tap(-0.75, -1.75), tap(0.25, 0.25))            # negative coordinates


def test_cornell_golden_regression():
    g = np.load(os.path.join(GOLD, "cornell_64.npz"))
    ps = S.cornell().pack()
    osc = orc.OracleScene(ps, True)
    cfg = S.render_config(64, 64, 2, 3); cam = S.camera_default(64, 64, 5)
    for use_bvh in (True, False):
        img, ids, rays = osc.render(cfg, cam, use_bvh)
        assert bits_equal(img, g["img"]).all()
        assert (ids[0] == g["inst"]).all() and (ids[1] == g["prim"]).all() and (ids[2] == g["tri"]).all()
        assert rays == int(g["rays"])


def test_no_pixel_jitter_and_seed_threading():
    """Q1/Q12: all samples share the primary ray; S samples != S frames of 1 sample (the seed is
    threaded through the samples)."""
    ps = S.cornell().pack()
    osc = orc.OracleScene(ps, True)
    cam = S.camera_default(32, 32, 0)
    a, ids_a, _ = osc.render(S.render_config(32, 32, 1, 3), cam)
    b, ids_b, _ = osc.render(S.render_config(32, 32, 4, 3), cam)
    assert all((x == y).all() for x, y in zip(ids_a, ids_b))
    assert not bits_equal(a, b).all()


def test_bounce_count_is_segment_count():
    """Q10: NUM_BOUNCES = 1 traces primaries only: emissive + sky, no BSDF sampling."""
    ps = S.cornell().pack()
    osc = orc.OracleScene(ps, True)
    cam = S.camera_default(32, 32, 0)
    img, ids, rays = osc.render(S.render_config(32, 32, 3, 1), cam)
    assert rays == 32 * 32 * 3
    miss = ids[0] == 0xFFFFFFFF
    assert np.allclose(img[~miss][:, :3].max(), 10.0) or (img[~miss][:, :3] <= 10.0).all()


@pytest.mark.parametrize("scene,W,H,spp,seg", [("cornell", 96, 96, 3, 2), ("soup", 96, 64, 2, 3), ("atrium", 160, 90, 2, 2)])
def test_zero_throughput_last_segments_need_no_trace(oracle_lib, scene, W, H, spp, seg):
    """The last segment of a sample draws no random numbers and only adds atten * (emission | sky), both finite: when atten is
    exactly (0, 0, 0) -- brdf() returns 0 for a shading normal that faces away from the viewer -- it adds exactly nothing.
    Skipping those rays (PT_FLAG_SKIP_DEAD_SEGMENTS) leaves every bit of the image alone and traces fewer rays."""
    if scene == "cornell":
        ps = S.cornell().pack()
    elif scene == "soup":
        ps = S.random_soup(17, 1500, 3, 6).pack()
    else:
        ps = S.atrium(1234, 20_000, 6, 64, (0.1, 1.0), 0.2, True).pack()
    osc = orc.OracleScene(ps, True)
    cam = S.camera_default(W, H, 0)
    a, _, rays_a = osc.render(S.render_config(W, H, spp, seg), cam, True)
    b, _, rays_b = osc.render(S.render_config(W, H, spp, seg, flags=1 << 10), cam, True)
    assert bits_equal(a, b).all() and rays_b <= rays_a
