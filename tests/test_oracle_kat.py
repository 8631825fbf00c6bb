"""Oracle pinned against (1) the known-answer vectors of SURVEY.md Appendix C, (2) golden vectors
produced by the reference's own brdf.glsl / rand.glsl compiled as C++ (tests/golden, made by
tools/make_golden.py) and (3) the live reference build when oracle/_ref exists."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import orc
from parity_util import bits_equal

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _seed(lib, prefix, x, y, f):
    s = (C.c_uint32 * 2)()
    getattr(lib, prefix + "_prng_seed")(x, y, f, s)
    return s


def _pcg(lib, prefix, s):
    o = (C.c_float * 2)()
    getattr(lib, prefix + "_pcg2d")(s, o)
    return np.float32(o[0]), np.float32(o[1])


# SURVEY.md Appendix C (derived from rand.glsl:76-94 by two independent routes)
APPENDIX_C = [
    ((0, 0, 0), (0x00000000, 0x00000000), (0x18e431a7, 0x055df4d1), (0.0972319618, 0.0209649093), (0xae860441, 0x69b5769c), (0.681732416, 0.412925154)),
    ((3, 7, 0), (0x448749e5, 0x2ad72673), (0x33c2987b, 0xf32032a2), (0.202188045, 0.949710071), (0x93bab5ce, 0xd19dd79b), (0.577067733, 0.818814754)),
    ((1919, 1079, 5), (0x269eec2d, 0xffe67ee6), (0x24af6a4e, 0x2228d776), (0.143301621, 0.133435696), (0x10af92b6, 0x73571d2d), (0.0651790351, 0.450547993)),
]


@pytest.mark.parametrize("case", APPENDIX_C)
def test_rng_known_answers(oracle_lib, case):
    (x, y, f), seed, st1, xi1, st2, xi2 = case
    s = _seed(oracle_lib, "orc", x, y, f)
    assert tuple(s) == seed
    a = _pcg(oracle_lib, "orc", s)
    assert tuple(s) == st1 and a == (np.float32(xi1[0]), np.float32(xi1[1]))
    b = _pcg(oracle_lib, "orc", s)
    assert tuple(s) == st2 and b == (np.float32(xi2[0]), np.float32(xi2[1]))


def test_pcg2d_can_return_one():
    # uint -> float is round-to-nearest: 0xFFFFFFFF * 2^-32 == 1.0f exactly (SURVEY a-10)
    assert np.float32(np.uint32(0xFFFFFFFF)) * np.float32(2.32830643654e-10) == np.float32(1.0)


def test_rng_golden_from_reference_glsl(oracle_lib):
    g = np.load(os.path.join(GOLD, "bsdf_ref_vectors.npz"))
    for i, (x, y, f) in enumerate(g["px"]):
        s = _seed(oracle_lib, "orc", int(x), int(y), int(f))
        assert tuple(s) == tuple(g["seeds"][i])
        for k in range(3):
            xi = _pcg(oracle_lib, "orc", s)
            assert tuple(s) == tuple(g["states"][i, k])
            assert xi == (g["xis"][i, k, 0], g["xis"][i, k, 1])


def test_bsdf_golden_from_reference_glsl(oracle_lib):
    g = np.load(os.path.join(GOLD, "bsdf_ref_vectors.npz"))
    L, pdf, f = orc.bsdf_batch(oracle_lib, g["N"], g["V"], g["kd"], g["f0"], g["rough"], g["xi"], "orc")
    assert bits_equal(L, g["L"]).all()
    assert bits_equal(pdf, g["pdf"]).all()
    assert bits_equal(f, g["f"]).all()


def test_bsdf_live_against_reference_glsl(oracle_lib, ref_glsl):
    """10^6 random records: the restatement equals the reference shader code bit for bit."""
    from tools.make_golden import bsdf_inputs
    N, V, kd, f0, rough, xi = bsdf_inputs(99, 1_000_000)
    a = orc.bsdf_batch(oracle_lib, N, V, kd, f0, rough, xi, "orc")
    b = orc.bsdf_batch(ref_glsl, N, V, kd, f0, rough, xi, "ref")
    for u, v in zip(a, b):
        assert bits_equal(u, v).all()


def test_rng_live_against_reference_glsl(oracle_lib, ref_glsl):
    rng = np.random.default_rng(5)
    for _ in range(2000):
        x, y, f = int(rng.integers(0, 8192)), int(rng.integers(0, 8192)), int(rng.integers(0, 2 ** 32))
        s1, s2 = _seed(oracle_lib, "orc", x, y, f), _seed(ref_glsl, "ref", x, y, f)
        assert tuple(s1) == tuple(s2)
        assert _pcg(oracle_lib, "orc", s1) == _pcg(ref_glsl, "ref", s2) and tuple(s1) == tuple(s2)


def test_float_spot_values(oracle_lib):
    # SURVEY.md Appendix C floating-point spot values (±2 ulp: they were produced with glibc sinf/cosf)
    out = (C.c_float * 3)()
    oracle_lib.orc_fresnel_schlick((C.c_float * 3)(0.04, 0.04, 0.04), (C.c_float * 3)(1, 1, 1), C.c_float(0.5), out)
    assert abs(out[0] - 0.0700000003) < 1e-8
    m = (C.c_float * 9)()
    oracle_lib.orc_shading_space((C.c_float * 3)(0, 0, 1), m)
    assert list(m) == [1, 0, 0, 0, 1, 0, 0, 0, 1]
    oracle_lib.orc_shading_space((C.c_float * 3)(0, 0, -1), m)
    assert list(m) == [1, 0, 0, 0, -1, 0, 0, 0, -1]
    n = np.array([0.1, 0.2, 1.0]); n /= np.linalg.norm(n)
    v = np.array([0.3, -0.1, 0.9]); v /= np.linalg.norm(v)
    L, pdf, f = orc.bsdf_batch(oracle_lib, n[None].astype(np.float32), v[None].astype(np.float32), np.array([[0.6, 0.5, 0.4]], np.float32),
                               np.array([[0.04, 0.04, 0.04]], np.float32), np.array([0.3], np.float32),
                               np.array([[0.202188045, 0.949710071]], np.float32))
    assert np.allclose(L[0], [-0.783413112, 0.591820419, 0.189769357], rtol=0, atol=2e-6)
    assert abs(pdf[0] - 0.0775616914) < 2e-7
    assert np.allclose(f[0], [0.205958679, 0.176263198, 0.146567732], rtol=0, atol=2e-7)


def test_shading_space_orthonormal(oracle_lib):
    rng = np.random.default_rng(0)
    for _ in range(200):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        m = (C.c_float * 9)()
        oracle_lib.orc_shading_space((C.c_float * 3)(*n), m)
        M = np.array(list(m)).reshape(3, 3)
        assert np.allclose(M @ M.T, np.eye(3), atol=2e-6)


def test_furnace_bound(oracle_lib):
    """White diffuse furnace: the Monte-Carlo estimate of brdf*cos/pdf stays <= 1 (+ noise)."""
    n = 200000
    rng = np.random.default_rng(1)
    N = np.tile(np.array([[0, 0, 1]], np.float32), (n, 1))
    V = np.tile(np.array([[0.3, 0.2, 0.9327379]], np.float32), (n, 1))
    kd = np.ones((n, 3), np.float32); f0 = np.full((n, 3), 0.02, np.float32)
    rough = np.full(n, 0.5, np.float32); xi = rng.random((n, 2)).astype(np.float32)
    L, pdf, f = orc.bsdf_batch(oracle_lib, N, V, kd, f0, rough, xi, "orc")
    cos = L[:, 2]
    ok = (cos > 0) & (pdf > 0)
    w = np.where(ok, f[:, 0] * cos / np.where(ok, pdf, 1), 0)
    assert np.isfinite(w).all()
    assert w.mean() < 1.05


# ------------------------------------------------------------------ the whole raygen loop, pinned to the reference's shader sources
# oracle/_ref/libref_glsl.so holds ray_gen.rgen main(), closest_hit.rchit, miss.rmiss, getTriangleData (triangle.glsl:46-112),
# getMaterialData (material.glsl:31-69), brdf.glsl and rand.glsl compiled as C++ from /root/reference (oracle/ref_glsl/rewrite.py).
def _ref_frame_cases():
    from tools.make_golden import REF_FRAMES
    return sorted(REF_FRAMES)


@pytest.mark.parametrize("name", _ref_frame_cases())
def test_oracle_frames_equal_golden_frames_of_the_reference_raygen(oracle_lib, name):
    """orc_render (the hand restatement) against frames rendered by the reference's own ray_gen.rgen main()."""
    from tools import scene as S
    from tools.make_golden import REF_FRAMES
    g = np.load(os.path.join(GOLD, "raygen_ref_frames.npz"))
    builder, W, H, spp, seg, frame = REF_FRAMES[name]
    img, _, rays = orc.OracleScene(builder().pack(), True).render(S.render_config(W, H, spp, seg), S.camera_default(W, H, frame), True)
    eq = bits_equal(img, g[name])
    assert eq.all(), f"{int((~eq.all(axis=2)).sum())} of {W * H} pixels differ from the reference raygen"
    assert rays == int(g[name + "_rays"])           # one traceRayEXT per executed segment (ray_gen.rgen:51)


@pytest.mark.parametrize("case", ["cornell", "soup-bruteforce", "atrium", "instanced", "tiles"])
def test_raygen_live_against_reference_glsl(oracle_lib, ref_glsl, case):
    """Whole frames through the reference's main(), one invocation per pixel, against orc_render — bit for bit."""
    from tools import scene as S
    if not hasattr(ref_glsl, "ref_render"):
        pytest.skip("prebuilt oracle/_ref predates the full-pipeline build")
    builder, W, H, spp, seg, frame, use_bvh, part = {
        "cornell": (lambda: S.cornell(), 160, 120, 3, 5, 7, True, (0, 1)),
        "soup-bruteforce": (lambda: S.random_soup(5, 600, 3, 6), 96, 54, 2, 4, 1, False, (0, 1)),
        "atrium": (lambda: S.atrium(1234, 20000, 6, 64), 240, 135, 2, 6, 2, True, (0, 1)),
        "instanced": (lambda: S.instanced(1234, 4, 64), 192, 108, 3, 4, 0, True, (0, 1)),
        "tiles": (lambda: S.atrium(1234, 20000, 6, 64), 200, 100, 1, 2, 3, True, (2, 3)),
    }[case]
    osc = orc.OracleScene(builder().pack(), True)
    cfg = S.render_config(W, H, spp, seg, rank=part[0], world=part[1], partition=1 if part[1] > 1 else 0)
    cam = S.camera_default(W, H, frame)
    a, _, rays_a = osc.render(cfg, cam, use_bvh)
    b, rays_b = osc.ref_render(cfg, cam, use_bvh)
    eq = bits_equal(a, b)
    assert eq.all(), f"{int((~eq.all(axis=2)).sum())} of {W * H} pixels differ"
    assert rays_a == rays_b and rays_a > W * H // part[1]


def test_hit_shading_live_against_reference_glsl(oracle_lib, ref_glsl):
    """10^5 random hits through the reference's closest_hit.rchit + getTriangleData + getMaterialData against
    oracle.cpp:get_triangle_data / get_material_data: uv, normal, tangent, geometric normal, material index, albedo,
    shading normal, emission, roughness, metallic — every word."""
    from parity_util import random_rays
    from tools import scene as S
    if not hasattr(ref_glsl, "ref_shade_hits"):
        pytest.skip("prebuilt oracle/_ref predates the full-pipeline build")
    total = 0
    for builder, extent in ((lambda: S.atrium(1234, 20000, 6, 64), 6.0), (lambda: S.instanced(1234, 4, 64), 30.0),
                            (lambda: S.random_soup(3, 3000, 3, 7), 8.0), (lambda: S.cornell(), 2.0)):
        osc = orc.OracleScene(builder().pack(), True)
        hits = osc.trace(random_rays(17, 60_000, extent, orc.PtRay_dtype))
        hits = hits[hits["instance"] != 0xFFFFFFFF]
        a, b = osc.shade_hits(hits), osc.ref_shade_hits(hits)
        assert len(hits) > 1000 and bits_equal(a, b).all()
        total += len(hits)
    assert total >= 100_000


def test_stated_tolerance_between_two_conforming_readings_of_the_reference(oracle_lib, ref_glsl):
    """north_star: "radiance must agree within a stated per-pixel relative tolerance, with converged-image mean error <= 1 %".
    Against the reference's shaders under the arithmetic contract (glsl_compat.h [contract] items) the oracle and the CUDA
    path are bit-exact — tolerance 0.  GLSL leaves normalize() and sin()/cos() precision open, so a second conforming
    reading (normalize = v / length(v), libm sinf/cosf; libref_glsl_alt.so) is rendered too and the drift between the two is
    the tolerance that any comparison with a real Vulkan driver has to allow.  Measured here and asserted with margin:
      * BASELINE's metric shape (1 spp, 2 segments): >= 99.8 % of pixels within 1e-4 relative, every image mean within 0.1 %;
      * deep paths (4 spp, 5 segments): image mean within 1 % — per-pixel agreement is NOT claimed there: path tracing
        amplifies a 1-ulp difference at every glossy bounce (atrium: 70 % of pixels within 1e-4, Cornell 99.8 %)."""
    from tools import scene as S
    if orc.ref_alt_lib() is None or not hasattr(ref_glsl, "ref_render"):
        pytest.skip("oracle/_ref/libref_glsl_alt.so not available")
    for builder, W, H in ((lambda: S.cornell(), 192, 192), (lambda: S.atrium(1234, 20000, 6, 64), 320, 180), (lambda: S.instanced(1234, 4, 64), 320, 180)):
        osc = orc.OracleScene(builder().pack(), True)
        for spp, seg in ((1, 2), (4, 5)):
            cfg, cam = S.render_config(W, H, spp, seg), S.camera_default(W, H, 3)
            a, _ = osc.ref_render(cfg, cam)
            b, _ = osc.ref_render(cfg, cam, alt=True)
            assert (np.isnan(a) == np.isnan(b)).all()
            a3, b3 = np.nan_to_num(a[..., :3].astype(np.float64)), np.nan_to_num(b[..., :3].astype(np.float64))
            rel = np.abs(a3 - b3).max(axis=2) / np.maximum(1e-3, np.maximum(a3.max(axis=2), b3.max(axis=2)))
            mean_err = abs(a3.mean() - b3.mean()) / a3.mean()
            if seg == 2:
                assert (rel <= 1e-4).mean() >= 0.998 and mean_err <= 1e-3
            else:
                assert mean_err <= 1e-2
