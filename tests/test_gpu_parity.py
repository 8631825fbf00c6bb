"""GPU parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle, the
golden fixtures and size-independent properties at full BASELINE.json sizes.

Bar: primary-hit ids bit-exact; radiance bit-exact too (the arithmetic contract of DESIGN.md makes
the fp32 expression trees identical: no FMA contraction, IEEE sqrt/div, one shared definition of sin/cos);
NaNs compare equal to NaNs.  north_star allows <=0.01 % edge-tie pixels and a radiance tolerance; the
tests hold the stricter bar and say so when they relax it."""
import functools
import os

import numpy as np
import pytest

import b200pt
from oracle import orc
from parity_util import bits_equal, hits_mismatch, random_rays
from tools import scene as S

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def render(ps, W, H, spp, seg, frame=0, **kw):
    r = b200pt.Renderer(ps, W, H, spp, seg, **kw)
    cam = S.camera_default(r.width * max(1, kw.get("divider", 1)), r.height * max(1, kw.get("divider", 1)), frame)
    r.record(cam)
    r.sync()
    out = r.image(), r.ids(), r.stats()
    r.deinit()
    return out


def assert_matches_oracle(ps, W, H, spp, seg, frame=0, use_bvh=True, **kw):
    img, ids, st = render(ps, W, H, spp, seg, frame, **kw)
    oimg, oids, orays = orc.OracleScene(ps, True).render(S.render_config(W, H, spp, seg), S.camera_default(W, H, frame), use_bvh)
    for name, a, b in zip(("instance", "primitive", "triangle"), ids, oids):
        assert (a == b).all(), f"{name} ids differ in {int((a != b).sum())} pixels"
    eq = bits_equal(img, oimg)
    assert eq.all(), f"{int((~eq.all(axis=2)).sum())} of {W * H} pixels differ from the oracle"
    return st, orays


# ------------------------------------------------------------------ shading restatement vs reference GLSL vectors
def test_bsdf_kernel_matches_reference_glsl_vectors():
    g = np.load(os.path.join(GOLD, "bsdf_ref_vectors.npz"))
    L, pdf, f = b200pt.debug_bsdf(g["N"], g["V"], g["kd"], g["f0"], g["rough"], g["xi"])
    assert bits_equal(L, g["L"]).all() and bits_equal(pdf, g["pdf"]).all() and bits_equal(f, g["f"]).all()


def test_bsdf_kernel_matches_oracle_on_a_million_records():
    from tools.make_golden import bsdf_inputs
    N, V, kd, f0, rough, xi = bsdf_inputs(4242, 1_000_000)
    a = b200pt.debug_bsdf(N, V, kd, f0, rough, xi)
    b = orc.bsdf_batch(orc.lib(), N, V, kd, f0, rough, xi, "orc")
    for name, u, v in zip(("L", "pdf", "f"), a, b):
        bad = ~bits_equal(u, v)
        # exact: sin/cos are include/b200pt_contract.h on both sides (explicit double arithmetic, no library call)
        assert bad.sum() == 0, f"{name}: {int(bad.sum())} words differ"


def test_rng_kernel_matches_reference_glsl_vectors():
    g = np.load(os.path.join(GOLD, "bsdf_ref_vectors.npz"))
    for k in range(3):
        state, xi = b200pt.debug_rng(g["px"], k + 1)
        assert (state == g["states"][:, k]).all() and bits_equal(xi, g["xis"][:, k]).all()
    state, _ = b200pt.debug_rng(g["px"], 0)
    assert (state == g["seeds"]).all()


def test_texture_filter_matches_oracle():
    import ctypes as C
    rng = np.random.default_rng(0)
    for (w, h) in ((2, 2), (5, 3), (64, 64), (1, 7)):
        tex = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        uv = ((rng.random((4000, 2)) - 0.3) * 3).astype(np.float32)
        uv[:8] = [[0, 0], [1, 1], [0.5, 0.5], [-1, -1], [1e6, -1e6], [0.25, 0.75], [1 / w, 1 / h], [3e9, 1]]
        got = b200pt.debug_texture(tex, uv)
        out = (C.c_float * 4)()
        ref = np.empty_like(got)
        for i, (u, v) in enumerate(uv):
            orc.lib().orc_sample_texture(tex.ctypes.data_as(C.c_void_p), w, h, C.c_float(u), C.c_float(v), out)
            ref[i] = list(out)
        assert bits_equal(got, ref).all()


def test_unorm8_conversion_is_exact():
    """texel / 255.0 is evaluated without the IEEE division sequence; it must equal it for all 256 bytes."""
    tex = np.zeros((1, 256, 4), np.uint8)
    v = np.arange(256, dtype=np.uint8)
    tex[0, :, 0] = v; tex[0, :, 1] = v[::-1]; tex[0, :, 2] = v ^ 0x55; tex[0, :, 3] = 255 - (v >> 1)
    uv = np.stack([(np.arange(256) + 0.5) / 256.0, np.full(256, 0.5)], axis=1).astype(np.float32)   # texel centres: weights 1,0,0,0
    got = b200pt.debug_texture(tex, uv)
    want = tex[0].astype(np.float32) / np.float32(255.0)
    assert bits_equal(got, want).all()


# ------------------------------------------------------------------ traversal kernel vs brute force
# builder stages: default = binned SAH + cost-optimal collapse; LBVH = Morton radix tree + greedy collapse (fast build)
LBVH, OPT, SAH, PLOC = b200pt.PT_FLAG_BVH_LBVH, b200pt.PT_FLAG_BVH_OPTIMAL_COLLAPSE, b200pt.PT_FLAG_BVH_SAH, b200pt.PT_FLAG_BVH_PLOC


@functools.lru_cache(maxsize=None)
def _brute_force_case(seed, ntri, nmesh, ninst):
    """Scene, rays and the oracle's brute-force hits (every ray against every triangle, no hierarchy): they do not
    depend on the builder / kernel flags under test, so they are computed once per scene."""
    ps = S.random_soup(seed, ntri, nmesh, ninst).pack()
    rays = random_rays(seed, 100000, dtype=b200pt.PtRay_dtype)
    # axis-parallel and zero-component directions, rays starting on geometry
    rays["dir"][:300] = np.tile(np.eye(3, dtype=np.float32), (100, 1))
    rays["dir"][300:600] = -np.tile(np.eye(3, dtype=np.float32), (100, 1))
    ho = orc.OracleScene(ps, False).trace(rays, use_bvh=False)
    rays.setflags(write=False); ho.setflags(write=False)
    return ps, rays, ho


@pytest.mark.parametrize("flags", [0, b200pt.PT_FLAG_TRACE_WALKER, b200pt.PT_FLAG_TRACE_BVH2, LBVH, LBVH | OPT, SAH, LBVH | b200pt.PT_FLAG_TRACE_WALKER,
                                   LBVH | b200pt.PT_FLAG_TRACE_BVH2, PLOC, PLOC | b200pt.PT_FLAG_TRACE_BVH2])
@pytest.mark.parametrize("seed,ntri,nmesh,ninst", [(3, 3000, 3, 7), (5, 3, 3, 4), (11, 20000, 2, 40), (12, 64, 5, 1)])
def test_traversal_matches_brute_force(flags, seed, ntri, nmesh, ninst):
    ps, rays, ho = _brute_force_case(seed, ntri, nmesh, ninst)
    r = b200pt.Renderer(ps, 32, 32, 1, 1, flags=flags)
    hg = r.trace(rays)
    bad = hits_mismatch(hg, ho)
    assert not bad.any(), f"{int(bad.sum())} of {len(rays)} rays differ from brute force"
    r.deinit()


@pytest.mark.parametrize("flags", [0, LBVH, SAH, PLOC])
def test_duplicated_geometry_builds_and_keeps_the_tie_break(flags):
    """6000 coincident triangles (two primitives, three instances at the same place): the builders must
    terminate on runs of identical boxes, and the closest hit is the lexicographically smallest
    (t, instance, primitive, triangle)."""
    sb = S.SceneBuilder(); sb.add_material()
    P = np.tile(np.array([[-1, -1, 2], [-1, 1, 2], [1, -1, 2]], np.float32), (1000, 1))
    N = np.tile(np.array([[0, 0, -1]], np.float32), (3000, 1)); T = np.tile(np.array([[1, 0, 0, 1]], np.float32), (3000, 1))
    UV = np.zeros((3000, 2), np.float32); I = np.arange(3000, dtype=np.uint32)
    s0 = S.Soup(); s0.add(P, N, T, UV, I.reshape(-1, 3)); s1 = S.Soup(); s1.add(P, N, T, UV, I.reshape(-1, 3))
    mesh = sb.add_mesh([s0.prim(0), s1.prim(0)])
    for _ in range(3):
        sb.add_node(mesh)
    ps = sb.pack()
    r = b200pt.Renderer(ps, 32, 32, 1, 1, flags=flags)
    rays = np.zeros(64, b200pt.PtRay_dtype)
    rays["origin"] = [-0.5, -0.25, 0]; rays["dir"] = [0, 0, 1]; rays["tmin"] = 1e-3; rays["tmax"] = 1e4
    h = r.trace(rays)
    ho = orc.OracleScene(ps, False).trace(rays, use_bvh=False)
    assert not hits_mismatch(h, ho).any()
    assert (h["instance"] == 0).all() and (h["primitive_index"] == 0).all() and (h["triangle_index"] == 0).all() and (h["t"] == 2.0).all()
    r.deinit()


def test_trace_edge_cases():
    ps = S.cornell().pack()
    r = b200pt.Renderer(ps, 32, 32, 1, 1)
    assert len(r.trace(np.zeros(0, b200pt.PtRay_dtype))) == 0
    rays = np.zeros(4, b200pt.PtRay_dtype)
    rays["origin"] = [0.9, 0.9, 0]; rays["dir"] = [0, 0, 1]       # sees only the back wall
    rays["tmin"] = [1e-3, 3.0, 1e-3, 2.9]; rays["tmax"] = [1e4, 1e4, 3.0, 3.1]     # back wall at z = 3: open interval
    h = r.trace(rays)
    ho = orc.OracleScene(ps, False).trace(rays, use_bvh=False)
    assert not hits_mismatch(h, ho).any()
    assert h["t"][1] == np.float32(1e32) and h["t"][2] == np.float32(1e32) and h["t"][3] == 3.0 and h["instance"][1] == 0xFFFFFFFF
    r.deinit()


def test_empty_scene_renders_sky():
    sb = S.SceneBuilder(); sb.add_material()
    ps = sb.pack()
    img, ids, st = render(ps, 64, 32, 2, 3)
    oimg, oids, _ = orc.OracleScene(ps, True).render(S.render_config(64, 32, 2, 3), S.camera_default(64, 32, 0), True)
    assert bits_equal(img, oimg).all() and (ids[0] == 0xFFFFFFFF).all()


# ------------------------------------------------------------------ the trace loop vs the oracle
def test_cornell_golden_fixture():
    g = np.load(os.path.join(GOLD, "cornell_64.npz"))
    img, ids, st = render(S.cornell().pack(), 64, 64, 2, 3, frame=5)
    assert bits_equal(img, g["img"]).all()
    assert (ids[0] == g["inst"]).all() and (ids[1] == g["prim"]).all() and (ids[2] == g["tri"]).all()


@pytest.mark.parametrize("name", ["atrium", "cornell", "instanced", "soup"])
def test_frames_equal_golden_frames_of_the_reference_raygen(name):
    """tests/golden/raygen_ref_frames.npz: whole frames rendered by the reference's own ray_gen.rgen main()
    (+ closest_hit.rchit, miss.rmiss, triangle/material/brdf/rand.glsl) compiled as C++ (tools/make_golden.py);
    the CUDA image equals them bit for bit, and so does the executed-ray count."""
    from tools.make_golden import REF_FRAMES
    g = np.load(os.path.join(GOLD, "raygen_ref_frames.npz"))
    builder, W, H, spp, seg, frame = REF_FRAMES[name]
    img, ids, st = render(builder().pack(), W, H, spp, seg, frame, flags=b200pt.PT_FLAG_NO_PRIMARY_REUSE)
    eq = bits_equal(img, g[name])
    assert eq.all(), f"{int((~eq.all(axis=2)).sum())} of {W * H} pixels differ from the reference raygen"
    assert st.rays_traced == int(g[name + "_rays"])     # without primary-hit reuse: one query per traceRayEXT of the reference
    img2, _, st2 = render(builder().pack(), W, H, spp, seg, frame)
    assert bits_equal(img2, g[name]).all() and st2.rays_traced < st.rays_traced


def test_config_c1_cornell_512():
    st, orays = assert_matches_oracle(S.cornell().pack(), 512, 512, 1, 2, use_bvh=False)
    assert st.rays_traced == orays          # 1 spp: no primary reuse, identical ray counts


@pytest.mark.parametrize("W,H,spp,seg,frame", [(96, 64, 4, 5, 0), (100, 60, 2, 3, 3), (33, 9, 3, 1, 1), (31, 7, 1, 8, 2)])
def test_cornell_sample_and_bounce_loops(W, H, spp, seg, frame):
    assert_matches_oracle(S.cornell().pack(), W, H, spp, seg, frame)


def test_random_soup_with_scaled_rotated_instances():
    # exercises the literal A^-1 "normal matrix" (quirk Q3/Q4) under rotation + non-uniform scale
    assert_matches_oracle(S.random_soup(3, 3000, 3, 7).pack(), 160, 90, 2, 4, use_bvh=False)


def test_instancing_scene_small():
    assert_matches_oracle(S.instanced(1234, 4, 64).pack(), 320, 180, 2, 4)


def test_material_sorted_shading_is_bit_identical():
    """PT_FLAG_SORT_BY_MATERIAL reorders the shading queue only: same image, same ids, same ray count."""
    ps = S.atrium(1234, 20000, 6, 64).pack()
    a_img, a_ids, a_st = render(ps, 200, 120, 3, 4, 1)
    b_img, b_ids, b_st = render(ps, 200, 120, 3, 4, 1, flags=b200pt.PT_FLAG_SORT_BY_MATERIAL)
    assert bits_equal(a_img, b_img).all() and all((x == y).all() for x, y in zip(a_ids, b_ids))
    assert a_st.rays_traced == b_st.rays_traced and b_st.kernel_launches > a_st.kernel_launches
    assert_matches_oracle(ps, 200, 120, 3, 4, 1, flags=b200pt.PT_FLAG_SORT_BY_MATERIAL)


def test_interleaved_material_textures_are_bit_identical():
    """Maps of one material that share an extent are interleaved per texel on the device (MatRec::atlas);
    materials with differently sized maps, or a single map, keep separate textures.  Same image either way."""
    sb = S.atrium(1234, 20000, 6, 64)
    rng = np.random.default_rng(9)
    small = sb.add_texture(rng.integers(0, 256, (16, 32, 4), dtype=np.uint8))      # 32x16: not the 64x64 of the others
    sb.materials[1]["normal_tex"] = small                                          # mixed extents -> fallback path
    sb.materials[2]["mr_tex"] = None; sb.materials[2]["normal_tex"] = None         # a single map -> fallback path
    ps = sb.pack()
    a_img, a_ids, _ = render(ps, 200, 120, 2, 4, 2)
    b_img, b_ids, _ = render(ps, 200, 120, 2, 4, 2, flags=b200pt.PT_FLAG_NO_TEXTURE_INTERLEAVE)
    assert bits_equal(a_img, b_img).all() and all((x == y).all() for x, y in zip(a_ids, b_ids))
    assert_matches_oracle(ps, 200, 120, 2, 4, 2)


def test_identity_and_transformed_instances_mix():
    """Identity instances are entered without re-deriving the ray frame (kInstIdentity); leaving a
    transformed instance must restore the world frame before the next identity instance is entered."""
    sb = S.random_soup(21, 1500, 3, 0)
    for k in range(3):
        sb.add_node(k)                                            # identity
        sb.add_node(k, translation=(0.3 * k - 0.2, 0.1, 0.4))     # pure translation: not identity
        sb.add_node(k, rotation=(0.0, 0.38268343, 0.0, 0.92387953), scale=(1.0, 0.7, 1.2))
    ps = sb.pack()
    assert_matches_oracle(ps, 160, 90, 2, 4, use_bvh=False)
    r = b200pt.Renderer(ps, 32, 32, 1, 1)
    rays = random_rays(5, 100000, dtype=b200pt.PtRay_dtype)
    assert not hits_mismatch(r.trace(rays), orc.OracleScene(ps, False).trace(rays, use_bvh=False)).any()
    r.deinit()


@pytest.fixture(scope="module")
def sponza():
    return S.sponza_scale().pack()


def test_config_c2_sponza_scale_1080p_vs_oracle(sponza):
    """BASELINE.json configs[1] at full size, against the multi-threaded oracle."""
    st, orays = assert_matches_oracle(sponza, 1920, 1080, 1, 2)
    assert st.rays_traced == orays and st.num_triangles == sponza.num_triangles


def assert_tile_sample_matches_oracle(ps, W, H, spp, seg, rank, world, frame=0):
    """Config-size parity through the tile partition as a sampler: the GPU renders the interleaved tiles
    `tile % world == rank` of the full-size frame (bit-identical to the same pixels of the whole frame by
    test_tile_partition_sums_to_single_gpu_image), the oracle renders the same (rank, world) share
    (oracle.cpp:orc_render) — ids and radiance bit-exact on every owned pixel, untouched elsewhere."""
    T = b200pt.PT_PARTITION_TILES
    img, ids, st = render(ps, W, H, spp, seg, frame, rank=rank, world=world, partition=T)
    cfg = S.render_config(W, H, spp, seg, rank=rank, world=world, partition=T)
    oimg, oids, orays = orc.OracleScene(ps, True).render(cfg, S.camera_default(W, H, frame), True)
    ty, tx = np.meshgrid(np.arange(H) // 8, np.arange(W) // 32, indexing="ij")
    owned = (ty * ((W + 31) // 32) + tx) % world == rank
    assert owned.sum() >= W * H // world - 32 * 8 * 2
    assert (oimg[owned][:, 3] == 1.0).all() and (oimg[~owned] == 0).all()
    for name, a, b in zip(("instance", "primitive", "triangle"), ids, oids):
        assert (a == b).all(), f"{name} ids differ in {int((a != b).sum())} pixels"
    eq = bits_equal(img, oimg)
    assert eq.all(), f"{int((~eq.all(axis=2)).sum())} of {int(owned.sum())} owned pixels differ from the oracle"
    return st, orays


def test_config_c3_4k_64spp_8seg_tile_sample_vs_oracle(sponza):
    """BASELINE.json configs[2] (3840x2160, 64 spp, 8 segments): tile share 17 of 64 against the oracle."""
    st, orays = assert_tile_sample_matches_oracle(sponza, 3840, 2160, 64, 8, 17, 64)
    assert st.rays_traced < orays           # the primary hit is traced once for the 64 samples (quirk Q1)


def test_config_c4_instanced_1000_tile_sample_vs_oracle():
    """BASELINE.json configs[3] (1 000 rotated instances of a 10 080-triangle mesh, 1080p, 4 spp, 4 segments): tile share 1 of 4."""
    ps = S.instanced(1234).pack()
    assert len(ps.instances) == 1000 and ps.num_effective_triangles > 9_000_000
    assert_tile_sample_matches_oracle(ps, 1920, 1080, 4, 4, 1, 4)


def test_config_c5_heavy_material_tile_sample_vs_oracle():
    """BASELINE.json configs[4] (~1 M triangles, 48 maps of 4096^2, 1080p, 16 spp, 6 segments): tile share 5 of 8."""
    ps = S.heavy_material(1234).pack()
    assert ps.num_triangles > 900_000 and max(t.shape[1] for t in ps.textures) == 4096
    assert_tile_sample_matches_oracle(ps, 1920, 1080, 16, 6, 5, 8)


def test_configs_c1_c2_full_size_vs_the_reference_shaders_live(sponza):
    """BASELINE.json configs[0] and configs[1] at full size against the REFERENCE's own shader sources compiled as C++
    (oracle/_ref/libref_glsl.so: one ray_gen.rgen main() per pixel, closest_hit.rchit / miss.rmiss behind traceRayEXT) —
    radiance bit for bit, and as many executed queries as the reference issues traceRayEXT calls."""
    if orc.ref_lib() is None or not hasattr(orc.ref_lib(), "ref_render"):
        pytest.skip("oracle/_ref/libref_glsl.so not available (built where /root/reference exists; it travels with the tree)")
    for ps, W, H in ((S.cornell().pack(), 512, 512), (sponza, 1920, 1080)):
        img, _, st = render(ps, W, H, 1, 2)
        rimg, rrays = orc.OracleScene(ps, True).ref_render(S.render_config(W, H, 1, 2), S.camera_default(W, H, 0), True)
        eq = bits_equal(img, rimg)
        assert eq.all(), f"{int((~eq.all(axis=2)).sum())} of {W * H} pixels differ from the reference shaders"
        assert st.rays_traced == rrays


def test_c2_reference_walker_and_no_reuse_equivalence(sponza):
    """Size-independent properties at 1080p: the BVH2 reference walker, and re-tracing the primary
    ray for every sample, give the same image as the BVH8 kernel with primary-hit reuse."""
    a, ia, sa = render(sponza, 1920, 1080, 3, 3, frame=1)
    b, ib, sb_ = render(sponza, 1920, 1080, 3, 3, frame=1, flags=b200pt.PT_FLAG_TRACE_BVH2)
    w, iw, sw = render(sponza, 1920, 1080, 3, 3, frame=1, flags=b200pt.PT_FLAG_TRACE_WALKER)
    assert bits_equal(a, w).all() and all((x == y).all() for x, y in zip(ia, iw))
    c, ic, sc = render(sponza, 1920, 1080, 3, 3, frame=1, flags=b200pt.PT_FLAG_NO_PRIMARY_REUSE)
    assert bits_equal(a, b).all() and bits_equal(a, c).all()
    assert all((x == y).all() and (x == z).all() for x, y, z in zip(ia, ib, ic))
    assert sc.rays_traced > sa.rays_traced        # reuse saves (S-1) primary rays per pixel


def test_tile_partition_sums_to_single_gpu_image(sponza):
    """N-rank tile partitioning emulated on one GPU (ranks run one after the other): the sum of the
    per-rank images equals the 1-rank image bit for bit, and every pixel is owned once."""
    W, H = 1000, 500    # ragged edge tiles (1000 % 32 != 0, 500 % 8 != 0)
    full, _, _ = render(sponza, W, H, 2, 3, frame=2)
    for world in (2, 3, 8):
        acc = np.zeros_like(full); owned = np.zeros((H, W), np.int32)
        for rank in range(world):
            img, _, _ = render(sponza, W, H, 2, 3, frame=2, rank=rank, world=world, partition=b200pt.PT_PARTITION_TILES)
            acc += img; owned += (img[..., 3] != 0)
        assert (owned == 1).all()
        assert bits_equal(acc, full).all()


def test_ray_counts_and_stats(sponza):
    img, ids, st = render(sponza, 640, 360, 1, 2)
    hit = int((ids[0] != 0xFFFFFFFF).sum())
    assert 640 * 360 <= st.rays_traced <= 640 * 360 + hit
    assert hit <= st.rays_hit < st.rays_traced
    _, ids1, st1 = render(sponza, 640, 360, 1, 1)                 # one segment: the hits are the primary hits
    assert st1.rays_traced == 640 * 360 and st1.rays_hit == int((ids1[0] != 0xFFFFFFFF).sum())
    assert st.kernel_launches == 1 + 2 * 2 and st.bvh_nodes > 0 and st.ms_build > 0


# ------------------------------------------------------------------ outputs and the rest of the RaytracingPass surface
def test_bgra8_is_the_reference_storage_format():
    ps = S.cornell().pack()
    r = b200pt.Renderer(ps, 128, 64, 2, 3)
    r.record(S.camera_default(128, 64, 0)); r.sync()
    img = r.image(); b = r.image_bgra8()
    q = (np.clip(np.nan_to_num(img, nan=0.0), 0, 1) * np.float32(255.0) + np.float32(0.5)).astype(np.uint8)
    assert np.array_equal(b[..., 0], q[..., 2]) and np.array_equal(b[..., 1], q[..., 1]) and np.array_equal(b[..., 2], q[..., 0])
    assert (b[..., 3] == 255).all()
    r.deinit()


def test_divider_blit_samples_texel_centres_on_non_divisible_extents():
    """vkCmdBlitImage with VK_FILTER_NEAREST (RaytracingPass.zig:599-635) reads the source texel under the destination texel's
    centre, floor((x + 0.5) * sw / dw); with 1921 / 2 -> 960 that differs from floor(x * sw / dw) from x = 2 on."""
    ps = S.cornell().pack()
    r = b200pt.Renderer(ps, 1921, 101, 1, 1, divider=2)
    assert (r.width, r.height) == (960, 50)
    r.record(S.camera_default(1921, 101, 0)); r.sync()
    small, up = r.image(), r.image_upscaled()
    assert up.shape == (101, 1921, 4)
    sx = np.minimum(959, ((2 * np.arange(1921) + 1) * 960) // (2 * 1921))
    sy = np.minimum(49, ((2 * np.arange(101) + 1) * 50) // (2 * 101))
    assert sx[2] == 1 and (2 * 960) // 1921 == 0                                    # the case the advisor named
    assert np.array_equal(up, small[sy][:, sx])
    r.deinit()


def test_russian_roulette_extension_matches_the_oracle_and_is_unbiased():
    """PT_FLAG_RUSSIAN_ROULETTE (north_star (c); SURVEY Q11: opt-in, default off, excluded from parity with the reference).
    GPU == oracle bit for bit with the flag on both sides; fewer rays; the mean over frames stays within noise."""
    ps = S.cornell().pack()
    W, H, spp, seg = 96, 96, 4, 8
    RR = b200pt.PT_FLAG_RUSSIAN_ROULETTE
    img, ids, st = render(ps, W, H, spp, seg, 1, flags=RR)
    cfg = S.render_config(W, H, spp, seg, flags=RR)
    oimg, oids, orays = orc.OracleScene(ps, True).render(cfg, S.camera_default(W, H, 1), True)
    assert bits_equal(img, oimg).all() and all((a == b).all() for a, b in zip(ids, oids))
    plain, _, st0 = render(ps, W, H, spp, seg, 1)
    assert st.rays_traced < st0.rays_traced and not bits_equal(img, plain).all()
    acc = {0: np.zeros(3), RR: np.zeros(3)}
    for flags in acc:
        r = b200pt.Renderer(ps, W, H, 16, seg, flags=flags)
        for f in range(8):
            r.record(S.camera_default(W, H, f)); r.sync()
            acc[flags] += np.nan_to_num(r.image()[..., :3]).reshape(-1, 3).mean(axis=0)
        r.deinit()
    assert np.allclose(acc[RR], acc[0], rtol=0.02)


def test_hw_bilinear_texture_objects_within_the_stated_tolerance():
    """PT_FLAG_HW_BILINEAR (A/B switch, north_star (c) / SURVEY H4): material maps through cudaTextureObject_t.  The texture
    units filter with 1.8 fixed-point weights, so the image is not bit-identical to the reference arithmetic (four texel
    fetches + fp32 lerp); stated tolerance at BASELINE's metric shape (1 spp, 2 segments): >= 97 % of the pixels within
    2e-2 relative, image mean within 0.5 %; ids are untouched (primary hits do not read textures)."""
    ps = S.atrium(1234, 20000, 6, 64).pack()
    W, H = 640, 360
    a, ia, _ = render(ps, W, H, 1, 2, 1)
    b, ib, _ = render(ps, W, H, 1, 2, 1, flags=b200pt.PT_FLAG_HW_BILINEAR)
    assert all((x == y).all() for x, y in zip(ia, ib))
    a3, b3 = np.nan_to_num(a[..., :3].astype(np.float64)), np.nan_to_num(b[..., :3].astype(np.float64))
    rel = np.abs(a3 - b3).max(axis=2) / np.maximum(1e-3, np.maximum(a3.max(axis=2), b3.max(axis=2)))
    print("hw bilinear: within 2e-2:", (rel <= 2e-2).mean(), "within 1e-3:", (rel <= 1e-3).mean(), "mean err:", abs(a3.mean() - b3.mean()) / a3.mean())
    assert not bits_equal(a, b).all()
    assert (rel <= 2e-2).mean() >= 0.97 and abs(a3.mean() - b3.mean()) / a3.mean() <= 5e-3


def test_divider_upscale_and_resize():
    ps = S.cornell().pack()
    r = b200pt.Renderer(ps, 256, 128, 1, 2, divider=2)
    assert (r.width, r.height) == (128, 64)                  # storage image = extent / divider (RaytracingPass.zig:149-152)
    cam = S.camera_default(256, 128, 0)                      # aspect from the window extent (main.zig:222)
    r.record(cam); r.sync()
    small = r.image(); up = r.image_upscaled()
    assert up.shape == (128, 256, 4)
    assert np.array_equal(up, np.repeat(np.repeat(small, 2, axis=0), 2, axis=1))     # nearest blit (:599-635)
    osc = orc.OracleScene(ps, True)
    oimg, _, _ = osc.render(S.render_config(256, 128, 1, 2, divider=2), cam, True)
    assert bits_equal(small, oimg).all()
    r.resize(96, 48)                                          # resize drops the divider (quirk Q13)
    assert (r.width, r.height) == (96, 48)
    cam2 = S.camera_default(96, 48, 1)
    r.record(cam2); r.sync()
    oimg2, _, _ = osc.render(S.render_config(96, 48, 1, 2), cam2, True)
    assert bits_equal(r.image(), oimg2).all()
    r.deinit()


def test_progressive_accumulation_converges_to_oracle_mean():
    """Converged-image criterion of north_star: mean over frames within 1 % of the oracle's mean
    (here exactly equal per frame, so the means agree to rounding)."""
    ps = S.cornell().pack()
    W, H, frames = 64, 64, 16
    r = b200pt.Renderer(ps, W, H, 1, 3)
    r.set_accumulate(True)
    osc = orc.OracleScene(ps, True)
    acc = np.zeros((H, W, 4), np.float32)
    for f in range(frames):
        cam = S.camera_default(W, H, f)
        r.record(cam)
        acc += osc.render(S.render_config(W, H, 1, 3), cam, True)[0]
    mean = r.accum_mean()
    assert bits_equal(mean, acc * np.float32(1.0 / frames)).all()
    assert abs(np.nanmean(mean[..., :3]) / np.nanmean((acc / frames)[..., :3]) - 1) < 0.01
    r.deinit()


def test_render_is_asynchronous_and_repeatable():
    ps = S.cornell().pack()
    r = b200pt.Renderer(ps, 200, 100, 2, 4)
    cams = [S.camera_default(200, 100, f) for f in range(3)]
    for c in cams:                       # queued back to back, no sync in between
        r.record(c)
    r.sync()
    last = r.image()
    r.record(cams[2]); r.sync()
    assert bits_equal(last, r.image()).all()
    r.deinit()


def test_pipelined_readback_returns_every_frame_intact():
    """pt_read_rgba32f_overlapped: the copy of frame k runs beside frame k+1; no frame may be torn."""
    import torch
    ps = S.atrium(1234, 20000, 6, 64).pack()
    W, H = 640, 360
    r = b200pt.Renderer(ps, W, H, 1, 3)
    want = []
    for f in range(4):
        r.record(S.camera_default(W, H, f)); r.sync(); want.append(r.image().copy())
    bufs = [torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True) for _ in range(4)]
    for rep in range(3):
        for f in range(4):
            r.record(S.camera_default(W, H, f))
            b200pt.check(b200pt.lib().pt_read_rgba32f_overlapped(r.handle, bufs[f].data_ptr()))
        r.sync()
        for f in range(4):
            assert bits_equal(bufs[f].numpy(), want[f]).all(), f"frame {f} torn in repetition {rep}"
    r.deinit()


def test_invalid_scene_is_rejected():
    import ctypes as C
    ps = S.cornell().pack()
    d = ps.desc()
    bad = (S.PtPrimitive * len(ps.primitives))(*ps.primitives)
    bad[0].positions_offset = ps.triangle_data.size            # out of the blob
    d.primitives = C.cast(bad, C.POINTER(S.PtPrimitive))

    class Fake:
        def desc(self):
            return d
    with pytest.raises(b200pt.PtError) as e:
        b200pt.Renderer(Fake(), 32, 32)
    assert e.value.status == -1


def test_traversal_byte_counter(sponza):
    r = b200pt.Renderer(sponza, 640, 360, 1, 2)
    r.record(S.camera_default(640, 360, 0)); r.sync()
    before = r.image()
    nodes, tris, rays = r.count_traversal()
    assert rays == r.stats().rays_traced and nodes > rays and tris > 0
    assert bits_equal(before, r.image()).all()      # the counting walker renders the same frame
    r.deinit()


def test_lean_primary_records_survive_trace_mode_switches(sponza):
    """K-raygen leaves out the ray records nobody reads (origin: launch constant of the flat K-trace; direction with S == 1).
    The counting walker reads them, so a frame rendered through it must write them again, and the next frame of the
    persistent kernel must not depend on what the walker left behind; the walker-only renderer gives the same image."""
    cam = S.camera_default(320, 200, 0)
    for spp in (1, 3):
        r = b200pt.Renderer(sponza, 320, 200, spp, 3)
        r.record(cam); r.sync(); first = r.image()
        r.count_traversal()
        assert bits_equal(first, r.image()).all()
        r.record(S.camera_default(320, 200, 1)); r.sync()       # another frame in between
        r.record(cam); r.sync()
        assert bits_equal(first, r.image()).all()
        r.deinit()
        w = b200pt.Renderer(sponza, 320, 200, spp, 3, flags=b200pt.PT_FLAG_TRACE_WALKER)
        w.record(cam); w.sync()
        assert bits_equal(first, w.image()).all()
        w.deinit()


def test_trace_rays_rejects_counts_beyond_the_slot_index_range():
    """K-trace keeps a flag in bit 31 of a ray's slot index, so pt_trace_rays takes fewer than 2^31 rays (include/b200pt.h);
    the count is checked before either buffer is touched."""
    import ctypes as C
    r = b200pt.Renderer(S.cornell().pack(), 16, 16, 1, 1)
    dummy = np.zeros(1, b200pt.PtRay_dtype)
    out = np.zeros(1, b200pt.PtHit_dtype)
    status = b200pt.lib().pt_trace_rays(r.handle, dummy.ctypes.data, C.c_size_t(1 << 31), out.ctypes.data)
    assert status == -1 and b"too many rays" in b200pt.lib().pt_last_error()      # PT_ERR_INVALID_ARGUMENT
    assert len(r.trace(random_rays(7, 64, dtype=b200pt.PtRay_dtype))) == 64     # the renderer is still usable
    r.deinit()


# ------------------------------------------------------------------ animated instances / camera replay
def _moved_instances(ps, seed):
    """A copy of `ps` whose instances got new rigid transforms (rows of M), plus the (n, 16) matrix array."""
    import copy
    import ctypes as C
    rng = np.random.default_rng(seed)
    n = len(ps.instances)
    T = np.zeros((n, 16), np.float32)
    for i in range(n):
        M = np.array(list(ps.instances[i].transform), np.float32).reshape(4, 4)
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        R = S._quat_to_mat4(q[3], q[0], q[1], q[2]).astype(np.float32) if i % 5 else np.eye(4, dtype=np.float32)
        R[:3, 3] = (rng.random(3) - 0.5).astype(np.float32) * 0.8
        T[i] = S._matmul_f32(R, M).reshape(16)
    if n > 2:                                   # one instance becomes the exact identity, another leaves it
        T[1] = np.eye(4, dtype=np.float32).reshape(16)
    ps1 = copy.copy(ps)
    ps1.instances = (S.PtInstance * n)()
    for i in range(n):
        ps1.instances[i].mesh_index = ps.instances[i].mesh_index
        ps1.instances[i].transform = (C.c_float * 16)(*T[i].tolist())
    return ps1, T


def test_animated_instances_equal_a_fresh_renderer_and_the_oracle():
    ps0 = S.instanced(1234, 4, 64).pack()
    W, H, spp, seg = 320, 180, 2, 3
    ps1, T1 = _moved_instances(ps0, 3)
    T0 = np.array([list(i.transform) for i in ps0.instances], np.float32)
    r = b200pt.Renderer(ps0, W, H, spp, seg)
    cam = S.camera_default(W, H, 0)
    r.record(cam); r.sync(); img0, ids0 = r.image().copy(), [a.copy() for a in r.ids()]
    ms = r.set_instance_transforms(T1)                           # Tlas.init again, on the device
    assert 0.0 < ms < 50.0
    r.record(cam); r.sync(); img1, ids1 = r.image().copy(), [a.copy() for a in r.ids()]
    fresh_img, fresh_ids, _ = render(ps1, W, H, spp, seg)
    assert bits_equal(img1, fresh_img).all() and all((a == b).all() for a, b in zip(ids1, fresh_ids))
    oimg, oids, _ = orc.OracleScene(ps1, True).render(S.render_config(W, H, spp, seg), cam, True)
    assert bits_equal(img1, oimg).all() and all((a == b).all() for a, b in zip(ids1, oids))
    assert not bits_equal(img1, img0).all()                      # the scene did move
    # traversal through the rebuilt top level: random rays, every kernel variant agrees with the oracle
    rays = random_rays(11, 50_000, 30.0, b200pt.PtRay_dtype)
    assert not hits_mismatch(r.trace(rays), orc.OracleScene(ps1, True).trace(rays)).any()
    r.set_instance_transforms(T0)                                # and back
    r.record(cam); r.sync()
    assert bits_equal(r.image(), img0).all() and all((a == b).all() for a, b in zip(r.ids(), ids0))
    with pytest.raises(b200pt.PtError):
        r.set_instance_transforms(T0[:-1])                       # instance count is fixed by the scene
    r.deinit()


def test_cli_replays_a_camera_path(tmp_path):
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vulkan-pathtracer_b200", "b200pt")
    sb = S.cornell(1234)
    cwd = os.getcwd(); os.chdir(tmp_path)
    try:
        path = sb.write_gltf(str(tmp_path), "cornell", glb=True)
    finally:
        os.chdir(cwd)
    lines = ["0.0 - 0 0 0", "0.05 w 0 0 0", "0.05 w+left_shift 1 12 -6", "0.05 d+space 1 -3 2"]
    (tmp_path / "path.txt").write_text("# dt keys mouse dx dy\n" + "\n".join(lines) + "\n")
    out = tmp_path / "out.pfm"
    r = subprocess.run([exe, "-s", path, "-x", "96", "-y", "64", "-c", "2", "-b", "3", "--frames", "4", "--camera-path",
                        str(tmp_path / "path.txt"), "-o", str(out), "--quiet"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr
    cam = b200pt.Camera(np.float32(np.pi * 0.25), np.float32(96) / np.float32(64))
    for ln in lines:
        dt, keys, mouse, dx, dy = ln.split()
        cam.update(keys=() if keys == "-" else keys.split("+"), mouse_left=mouse == "1", cursor_delta=(int(dx), int(dy)), delta_time=float(dt))
    ls = b200pt.LoadedScene(path)
    ren = b200pt.Renderer(ls, 96, 64, 2, 3)
    ren.record(cam.pack(3)); ren.sync()
    want = ren.image()[..., :3]
    with open(out, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(v) for v in f.readline().split()); f.readline()
        got = np.frombuffer(f.read(), "<f4").reshape(h, w, 3)[::-1]
    assert bits_equal(got, want).all()
    assert cam.state.position[2] > 0.7 and cam.state.yaw != 0.0   # the path did move the camera
    ren.deinit()


def test_single_instance_scene_skips_the_top_level_only_when_identity():
    """One identity instance starts traversal at its BLAS root (DeviceScene::start_root); a moved one does not."""
    import copy
    import ctypes as C
    ps = S.random_soup(33, 2500, 1, 1).pack()                    # one instance, translated: general path
    W, H, spp, seg = 160, 90, 2, 3
    cam = S.camera_default(W, H, 1, position=(0.0, 0.0, -6.0))
    eye = np.eye(4, dtype=np.float32).reshape(1, 16)
    ps_id = copy.copy(ps)
    ps_id.instances = (S.PtInstance * 1)()
    ps_id.instances[0].mesh_index = ps.instances[0].mesh_index
    ps_id.instances[0].transform = (C.c_float * 16)(*eye[0].tolist())
    T0 = np.array([list(ps.instances[0].transform)], np.float32)
    rays = random_rays(3, 100_000, 6.0, b200pt.PtRay_dtype)
    r = b200pt.Renderer(ps, W, H, spp, seg)
    for scene, T in ((ps, None), (ps_id, eye), (ps, T0)):
        if T is not None:
            r.set_instance_transforms(T)
        osc = orc.OracleScene(scene, True)
        r.record(cam); r.sync()
        oimg, oids, _ = osc.render(S.render_config(W, H, spp, seg), cam, True)
        assert bits_equal(r.image(), oimg).all() and all((a == b).all() for a, b in zip(r.ids(), oids))
        assert not hits_mismatch(r.trace(rays), osc.trace(rays)).any()
    r.deinit()
    fresh_img, fresh_ids, _ = render(ps_id, W, H, spp, seg, 1)   # created as identity from the start (camera at 0)
    oimg, oids, _ = orc.OracleScene(ps_id, True).render(S.render_config(W, H, spp, seg), S.camera_default(W, H, 1), True)
    assert bits_equal(fresh_img, oimg).all() and all((a == b).all() for a, b in zip(fresh_ids, oids))


@pytest.mark.parametrize("builder", [0, LBVH])
def test_builder_is_deterministic_across_builds(builder):
    """Same scene, repeated builds with other work in between: identical digests of every builder stage (the SAH stage
    derives node ids from split positions and partitions stably, so it does not depend on atomic order either)."""
    ps = S.random_soup(3, 3000, 3, 7).pack()
    pc = S.cornell().pack()
    seen = set()
    for k in range(12):
        if k % 3 == 0:
            rc = b200pt.Renderer(pc, 96, 64, 1 + k % 2, 2); rc.record(S.camera_default(96, 64, k)); rc.sync(); rc.deinit()
        r = b200pt.Renderer(ps, 100, 60, 1, 2, flags=builder | (0, b200pt.PT_FLAG_TRACE_BVH2)[k & 1])
        seen.add(r.bvh_digest())
        r.deinit()
    assert len(seen) == 1, f"{len(seen)} different acceleration structures from the same input"


def test_image_does_not_depend_on_the_hierarchy():
    """The closest-hit rule is order-independent, so every builder gives the same image, ids and ray count."""
    ps = S.atrium(1234, 20000, 6, 64).pack()
    ref = None
    for flags in (0, LBVH, LBVH | OPT, SAH, PLOC):
        img, ids, st = render(ps, 256, 144, 2, 3, 1, flags=flags)
        if ref is None:
            ref = (img, ids, st.rays_traced)
        assert bits_equal(img, ref[0]).all() and all((a == b).all() for a, b in zip(ids, ref[1])) and st.rays_traced == ref[2]


def test_zero_throughput_last_segments_are_skipped_without_changing_a_bit():
    """PT_FLAG_SKIP_DEAD_SEGMENTS: same image and ids, fewer executed queries — and exactly as many as the oracle under the
    same flag (the saving is proven in the restatement too, tests/test_oracle_trace.py)."""
    SK = b200pt.PT_FLAG_SKIP_DEAD_SEGMENTS
    for ps, W, H, spp, seg in ((S.atrium(1234, 20000, 6, 64).pack(), 320, 180, 2, 3), (S.cornell().pack(), 128, 128, 3, 2)):
        a, ia, sa = render(ps, W, H, spp, seg, 2)
        b, ib, sb_ = render(ps, W, H, spp, seg, 2, flags=SK)
        assert bits_equal(a, b).all() and all((x == y).all() for x, y in zip(ia, ib))
        assert sb_.rays_traced <= sa.rays_traced
        c, _, sc = render(ps, W, H, spp, seg, 2, flags=SK | b200pt.PT_FLAG_NO_PRIMARY_REUSE)
        _, _, orays = orc.OracleScene(ps, True).render(S.render_config(W, H, spp, seg, flags=SK), S.camera_default(W, H, 2), True)
        assert bits_equal(a, c).all() and sc.rays_traced == orays
    assert sb_.rays_traced <= sa.rays_traced


def test_sibling_context_shares_the_scene_and_keeps_two_frames_in_flight():
    """pt_create_sibling: same device scene, own stream / path state / image; frames alternate between the two."""
    ps = S.atrium(1234, 20000, 6, 64).pack()
    W, H, spp, seg = 320, 180, 2, 3
    a = b200pt.Renderer(ps, W, H, spp, seg)
    b = a.sibling()
    want = []
    for f in range(4):
        a.record(S.camera_default(W, H, f)); a.sync(); want.append(a.image().copy())
    # two frames in flight: submit on a and b back to back, read each after its own fence
    for f in (0, 2):
        a.record(S.camera_default(W, H, f)); b.record(S.camera_default(W, H, f + 1))
        a.sync(); b.sync()
        assert bits_equal(a.image(), want[f]).all() and bits_equal(b.image(), want[f + 1]).all()
    rays = random_rays(9, 20000, 10.0, b200pt.PtRay_dtype)
    assert not hits_mismatch(a.trace(rays), b.trace(rays)).any()
    with pytest.raises(b200pt.PtError):
        b.set_instance_transforms(np.eye(4, dtype=np.float32).reshape(1, 16))     # the scene belongs to the parent
    # the parent moves its (single, identity) instance: the sibling follows at its next frame
    T = np.eye(4, dtype=np.float32); T[0, 3] = 0.75; T[2, 3] = -0.5
    a.set_instance_transforms(T.reshape(1, 16))
    cam = S.camera_default(W, H, 5)
    a.record(cam); b.record(cam); a.sync(); b.sync()
    assert bits_equal(a.image(), b.image()).all() and not bits_equal(a.image(), want[0]).all()
    b.deinit(); a.deinit()


def test_create_from_gltf_equals_load_then_create(tmp_path):
    sb = S.cornell(1234)
    cwd = os.getcwd(); os.chdir(tmp_path)
    try:
        path = sb.write_gltf(str(tmp_path), "cornell", glb=True)
    finally:
        os.chdir(cwd)
    a = b200pt.Renderer.from_gltf(path, 96, 64, 2, 3)
    b = b200pt.Renderer(sb.pack(), 96, 64, 2, 3)
    cam = S.camera_default(96, 64, 1)
    a.record(cam); b.record(cam); a.sync(); b.sync()
    assert bits_equal(a.image(), b.image()).all() and all((x == y).all() for x, y in zip(a.ids(), b.ids()))
    a.deinit(); b.deinit()
