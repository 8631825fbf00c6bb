"""The host data contract of Scene.load (Scene.zig:160-546) as restated by tools/scene.py."""
import numpy as np

from tools import scene as S


def test_triangle_data_sections_and_padding():
    ps = S.random_soup(1, 600, 3, 4).pack()
    prims = list(ps.primitives)
    sizes = [0] * 5
    largest = 0
    for p in prims:
        nv = p.max_vertex + 1
        isz = 4 if p.info & 0x80000000 else 2
        acc = [3 * p.triangle_count * isz, 12 * nv, 12 * nv, 16 * nv, 8 * nv]
        for k in range(5):
            sizes[k] += acc[k]
        largest = max(largest, max(acc))
    pad = 1 << (largest - 1).bit_length()                    # ceilPowerOfTwo (Scene.zig:209)
    offs, total = [], 0
    for k in range(5):
        offs.append(total)
        total += (sizes[k] + pad - 1) // pad * pad            # alignForward (Scene.zig:215)
    assert ps.triangle_data.size == total
    # first primitive starts every section; later ones are packed tightly in (mesh, primitive) order
    p0 = prims[0]
    assert [p0.indices_offset, p0.positions_offset, p0.normals_offset, p0.tangents_offset, p0.uvs_offset] == offs
    for a, b in zip(prims, prims[1:]):
        assert b.positions_offset == a.positions_offset + 12 * (a.max_vertex + 1)
        assert b.uvs_offset == a.uvs_offset + 8 * (a.max_vertex + 1)
        isz_b = 4 if b.info & 0x80000000 else 2
        assert b.indices_offset % isz_b == 0                  # Scene.zig:236-240
    # mesh ranges are consecutive
    assert ps.meshes[0].start == 0 and all(ps.meshes[i].end == ps.meshes[i + 1].start for i in range(len(ps.meshes) - 1))


def test_material_quantisation_truncates():
    sb = S.SceneBuilder()
    sb.add_material((0.5, 0.999, 1.0), roughness=0.3, metallic=0.7, emissive=(0.2, 0.0, 1.0), albedo_tex=None)
    s = S.Soup(); S.grid_patch(s, (0, 0, 1), (1, 0, 0), (0, 1, 0), 1, 1)
    sb.add_node(sb.add_mesh([s.prim(0)]))
    m = sb.pack().materials[0]
    assert m.albedo_factor == (127 | 254 << 8 | 255 << 16)                  # @intFromFloat(x*255) truncates, alpha 0
    assert m.metal_roughness_factor == (int(np.float32(0.3) * np.float32(255)) << 8 | int(np.float32(0.7) * np.float32(255)) << 16)
    assert m.emissive_factor == (51 | 0 << 8 | 255 << 16)
    assert m.albedo_texture_index == 0xFFFFFFFF and m.normal_texture_index == 0xFFFFFFFF


def test_instances_lifo_dfs_and_mesh_only_trs():
    sb = S.SceneBuilder()
    sb.add_material()
    s = S.Soup(); S.grid_patch(s, (0, 0, 1), (1, 0, 0), (0, 1, 0), 1, 1)
    mesh = sb.add_mesh([s.prim(0)])
    root = sb.add_node(None, translation=(100, 0, 0))             # mesh-less: its TRS is ignored (quirk Q7)
    a = sb.add_node(mesh, translation=(1, 0, 0), parent=root)
    b = sb.add_node(mesh, translation=(2, 0, 0), parent=root)
    c = sb.add_node(mesh, translation=(0, 5, 0), parent=a)
    inst = sb.pack().instances
    T = [np.array(i.transform).reshape(4, 4) for i in inst]
    # LIFO: children pushed a, b -> popped b first, then a, then a's child c
    assert [tuple(t[:3, 3]) for t in T] == [(2, 0, 0), (1, 0, 0), (1, 5, 0)]
    assert all(np.array_equal(t[:3, :3], np.eye(3)) for t in T)


def test_configs_have_expected_scale():
    c1 = S.cornell().pack()
    assert 30 <= c1.num_triangles <= 50 and len(c1.instances) == 3
    emissive = [m for m in c1.materials if m.emissive_factor]
    assert len(emissive) == 1
    c2 = S.sponza_scale(textures=False).pack()
    assert abs(c2.num_triangles - 260_000) <= 0.01 * 260_000
    assert len(c2.materials) == 26    # 25 PBR + the emissive quads
    assert any(p.info & 0x80000000 for p in c2.primitives) and any(not (p.info & 0x80000000) for p in c2.primitives)


def test_instanced_config_counts():
    c4 = S.instanced(1234, 10, 32).pack()
    assert len(c4.instances) == 1000
    assert 9_000_000 <= c4.num_effective_triangles <= 11_000_000
