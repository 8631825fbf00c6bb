"""pt_scene_load == Scene.load (Scene.zig:79-546): the native glTF loader must reproduce, byte for byte,
the host arrays tools/scene.py::SceneBuilder.pack() states for the same scene, from .gltf (+ .bin + .png)
and from .glb; error names mirror the Zig error set.  CPU only (no CUDA call is made)."""
import ctypes as C
import io
import json
import os
import struct

import numpy as np
import pytest

import b200pt
from tools import scene as S


def _as_bytes(ptr, n):
    return bytes(C.cast(ptr, C.POINTER(C.c_uint8 * n)).contents) if n else b""


def _compare(desc, ps):
    assert desc.triangle_data_len == ps.triangle_data.size
    assert _as_bytes(desc.triangle_data, desc.triangle_data_len) == ps.triangle_data.tobytes()
    assert desc.num_primitives == len(ps.primitives)
    for i, p in enumerate(ps.primitives):
        q = desc.primitives[i]
        for f, _ in S.PtPrimitive._fields_:
            assert getattr(q, f) == getattr(p, f), (i, f)
    assert desc.num_meshes == len(ps.meshes)
    assert [(desc.meshes[i].start, desc.meshes[i].end) for i in range(desc.num_meshes)] == [(m.start, m.end) for m in ps.meshes]
    assert desc.num_materials == len(ps.materials)
    for i, m in enumerate(ps.materials):
        for f, _ in S.PtMaterial._fields_:
            assert getattr(desc.materials[i], f) == getattr(m, f), (i, f)
    assert desc.num_textures == len(ps.textures)
    for i, t in enumerate(ps.textures):
        q = desc.textures[i]
        assert (q.width, q.height) == (t.shape[1], t.shape[0])
        assert _as_bytes(q.rgba, t.size) == t.tobytes()
    assert desc.num_instances == len(ps.instances)
    for i, inst in enumerate(ps.instances):
        assert desc.instances[i].mesh_index == inst.mesh_index
        a = np.array(desc.instances[i].transform[:], np.float32).view(np.uint32)
        b = np.array(inst.transform[:], np.float32).view(np.uint32)
        assert (a == b).all(), f"instance {i} transform bits differ"


def _scene_with_everything():
    """Textured materials, u16 + u32 primitives, a node hierarchy with rotation and scale, a mesh-less parent."""
    sb = S.random_soup(11, 900, 3, 5)
    rng = np.random.default_rng(5)
    t0 = sb.add_texture(rng.integers(0, 256, (16, 8, 4), dtype=np.uint8))
    t1 = sb.add_texture(rng.integers(0, 256, (5, 7, 4), dtype=np.uint8))
    m = sb.add_material((0.25, 0.5, 0.999), 0.37, 0.61, emissive=(0.1, 0.2, 0.3), albedo_tex=t0, mr_tex=t1, emissive_tex=t0, normal_tex=t1)
    s = S.Soup(); S.grid_patch(s, (0, 0, 2), (1, 0, 0), (0, 1, 0), 3, 2)
    mesh = sb.add_mesh([s.prim(m), s.prim(m, force_u32=True)])
    parent = sb.add_node(None, translation=(9, 9, 9))
    a = sb.add_node(mesh, translation=(0.5, -0.25, 1.0), rotation=(0.1, 0.2, 0.3, 0.9), scale=(1.5, 0.5, 2.0), parent=parent)
    sb.add_node(mesh, translation=(0.1, 0.2, 0.3), rotation=(0.0, 0.7071, 0.0, 0.7071), parent=a)
    return sb


@pytest.mark.parametrize("glb", [False, True])
def test_loader_matches_packed_scene(tmp_path, monkeypatch, glb):
    monkeypatch.chdir(tmp_path)        # URIs resolve against the cwd (Scene.zig:125,414 — quirk Q9)
    sb = _scene_with_everything()
    path = sb.write_gltf(str(tmp_path / "assets"), "everything", glb=glb)
    ls = b200pt.LoadedScene(path)
    _compare(ls.desc(), sb.pack())
    ls.close()


def test_loader_cornell_and_uri_relative_to_cwd(tmp_path, monkeypatch):
    sb = S.cornell()
    monkeypatch.chdir(tmp_path)
    path = sb.write_gltf(str(tmp_path / "sub" / "dir"), "cornell")
    ls = b200pt.LoadedScene(os.path.relpath(path))
    _compare(ls.desc(), sb.pack())
    ls.close()
    # from another cwd the relative buffer uri no longer resolves: the reference fails the same way
    other = tmp_path / "elsewhere"; other.mkdir()
    monkeypatch.chdir(other)
    with pytest.raises(b200pt.PtError) as e:
        b200pt.LoadedScene(path)
    assert e.value.name == "IoError"


def _minimal_doc():
    """One triangle, everything in one .bin; returns (doc, bin bytes)."""
    idx = np.array([0, 1, 2], np.uint16).tobytes() + b"\0\0"
    P = np.array([[0, 0, 1], [1, 0, 1], [0, 1, 1]], np.float32).tobytes()
    N = np.array([[0, 0, -1]] * 3, np.float32).tobytes()
    T = np.array([[1, 0, 0, 1]] * 3, np.float32).tobytes()
    UV = np.array([[0, 0], [1, 0], [0, 1]], np.float32).tobytes()
    blob = idx + P + N + T + UV
    views, off = [], 0
    for b in (idx[:6], P, N, T, UV):
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(b)})
        off += len(b) + (2 if b is idx[:6] or len(b) == 6 else 0)
    acc = [{"bufferView": 0, "componentType": 5123, "count": 3, "type": "SCALAR"},
           {"bufferView": 1, "componentType": 5126, "count": 3, "type": "VEC3"},
           {"bufferView": 2, "componentType": 5126, "count": 3, "type": "VEC3"},
           {"bufferView": 3, "componentType": 5126, "count": 3, "type": "VEC4"},
           {"bufferView": 4, "componentType": 5126, "count": 3, "type": "VEC2"}]
    doc = {"asset": {"version": "2.0"}, "scenes": [{"nodes": [0]}], "nodes": [{"mesh": 0}],
           "meshes": [{"primitives": [{"attributes": {"POSITION": 1, "NORMAL": 2, "TANGENT": 3, "TEXCOORD_0": 4}, "indices": 0, "material": 0}]}],
           "materials": [{}], "accessors": acc, "bufferViews": views, "buffers": [{"uri": "min.bin", "byteLength": len(blob)}]}
    return doc, blob


def _write(tmp_path, doc, blob, name="min.gltf"):
    (tmp_path / "min.bin").write_bytes(blob)
    (tmp_path / name).write_text(json.dumps(doc))
    return str(tmp_path / name)


def test_loader_defaults_follow_gltf_spec(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    doc, blob = _minimal_doc()
    ls = b200pt.LoadedScene(_write(tmp_path, doc, blob))
    d = ls.desc()
    m = d.materials[0]
    # absent factors: baseColor 1, metallic 1, roughness 1, emissive 0; absent textures: 0xFFFFFFFF
    assert m.albedo_factor == 0x00FFFFFF and m.metal_roughness_factor == (255 << 8 | 255 << 16) and m.emissive_factor == 0
    assert m.albedo_texture_index == m.normal_texture_index == 0xFFFFFFFF
    assert d.num_instances == 1 and list(d.instances[0].transform[:]) == [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]
    p = d.primitives[0]
    assert (p.triangle_count, p.max_vertex, p.info) == (1, 2, 0)
    # sections padded to ceilPow2(largest accessor = 48 bytes) = 64
    assert d.triangle_data_len == 5 * 64
    assert [p.indices_offset, p.positions_offset, p.normals_offset, p.tangents_offset, p.uvs_offset] == [0, 64, 128, 192, 256]
    ls.close()


ERRORS = [
    ("GltfNoTangents", lambda d: d["meshes"][0]["primitives"][0]["attributes"].pop("TANGENT")),
    ("GltfNoNormals", lambda d: d["meshes"][0]["primitives"][0]["attributes"].pop("NORMAL")),
    ("GltfNoPositions", lambda d: d["meshes"][0]["primitives"][0]["attributes"].pop("POSITION")),
    ("GltfNoTextureCoordinates", lambda d: d["meshes"][0]["primitives"][0]["attributes"].pop("TEXCOORD_0")),
    ("GltfNoIndices", lambda d: d["meshes"][0]["primitives"][0].pop("indices")),
    ("NoMaterial", lambda d: d["meshes"][0]["primitives"][0].pop("material")),
    ("GltfNotATriangleTopology", lambda d: d["meshes"][0]["primitives"][0].update(mode=1)),
    ("InvalidAttributeType", lambda d: d["accessors"][3].update(type="VEC3")),
    ("InvalidIndexType", lambda d: d["accessors"][0].update(componentType=5121)),
    ("NoBufferView", lambda d: d["accessors"][1].pop("bufferView")),
    ("NoScene", lambda d: d.update(scenes=[])),
    ("TooManyScenes", lambda d: d.update(scenes=[{"nodes": [0]}, {"nodes": [0]}])),
    ("IoError", lambda d: d["accessors"][4].update(count=300)),          # error.EndOfFile: short read
]


@pytest.mark.parametrize("name,mutate", ERRORS, ids=[e[0] for e in ERRORS])
def test_loader_error_names(tmp_path, monkeypatch, name, mutate):
    monkeypatch.chdir(tmp_path)
    doc, blob = _minimal_doc()
    mutate(doc)
    with pytest.raises(b200pt.PtError) as e:
        b200pt.LoadedScene(_write(tmp_path, doc, blob))
    assert e.value.name == name, e.value


def test_loader_rejects_extension_and_bad_json(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    doc, blob = _minimal_doc()
    with pytest.raises(b200pt.PtError) as e:
        b200pt.LoadedScene(_write(tmp_path, doc, blob, "min.obj"))
    assert e.value.name == "InvalidFileExtension"
    (tmp_path / "bad.gltf").write_text('{"asset": ')
    with pytest.raises(b200pt.PtError) as e:
        b200pt.LoadedScene(str(tmp_path / "bad.gltf"))
    assert e.value.name == "GltfParseError"
    with pytest.raises(b200pt.PtError) as e:
        b200pt.LoadedScene(str(tmp_path / "missing.glb"))
    assert e.value.name == "IoError"


def _png_variants():
    from PIL import Image
    rng = np.random.default_rng(3)
    rgba = rng.integers(0, 256, (9, 13, 4), dtype=np.uint8)
    out = []
    for mode in ("RGBA", "RGB", "L", "LA", "P"):
        im = Image.fromarray(rgba, "RGBA").convert(mode)
        buf = io.BytesIO(); im.save(buf, format="PNG")
        out.append((mode, buf.getvalue(), np.asarray(im.convert("RGBA"))))
    im16 = Image.fromarray((rgba[..., 0].astype(np.uint16) * 257), "I;16")
    buf = io.BytesIO(); im16.save(buf, format="PNG")
    g = rgba[..., 0]
    out.append(("I;16", buf.getvalue(), np.stack([g, g, g, np.full_like(g, 255)], -1)))
    im1 = Image.fromarray(rgba[..., 0] > 127)
    buf = io.BytesIO(); im1.save(buf, format="PNG")
    b = np.where(rgba[..., 0] > 127, 255, 0).astype(np.uint8)
    out.append(("1", buf.getvalue(), np.stack([b, b, b, np.full_like(b, 255)], -1)))
    return out


def test_png_decoder_all_colour_types(tmp_path, monkeypatch):
    """Textures arrive as RGBA8 whatever the PNG colour type / bit depth (stb_image's req_comp = 4)."""
    monkeypatch.chdir(tmp_path)
    for mode, png, expect in _png_variants():
        doc, blob = _minimal_doc()
        (tmp_path / "t.png").write_bytes(png)
        doc["images"] = [{"uri": "t.png"}]; doc["textures"] = [{"source": 0}]
        doc["materials"] = [{"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}}]
        ls = b200pt.LoadedScene(_write(tmp_path, doc, blob))
        t = ls.desc().textures[0]
        got = np.frombuffer(_as_bytes(t.rgba, t.width * t.height * 4), np.uint8).reshape(t.height, t.width, 4)
        assert (t.width, t.height) == (expect.shape[1], expect.shape[0]), mode
        assert (got == expect).all(), mode
        ls.close()


def test_png_interlaced_matches(tmp_path, monkeypatch):
    """Adam7: hand-built interlaced PNG of a known image (PIL cannot write interlaced files)."""
    import zlib
    monkeypatch.chdir(tmp_path)
    rng = np.random.default_rng(8)
    img = rng.integers(0, 256, (11, 10, 3), dtype=np.uint8)
    passes = [(0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)]
    raw = b""
    for x0, y0, dx, dy in passes:
        sub = img[y0::dy, x0::dx]
        if sub.size == 0:
            continue
        for row in sub:
            raw += b"\0" + row.tobytes()

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body))
    png = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", 10, 11, 8, 2, 0, 0, 1)) + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")
    doc, blob = _minimal_doc()
    (tmp_path / "i.png").write_bytes(png)
    doc["images"] = [{"uri": "i.png"}]; doc["textures"] = [{"source": 0}]
    ls = b200pt.LoadedScene(_write(tmp_path, doc, blob))
    t = ls.desc().textures[0]
    got = np.frombuffer(_as_bytes(t.rgba, t.width * t.height * 4), np.uint8).reshape(t.height, t.width, 4)
    assert (got[..., :3] == img).all() and (got[..., 3] == 255).all()
    ls.close()


def test_jpeg_baseline_close_to_pil(tmp_path, monkeypatch):
    """JPEG decoders differ by a few LSB (IDCT / up-sampling); the contract is 'close', not exact."""
    from PIL import Image
    monkeypatch.chdir(tmp_path)
    y, x = np.mgrid[0:48, 0:40]
    img = np.stack([(x * 6) % 256, (y * 5) % 256, ((x + y) * 3) % 256], -1).astype(np.uint8)
    for sub in (0, 2):           # 4:4:4 and 4:2:0
        buf = io.BytesIO(); Image.fromarray(img, "RGB").save(buf, format="JPEG", quality=92, subsampling=sub)
        ref = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB")).astype(int)
        doc, blob = _minimal_doc()
        (tmp_path / "t.jpg").write_bytes(buf.getvalue())
        doc["images"] = [{"uri": "t.jpg"}]; doc["textures"] = [{"source": 0}]
        ls = b200pt.LoadedScene(_write(tmp_path, doc, blob))
        t = ls.desc().textures[0]
        got = np.frombuffer(_as_bytes(t.rgba, t.width * t.height * 4), np.uint8).reshape(t.height, t.width, 4).astype(int)
        assert (got[..., 3] == 255).all()
        diff = np.abs(got[..., :3] - ref)
        assert diff.mean() < (1.0 if sub == 0 else 6.0), (sub, diff.mean(), diff.max())
        ls.close()
