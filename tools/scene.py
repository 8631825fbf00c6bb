"""Synthetic glTF scenes for the five BASELINE.json configs, and the reference's host data contract.

* `SceneBuilder` is an in-memory glTF-like scene (meshes -> primitives, materials, textures, node
  graph).  `write_gltf()` emits a spec-clean `.gltf + .bin + .png` (or `.glb`) that satisfies the
  reference loader's requirements (indexed triangles, POSITION/NORMAL/TANGENT/TEXCOORD_0 float,
  u16/u32 indices, a material on every primitive, exactly one scene —
  /root/reference/src/Scene.zig:176-182,251-256,495-496).
* `SceneBuilder.pack()` restates `Scene.load` (Scene.zig:79-546) on the in-memory scene and returns
  a `PackedScene` whose numpy arrays have exactly the reference's host layout
  (`triangle_data` sections, `Primitive`, `Mesh`, `Material`, `Texture`, `Instance`); its `desc()`
  is the `PtSceneDesc` of include/b200pt.h.
* `camera_default()` restates Camera.new + updateOrientation + updateVectors (Camera.zig:25-46,86-102)
  and the push-constant packing (RaytracingPass.zig:563-569).

Used by tests, bench.py and the CLI examples; pure numpy (PIL only to write PNGs).
"""
import ctypes as C
import json
import math
import os
import struct

import numpy as np

INVALID_TEX = 0xFFFFFFFF


# --------------------------------------------------------------------------------------------
# ctypes mirrors of include/b200pt.h
# --------------------------------------------------------------------------------------------
class PtPrimitive(C.Structure):
    _fields_ = [("indices_offset", C.c_uint64), ("positions_offset", C.c_uint64), ("normals_offset", C.c_uint64),
                ("tangents_offset", C.c_uint64), ("uvs_offset", C.c_uint64), ("max_vertex", C.c_uint32),
                ("triangle_count", C.c_uint32), ("info", C.c_uint32), ("_pad", C.c_uint32)]


class PtMesh(C.Structure):
    _fields_ = [("start", C.c_uint32), ("end", C.c_uint32)]


class PtMaterial(C.Structure):
    _fields_ = [("albedo_factor", C.c_uint32), ("metal_roughness_factor", C.c_uint32), ("emissive_factor", C.c_uint32),
                ("albedo_texture_index", C.c_uint32), ("metal_roughness_texture_index", C.c_uint32),
                ("emissive_texture_index", C.c_uint32), ("normal_texture_index", C.c_uint32)]


class PtTexture(C.Structure):
    _fields_ = [("rgba", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32)]


class PtInstance(C.Structure):
    _fields_ = [("mesh_index", C.c_uint64), ("transform", C.c_float * 16)]


class PtSceneDesc(C.Structure):
    _fields_ = [("triangle_data", C.c_void_p), ("triangle_data_len", C.c_size_t),
                ("primitives", C.POINTER(PtPrimitive)), ("num_primitives", C.c_size_t),
                ("meshes", C.POINTER(PtMesh)), ("num_meshes", C.c_size_t),
                ("materials", C.POINTER(PtMaterial)), ("num_materials", C.c_size_t),
                ("textures", C.POINTER(PtTexture)), ("num_textures", C.c_size_t),
                ("instances", C.POINTER(PtInstance)), ("num_instances", C.c_size_t)]


class PtRenderConfig(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("num_samples", C.c_uint32), ("num_bounces", C.c_uint32),
                ("divider", C.c_uint32), ("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("partition", C.c_uint32), ("flags", C.c_uint32)]


class PtCamera(C.Structure):
    _fields_ = [("position", C.c_float * 4), ("horizontal", C.c_float * 4), ("vertical", C.c_float * 4),
                ("forward", C.c_float * 4), ("frame_count", C.c_uint32)]


def render_config(width=1920, height=1080, num_samples=1, num_bounces=2, divider=1, device=0, rank=0, world=1,
                  partition=0, flags=0):
    """CLI defaults of /root/reference/src/main.zig:29-67."""
    return PtRenderConfig(width, height, num_samples, num_bounces, divider, device, rank, world, partition, flags)


def camera_default(width, height, frame_count=0, position=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0, fov=math.pi * 0.25):
    """Camera.new(fov = pi/4, aspect = w/h, pos) (main.zig:220-224) -> push constants."""
    f32 = np.float32
    aspect = f32(width) / f32(height)

    def quat_axis(deg, axis):  # zalgebra Quat.fromAxis(degrees, axis)
        r = math.radians(deg) * 0.5
        s = math.sin(r)
        return np.array([math.cos(r), axis[0] * s, axis[1] * s, axis[2] * s])

    def qmul(a, b):
        w1, x1, y1, z1 = a
        w2, x2, y2, z2 = b
        return np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                         w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])

    def rot(q, v):
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                      [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                      [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
        return R @ np.asarray(v, dtype=np.float64)

    q = qmul(quat_axis(yaw, (0, 1, 0)), quat_axis(pitch, (1, 0, 0)))       # Camera.zig:87-89
    forward = rot(q, (0, 0, 1)).astype(f32)                                 # :91  Vec3.forward() = +z
    left = rot(q, (-1, 0, 0)).astype(f32)                                   # :92  Vec3.left() = -x
    up = np.cross(left, forward).astype(f32)                                # :93
    vh = f32(2.0) * f32(math.tan(float(f32(fov) * f32(0.5))))               # :97
    vw = f32(vh * aspect)                                                   # :98
    cam = PtCamera()
    cam.position[:] = [position[0], position[1], position[2], 1.0]
    cam.horizontal[:] = list((left * vw).astype(f32)) + [0.0]               # :100
    cam.vertical[:] = list((up * vh).astype(f32)) + [0.0]                   # :101
    cam.forward[:] = list(forward) + [0.0]
    cam.frame_count = frame_count
    return cam


# --------------------------------------------------------------------------------------------
# in-memory glTF-like scene
# --------------------------------------------------------------------------------------------
class Prim:
    def __init__(self, indices, P, N, T, UV, material, force_u32=False):
        self.P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        self.N = np.ascontiguousarray(N, np.float32).reshape(-1, 3)
        self.T = np.ascontiguousarray(T, np.float32).reshape(-1, 4)
        self.UV = np.ascontiguousarray(UV, np.float32).reshape(-1, 2)
        nv = len(self.P)
        assert len(self.N) == nv and len(self.T) == nv and len(self.UV) == nv
        idx = np.asarray(indices).reshape(-1)
        assert idx.size % 3 == 0 and (idx.size == 0 or int(idx.max()) < nv)
        self.u32 = bool(force_u32 or nv > 65535)
        self.indices = np.ascontiguousarray(idx, np.uint32 if self.u32 else np.uint16)
        self.material = int(material)

    @property
    def num_tris(self):
        return self.indices.size // 3


class Node:
    def __init__(self, mesh=None, translation=(0, 0, 0), rotation=(0, 0, 0, 1), scale=(1, 1, 1), children=()):
        self.mesh = mesh
        self.translation = tuple(float(v) for v in translation)
        self.rotation = tuple(float(v) for v in rotation)  # x, y, z, w (glTF order)
        self.scale = tuple(float(v) for v in scale)
        self.children = list(children)


def _quat_to_mat4(w, x, y, z):
    """zalgebra Quat.toMat4 (normalises first); f32 arithmetic."""
    f = np.float32
    w, x, y, z = f(w), f(x), f(y), f(z)
    n = f(np.sqrt(w * w + x * x + y * y + z * z))
    w, x, y, z = w / n, x / n, y / n, z / n
    xx, yy, zz = x * x, y * y, z * z
    xy, xz, yz = x * y, x * z, y * z
    wx, wy, wz = w * x, w * y, w * z
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = f(1) - f(2) * (yy + zz); m[0, 1] = f(2) * (xy - wz); m[0, 2] = f(2) * (xz + wy)
    m[1, 0] = f(2) * (xy + wz); m[1, 1] = f(1) - f(2) * (xx + zz); m[1, 2] = f(2) * (yz - wx)
    m[2, 0] = f(2) * (xz - wy); m[2, 1] = f(2) * (yz + wx); m[2, 2] = f(1) - f(2) * (xx + yy)
    return m


def _matmul_f32(a, b):
    """4x4 product, f32, k-ascending accumulation (zalgebra Mat4.mul loops)."""
    out = np.zeros((4, 4), np.float32)
    for r in range(4):
        for c in range(4):
            s = np.float32(0)
            for k in range(4):
                s = np.float32(s + np.float32(a[r, k] * b[k, c]))
            out[r, c] = s
    return out


class PackedScene:
    """Host arrays in the reference layout (Scene.zig:68-77)."""

    def __init__(self):
        self.triangle_data = None
        self.primitives = None  # ctypes array of PtPrimitive
        self.meshes = None
        self.materials = None
        self.textures = []      # list of (H,W,4) uint8 arrays
        self.instances = None
        self._tex_structs = None

    def desc(self):
        d = PtSceneDesc()
        d.triangle_data = self.triangle_data.ctypes.data
        d.triangle_data_len = self.triangle_data.size
        d.primitives = C.cast(self.primitives, C.POINTER(PtPrimitive)); d.num_primitives = len(self.primitives)
        d.meshes = C.cast(self.meshes, C.POINTER(PtMesh)); d.num_meshes = len(self.meshes)
        d.materials = C.cast(self.materials, C.POINTER(PtMaterial)); d.num_materials = len(self.materials)
        self._tex_structs = (PtTexture * max(1, len(self.textures)))()
        for i, t in enumerate(self.textures):
            assert t.dtype == np.uint8 and t.flags["C_CONTIGUOUS"] and t.shape[2] == 4
            self._tex_structs[i] = PtTexture(t.ctypes.data, t.shape[1], t.shape[0])
        d.textures = C.cast(self._tex_structs, C.POINTER(PtTexture)); d.num_textures = len(self.textures)
        d.instances = C.cast(self.instances, C.POINTER(PtInstance)); d.num_instances = len(self.instances)
        return d

    @property
    def num_triangles(self):
        return sum(p.triangle_count for p in self.primitives)

    @property
    def num_effective_triangles(self):
        per_mesh = [sum(self.primitives[p].triangle_count for p in range(m.start, m.end)) for m in self.meshes]
        return sum(per_mesh[i.mesh_index] for i in self.instances)


def _ceil_pow2(n):
    return 1 if n <= 1 else 1 << (int(n) - 1).bit_length()


def _align(n, a):
    return (n + a - 1) // a * a


class SceneBuilder:
    def __init__(self):
        self.meshes = []     # list[list[Prim]]
        self.materials = []  # dicts
        self.textures = []   # (H,W,4) uint8
        self.nodes = []
        self.roots = []

    # -- authoring -------------------------------------------------------------------------
    def add_texture(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.uint8)
        assert rgba.ndim == 3 and rgba.shape[2] == 4
        self.textures.append(rgba)
        return len(self.textures) - 1

    def add_material(self, base_color=(1, 1, 1), roughness=1.0, metallic=0.0, emissive=(0, 0, 0), albedo_tex=None,
                     mr_tex=None, emissive_tex=None, normal_tex=None):
        self.materials.append(dict(base_color=tuple(float(v) for v in base_color), roughness=float(roughness),
                                   metallic=float(metallic), emissive=tuple(float(v) for v in emissive),
                                   albedo_tex=albedo_tex, mr_tex=mr_tex, emissive_tex=emissive_tex, normal_tex=normal_tex))
        return len(self.materials) - 1

    def add_mesh(self, prims):
        self.meshes.append(list(prims))
        return len(self.meshes) - 1

    def add_node(self, mesh=None, translation=(0, 0, 0), rotation=(0, 0, 0, 1), scale=(1, 1, 1), parent=None):
        self.nodes.append(Node(mesh, translation, rotation, scale))
        i = len(self.nodes) - 1
        if parent is None:
            self.roots.append(i)
        else:
            self.nodes[parent].children.append(i)
        return i

    # -- Scene.load restated ---------------------------------------------------------------
    def pack(self):
        ps = PackedScene()
        # loadMeshes (Scene.zig:160-352)
        prims = [p for m in self.meshes for p in m]
        sizes = [0] * 5
        largest = 0
        for p in prims:
            for k, a in enumerate((p.indices, p.P, p.N, p.T, p.UV)):
                sizes[k] += a.nbytes
                largest = max(largest, a.nbytes)
        largest = _ceil_pow2(largest)                                   # :209
        offsets, total = [], 0
        for k in range(5):                                              # :211-217
            offsets.append(total)
            sizes[k] = _align(sizes[k], largest)
            total += sizes[k]
        blob = np.zeros(total, np.uint8)
        cur = list(offsets)
        P = (PtPrimitive * max(1, len(prims)))()
        for i, p in enumerate(prims):                                   # :228-333
            cur[0] = _align(cur[0], p.indices.itemsize)                 # :236-240
            P[i] = PtPrimitive(cur[0], cur[1], cur[2], cur[3], cur[4], len(p.P) - 1, p.num_tris,
                               (p.material & 0xFFFFFF) | (0x80000000 if p.u32 else 0), 0)
            for k, a in enumerate((p.indices, p.P, p.N, p.T, p.UV)):
                blob[cur[k]:cur[k] + a.nbytes] = a.view(np.uint8).reshape(-1)
                cur[k] += a.nbytes
        ps.triangle_data = blob
        ps.primitives = (PtPrimitive * len(prims))(*P[:len(prims)]) if prims else (PtPrimitive * 0)()
        M = (PtMesh * len(self.meshes))()
        s = 0
        for i, m in enumerate(self.meshes):                             # :338-347
            M[i] = PtMesh(s, s + len(m)); s += len(m)
        ps.meshes = M
        # loadTextures (:354-446): one per glTF texture, RGBA8
        ps.textures = list(self.textures)
        # loadMaterials (:448-485): trunc(x*255) in f32
        def q(x):
            return int(np.float32(x) * np.float32(255.0)) & 0xFF
        mats = (PtMaterial * len(self.materials))()
        for i, m in enumerate(self.materials):
            bc, em = m["base_color"], m["emissive"]
            t = lambda v: INVALID_TEX if v is None else int(v)
            mats[i] = PtMaterial(q(bc[0]) | q(bc[1]) << 8 | q(bc[2]) << 16,
                                 q(m["roughness"]) << 8 | q(m["metallic"]) << 16,
                                 q(em[0]) | q(em[1]) << 8 | q(em[2]) << 16,
                                 t(m["albedo_tex"]), t(m["mr_tex"]), t(m["emissive_tex"]), t(m["normal_tex"]))
        ps.materials = mats
        # loadScene (:487-546): LIFO DFS per root; TRS only on mesh nodes; M = S*((T*P)*R) (zalgebra
        # translate/scale pre-multiply — UNVERIFIED); stored transposed = rows of M.
        inst = []
        for root in self.roots:
            stack = [(root, np.eye(4, dtype=np.float32))]
            while stack:
                ni, pm = stack.pop()
                n = self.nodes[ni]
                if n.mesh is not None:
                    T = np.eye(4, dtype=np.float32); T[:3, 3] = n.translation
                    m = _matmul_f32(T, pm)
                    x, y, z, w = n.rotation
                    m = _matmul_f32(m, _quat_to_mat4(w, x, y, z))
                    S = np.eye(4, dtype=np.float32); S[0, 0], S[1, 1], S[2, 2] = n.scale
                    m = _matmul_f32(S, m)
                    inst.append((n.mesh, m.copy()))
                else:
                    m = pm
                for c in n.children:
                    stack.append((c, m))
        I = (PtInstance * len(inst))()
        for i, (mesh, m) in enumerate(inst):
            I[i].mesh_index = mesh
            I[i].transform[:] = [float(v) for v in m.reshape(-1)]
        ps.instances = I
        return ps

    # -- glTF writer -------------------------------------------------------------------------
    def write_gltf(self, directory, name, glb=False, relative_to_cwd=True):
        """Writes <directory>/<name>.gltf (+ .bin, .png) or <name>.glb.  URIs are written relative to
        the current working directory when `relative_to_cwd` (the reference opens them relative
        to the cwd, Scene.zig:125,414 — quirk Q9), else relative to the file."""
        from PIL import Image
        import io
        os.makedirs(directory, exist_ok=True)
        chunks, views, accessors = [], [], []
        off = 0

        def add_view(data, target=None):
            nonlocal off
            pad = (-off) % 4
            if pad:
                chunks.append(b"\0" * pad); off += pad
            v = {"buffer": 0, "byteOffset": off, "byteLength": len(data)}
            if target:
                v["target"] = target
            views.append(v); chunks.append(data); off += len(data)
            return len(views) - 1

        def add_acc(arr, ctype, typ, target, minmax=False):
            v = add_view(arr.tobytes(), target)
            a = {"bufferView": v, "componentType": ctype, "count": int(arr.shape[0] if arr.ndim > 1 else arr.size), "type": typ}
            if minmax:
                a["min"] = [float(x) for x in arr.min(axis=0)]; a["max"] = [float(x) for x in arr.max(axis=0)]
            accessors.append(a)
            return len(accessors) - 1

        gmeshes = []
        for m in self.meshes:
            gp = []
            for p in m:
                ia = add_acc(p.indices, 5125 if p.u32 else 5123, "SCALAR", 34963)
                pa = add_acc(p.P, 5126, "VEC3", 34962, True)
                na = add_acc(p.N, 5126, "VEC3", 34962)
                ta = add_acc(p.T, 5126, "VEC4", 34962)
                ua = add_acc(p.UV, 5126, "VEC2", 34962)
                gp.append({"attributes": {"POSITION": pa, "NORMAL": na, "TANGENT": ta, "TEXCOORD_0": ua}, "indices": ia,
                           "material": p.material, "mode": 4})
            gmeshes.append({"primitives": gp})
        images = []
        for i, t in enumerate(self.textures):
            buf = io.BytesIO(); Image.fromarray(t, "RGBA").save(buf, format="PNG", compress_level=1)
            if glb:
                images.append({"bufferView": add_view(buf.getvalue()), "mimeType": "image/png"})
            else:
                fn = f"{name}_tex{i}.png"
                with open(os.path.join(directory, fn), "wb") as f:
                    f.write(buf.getvalue())
                uri = os.path.relpath(os.path.join(directory, fn)) if relative_to_cwd else fn
                images.append({"uri": uri})
        gmats = []
        for m in self.materials:
            pbr = {"baseColorFactor": list(m["base_color"]) + [1.0], "metallicFactor": m["metallic"], "roughnessFactor": m["roughness"]}
            g = {"pbrMetallicRoughness": pbr, "emissiveFactor": list(m["emissive"])}
            if m["albedo_tex"] is not None: pbr["baseColorTexture"] = {"index": m["albedo_tex"]}
            if m["mr_tex"] is not None: pbr["metallicRoughnessTexture"] = {"index": m["mr_tex"]}
            if m["emissive_tex"] is not None: g["emissiveTexture"] = {"index": m["emissive_tex"]}
            if m["normal_tex"] is not None: g["normalTexture"] = {"index": m["normal_tex"]}
            gmats.append(g)
        gnodes = []
        for n in self.nodes:
            g = {}
            if n.mesh is not None: g["mesh"] = n.mesh
            if n.translation != (0.0, 0.0, 0.0): g["translation"] = list(n.translation)
            if n.rotation != (0.0, 0.0, 0.0, 1.0): g["rotation"] = list(n.rotation)
            if n.scale != (1.0, 1.0, 1.0): g["scale"] = list(n.scale)
            if n.children: g["children"] = list(n.children)
            gnodes.append(g)
        doc = {"asset": {"version": "2.0", "generator": "b200pt tools/scene.py"}, "scene": 0, "scenes": [{"nodes": list(self.roots)}],
               "nodes": gnodes, "meshes": gmeshes, "materials": gmats, "accessors": accessors, "bufferViews": views}
        if self.textures:
            doc["images"] = images
            doc["textures"] = [{"source": i} for i in range(len(self.textures))]
        blob = b"".join(chunks)
        if glb:
            doc["buffers"] = [{"byteLength": len(blob)}]
            js = json.dumps(doc, separators=(",", ":")).encode()
            js += b" " * ((-len(js)) % 4)
            blob += b"\0" * ((-len(blob)) % 4)
            path = os.path.join(directory, name + ".glb")
            with open(path, "wb") as f:
                f.write(struct.pack("<III", 0x46546C67, 2, 12 + 8 + len(js) + 8 + len(blob)))
                f.write(struct.pack("<II", len(js), 0x4E4F534A)); f.write(js)
                f.write(struct.pack("<II", len(blob), 0x004E4942)); f.write(blob)
            return path
        binfn = name + ".bin"
        with open(os.path.join(directory, binfn), "wb") as f:
            f.write(blob)
        doc["buffers"] = [{"byteLength": len(blob), "uri": os.path.relpath(os.path.join(directory, binfn)) if relative_to_cwd else binfn}]
        path = os.path.join(directory, name + ".gltf")
        with open(path, "w") as f:
            json.dump(doc, f)
        return path


# --------------------------------------------------------------------------------------------
# geometry helpers (all produce front faces = CCW seen from the side the normal points to)
# --------------------------------------------------------------------------------------------
class Soup:
    """Triangle soup with attributes; concatenated per material into primitives."""

    def __init__(self):
        self.P, self.N, self.T, self.UV, self.I = [], [], [], [], []
        self.nv = 0

    def add(self, P, N, T, UV, I):
        P = np.asarray(P, np.float32).reshape(-1, 3)
        self.P.append(P); self.N.append(np.asarray(N, np.float32).reshape(-1, 3))
        self.T.append(np.asarray(T, np.float32).reshape(-1, 4)); self.UV.append(np.asarray(UV, np.float32).reshape(-1, 2))
        self.I.append(np.asarray(I, np.int64).reshape(-1) + self.nv)
        self.nv += len(P)

    @property
    def num_tris(self):
        return sum(i.size for i in self.I) // 3

    def prim(self, material, force_u32=False):
        return Prim(np.concatenate(self.I), np.concatenate(self.P), np.concatenate(self.N), np.concatenate(self.T),
                    np.concatenate(self.UV), material, force_u32)


def grid_patch(soup, origin, eu, ev, nu, nv, uv_scale=(1.0, 1.0), displace=None, facing=None):
    """Planar nu x nv quad grid spanned by eu, ev from origin; normal = normalize(eu x ev).  When
    `facing` is given the two edges are swapped if needed so that the normal has a positive
    component along it."""
    origin, eu, ev = (np.asarray(a, np.float64) for a in (origin, eu, ev))
    if facing is not None and np.dot(np.cross(eu, ev), np.asarray(facing, np.float64)) < 0:
        eu, ev, nu, nv, uv_scale = ev, eu, nv, nu, (uv_scale[1], uv_scale[0])
    n = np.cross(eu, ev); n /= np.linalg.norm(n)
    t = eu / np.linalg.norm(eu)
    s, r = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="xy")
    P = origin + s[..., None] * eu + r[..., None] * ev
    if displace is not None:
        P = P + n * displace(s, r)[..., None]
    P = P.reshape(-1, 3)
    N = np.broadcast_to(n, P.shape)
    w = 1.0 if np.dot(np.cross(n, t), ev) > 0 else -1.0
    T = np.concatenate([np.broadcast_to(t, P.shape), np.full((len(P), 1), w)], axis=1)
    UV = np.stack([s.reshape(-1) * uv_scale[0], r.reshape(-1) * uv_scale[1]], axis=1)
    i = (np.arange(nv)[:, None] * (nu + 1) + np.arange(nu)[None, :]).reshape(-1)
    I = np.stack([i, i + 1, i + nu + 2, i, i + nu + 2, i + nu + 1], axis=1)
    soup.add(P, N, T, UV, I)


def box(soup, lo, hi, outward=True, uv_scale=(1.0, 1.0), n=1):
    """Axis-aligned box; faces point outward (or inward for rooms)."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    d = hi - lo
    ex, ey, ez = np.array([d[0], 0, 0]), np.array([0, d[1], 0]), np.array([0, 0, d[2]])
    faces = [  # (origin, eu, ev) with eu x ev = outward normal
        (lo + ex, ez, ey), (lo, ey, ez),          # +x, -x
        (lo + ey, ez, ex), (lo, ex, ez),          # +y, -y
        (lo + ez, ex, ey), (lo, ey, ex),          # +z, -z
    ]
    normals = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    for (o, eu, ev), nrm in zip(faces, normals):
        f = np.asarray(nrm, np.float64) * (1.0 if outward else -1.0)
        grid_patch(soup, o, eu, ev, n, n, uv_scale, facing=f)


def revolve(soup, center, profile_r, profile_y, segments, uv_scale=(1.0, 1.0), bump=None):
    """Surface of revolution around +y through `center`; profile radii/heights (bottom to top),
    outward-facing.  `bump(theta, v) -> radial displacement` makes every triangle unique."""
    center = np.asarray(center, np.float64)
    profile_r, profile_y = np.asarray(profile_r, np.float64), np.asarray(profile_y, np.float64)
    rings = len(profile_r)
    th = np.linspace(0.0, 2.0 * np.pi, segments + 1)
    TH, R = np.meshgrid(th, profile_r, indexing="xy")
    _, Y = np.meshgrid(th, profile_y, indexing="xy")
    V = np.broadcast_to(np.linspace(0, 1, rings)[:, None], TH.shape)
    if bump is not None:
        R = R + bump(TH, V)
    P = np.stack([center[0] + R * np.cos(TH), center[1] + Y, center[2] + R * np.sin(TH)], axis=-1)
    # normals from the profile slope (analytic up to the bump)
    dr = np.gradient(profile_r); dy = np.gradient(profile_y)
    nl = np.sqrt(dr * dr + dy * dy) + 1e-20
    nr, ny = (dy / nl)[:, None], (-dr / nl)[:, None]
    N = np.stack([nr * np.cos(TH), np.broadcast_to(ny, TH.shape), nr * np.sin(TH)], axis=-1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    Tg = np.stack([-np.sin(TH), np.zeros_like(TH), np.cos(TH)], axis=-1)
    # bitangent should follow +v (up): cross(N, T) = up for this parametrisation?  pick w accordingly
    up = np.stack([np.zeros_like(TH), np.ones_like(TH), np.zeros_like(TH)], axis=-1)
    w = np.where(np.einsum("...k,...k", np.cross(N, Tg), up) >= 0, 1.0, -1.0)
    T = np.concatenate([Tg, w[..., None]], axis=-1)
    UV = np.stack([TH / (2 * np.pi) * uv_scale[0], V * uv_scale[1]], axis=-1)
    i = (np.arange(rings - 1)[:, None] * (segments + 1) + np.arange(segments)[None, :]).reshape(-1)
    s1 = segments + 1
    # outward-facing: (i, i+s1, i+s1+1), (i, i+s1+1, i+1)
    I = np.stack([i, i + s1, i + s1 + 1, i, i + s1 + 1, i + 1], axis=1)
    soup.add(P.reshape(-1, 3), N.reshape(-1, 3), T.reshape(-1, 4), UV.reshape(-1, 2), I)


def uv_sphere(soup, center, radius, segments, rings, uv_scale=(1.0, 1.0), bump=None):
    """Outward-facing UV sphere (poles slightly cut so no degenerate triangles)."""
    phi = np.linspace(-0.5 * np.pi + 0.02, 0.5 * np.pi - 0.02, rings)
    revolve(soup, center, radius * np.cos(phi), radius * np.sin(phi), segments, uv_scale, bump)


# --------------------------------------------------------------------------------------------
# procedural textures (value noise; deterministic in (seed, id))
# --------------------------------------------------------------------------------------------
def _value_noise(rng, size, cells):
    g = rng.random((cells + 1, cells + 1), dtype=np.float32)
    g[-1, :] = g[0, :]; g[:, -1] = g[:, 0]   # tileable (REPEAT addressing)
    x = np.linspace(0, cells, size, endpoint=False, dtype=np.float32)
    i = x.astype(np.int32); f = x - i
    f = f * f * (3 - 2 * f)
    a = g[i][:, i]; b = g[i][:, i + 1]; c = g[i + 1][:, i]; d = g[i + 1][:, i + 1]
    fx, fy = f[None, :], f[:, None]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def fbm(rng, size, octaves=4, base_cells=4):
    out = np.zeros((size, size), np.float32); amp, tot = 1.0, 0.0
    for o in range(octaves):
        cells = min(size // 2, base_cells << o)
        out += amp * _value_noise(rng, size, cells); tot += amp; amp *= 0.5
    return out / tot


def make_pbr_textures(seed, size, tint, rough_range=(0.2, 1.0), metal=False):
    """-> (albedo, metal_rough, normal) RGBA8 textures."""
    rng = np.random.default_rng(seed)
    h = fbm(rng, size, 5, 4)
    detail = fbm(rng, size, 3, 32)
    checker = ((np.arange(size)[:, None] // max(1, size // 8) + np.arange(size)[None, :] // max(1, size // 8)) & 1).astype(np.float32)
    lum = 0.55 + 0.35 * h + 0.1 * detail - 0.08 * checker
    alb = np.clip(lum[..., None] * np.asarray(tint, np.float32)[None, None, :], 0, 1)
    albedo = np.empty((size, size, 4), np.uint8)
    albedo[..., :3] = (alb * 255.0 + 0.5).astype(np.uint8); albedo[..., 3] = 255
    r = rough_range[0] + (rough_range[1] - rough_range[0]) * detail
    mr = np.empty((size, size, 4), np.uint8)
    mr[..., 0] = 255
    mr[..., 1] = (np.clip(r, 0, 1) * 255.0 + 0.5).astype(np.uint8)
    mr[..., 2] = 255 if metal else 0
    if metal:
        mr[..., 2] = (np.clip(0.6 + 0.4 * h, 0, 1) * 255.0 + 0.5).astype(np.uint8)
    mr[..., 3] = 255
    hh = h + 0.3 * detail
    strength = 2.0 * size / 256.0
    dx = (np.roll(hh, -1, axis=1) - np.roll(hh, 1, axis=1)) * strength * 8
    dy = (np.roll(hh, -1, axis=0) - np.roll(hh, 1, axis=0)) * strength * 8
    n = np.stack([-dx, -dy, np.ones_like(dx)], axis=-1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    normal = np.empty((size, size, 4), np.uint8)
    normal[..., :3] = ((n * 0.5 + 0.5) * 255.0 + 0.5).astype(np.uint8); normal[..., 3] = 255
    return albedo, mr, normal


# --------------------------------------------------------------------------------------------
# C1 — Cornell box  (~40 tris, 1 emissive quad, 512x512, 1 spp, 2 segments)
# --------------------------------------------------------------------------------------------
def cornell(seed=1234):
    sb = SceneBuilder()
    rng = np.random.default_rng(seed)
    checker = np.zeros((64, 64, 4), np.uint8)
    c = ((np.arange(64)[:, None] // 8 + np.arange(64)[None, :] // 8) & 1).astype(np.uint8)
    checker[..., 0] = 140 + 100 * c; checker[..., 1] = 140 + 100 * c; checker[..., 2] = 130 + 100 * c; checker[..., 3] = 255
    t_checker = sb.add_texture(checker)
    _, _, nrm = make_pbr_textures(seed + 1, 64, (1, 1, 1))
    t_normal = sb.add_texture(nrm)
    m_white = sb.add_material((0.73, 0.73, 0.73), 1.0, 0.0)
    m_red = sb.add_material((0.65, 0.05, 0.05), 1.0, 0.0)
    m_green = sb.add_material((0.12, 0.45, 0.15), 1.0, 0.0)
    m_floor = sb.add_material((1.0, 1.0, 1.0), 0.9, 0.0, albedo_tex=t_checker, normal_tex=t_normal)
    m_light = sb.add_material((0.8, 0.8, 0.8), 1.0, 0.0, emissive=(1.0, 1.0, 1.0))
    m_metal = sb.add_material((0.9, 0.85, 0.6), 0.2, 1.0)
    # room: x,y in [-1,1], z in [1,3]; open towards the camera (-z side)
    def wall(o, eu, ev, facing, mat, uvs=(1, 1)):
        s = Soup(); grid_patch(s, o, eu, ev, 1, 1, uvs, facing=facing); return s.prim(mat)
    room = [
        wall((-1, -1, 3), (2, 0, 0), (0, 2, 0), (0, 0, -1), m_white),           # back
        wall((-1, -1, 1), (0, 0, 2), (0, 2, 0), (1, 0, 0), m_red),              # left (x=-1)
        wall((1, -1, 3), (0, 0, -2), (0, 2, 0), (-1, 0, 0), m_green),           # right (x=+1)
        wall((-1, -1, 1), (2, 0, 0), (0, 0, 2), (0, 1, 0), m_floor, (4, 4)),    # floor
        wall((-1, 1, 3), (2, 0, 0), (0, 0, -2), (0, -1, 0), m_white),           # ceiling
        wall((-0.35, 0.995, 2.35), (0.7, 0, 0), (0, 0, -0.7), (0, -1, 0), m_light),  # light
    ]
    mesh_room = sb.add_mesh(room)
    s1 = Soup(); box(s1, (-0.3, -0.6, -0.3), (0.3, 0.6, 0.3))
    s2 = Soup(); box(s2, (-0.3, -0.3, -0.3), (0.3, 0.3, 0.3))
    mesh_tall = sb.add_mesh([s1.prim(m_white)])
    mesh_cube = sb.add_mesh([s2.prim(m_metal)])
    sb.add_node(mesh_room)
    a = math.radians(18.0)
    sb.add_node(mesh_tall, translation=(-0.35, -0.4, 2.4), rotation=(0, math.sin(a / 2), 0, math.cos(a / 2)))
    a = math.radians(-17.0)
    sb.add_node(mesh_cube, translation=(0.4, -0.7, 1.7), rotation=(0, math.sin(a / 2), 0, math.cos(a / 2)))
    return sb


# --------------------------------------------------------------------------------------------
# C2 / C3 / C5 — Sponza-scale atrium (one big mesh, one primitive per material, a few props)
# --------------------------------------------------------------------------------------------
def atrium(seed=1234, target_tris=260_000, n_materials=25, tex_size=1024, rough_range=(0.1, 1.0), p_metal=0.2,
           textures=True):
    rng = np.random.default_rng(seed)
    sb = SceneBuilder()
    mats = []
    params = []
    for i in range(n_materials):
        metal = bool(rng.random() < p_metal)
        tint = 0.35 + 0.6 * rng.random(3)
        rough = rough_range[0] + (rough_range[1] - rough_range[0]) * rng.random()
        params.append((metal, tint, rough))
    if textures:
        # every material's maps come from their own generator (seed * 1000 + i): independent, so they are made in parallel
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max(1, min(n_materials, (os.cpu_count() or 4)))) as ex:
            maps = list(ex.map(lambda i: make_pbr_textures(seed * 1000 + i, tex_size, params[i][1], (0.3, 1.0), params[i][0]), range(n_materials)))
    for i, (metal, tint, rough) in enumerate(params):
        if textures:
            a, mr, n = maps[i]
            ta, tm, tn = sb.add_texture(a), sb.add_texture(mr), sb.add_texture(n)
            mats.append(sb.add_material((1, 1, 1), rough, 1.0 if metal else 0.0, albedo_tex=ta, mr_tex=tm, normal_tex=tn))
        else:
            mats.append(sb.add_material(tint, rough, 1.0 if metal else 0.0))
    m_light = sb.add_material((1, 1, 1), 1.0, 0.0, emissive=(1.0, 0.95, 0.85))
    soups = {m: Soup() for m in mats + [m_light]}
    pick = lambda k: mats[k % len(mats)]

    # dimensions: nave along +z, camera at the origin 2 m above the floor
    X0, X1, XW = -3.5, 3.5, 6.5          # colonnade lines, outer walls
    Z0, Z1 = -3.0, 33.0
    YF, Y1, Y2 = -2.0, 2.5, 7.0          # floor, first gallery, top
    n_bays = 12
    bay = (Z1 - Z0) / n_bays

    # a global tessellation factor reached by bisection on the triangle budget
    def build(f, soups):
        g = lambda n: max(1, int(round(n * f)))
        # floor + gallery floors + walls (displaced a little so tessellation is real geometry)
        wob = lambda a, b: 0.01 * np.sin(a * 37.0) * np.cos(b * 29.0)
        grid_patch(soups[pick(0)], (-XW, YF, Z0), (0, 0, Z1 - Z0), (2 * XW, 0, 0), g(90), g(36), (18, 6.5), wob, facing=(0, 1, 0))  # floor (+y)
        grid_patch(soups[pick(1)], (-XW, YF, Z1), (2 * XW, 0, 0), (0, Y2 - YF, 0), g(26), g(18), (6, 4), wob, facing=(0, 0, -1))  # far wall (-z)
        grid_patch(soups[pick(2)], (XW, YF, Z0), (-2 * XW, 0, 0), (0, Y2 - YF, 0), g(26), g(18), (6, 4), wob, facing=(0, 0, 1))  # near wall (+z)
        grid_patch(soups[pick(3)], (-XW, YF, Z0), (0, 0, Z1 - Z0), (0, Y2 - YF, 0), g(72), g(18), (18, 4), wob, facing=(1, 0, 0))  # left wall (+x)
        grid_patch(soups[pick(4)], (XW, YF, Z1), (0, 0, -(Z1 - Z0)), (0, Y2 - YF, 0), g(72), g(18), (18, 4), wob, facing=(-1, 0, 0))  # right wall (-x)
        for side, xa, xb in ((-1, -XW, X0), (1, X1, XW)):
            # gallery slab: top (+y) and underside (-y)
            grid_patch(soups[pick(5)], (xa, Y1, Z0), (0, 0, Z1 - Z0), (xb - xa, 0, 0), g(72), g(6), (18, 2), wob, facing=(0, 1, 0))
            grid_patch(soups[pick(6)], (xa, Y1 - 0.3, Z0), (xb - xa, 0, 0), (0, 0, Z1 - Z0), g(6), g(72), (2, 18), wob, facing=(0, -1, 0))
            # roof strips over the galleries (underside visible), nave stays open to the sky
            grid_patch(soups[pick(7)], (xa, Y2, Z0), (xb - xa, 0, 0), (0, 0, Z1 - Z0), g(6), g(72), (2, 18), wob, facing=(0, -1, 0))
            # slab edge facing the nave
            xe = xb if side < 0 else xa
            eu = (0, 0, Z1 - Z0) if side < 0 else (0, 0, -(Z1 - Z0))
            o = (xe, Y1 - 0.3, Z0) if side < 0 else (xe, Y1 - 0.3, Z1)
            grid_patch(soups[pick(8)], o, eu, (0, 0.3, 0), g(72), 1, (18, 0.2), facing=(-side, 0, 0))
        # columns: two storeys x two sides x (n_bays+1)
        k = 0
        for storey, (ya, yb, r) in enumerate(((YF, Y1 - 0.3, 0.32), (Y1, Y2, 0.24))):
            for xs in (X0, X1):
                for b in range(n_bays + 1):
                    zc = Z0 + b * bay
                    rings = g(34)
                    v = np.linspace(0, 1, rings)
                    pr = r * (1.0 - 0.12 * v + 0.35 * np.exp(-((v - 0.02) / 0.03) ** 2) + 0.3 * np.exp(-((v - 0.98) / 0.03) ** 2))
                    py = ya + (yb - ya) * v
                    ph = rng.random() * 6.28
                    flutes = 16
                    bump = lambda TH, V, ph=ph, r=r: 0.035 * r * np.cos(flutes * TH + ph) * np.sin(np.pi * V) ** 0.5
                    revolve(soups[pick(9 + (k % 4))], (xs, 0.0, zc), pr, py, g(44), (3, 6), bump)
                    k += 1
        # arches between first-storey columns: half tori approximated by bent tubes
        for xs in (X0, X1):
            for b in range(n_bays):
                zc = Z0 + (b + 0.5) * bay
                segs, sides = g(20), g(10)
                a = np.linspace(0, np.pi, segs + 1)
                R = bay * 0.5 - 0.05
                cy, cz = (Y1 - 0.3) - R - 0.05 + 0.0, zc
                # tube around the arc (in the y-z plane at x = xs)
                th = np.linspace(0, 2 * np.pi, sides + 1)
                A, TH = np.meshgrid(a, th, indexing="ij")
                rr = 0.14
                cyl_y = cy + (R + rr * np.cos(TH)) * np.sin(A)
                cyl_z = cz - (R + rr * np.cos(TH)) * np.cos(A)
                cyl_x = xs + rr * np.sin(TH)
                P = np.stack([cyl_x, cyl_y, cyl_z], axis=-1)
                N = np.stack([np.sin(TH), np.cos(TH) * np.sin(A), -np.cos(TH) * np.cos(A)], axis=-1)
                Tg = np.stack([np.zeros_like(A), np.cos(A), np.sin(A)], axis=-1)
                T = np.concatenate([Tg, np.ones_like(A)[..., None]], axis=-1)
                UV = np.stack([A / np.pi * 4, TH / (2 * np.pi)], axis=-1)
                i = (np.arange(segs)[:, None] * (sides + 1) + np.arange(sides)[None, :]).reshape(-1)
                s1 = sides + 1
                I = np.stack([i, i + s1, i + s1 + 1, i, i + s1 + 1, i + 1], axis=1)
                # orientation check on the first triangle; flip if needed
                p0, p1, p2 = P.reshape(-1, 3)[I[0, 0]], P.reshape(-1, 3)[I[0, 1]], P.reshape(-1, 3)[I[0, 2]]
                if np.dot(np.cross(p1 - p0, p2 - p0), N.reshape(-1, 3)[I[0, 0]]) < 0:
                    I = I[:, [0, 2, 1, 3, 5, 4]]
                soups[pick(13 + (b % 3))].add(P.reshape(-1, 3), N.reshape(-1, 3), T.reshape(-1, 4), UV.reshape(-1, 2), I)
        # props in the nave: spheres and vases (metallic/glossy variety)
        for j in range(10):
            zc = 3.0 + j * 2.9
            xs = (-1.4 if j % 2 else 1.4) + 0.3 * math.sin(j * 1.7)
            rad = 0.45 + 0.15 * math.sin(j * 2.3)
            ph = j * 0.7
            bump = lambda TH, V, ph=ph, rad=rad: 0.05 * rad * np.sin(9 * TH + ph) * np.sin(7 * np.pi * V)
            if j % 3 == 0:
                v = np.linspace(0, 1, g(40))
                pr = rad * (0.35 + 0.65 * np.sin(np.pi * np.clip(v * 0.9 + 0.05, 0, 1)) ** 1.5)
                revolve(soups[pick(16 + j % 9)], (xs, 0, zc), pr, YF + 1.6 * v, g(56), (2, 2), bump)
            else:
                uv_sphere(soups[pick(16 + j % 9)], (xs, YF + rad, zc), rad, g(56), g(40), (2, 2), bump)
        # two emissive quads hanging under the roof strips (normal -y)
        grid_patch(soups[m_light], (-XW + 0.5, Y2 - 0.05, 6.0), (2.0, 0, 0), (0, 0, 8.0), 1, 1, facing=(0, -1, 0))
        grid_patch(soups[m_light], (X1 + 0.5, Y2 - 0.05, 16.0), (2.0, 0, 0), (0, 0, 8.0), 1, 1, facing=(0, -1, 0))
        return sum(s.num_tris for s in soups.values())

    lo, hi = 0.05, 8.0
    state = rng.bit_generator.state
    for _ in range(18):
        mid = math.sqrt(lo * hi)
        rng.bit_generator.state = state
        n = build(mid, {m: Soup() for m in soups})
        if n < target_tris:
            lo = mid
        else:
            hi = mid
        if abs(n - target_tris) <= 0.005 * target_tris:
            break
    rng.bit_generator.state = state
    build(mid, soups)
    # both index widths occur (like real assets): big primitives use u32 indices, small ones u16
    prims = [soups[m].prim(m, force_u32=soups[m].nv > 20000) for m in mats + [m_light] if soups[m].nv > 0]
    mesh = sb.add_mesh(prims)
    sb.add_node(mesh)
    return sb


def sponza_scale(seed=1234, textures=True, tex_size=1024):
    """C2/C3: ~260 k triangles, 25 textured PBR materials."""
    return atrium(seed, 260_000, 25, tex_size, (0.1, 1.0), 0.2, textures)


def heavy_material(seed=1234, tex_size=4096, target_tris=1_000_000):
    """C5: ~1 M triangles, 16 glossy/metallic materials with 4096^2 albedo/metal-rough/normal maps."""
    return atrium(seed, target_tris, 16, tex_size, (0.05, 0.5), 0.6, True)


# --------------------------------------------------------------------------------------------
# C4 — instancing stress: 1000 instances of one ~10k-triangle mesh
# --------------------------------------------------------------------------------------------
def instanced(seed=1234, n_side=10, tex_size=256):
    rng = np.random.default_rng(seed)
    sb = SceneBuilder()
    a, mr, n = make_pbr_textures(seed, tex_size, (0.8, 0.7, 0.6), (0.2, 0.8), False)
    ta, tm, tn = sb.add_texture(a), sb.add_texture(mr), sb.add_texture(n)
    m0 = sb.add_material((1, 1, 1), 0.6, 0.0, albedo_tex=ta, mr_tex=tm, normal_tex=tn)
    m1 = sb.add_material((0.95, 0.8, 0.5), 0.25, 1.0, normal_tex=tn)
    m_light = sb.add_material((1, 1, 1), 1.0, 0.0, emissive=(1.0, 0.6, 0.3))
    bump = lambda TH, V: 0.12 * np.sin(5 * TH) * np.sin(6 * np.pi * V) + 0.05 * np.cos(11 * TH + 3 * V)
    s_body, s_band = Soup(), Soup()
    # 72 x 70 x 2 = 10080 triangles, split into two primitives (two geometries in the BLAS)
    phi = np.linspace(-0.5 * np.pi + 0.02, 0.5 * np.pi - 0.02, 71)
    lo, hi = phi[:36], phi[35:]
    revolve(s_body, (0, 0, 0), np.cos(lo), np.sin(lo), 72, (2, 1), bump)
    revolve(s_band, (0, 0, 0), np.cos(hi), np.sin(hi), 72, (2, 1), bump)
    mesh = sb.add_mesh([s_body.prim(m0), s_band.prim(m1)])
    s_l = Soup(); box(s_l, (-0.3, -0.3, -0.3), (0.3, 0.3, 0.3))
    mesh_light = sb.add_mesh([s_l.prim(m_light)])
    k = 0
    for ix in range(n_side):
        for iy in range(n_side):
            for iz in range(n_side):
                jitter = (rng.random(3) - 0.5) * 1.2
                pos = np.array([(ix - (n_side - 1) / 2) * 3.0, (iy - (n_side - 1) / 2) * 3.0, 5.0 + iz * (55.0 / max(1, n_side - 1))]) + jitter
                q = rng.normal(size=4); q /= np.linalg.norm(q)
                sb.add_node(mesh_light if (k % 97 == 0) else mesh, translation=pos, rotation=(q[0], q[1], q[2], q[3]))
                k += 1
    return sb


# --------------------------------------------------------------------------------------------
# random scenes for traversal / tie-break tests
# --------------------------------------------------------------------------------------------
def random_soup(seed, n_tris=2000, n_meshes=3, n_instances=6, extent=4.0, tri_size=0.6, with_scale=True):
    """Random triangles in several meshes, instanced with random rotation/translation(/scale)."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder()
    m = [sb.add_material(rng.random(3), 0.2 + 0.8 * rng.random(), float(rng.random() < 0.3), emissive=(rng.random(3) * (rng.random() < 0.2)))
         for _ in range(4)]
    for mi in range(n_meshes):
        prims = []
        for pi in range(1 + mi % 3):
            n = max(1, n_tris // (n_meshes * (1 + mi % 3)))
            c = (rng.random((n, 1, 3)) - 0.5) * extent
            P = (c + (rng.random((n, 3, 3)) - 0.5) * tri_size).reshape(-1, 3).astype(np.float32)
            if pi == 0 and mi == 0:
                P[: 3 * (n // 8)] = np.round(P[: 3 * (n // 8)] * 4) / 4   # axis-aligned / coplanar / degenerate cases
            nrm = np.cross(P[1::3] - P[0::3], P[2::3] - P[0::3])
            ln = np.linalg.norm(nrm, axis=1, keepdims=True)
            nrm = np.where(ln > 0, nrm / np.maximum(ln, 1e-30), np.array([[0, 0, 1.0]]))
            N = np.repeat(nrm, 3, axis=0)
            e = P[1::3] - P[0::3]
            le = np.linalg.norm(e, axis=1, keepdims=True)
            e = np.where(le > 0, e / np.maximum(le, 1e-30), np.array([[1.0, 0, 0]]))
            T = np.concatenate([np.repeat(e, 3, axis=0), np.ones((3 * n, 1))], axis=1)
            UV = rng.random((3 * n, 2))
            prims.append(Prim(np.arange(3 * n), P, N, T, UV, m[(mi + pi) % 4], force_u32=(pi % 2 == 1)))
        sb.add_mesh(prims)
    for i in range(n_instances):
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        if i == 0:
            q = np.array([0, 0, 0, 1.0])
        s = (0.5 + rng.random(3) * 1.5) if (with_scale and i % 3 == 2) else (1, 1, 1)
        sb.add_node(i % n_meshes, translation=(rng.random(3) - 0.5) * extent + np.array([0, 0, extent * 1.5]), rotation=q, scale=s)
    return sb


CONFIGS = {
    # name: (builder, width, height, samples, segments)
    "c1": (cornell, 512, 512, 1, 2),
    "c2": (sponza_scale, 1920, 1080, 1, 2),
    "c3": (sponza_scale, 3840, 2160, 64, 8),
    "c4": (instanced, 1920, 1080, 4, 4),
    "c5": (heavy_material, 1920, 1080, 16, 6),
    # one rank's share of C3 on 8 GPUs (an eighth of the pixels): small wavefronts, same iteration count
    "c3_8th": (sponza_scale, 1360, 760, 64, 8),
}
