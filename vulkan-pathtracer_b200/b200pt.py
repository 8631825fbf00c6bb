"""Host-side mirror of the reference's RaytracingPass interface over libb200pt.so (ctypes).

    RaytracingPass.init(...)    -> Renderer(scene, width, height, num_samples, num_bounces, divider)
    RaytracingPass.record(...)  -> Renderer.record(camera)   (asynchronous, like a command buffer)
    frame fence                 -> Renderer.sync()
    RaytracingPass.resize(...)  -> Renderer.resize(w, h)
    RaytracingPass.deinit()     -> Renderer.deinit()

(/root/reference/src/RaytracingPass.zig:87-97, 520-528, 677-683, 202).  There is NO CPU fallback:
importing works anywhere (so the symbol table can be checked), but creating a Renderer without the
CUDA library or without a GPU raises.
"""
import ctypes as C
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from tools.scene import (PtCamera, PtInstance, PtMaterial, PtMesh, PtPrimitive, PtRenderConfig,  # noqa: E402,F401
                         PtSceneDesc, PtTexture)

# B200PT_LIB selects an alternative build of the same library (A/B kernel experiments)
LIB_PATH = os.environ.get("B200PT_LIB") or os.path.join(_HERE, "libb200pt.so")

PT_PARTITION_NONE, PT_PARTITION_TILES = 0, 1
PT_FLAG_SORT_BY_MATERIAL, PT_FLAG_TRACE_BVH2, PT_FLAG_NO_PRIMARY_REUSE, PT_FLAG_TRACE_WALKER, PT_FLAG_BVH_PLOC, PT_FLAG_NO_TEXTURE_INTERLEAVE = 1, 2, 4, 8, 16, 32
PT_FLAG_BVH_SAH, PT_FLAG_BVH_OPTIMAL_COLLAPSE, PT_FLAG_RUSSIAN_ROULETTE, PT_FLAG_BVH_LBVH, PT_FLAG_SKIP_DEAD_SEGMENTS, PT_FLAG_HW_BILINEAR = 64, 128, 256, 512, 1024, 2048

PtHit_dtype = np.dtype([("t", "<f4"), ("b1", "<f4"), ("b2", "<f4"), ("instance", "<u4"), ("primitive_index", "<u4"),
                        ("triangle_index", "<u4")])
NOALLOC_MAX_RAYS = 131072   # PT_NOALLOC_MAX_RAYS
PtRay_dtype = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4")])


class PtStats(C.Structure):
    _fields_ = [("rays_traced", C.c_uint64), ("rays_hit", C.c_uint64), ("wavefront_iterations", C.c_uint32),
                ("kernel_launches", C.c_uint32), ("ms_trace", C.c_float), ("ms_shade", C.c_float), ("ms_other", C.c_float),
                ("ms_total", C.c_float), ("bvh_nodes", C.c_uint64), ("bvh_bytes", C.c_uint64), ("num_triangles", C.c_uint64),
                ("ms_build", C.c_float)]


class PtCameraState(C.Structure):
    """Camera (Camera.zig:10-23): angles in degrees, rotation (w, x, y, z)."""
    _fields_ = [("fov", C.c_float), ("aspect_ratio", C.c_float), ("position", C.c_float * 3), ("rotation", C.c_float * 4),
                ("pitch", C.c_float), ("yaw", C.c_float), ("forward", C.c_float * 3), ("left", C.c_float * 3),
                ("up", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3)]


class PtInput(C.Structure):
    """One frame of Input + Timer as Camera.update reads them (Input.zig:16-19, Timer.zig:6)."""
    _fields_ = [("keys", C.c_uint32), ("mouse_left", C.c_uint32), ("cursor_delta_x", C.c_int32),
                ("cursor_delta_y", C.c_int32), ("delta_time", C.c_float)]


PT_KEY_W, PT_KEY_S, PT_KEY_D, PT_KEY_A, PT_KEY_SPACE, PT_KEY_LEFT_CTRL, PT_KEY_LEFT_SHIFT = 1, 2, 4, 8, 16, 32, 64
KEY_NAMES = {"w": PT_KEY_W, "s": PT_KEY_S, "d": PT_KEY_D, "a": PT_KEY_A, "space": PT_KEY_SPACE,
             "left_ctrl": PT_KEY_LEFT_CTRL, "left_shift": PT_KEY_LEFT_SHIFT}


class PtError(RuntimeError):
    def __init__(self, status, name, message):
        super().__init__(f"{name} ({status}): {message}")
        self.status, self.name, self.message = status, name, message


_lib = None


def lib():
    """Load libb200pt.so (fails loudly when it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PtError(-2, "CudaError", f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
                          "(make -C vulkan-pathtracer_b200/csrc); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.pt_last_error.restype = C.c_char_p
        L.pt_status_name.restype = C.c_char_p
        L.pt_status_name.argtypes = [C.c_int]
        L.pt_version.restype = C.c_char_p
        L.pt_create.argtypes = [C.POINTER(PtSceneDesc), C.POINTER(PtRenderConfig), C.POINTER(C.c_void_p)]
        L.pt_destroy.argtypes = [C.c_void_p]
        L.pt_destroy.restype = None
        L.pt_render.argtypes = [C.c_void_p, C.POINTER(PtCamera), C.c_void_p]
        L.pt_sync.argtypes = [C.c_void_p]
        L.pt_resize.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.pt_render_extent.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        for n in ("pt_read_rgba32f", "pt_read_bgra8", "pt_read_rgba32f_upscaled"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_void_p]
        for n in ("pt_read_rgba32f_async", "pt_read_bgra8_async"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.pt_device_image.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.pt_read_ids.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pt_set_accumulate.argtypes = [C.c_void_p, C.c_int]
        L.pt_reset_accum.argtypes = [C.c_void_p]
        L.pt_device_accum.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
        L.pt_read_accum_mean.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.pt_set_profiling.argtypes = [C.c_void_p, C.c_int]
        L.pt_get_stats.argtypes = [C.c_void_p, C.POINTER(PtStats)]
        L.pt_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.pt_trace_rays_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.pt_count_traversal.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.pt_read_rgba32f_overlapped.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_ipc_export_image.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_ipc_open_root_image.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_camera_new.argtypes = [C.c_float, C.c_float, C.POINTER(C.c_float * 3), C.c_float, C.c_float, C.c_uint32,
                                    C.POINTER(PtCamera)]
        L.pt_camera_state_new.argtypes = [C.c_float, C.c_float, C.POINTER(C.c_float * 3), C.POINTER(PtCameraState)]
        L.pt_camera_update.argtypes = [C.POINTER(PtCameraState), C.POINTER(PtInput)]
        L.pt_camera_update_aspect_ratio.argtypes = [C.POINTER(PtCameraState), C.c_float]
        L.pt_camera_pack.argtypes = [C.POINTER(PtCameraState), C.c_uint32, C.POINTER(PtCamera)]
        L.pt_create_from_gltf.argtypes = [C.c_char_p, C.POINTER(PtRenderConfig), C.POINTER(C.c_void_p)]
        L.pt_create_sibling.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.pt_debug_bvh_digest.argtypes = [C.c_void_p, C.POINTER(C.c_uint64 * 6)]
        L.pt_set_instance_transforms.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_float)]
        L.pt_write_pfm.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.pt_write_exr.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        if hasattr(L, "pt_comm_init"):      # absent only from A/B variant builds that predate the multi-GPU surface (B200PT_LIB)
            L.pt_comm_available.argtypes = [C.POINTER(C.c_int)]
            L.pt_comm_unique_id.argtypes = [C.c_void_p]
            L.pt_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
            L.pt_comm_init_all.argtypes = [C.POINTER(C.c_void_p), C.c_int]
            L.pt_comm_rank.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
            L.pt_reduce_image.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
            L.pt_reduce_accum.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
            L.pt_comm_destroy.argtypes = [C.c_void_p]
            L.pt_add_accum.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_scene_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.pt_scene_desc.argtypes = [C.c_void_p, C.POINTER(PtSceneDesc)]
        L.pt_scene_free.argtypes = [C.c_void_p]
        L.pt_scene_free.restype = None
        _lib = L
    return _lib


def check(status):
    if status != 0:
        L = lib()
        raise PtError(status, L.pt_status_name(status).decode(), (L.pt_last_error() or b"").decode())


def comm_available():
    """NCCL version code when the library could open NCCL, else 0."""
    v = C.c_int(0)
    return v.value if lib().pt_comm_available(C.byref(v)) else 0


def comm_unique_id():
    """ncclGetUniqueId: 128 bytes to ship from rank 0 to every other rank."""
    buf = (C.c_uint8 * 128)()
    check(lib().pt_comm_unique_id(buf))
    return bytes(buf)


def comm_init_all(renderers):
    """One process, one renderer per device: renderer i becomes rank i (ncclCommInitAll)."""
    arr = (C.c_void_p * len(renderers))(*[r.handle for r in renderers])
    check(lib().pt_comm_init_all(arr, len(renderers)))


class comm_group:
    """ncclGroupStart / ncclGroupEnd around the per-rank reduce calls that ONE thread issues for several devices."""

    def __enter__(self):
        check(lib().pt_comm_group_start())

    def __exit__(self, *exc):
        check(lib().pt_comm_group_end())
        return False


def device_count():
    return lib().pt_device_count()


def camera_new(fov, aspect_ratio, position=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0, frame_count=0):
    """Camera.new(fov, aspect, position) (Camera.zig:25-46) through the C ABI."""
    cam = PtCamera()
    pos = (C.c_float * 3)(*position)
    check(lib().pt_camera_new(fov, aspect_ratio, C.byref(pos), yaw, pitch, frame_count, C.byref(cam)))
    return cam


class Camera:
    """The reference's Camera value (Camera.zig) for replaying recorded input: new / update / updateAspectRatio."""

    def __init__(self, fov, aspect_ratio, position=(0.0, 0.0, 0.0)):
        self.state = PtCameraState()
        pos = (C.c_float * 3)(*position)
        check(lib().pt_camera_state_new(fov, aspect_ratio, C.byref(pos), C.byref(self.state)))

    def update(self, keys=(), mouse_left=False, cursor_delta=(0, 0), delta_time=0.0):
        """Camera.update(input, timer) (Camera.zig:48-79); `keys` are names from KEY_NAMES or a PT_KEY_* mask."""
        mask = keys if isinstance(keys, int) else sum(KEY_NAMES[k] for k in set(keys))
        inp = PtInput(mask, 1 if mouse_left else 0, int(cursor_delta[0]), int(cursor_delta[1]), delta_time)
        check(lib().pt_camera_update(C.byref(self.state), C.byref(inp)))
        return self

    def update_aspect_ratio(self, aspect_ratio):
        check(lib().pt_camera_update_aspect_ratio(C.byref(self.state), aspect_ratio))
        return self

    def pack(self, frame_count=0):
        """The push constants of one frame (RaytracingPass.zig:563-569)."""
        cam = PtCamera()
        check(lib().pt_camera_pack(C.byref(self.state), frame_count, C.byref(cam)))
        return cam


class LoadedScene:
    """Scene.load(path) (Scene.zig:79) through the C ABI; owns the host arrays."""

    def __init__(self, path):
        self.handle = C.c_void_p()
        check(lib().pt_scene_load(os.fsencode(path), C.byref(self.handle)))
        self._desc = PtSceneDesc()
        check(lib().pt_scene_desc(self.handle, C.byref(self._desc)))

    def desc(self):
        return self._desc

    def close(self):
        if self.handle:
            lib().pt_scene_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Renderer:
    """RaytracingPass over CUDA.  `scene` is anything with a .desc() -> PtSceneDesc
    (tools.scene.PackedScene or LoadedScene)."""

    def __init__(self, scene, width=1920, height=1080, num_samples=1, num_bounces=2, divider=1, device=0, rank=0, world=1,
                 partition=PT_PARTITION_NONE, flags=0):
        self.handle = C.c_void_p()
        self.cfg = PtRenderConfig(width, height, num_samples, num_bounces, divider, device, rank, world, partition, flags)
        desc = scene.desc()
        check(lib().pt_create(C.byref(desc), C.byref(self.cfg), C.byref(self.handle)))
        w, h = C.c_uint32(), C.c_uint32()
        check(lib().pt_render_extent(self.handle, C.byref(w), C.byref(h)))
        self.width, self.height = w.value, h.value

    @classmethod
    def from_gltf(cls, path, width=1920, height=1080, num_samples=1, num_bounces=2, divider=1, device=0, flags=0):
        """RaytracingPass.init with a scene path (RaytracingPass.zig:87-103)."""
        self = cls.__new__(cls)
        self.handle = C.c_void_p()
        self.cfg = PtRenderConfig(width, height, num_samples, num_bounces, divider, device, 0, 1, PT_PARTITION_NONE, flags)
        check(lib().pt_create_from_gltf(os.fsencode(path), C.byref(self.cfg), C.byref(self.handle)))
        w, h = C.c_uint32(), C.c_uint32()
        check(lib().pt_render_extent(self.handle, C.byref(w), C.byref(h)))
        self.width, self.height = w.value, h.value
        return self

    def sibling(self):
        """A second frame context over the same device scene (frames in flight, main.zig:209-217,255-259): own stream,
        path state and image.  Deinit it before this renderer."""
        other = Renderer.__new__(Renderer)
        other.handle = C.c_void_p()
        other.cfg = self.cfg
        check(lib().pt_create_sibling(self.handle, C.byref(other.handle)))
        other.width, other.height = self.width, self.height
        return other

    # -- RaytracingPass.record / fence / resize / deinit
    def record(self, camera, stream=None):
        check(lib().pt_render(self.handle, C.byref(camera), stream))

    def sync(self):
        check(lib().pt_sync(self.handle))

    def resize(self, width, height):
        check(lib().pt_resize(self.handle, width, height))
        w, h = C.c_uint32(), C.c_uint32()
        check(lib().pt_render_extent(self.handle, C.byref(w), C.byref(h)))
        self.width, self.height = w.value, h.value

    def deinit(self):
        if self.handle:
            lib().pt_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.deinit()
        except Exception:
            pass

    # -- outputs
    def image(self):
        out = np.empty((self.height, self.width, 4), np.float32)
        check(lib().pt_read_rgba32f(self.handle, out.ctypes.data))
        return out

    def image_bgra8(self):
        out = np.empty((self.height, self.width, 4), np.uint8)
        check(lib().pt_read_bgra8(self.handle, out.ctypes.data))
        return out

    def image_upscaled(self):
        out = np.empty((self.cfg.height, self.cfg.width, 4), np.float32)
        check(lib().pt_read_rgba32f_upscaled(self.handle, out.ctypes.data))
        return out

    def ids(self):
        a = [np.empty((self.height, self.width), np.uint32) for _ in range(3)]
        check(lib().pt_read_ids(self.handle, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data))
        return tuple(a)

    def device_image_ptr(self):
        p = C.c_void_p()
        check(lib().pt_device_image(self.handle, C.byref(p)))
        return p.value

    def set_accumulate(self, on):
        check(lib().pt_set_accumulate(self.handle, 1 if on else 0))

    def reset_accum(self):
        check(lib().pt_reset_accum(self.handle))

    def device_accum_ptr(self):
        p, n = C.c_void_p(), C.c_uint32()
        check(lib().pt_device_accum(self.handle, C.byref(p), C.byref(n)))
        return p.value, n.value

    def accum_mean(self, total_frames=0):
        out = np.empty((self.height, self.width, 4), np.float32)
        check(lib().pt_read_accum_mean(self.handle, out.ctypes.data, total_frames))
        return out

    # -- multi-GPU combine (NCCL inside the library; include/b200pt.h "multi-GPU combine")
    def comm_init(self, rank, world, unique_id):
        """One process per GPU: `unique_id` = the 128 bytes of comm_unique_id() made on rank 0."""
        buf = (C.c_uint8 * 128)(*bytes(unique_id))
        check(lib().pt_comm_init(self.handle, rank, world, buf))

    def reduce_image(self, root=0, stream=None):
        check(lib().pt_reduce_image(self.handle, root, stream))

    def reduce_accum(self, root=0, stream=None):
        check(lib().pt_reduce_accum(self.handle, root, stream))

    def add_accum(self, other):
        check(lib().pt_add_accum(self.handle, other.handle))

    def set_instance_transforms(self, transforms):
        """Animated instances: new (n, 16) row-major matrices, top level rebuilt on the device; returns the ms."""
        t = np.ascontiguousarray(transforms, np.float32).reshape(-1, 16)
        ms = C.c_float(0.0)
        check(lib().pt_set_instance_transforms(self.handle, t.ctypes.data, t.shape[0], C.byref(ms)))
        return ms.value

    def bvh_digest(self):
        d = (C.c_uint64 * 6)()
        check(lib().pt_debug_bvh_digest(self.handle, C.byref(d)))
        return tuple(d)

    def set_profiling(self, on):
        check(lib().pt_set_profiling(self.handle, 1 if on else 0))

    def stats(self):
        s = PtStats()
        check(lib().pt_get_stats(self.handle, C.byref(s)))
        return s

    def trace(self, rays):
        rays = np.ascontiguousarray(rays, dtype=PtRay_dtype)
        hits = np.empty(len(rays), PtHit_dtype)
        check(lib().pt_trace_rays(self.handle, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def trace_noalloc(self, device_rays, n, device_hits, stream=None):
        """pt_trace_rays_noalloc: DEVICE addresses (ints) of n PtRay / n PtHit records; enqueued on `stream` (None = the
        renderer's own, which sync() waits for).  No allocation; n <= NOALLOC_MAX_RAYS."""
        check(lib().pt_trace_rays_noalloc(self.handle, device_rays, n, device_hits, stream))

    def count_traversal(self):
        n, t, r = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().pt_count_traversal(self.handle, C.byref(n), C.byref(t), C.byref(r)))
        return n.value, t.value, r.value


def read_bandwidth(nbytes, iters=50, device=0):
    """(GB/s, SM count): read bandwidth of an nbytes device buffer (fits the L2 -> L2 bandwidth; large -> HBM)."""
    g, n = C.c_float(0.0), C.c_int(0)
    L = lib()
    L.pt_debug_read_bandwidth.argtypes = [C.c_int, C.c_size_t, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int)]
    check(L.pt_debug_read_bandwidth(device, nbytes, iters, C.byref(g), C.byref(n)))
    return g.value, n.value


def debug_bsdf(N, V, kd, f0, rough, xi, device=0):
    """sample_brdf / get_brdf_density / brdf evaluated by the CUDA restatement."""
    n = len(rough)
    a = [np.ascontiguousarray(x, np.float32) for x in (N, V, kd, f0, rough, xi)]
    L, pdf, f = np.empty((n, 3), np.float32), np.empty(n, np.float32), np.empty((n, 3), np.float32)
    fn = lib().pt_debug_bsdf
    fn.argtypes = [C.c_int, C.c_size_t] + [C.c_void_p] * 9
    check(fn(device, n, *[x.ctypes.data for x in a], L.ctypes.data, pdf.ctypes.data, f.ctypes.data))
    return L, pdf, f


def debug_rng(pixel_xy_frame, draws, device=0):
    p = np.ascontiguousarray(pixel_xy_frame, np.uint32).reshape(-1, 3)
    state, xi = np.empty((len(p), 2), np.uint32), np.empty((len(p), 2), np.float32)
    fn = lib().pt_debug_rng
    fn.argtypes = [C.c_int, C.c_size_t, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
    check(fn(device, len(p), p.ctypes.data, draws, state.ctypes.data, xi.ctypes.data))
    return state, xi


def debug_texture(rgba, uv, device=0):
    rgba = np.ascontiguousarray(rgba, np.uint8)
    uv = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    out = np.empty((len(uv), 4), np.float32)
    fn = lib().pt_debug_texture
    fn.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p]
    check(fn(device, rgba.ctypes.data, rgba.shape[1], rgba.shape[0], len(uv), uv.ctypes.data, out.ctypes.data))
    return out


def set_option(key, value):
    fn = lib().pt_set_option
    fn.argtypes = [C.c_char_p, C.c_int]
    check(fn(key.encode(), int(value)))


def write_pfm(path, rgba):
    rgba = np.ascontiguousarray(rgba, np.float32)
    check(lib().pt_write_pfm(os.fsencode(path), rgba.ctypes.data, rgba.shape[1], rgba.shape[0]))


def write_exr(path, rgba):
    rgba = np.ascontiguousarray(rgba, np.float32)
    check(lib().pt_write_exr(os.fsencode(path), rgba.ctypes.data, rgba.shape[1], rgba.shape[0]))
