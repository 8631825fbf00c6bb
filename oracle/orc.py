"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so) and of the reference's own shader
sources (ray_gen.rgen, closest_hit.rchit, miss.rmiss, triangle/material/brdf/rand/common.glsl) compiled
as C++ (oracle/_ref/libref_glsl.so).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
REF_PATH = os.path.join(HERE, "_ref", "libref_glsl.so")


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference exists) with oracle/Makefile."""
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "oracle.cpp")):
        subprocess.check_call(["make", "-C", HERE, "_build/liboracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/src/shaders"):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


import sys
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from tools.scene import (PtCamera, PtInstance, PtMaterial, PtMesh, PtPrimitive, PtRenderConfig,  # noqa: E402,F401
                         PtSceneDesc, PtTexture)

PtHit_dtype = np.dtype([("t", "<f4"), ("b1", "<f4"), ("b2", "<f4"), ("instance", "<u4"), ("primitive_index", "<u4"),
                        ("triangle_index", "<u4")])
PtRay_dtype = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4")])

_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_create.argtypes = [C.POINTER(PtSceneDesc), C.c_int]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_render.restype = C.c_uint64
        L.orc_render.argtypes = [C.c_void_p, C.POINTER(PtRenderConfig), C.POINTER(PtCamera), C.c_int, C.c_int,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_render_rows.restype = C.c_uint64
        L.orc_render_rows.argtypes = [C.c_void_p, C.POINTER(PtRenderConfig), C.POINTER(PtCamera), C.c_int, C.c_int,
                                      C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orc_intersect_tri.restype = C.c_int
        L.orc_hardware_threads.restype = C.c_int
        L.orc_shade_hits.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        _lib = L
    return _lib


def ref_lib():
    """The reference's brdf.glsl/rand.glsl as C++; None when it was never built (no /root/reference)."""
    global _ref
    if _ref is None and os.path.exists(REF_PATH):
        R = C.CDLL(REF_PATH)
        if hasattr(R, "ref_render"):
            R.ref_render.restype = C.c_uint64
            R.ref_render.argtypes = [C.POINTER(PtSceneDesc), C.POINTER(PtRenderConfig), C.POINTER(PtCamera), C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_int, C.c_void_p]
            R.ref_shade_hits.argtypes = [C.POINTER(PtSceneDesc), C.c_void_p, C.c_size_t, C.c_void_p]
        _ref = R
    return _ref


def ref_alt_lib():
    """The same reference shaders under the other legal precision reading (glsl_compat.h, GLSL_ALT_PRECISION)."""
    path = os.path.join(HERE, "_ref", "libref_glsl_alt.so")
    if not os.path.exists(path):
        return None
    R = C.CDLL(path)
    R.ref_render.restype = C.c_uint64
    R.ref_render.argtypes = [C.POINTER(PtSceneDesc), C.POINTER(PtRenderConfig), C.POINTER(PtCamera), C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    return R


def _fp(a):
    return a.ctypes.data_as(C.c_void_p)


def bsdf_batch(L, N, V, kd, f0, rough, xi, prefix="orc"):
    """Run sample_brdf -> get_brdf_density -> brdf on n records; returns (L, pdf, f)."""
    n = len(rough)
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (N, V, kd, f0, rough, xi)]
    Lo = np.empty((n, 3), np.float32)
    pdf = np.empty(n, np.float32)
    f = np.empty((n, 3), np.float32)
    fn = getattr(L, prefix + "_bsdf_batch")
    fn.argtypes = [C.c_size_t] + [C.c_void_p] * 9
    fn(n, *[_fp(a) for a in arrs], _fp(Lo), _fp(pdf), _fp(f))
    return Lo, pdf, f


class OracleScene:
    """Owns an oracle-side acceleration structure over a python scene (see tools/scene.py)."""

    def __init__(self, scene, with_bvh=True):
        self.scene = scene  # keeps the numpy arrays alive
        self.desc = scene.desc()
        self.handle = lib().orc_scene_create(C.byref(self.desc), 1 if with_bvh else 0)

    def close(self):
        if self.handle:
            lib().orc_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        self.close()

    def render(self, cfg, cam, use_bvh=True, nthreads=None, ids=True):
        W, H = cfg.width // max(1, cfg.divider), cfg.height // max(1, cfg.divider)
        rgba = np.zeros((H, W, 4), np.float32)
        inst = np.full((H, W), 0xFFFFFFFF, np.uint32) if ids else None
        prim = np.full((H, W), 0xFFFFFFFF, np.uint32) if ids else None
        tri = np.full((H, W), 0xFFFFFFFF, np.uint32) if ids else None
        nt = nthreads or lib().orc_hardware_threads()
        rays = lib().orc_render(self.handle, C.byref(cfg), C.byref(cam), 1 if use_bvh else 0, nt, _fp(rgba),
                                _fp(inst) if ids else None, _fp(prim) if ids else None, _fp(tri) if ids else None)
        return rgba, (inst, prim, tri), rays

    def render_rows(self, cfg, cam, y0, y1, row_step=1, nthreads=None):
        W, H = cfg.width // max(1, cfg.divider), cfg.height // max(1, cfg.divider)
        rgba = np.zeros((H, W, 4), np.float32)
        nt = nthreads or lib().orc_hardware_threads()
        rays = lib().orc_render_rows(self.handle, C.byref(cfg), C.byref(cam), 1, nt, y0, y1, row_step, _fp(rgba))
        return rgba, rays

    def shade_hits(self, hits):
        """getTriangleData + getMaterialData of every hit -> (n, 25) float32 (layout: orc_shade_hits)."""
        hits = np.ascontiguousarray(hits, dtype=PtHit_dtype)
        out = np.zeros((len(hits), 25), np.float32)
        lib().orc_shade_hits(self.handle, _fp(hits), len(hits), _fp(out))
        return out

    # -- the same two through the REFERENCE's own shader sources (oracle/_ref/libref_glsl.so); the acceleration-structure
    # query behind traceRayEXT is this oracle's closest_hit (the Vulkan driver's part has no source)
    def ref_render(self, cfg, cam, use_bvh=True, nthreads=None, alt=False):
        W, H = cfg.width // max(1, cfg.divider), cfg.height // max(1, cfg.divider)
        rgba = np.zeros((H, W, 4), np.float32)
        nt = nthreads or lib().orc_hardware_threads()
        fn = C.cast(lib().orc_closest_hit_one, C.c_void_p)
        rays = (ref_alt_lib() if alt else ref_lib()).ref_render(C.byref(self.desc), C.byref(cfg), C.byref(cam), fn, self.handle, 1 if use_bvh else 0, nt, _fp(rgba))
        return rgba, rays

    def ref_shade_hits(self, hits):
        hits = np.ascontiguousarray(hits, dtype=PtHit_dtype)
        out = np.zeros((len(hits), 25), np.float32)
        ref_lib().ref_shade_hits(C.byref(self.desc), _fp(hits), len(hits), _fp(out))
        return out

    def trace(self, rays, use_bvh=True, nthreads=None):
        rays = np.ascontiguousarray(rays, dtype=PtRay_dtype)
        hits = np.empty(len(rays), PtHit_dtype)
        nt = nthreads or lib().orc_hardware_threads()
        lib().orc_trace(self.handle, _fp(rays), len(rays), 1 if use_bvh else 0, nt, _fp(hits))
        return hits
