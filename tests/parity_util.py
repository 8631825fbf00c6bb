import numpy as np


def bits_equal(a, b):
    """Bitwise equality of fp32 arrays, NaN == NaN (NaN payloads differ between x86 and the GPU)."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def hits_mismatch(a, b):
    bad = np.zeros(len(a), bool)
    for f in ("instance", "primitive_index", "triangle_index"):
        bad |= a[f] != b[f]
    for f in ("t", "b1", "b2"):
        bad |= a[f].view(np.uint32) != b[f].view(np.uint32)
    return bad


def random_rays(seed, n, extent=8.0, dtype=None):
    rng = np.random.default_rng(seed)
    o = (rng.random((n, 3)) - 0.5) * extent * 2
    o[:, 2] -= extent * 0.2
    t = (rng.random((n, 3)) - 0.5) * extent + np.array([0, 0, extent * 0.75])
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros(n, dtype)
    rays["origin"] = o
    rays["dir"] = d
    rays["tmin"] = 1e-3
    rays["tmax"] = 1e4
    return rays
