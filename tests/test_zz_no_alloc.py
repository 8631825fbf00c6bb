"""pt_trace_rays_noalloc (include/b200pt.h): the closest-hit query with its scratch in a fixed __device__ array,
for callers that cannot call cudaMalloc.  Checked through the C ABI against the existing entry point,
pt_trace_rays, on the same rays: bit-exact (the suite's bar for hit records), on calls in a row that share the
scratch array.  Collected last on purpose (the file name).

The test owns the device buffers (torch is only the device-memory allocator here); the entry point is given their
addresses."""
import numpy as np
import pytest

import b200pt
from parity_util import hits_mismatch, random_rays
from tools import scene as S

pytestmark = pytest.mark.gpu
PT_ERR_INVALID_ARGUMENT = -1   # include/b200pt.h


def to_device(torch, rays):
    return torch.from_numpy(np.frombuffer(rays.tobytes(), np.uint8).copy()).cuda()


def from_device(t):
    return np.frombuffer(t.cpu().numpy().tobytes(), b200pt.PtHit_dtype)


@pytest.mark.parametrize("scene", ["instanced", "flat"])
def test_no_alloc_query_matches_the_allocating_one_on_calls_in_a_row(scene):
    import torch
    # instanced: 3 meshes under 7 transformed instances (general kernel); flat: one identity instance (kFlat kernel)
    ps = (S.random_soup(3, 3000, 3, 7) if scene == "instanced" else S.cornell()).pack()
    r = b200pt.Renderer(ps, 32, 32, 1, 1)
    batches = [random_rays(21, 100000, dtype=b200pt.PtRay_dtype),
               random_rays(22, 4097, dtype=b200pt.PtRay_dtype),        # fewer rays than the call before: stale slots must not leak
               random_rays(23, b200pt.NOALLOC_MAX_RAYS, dtype=b200pt.PtRay_dtype)]   # exactly the maximum
    batches[0]["dir"][:300] = np.tile(np.eye(3, dtype=np.float32), (100, 1))   # axis-parallel directions
    want = [r.trace(b) for b in batches]                                        # existing entry point (allocates per call)
    d_rays = [to_device(torch, b) for b in batches]
    d_hits = [torch.full((len(b) * b200pt.PtHit_dtype.itemsize,), 0xAB, dtype=torch.uint8, device="cuda") for b in batches]
    torch.cuda.synchronize()
    # first round: also the set-up call of this device (symbol address look-up, lazy kernel load)
    for b, dr, dh in zip(batches, d_rays, d_hits):
        r.trace_noalloc(dr.data_ptr(), len(b), dh.data_ptr())   # enqueued back to back on the renderer's stream, no sync between
    r.sync()
    for i, (w, dh) in enumerate(zip(want, d_hits)):
        bad = hits_mismatch(from_device(dh), w)
        assert not bad.any(), f"call {i}: {int(bad.sum())} of {len(w)} hits differ from pt_trace_rays"
    # second round in the reverse order, into fresh output buffers; device memory in use must not move
    for dh in d_hits:
        dh.fill_(0xCD)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for b, dr, dh in reversed(list(zip(batches, d_rays, d_hits))):
        r.trace_noalloc(dr.data_ptr(), len(b), dh.data_ptr())
    r.sync()
    free_after = torch.cuda.mem_get_info()[0]
    for i, (w, dh) in enumerate(zip(want, d_hits)):
        bad = hits_mismatch(from_device(dh), w)
        assert not bad.any(), f"second round, batch {i}: {int(bad.sum())} of {len(w)} hits differ from pt_trace_rays"
    assert free_after >= free_before, f"device memory in use grew by {free_before - free_after} bytes across no-allocation calls"
    r.deinit()


def test_no_alloc_query_on_a_caller_stream():
    import torch
    ps = S.random_soup(5, 500, 2, 3).pack()
    r = b200pt.Renderer(ps, 32, 32, 1, 1)
    rays = random_rays(31, 20000, dtype=b200pt.PtRay_dtype)
    want = r.trace(rays)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dr = to_device(torch, rays)
        dh = torch.zeros(len(rays) * b200pt.PtHit_dtype.itemsize, dtype=torch.uint8, device="cuda")
        r.trace_noalloc(dr.data_ptr(), len(rays), dh.data_ptr(), stream=st.cuda_stream)
    st.synchronize()
    assert not hits_mismatch(from_device(dh), want).any()
    r.deinit()


def test_no_alloc_query_rejects_what_it_cannot_hold():
    import torch
    ps = S.cornell().pack()
    r = b200pt.Renderer(ps, 32, 32, 1, 1)
    n = b200pt.NOALLOC_MAX_RAYS + 1
    dr = torch.zeros(n * b200pt.PtRay_dtype.itemsize, dtype=torch.uint8, device="cuda")
    dh = torch.full((n * b200pt.PtHit_dtype.itemsize,), 0x5A, dtype=torch.uint8, device="cuda")
    with pytest.raises(b200pt.PtError) as e:
        r.trace_noalloc(dr.data_ptr(), n, dh.data_ptr())
    assert e.value.status == PT_ERR_INVALID_ARGUMENT and "PT_NOALLOC_MAX_RAYS" in e.value.message
    r.sync()
    assert bool((dh == 0x5A).all()), "a rejected call must not launch anything"
    with pytest.raises(b200pt.PtError):                      # null buffers
        r.trace_noalloc(None, 16, None)
    r.trace_noalloc(None, 0, None)                           # empty batch: nothing to do
    r.deinit()
    walker = b200pt.Renderer(ps, 32, 32, 1, 1, flags=b200pt.PT_FLAG_TRACE_WALKER)
    with pytest.raises(b200pt.PtError) as e:
        walker.trace_noalloc(dr.data_ptr(), 16, dh.data_ptr())
    assert e.value.status == PT_ERR_INVALID_ARGUMENT
    walker.deinit()
