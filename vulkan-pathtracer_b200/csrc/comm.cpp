// comm.cpp — multi-GPU combine inside the library ("per-rank accumulation buffers are combined with an NCCL
// reduce"; the C ABI takes (rank, world, ncclUniqueId bytes)).  One communicator per rank (one process per GPU:
// pt_comm_init) or per device of one process (pt_comm_init_all); a renderer's siblings use their parent's.
//
// NCCL is bound at run time (dlopen of libnccl.so.2) instead of at link time: a host that already carries an NCCL in its
// address space (PyTorch bundles its own) must not get a second one, and a single-GPU user needs none at all.  If the
// library cannot be opened every entry point here fails with PT_ERR_NCCL — there is no fallback path.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstring>
#include <mutex>
#include <string>

#include "renderer.h"

using namespace pt;

namespace {

// the slice of nccl.h (2.x ABI, stable since 2.0) this file needs
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
enum { ncclSuccess = 0, ncclInProgress = 7 };
enum { ncclFloat32 = 7 };
enum { ncclSum = 0 };

struct Nccl {
    void* so = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*CommGetAsyncError)(ncclComm_t, int*) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string why;
};

Nccl& nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {getenv("B200PT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            if (!nm || !*nm) continue;
            n.so = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (n.so) break;
            n.why = dlerror();
        }
        if (!n.so) return;
        bool ok = true;
        auto sym = [&](const char* s) { void* p = dlsym(n.so, s); if (!p) { ok = false; n.why = std::string("missing symbol ") + s; } return p; };
        n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
        n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
        n.CommInitAll = reinterpret_cast<decltype(n.CommInitAll)>(sym("ncclCommInitAll"));
        n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
        n.CommGetAsyncError = reinterpret_cast<decltype(n.CommGetAsyncError)>(sym("ncclCommGetAsyncError"));
        n.Reduce = reinterpret_cast<decltype(n.Reduce)>(sym("ncclReduce"));
        n.AllReduce = reinterpret_cast<decltype(n.AllReduce)>(sym("ncclAllReduce"));
        n.GroupStart = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
        n.GroupEnd = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
        n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
        n.GetVersion = reinterpret_cast<decltype(n.GetVersion)>(sym("ncclGetVersion"));
        if (!ok) { dlclose(n.so); n.so = nullptr; }
    });
    return n;
}

int need_nccl() {
    Nccl& n = nccl();
    if (n.so) return PT_OK;
    set_error("NCCL is not available (libnccl.so.2: " + n.why + "); set B200PT_NCCL_LIB or install NCCL");
    return PT_ERR_NCCL;
}

#define PT_NCCL_CHECK(expr)                                                                          \
    do {                                                                                             \
        const int _r = (expr);                                                                       \
        if (_r != ncclSuccess) {                                                                     \
            set_error(std::string(#expr) + ": " + nccl().GetErrorString(_r));                        \
            return PT_ERR_NCCL;                                                                      \
        }                                                                                            \
    } while (0)

}  // namespace

struct PtComm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    cudaStream_t stream = nullptr;      // communication stream: reduces run beside the next frame's kernels
    cudaEvent_t ev_ready = nullptr;     // "the buffer to reduce is complete", recorded on the render stream
};

namespace {

PtComm* comm_of(PtRenderer* r) { return r->comm ? r->comm : (r->parent ? r->parent->comm : nullptr); }

int finish_init(PtRenderer* r, ncclComm_t c, int rank, int world) {
    PtComm* pc = new PtComm();
    pc->comm = c; pc->rank = rank; pc->world = world;
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    PT_CUDA_CHECK(cudaStreamCreateWithFlags(&pc->stream, cudaStreamNonBlocking));
    PT_CUDA_CHECK(cudaEventCreateWithFlags(&pc->ev_ready, cudaEventDisableTiming));
    r->comm = pc;
    return PT_OK;
}

// an asynchronous communicator error (a peer died, a link failed) must surface as a status, not as a hang
int poll_async(PtComm* pc) {
    int async = ncclSuccess;
    PT_NCCL_CHECK(nccl().CommGetAsyncError(pc->comm, &async));
    if (async != ncclSuccess && async != ncclInProgress) {
        set_error(std::string("NCCL asynchronous error: ") + nccl().GetErrorString(async));
        return PT_ERR_NCCL;
    }
    return PT_OK;
}

// sum of `count` floats over all ranks into root's buffer, in place.  stream == nullptr: on the communication stream,
// ordered after everything enqueued so far on the renderer's render stream; the renderer's next frame waits for it only
// before it first writes the image (render_impl), and reads of the image are ordered after it (follow_stream).
int reduce_buffer(PtRenderer* r, float* buf, size_t count, int root, void* user_stream) {
    int rc = need_nccl(); if (rc != PT_OK) return rc;
    PtComm* pc = comm_of(r);
    if (!pc) { set_error("no communicator: call pt_comm_init / pt_comm_init_all first"); return PT_ERR_INVALID_ARGUMENT; }
    if (root < 0 || root >= pc->world) { set_error("reduce: root out of range"); return PT_ERR_INVALID_ARGUMENT; }
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    if ((rc = poll_async(pc)) != PT_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(user_stream);
    if (!st) {
        cudaStream_t render = r->last_stream ? r->last_stream : r->stream;
        PT_CUDA_CHECK(cudaEventRecord(pc->ev_ready, render));
        PT_CUDA_CHECK(cudaStreamWaitEvent(pc->stream, pc->ev_ready, 0));
        st = pc->stream;
    }
    PT_NCCL_CHECK(nccl().Reduce(buf, buf, count, ncclFloat32, ncclSum, root, pc->comm, st));
    if (!user_stream) {
        if (!r->ev_reduce_done) PT_CUDA_CHECK(cudaEventCreateWithFlags(&r->ev_reduce_done, cudaEventDisableTiming));
        PT_CUDA_CHECK(cudaEventRecord(r->ev_reduce_done, st));
        r->reduce_pending = true;
        r->last_stream = st;            // read-backs of this renderer follow the reduce
    }
    return PT_OK;
}

}  // namespace

void comm_release(PtRenderer* r) {
    PtComm* pc = r->comm;
    if (!pc) return;
    cudaSetDevice(r->device);
    if (pc->stream) cudaStreamSynchronize(pc->stream);
    if (pc->comm && nccl().so) nccl().CommDestroy(pc->comm);
    if (pc->ev_ready) cudaEventDestroy(pc->ev_ready);
    if (pc->stream) cudaStreamDestroy(pc->stream);
    delete pc;
    r->comm = nullptr;
}

extern "C" {

int pt_comm_available(int* nccl_version) {
    if (need_nccl() != PT_OK) return 0;
    if (nccl_version) nccl().GetVersion(nccl_version);
    return 1;
}

int pt_comm_unique_id(uint8_t id[128]) {
    if (!id) { set_error("pt_comm_unique_id: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    int rc = need_nccl(); if (rc != PT_OK) return rc;
    ncclUniqueId u;
    PT_NCCL_CHECK(nccl().GetUniqueId(&u));
    std::memcpy(id, u.internal, 128);
    return PT_OK;
}

int pt_comm_init(PtRenderer* r, int rank, int world, const uint8_t id[128]) {
    if (!r || !id) { set_error("pt_comm_init: null argument"); return PT_ERR_INVALID_ARGUMENT; }
    if (world < 1 || rank < 0 || rank >= world) { set_error("pt_comm_init: invalid rank/world"); return PT_ERR_INVALID_ARGUMENT; }
    if (r->comm || r->parent) { set_error("pt_comm_init: already initialised, or a sibling (siblings use their parent's communicator)"); return PT_ERR_INVALID_ARGUMENT; }
    int rc = need_nccl(); if (rc != PT_OK) return rc;
    PT_CUDA_CHECK(cudaSetDevice(r->device));
    ncclUniqueId u;
    std::memcpy(u.internal, id, 128);
    ncclComm_t c = nullptr;
    PT_NCCL_CHECK(nccl().CommInitRank(&c, world, u, rank));
    return finish_init(r, c, rank, world);
}

int pt_comm_init_all(PtRenderer** renderers, int n) {
    if (!renderers || n < 1 || n > 64) { set_error("pt_comm_init_all: invalid argument"); return PT_ERR_INVALID_ARGUMENT; }
    int rc = need_nccl(); if (rc != PT_OK) return rc;
    int devs[64];
    ncclComm_t comms[64];
    for (int i = 0; i < n; ++i) {
        if (!renderers[i] || renderers[i]->comm || renderers[i]->parent) { set_error("pt_comm_init_all: null, sibling or already initialised renderer"); return PT_ERR_INVALID_ARGUMENT; }
        devs[i] = renderers[i]->device;
        for (int j = 0; j < i; ++j)
            if (devs[j] == devs[i]) { set_error("pt_comm_init_all: two renderers on one device"); return PT_ERR_INVALID_ARGUMENT; }
    }
    PT_NCCL_CHECK(nccl().CommInitAll(comms, n, devs));
    for (int i = 0; i < n; ++i)
        if ((rc = finish_init(renderers[i], comms[i], i, n)) != PT_OK) return rc;
    return PT_OK;
}

int pt_comm_rank(PtRenderer* r, int* rank, int* world) {
    PtComm* pc = r ? comm_of(r) : nullptr;
    if (!pc) { set_error("pt_comm_rank: no communicator"); return PT_ERR_INVALID_ARGUMENT; }
    if (rank) *rank = pc->rank;
    if (world) *world = pc->world;
    return PT_OK;
}

int pt_comm_group_start(void) { int rc = need_nccl(); if (rc != PT_OK) return rc; PT_NCCL_CHECK(nccl().GroupStart()); return PT_OK; }
int pt_comm_group_end(void) { int rc = need_nccl(); if (rc != PT_OK) return rc; PT_NCCL_CHECK(nccl().GroupEnd()); return PT_OK; }

int pt_reduce_image(PtRenderer* r, int root, void* cuda_stream) {
    if (!r) { set_error("pt_reduce_image: null renderer"); return PT_ERR_INVALID_ARGUMENT; }
    return reduce_buffer(r, r->d_image, (size_t)r->pm.width * r->pm.height * 4, root, cuda_stream);
}

int pt_reduce_accum(PtRenderer* r, int root, void* cuda_stream) {
    if (!r) { set_error("pt_reduce_accum: null renderer"); return PT_ERR_INVALID_ARGUMENT; }
    if (!r->d_accum) { set_error("pt_reduce_accum: accumulation is not enabled (pt_set_accumulate) or no frame was rendered"); return PT_ERR_INVALID_ARGUMENT; }
    return reduce_buffer(r, r->d_accum, (size_t)r->pm.width * r->pm.height * 4, root, cuda_stream);
}

int pt_comm_destroy(PtRenderer* r) {
    if (!r) { set_error("pt_comm_destroy: null renderer"); return PT_ERR_INVALID_ARGUMENT; }
    comm_release(r);
    return PT_OK;
}

}  // extern "C"
