// nd_sync.cu -- the multi-GPU outer sync as a bucketed NCCL pipeline (one process per GPU), and the
// single-GPU outer step with a host-resident master copy.  (The peer-memory and in-switch variants,
// which are faster on one NVSwitch node, live in nd_peer.cu.)
//
// Replaces the reference's per-tensor blocking `dist.all_reduce(AVG)` loop
// (nanodiloco/diloco/diloco.py:46-50) and its outer optimizer step
// (diloco.py:52-54) with a bucketed pipeline over NCCL / NVLink 5:
//
//   bucket b:  K1 pseudo-grad  ->  reduce-scatter(sum)  ->  K2 fused Nesterov on
//              this rank's slice  ->  all-gather of the new parameters  ->  master
//              refresh (device copy), each bucket on side stream b % n_streams so
//              RS(b+1) overlaps K2(b) and AG(b-1).
//
// NCCL is resolved at run time with dlopen so that the kernel library loads on a
// machine without NCCL (and on the CPU-only build container); the host process
// has normally loaded libnccl.so.2 already (torch does).
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "nd_common.cuh"
#include "nd_error.h"

namespace nd {

// ---- minimal NCCL surface (ABI-stable across 2.x) ---------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[ND_UNIQUE_ID_BYTES]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclFloat32 = 7, ncclBfloat16 = 9 };
enum { ncclSum = 0 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*ReduceScatter)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

static NcclApi g_nccl;

static int load_nccl() {
    if (g_nccl.handle) return ND_OK;
    const char* env = getenv("ND_NCCL_LIB");
    const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* nm : names) {
        if (!nm) continue;
        h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return set_error(ND_ENCCL, "cannot dlopen libnccl.so.2 (%s); set ND_NCCL_LIB", dlerror());
#define ND_SYM(field, name)                                                            \
    do {                                                                               \
        *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);                     \
        if (!g_nccl.field) return set_error(ND_ENCCL, "libnccl lacks symbol %s", name); \
    } while (0)
    ND_SYM(GetUniqueId, "ncclGetUniqueId");
    ND_SYM(CommInitRank, "ncclCommInitRank");
    ND_SYM(CommDestroy, "ncclCommDestroy");
    ND_SYM(ReduceScatter, "ncclReduceScatter");
    ND_SYM(AllGather, "ncclAllGather");
    ND_SYM(GetErrorString, "ncclGetErrorString");
#undef ND_SYM
    g_nccl.handle = h;
    return ND_OK;
}

#define ND_NCCL(call)                                                                              \
    do {                                                                                           \
        int r_ = (call);                                                                           \
        if (r_ != ncclSuccess)                                                                     \
            return set_error(ND_ENCCL, "%s failed: %s", #call, g_nccl.GetErrorString(r_));         \
    } while (0)

}  // namespace nd

using namespace nd;

struct nd_sync_ctx {
    ncclComm_t comm = nullptr;
    int world = 1, rank = 0;
    int64_t bucket = 0;  // elements per bucket, multiple of 4*world
    int n_streams = 1;
    std::vector<cudaStream_t> streams;
    std::vector<cudaEvent_t> done;  // one per side stream
    cudaEvent_t fork = nullptr, t0 = nullptr, t1 = nullptr;
    bool timed = false;
};

static int sync_ctx_streams(nd_sync_ctx* c) {
    c->streams.assign(c->n_streams, nullptr);
    c->done.assign(c->n_streams, nullptr);
    for (int i = 0; i < c->n_streams; ++i) {
        ND_CUDA(cudaStreamCreateWithFlags(&c->streams[i], cudaStreamNonBlocking));
        ND_CUDA(cudaEventCreateWithFlags(&c->done[i], cudaEventDisableTiming));
    }
    ND_CUDA(cudaEventCreateWithFlags(&c->fork, cudaEventDisableTiming));
    ND_CUDA(cudaEventCreate(&c->t0));
    ND_CUDA(cudaEventCreate(&c->t1));
    return ND_OK;
}

extern "C" {

int nd_sync_ctx_destroy(nd_sync_ctx* c);

int nd_nccl_unique_id(char out_host[ND_UNIQUE_ID_BYTES]) {
    if (!out_host) return set_error(ND_EINVAL, "nd_nccl_unique_id: null output");
    int rc = load_nccl();
    if (rc) return rc;
    ncclUniqueId id;
    ND_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(out_host, id.internal, ND_UNIQUE_ID_BYTES);
    return ND_OK;
}

int nd_sync_ctx_create(const char unique_id_host[ND_UNIQUE_ID_BYTES], int world, int rank, int64_t bucket_elems,
                       int n_streams, nd_sync_ctx** out) {
    if (!out) return set_error(ND_EINVAL, "nd_sync_ctx_create: null out");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return set_error(ND_EINVAL, "nd_sync_ctx_create: bad world/rank");
    if (bucket_elems < 1 || n_streams < 1 || n_streams > 16)
        return set_error(ND_EINVAL, "nd_sync_ctx_create: bad bucket_elems/n_streams");
    if (world > 1 && !unique_id_host) return set_error(ND_EINVAL, "nd_sync_ctx_create: null unique id");
    nd_sync_ctx* c = new nd_sync_ctx();
    c->world = world;
    c->rank = rank;
    const int64_t q = 4 * (int64_t)world;
    c->bucket = (bucket_elems + q - 1) / q * q;
    c->n_streams = n_streams;
    if (world > 1) {
        int rc = load_nccl();
        if (rc) { delete c; return rc; }
        ncclUniqueId id;
        memcpy(id.internal, unique_id_host, ND_UNIQUE_ID_BYTES);
        int r = g_nccl.CommInitRank(&c->comm, world, id, rank);
        if (r != ncclSuccess) {
            delete c;
            return set_error(ND_ENCCL, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
        }
    }
    if (int rc = sync_ctx_streams(c)) {   // streams, events: on failure release whatever exists so far
        nd_sync_ctx_destroy(c);
        return rc;
    }
    *out = c;
    return ND_OK;
}

int nd_sync_ctx_destroy(nd_sync_ctx* c) {
    if (!c) return ND_OK;
    for (auto s : c->streams) if (s) cudaStreamDestroy(s);
    for (auto e : c->done) if (e) cudaEventDestroy(e);
    if (c->fork) cudaEventDestroy(c->fork);
    if (c->t0) cudaEventDestroy(c->t0);
    if (c->t1) cudaEventDestroy(c->t1);
    if (c->comm) g_nccl.CommDestroy(c->comm);
    delete c;
    return ND_OK;
}

int64_t nd_sync_padded_elems(const nd_sync_ctx* c, int64_t n) {
    if (!c || n < 0) return -1;
    const int64_t q = 4 * (int64_t)c->world;
    return (n + q - 1) / q * q;
}

int64_t nd_sync_shard_elems(const nd_sync_ctx* c, int64_t n) {
    if (!c || n < 0) return -1;
    return nd_sync_padded_elems(c, n) / c->world;
}

int64_t nd_sync_bucket_elems(const nd_sync_ctx* c) { return c ? c->bucket : -1; }

int nd_outer_sync(nd_sync_ctx* c, float* local, float* master, float* mom_shard, float* pg, int64_t n_padded,
                  float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                  nd_stream_t stream_) {
    if (!c) return set_error(ND_ESTATE, "nd_outer_sync: null context");
    if (n_padded < 0 || n_padded % (4 * (int64_t)c->world) != 0)
        return set_error(ND_EINVAL, "nd_outer_sync: n_padded must be a multiple of 4*world (nd_sync_padded_elems)");
    if (n_padded && (!local || !master || !mom_shard || (c->world > 1 && !pg)))
        return set_error(ND_EINVAL, "nd_outer_sync: null pointer");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const int W = c->world;

    ND_CUDA(cudaEventRecord(c->t0, stream));
    if (W == 1) {
        // no exchange step: kernels (1)+(2) in one pass, 24 B/element
        void* locals[1] = {local};
        int rc = nd_outer_fused(master, mom_shard, locals, 1, ND_F32, n_padded, lr, mu, dampening, wd, nesterov,
                                first_step, nullptr, nullptr, stream_);
        if (rc) return rc;
        ND_CUDA(cudaEventRecord(c->t1, stream));
        c->timed = true;
        return ND_OK;
    }

    // fork: side streams wait for everything already queued on the caller's stream
    ND_CUDA(cudaEventRecord(c->fork, stream));
    for (int i = 0; i < c->n_streams; ++i) ND_CUDA(cudaStreamWaitEvent(c->streams[i], c->fork, 0));

    // Software-pipelined issue order: stage A(b+1) = {K1, reduce-scatter} is queued before
    // stage B(b) = {K2, all-gather, master refresh}, so that on the communicator RS(b+1)
    // is on the wire while K2(b) runs (NCCL executes one communicator's collectives in
    // issue order, whatever stream they are on).
    const int64_t nb = (n_padded + c->bucket - 1) / c->bucket;
    auto bucket_len = [&](int64_t b) {
        const int64_t off = b * c->bucket;
        return (n_padded - off < c->bucket) ? (n_padded - off) : c->bucket;  // multiple of 4*W
    };
    auto stage_a = [&](int64_t b) -> int {
        const int64_t off = b * c->bucket, len = bucket_len(b), slice = len / W;
        cudaStream_t s = c->streams[b % c->n_streams];
        float* pb = pg + off;
        int rc = nd_pseudo_grad(master + off, local + off, ND_F32, pb, len, reinterpret_cast<nd_stream_t>(s));
        if (rc) return rc;
        // the reduced slice lands in place of our own slice of the staging buffer
        ND_NCCL(g_nccl.ReduceScatter(pb, pb + (int64_t)c->rank * slice, (size_t)slice, ncclFloat32, ncclSum,
                                     c->comm, s));
        return ND_OK;
    };
    auto stage_b = [&](int64_t b) -> int {
        const int64_t off = b * c->bucket, len = bucket_len(b), slice = len / W;
        const int64_t mine = off + (int64_t)c->rank * slice;
        // every full bucket contributes bucket/W elements to this rank's shard
        const int64_t shard_off = b * (c->bucket / W);
        cudaStream_t s = c->streams[b % c->n_streams];
        int rc = nd_outer_nesterov(master + mine, mom_shard + shard_off, pg + mine, local + mine, ND_F32, slice,
                                   (float)W, lr, mu, dampening, wd, nesterov, first_step, nullptr, nullptr,
                                   reinterpret_cast<nd_stream_t>(s));
        if (rc) return rc;
        ND_NCCL(g_nccl.AllGather(local + mine, local + off, (size_t)slice, ncclFloat32, c->comm, s));
        // master replica refresh: theta_sync <- the gathered parameters (diloco.py:54)
        ND_CUDA(cudaMemcpyAsync(master + off, local + off, (size_t)len * sizeof(float), cudaMemcpyDeviceToDevice, s));
        return ND_OK;
    };
    if (nb > 0) {
        int rc = stage_a(0);
        if (rc) return rc;
    }
    for (int64_t b = 0; b < nb; ++b) {
        int rc;
        if (b + 1 < nb && (rc = stage_a(b + 1))) return rc;
        if ((rc = stage_b(b))) return rc;
    }
    // join
    for (int i = 0; i < c->n_streams; ++i) {
        ND_CUDA(cudaEventRecord(c->done[i], c->streams[i]));
        ND_CUDA(cudaStreamWaitEvent(stream, c->done[i], 0));
    }
    ND_CUDA(cudaEventRecord(c->t1, stream));
    c->timed = true;
    return ND_OK;
}

int nd_outer_step_offloaded(nd_sync_ctx* c, float* master_host, float* staging, int64_t chunk_elems, float* mom,
                            void* local, int local_dtype, int64_t n, float lr, float mu, float dampening, float wd,
                            int nesterov, int first_step, nd_stream_t stream_) {
    if (!c) return set_error(ND_ESTATE, "nd_outer_step_offloaded: null context");
    if (c->world != 1) return set_error(ND_ESTATE, "nd_outer_step_offloaded: world size must be 1");
    if (n < 0 || chunk_elems < 4 || chunk_elems % 4 != 0)
        return set_error(ND_EINVAL, "nd_outer_step_offloaded: chunk_elems must be a positive multiple of 4");
    if (n && (!master_host || !staging || !mom || !local))
        return set_error(ND_EINVAL, "nd_outer_step_offloaded: null pointer");
    if (local_dtype != ND_F32 && local_dtype != ND_BF16)
        return set_error(ND_EINVAL, "nd_outer_step_offloaded: unknown local_dtype");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const size_t lsz = local_dtype == ND_BF16 ? 2 : 4;
    ND_CUDA(cudaEventRecord(c->t0, stream));
    ND_CUDA(cudaEventRecord(c->fork, stream));
    for (int i = 0; i < c->n_streams; ++i) ND_CUDA(cudaStreamWaitEvent(c->streams[i], c->fork, 0));
    int64_t k = 0;
    for (int64_t off = 0; off < n; off += chunk_elems, ++k) {
        const int64_t len = (n - off < chunk_elems) ? (n - off) : chunk_elems;
        const int slot = (int)(k % c->n_streams);
        cudaStream_t s = c->streams[slot];  // stream order also protects the slot from reuse
        float* m_dev = staging + (int64_t)slot * chunk_elems;
        ND_CUDA(cudaMemcpyAsync(m_dev, master_host + off, (size_t)len * 4, cudaMemcpyHostToDevice, s));
        void* locals[1] = {static_cast<char*>(local) + (size_t)off * lsz};
        int rc = nd_outer_fused(m_dev, mom + off, locals, 1, local_dtype, len, lr, mu, dampening, wd, nesterov,
                                first_step, nullptr, nullptr, reinterpret_cast<nd_stream_t>(s));
        if (rc) return rc;
        ND_CUDA(cudaMemcpyAsync(master_host + off, m_dev, (size_t)len * 4, cudaMemcpyDeviceToHost, s));
    }
    for (int i = 0; i < c->n_streams; ++i) {
        ND_CUDA(cudaEventRecord(c->done[i], c->streams[i]));
        ND_CUDA(cudaStreamWaitEvent(stream, c->done[i], 0));
    }
    ND_CUDA(cudaEventRecord(c->t1, stream));
    c->timed = true;
    return ND_OK;
}

int nd_sync_last_ms(nd_sync_ctx* c, float* ms_out) {
    if (!c || !ms_out) return set_error(ND_EINVAL, "nd_sync_last_ms: null argument");
    if (!c->timed) return set_error(ND_ESTATE, "nd_sync_last_ms: no nd_outer_sync has run");
    ND_CUDA(cudaEventSynchronize(c->t1));
    ND_CUDA(cudaEventElapsedTime(ms_out, c->t0, c->t1));
    return ND_OK;
}

}  // extern "C"
