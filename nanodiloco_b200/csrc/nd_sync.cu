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
#include <utility>
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
    // optional (NCCL >= 2.19): buffers NCCL can map into the switch / use without staging copies
    int (*MemAlloc)(void**, size_t) = nullptr;
    int (*MemFree)(void*) = nullptr;
    int (*CommRegister)(ncclComm_t, void*, size_t, void**) = nullptr;
    int (*CommDeregister)(ncclComm_t, void*) = nullptr;
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
    *reinterpret_cast<void**>(&g_nccl.MemAlloc) = dlsym(h, "ncclMemAlloc");
    *reinterpret_cast<void**>(&g_nccl.MemFree) = dlsym(h, "ncclMemFree");
    *reinterpret_cast<void**>(&g_nccl.CommRegister) = dlsym(h, "ncclCommRegister");
    *reinterpret_cast<void**>(&g_nccl.CommDeregister) = dlsym(h, "ncclCommDeregister");
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
    ncclComm_t comm_ag = nullptr;     // optional second communicator: all-gathers stop queueing behind reduce-scatters
    std::vector<std::pair<ncclComm_t, void*>> registrations;   // (communicator, handle) of registered user buffers
    bool profile = false;             // record the per-bucket stage events below
    std::vector<cudaEvent_t> stage_events;   // kStageEvents per bucket
    int64_t profiled_buckets = 0;
    static constexpr int kStageEvents = 7;   // K1 start/end, RS end, K2 start/end, AG end, refresh end
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
    for (auto& r : c->registrations)
        if (g_nccl.CommDeregister) g_nccl.CommDeregister(r.first, r.second);
    for (auto e : c->stage_events) if (e) cudaEventDestroy(e);
    if (c->comm_ag) g_nccl.CommDestroy(c->comm_ag);
    if (c->comm) g_nccl.CommDestroy(c->comm);
    delete c;
    return ND_OK;
}

int nd_sync_ctx_add_ag_comm(nd_sync_ctx* c, const char unique_id_host[ND_UNIQUE_ID_BYTES]) {
    if (!c) return set_error(ND_ESTATE, "nd_sync_ctx_add_ag_comm: null context");
    if (c->world == 1 || c->comm_ag) return ND_OK;
    if (!unique_id_host) return set_error(ND_EINVAL, "nd_sync_ctx_add_ag_comm: null unique id");
    ncclUniqueId id;
    memcpy(id.internal, unique_id_host, ND_UNIQUE_ID_BYTES);
    ND_NCCL(g_nccl.CommInitRank(&c->comm_ag, c->world, id, c->rank));
    return ND_OK;
}

int nd_nccl_mem_alloc(size_t bytes, void** dev_ptr_out) {
    if (!dev_ptr_out || bytes == 0) return set_error(ND_EINVAL, "nd_nccl_mem_alloc: bad argument");
    if (int rc = load_nccl()) return rc;
    if (!g_nccl.MemAlloc) return set_error(ND_ENCCL, "this libnccl has no ncclMemAlloc (needs NCCL >= 2.19)");
    ND_NCCL(g_nccl.MemAlloc(dev_ptr_out, bytes));
    return ND_OK;
}

int nd_nccl_mem_free(void* dev_ptr) {
    if (!dev_ptr) return ND_OK;
    if (!g_nccl.MemFree) return set_error(ND_ENCCL, "this libnccl has no ncclMemFree");
    ND_NCCL(g_nccl.MemFree(dev_ptr));
    return ND_OK;
}

int nd_sync_register(nd_sync_ctx* c, void* buf, size_t bytes) {
    if (!c || !buf || bytes == 0) return set_error(ND_EINVAL, "nd_sync_register: bad argument");
    if (c->world == 1) return ND_OK;
    if (!g_nccl.CommRegister) return set_error(ND_ENCCL, "this libnccl has no ncclCommRegister (needs NCCL >= 2.19)");
    for (ncclComm_t comm : {c->comm, c->comm_ag}) {
        if (!comm) continue;
        void* h = nullptr;
        ND_NCCL(g_nccl.CommRegister(comm, buf, bytes, &h));
        c->registrations.emplace_back(comm, h);
    }
    return ND_OK;
}

int nd_sync_set_profiling(nd_sync_ctx* c, int on) {
    if (!c) return set_error(ND_ESTATE, "nd_sync_set_profiling: null context");
    c->profile = on != 0;
    return ND_OK;
}

int nd_sync_timeline(nd_sync_ctx* c, float* rows_out, int64_t cap_rows, int64_t* n_rows_out) {
    if (!c || !n_rows_out) return set_error(ND_EINVAL, "nd_sync_timeline: null argument");
    *n_rows_out = c->profiled_buckets;
    if (!rows_out || c->profiled_buckets == 0) return ND_OK;
    ND_CUDA(cudaEventSynchronize(c->t1));
    const int K = nd_sync_ctx::kStageEvents;
    for (int64_t b = 0; b < c->profiled_buckets && b < cap_rows; ++b)
        for (int k = 0; k < K; ++k)
            ND_CUDA(cudaEventElapsedTime(&rows_out[b * K + k], c->t0, c->stage_events[b * K + k]));
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
    NvtxRange range_all("nd_outer_sync (NCCL pipeline)");

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
    constexpr int KE = nd_sync_ctx::kStageEvents;
    c->profiled_buckets = 0;
    if (c->profile) {
        while ((int64_t)c->stage_events.size() < nb * KE) {
            cudaEvent_t e;
            ND_CUDA(cudaEventCreate(&e));
            c->stage_events.push_back(e);
        }
        c->profiled_buckets = nb;
    }
    auto mark = [&](int64_t b, int k, cudaStream_t s) -> int {
        if (c->profile) ND_CUDA(cudaEventRecord(c->stage_events[b * KE + k], s));
        return ND_OK;
    };
    auto bucket_len = [&](int64_t b) {
        const int64_t off = b * c->bucket;
        return (n_padded - off < c->bucket) ? (n_padded - off) : c->bucket;  // multiple of 4*W
    };
    auto stage_a = [&](int64_t b) -> int {
        const int64_t off = b * c->bucket, len = bucket_len(b), slice = len / W;
        cudaStream_t s = c->streams[b % c->n_streams];
        float* pb = pg + off;
        int rc;
        NvtxRange range("nd.sync.bucket: K1 pseudo-gradient + reduce-scatter");
        if ((rc = mark(b, 0, s))) return rc;
        rc = nd_pseudo_grad(master + off, local + off, ND_F32, pb, len, reinterpret_cast<nd_stream_t>(s));
        if (rc) return rc;
        if ((rc = mark(b, 1, s))) return rc;
        // the reduced slice lands in place of our own slice of the staging buffer
        ND_NCCL(g_nccl.ReduceScatter(pb, pb + (int64_t)c->rank * slice, (size_t)slice, ncclFloat32, ncclSum,
                                     c->comm, s));
        return mark(b, 2, s);
    };
    auto stage_b = [&](int64_t b) -> int {
        const int64_t off = b * c->bucket, len = bucket_len(b), slice = len / W;
        const int64_t mine = off + (int64_t)c->rank * slice;
        // every full bucket contributes bucket/W elements to this rank's shard
        const int64_t shard_off = b * (c->bucket / W);
        cudaStream_t s = c->streams[b % c->n_streams];
        int rc;
        NvtxRange range("nd.sync.bucket: K2 shard update + all-gather + master refresh");
        if ((rc = mark(b, 3, s))) return rc;
        rc = nd_outer_nesterov(master + mine, mom_shard + shard_off, pg + mine, local + mine, ND_F32, slice,
                               (float)W, lr, mu, dampening, wd, nesterov, first_step, nullptr, nullptr,
                               reinterpret_cast<nd_stream_t>(s));
        if (rc) return rc;
        if ((rc = mark(b, 4, s))) return rc;
        ND_NCCL(g_nccl.AllGather(local + mine, local + off, (size_t)slice, ncclFloat32,
                                 c->comm_ag ? c->comm_ag : c->comm, s));
        if ((rc = mark(b, 5, s))) return rc;
        // master replica refresh: theta_sync <- the slices other ranks own (diloco.py:54); ours was written by K2
        if (mine > off)
            ND_CUDA(cudaMemcpyAsync(master + off, local + off, (size_t)(mine - off) * sizeof(float),
                                    cudaMemcpyDeviceToDevice, s));
        if (mine + slice < off + len)
            ND_CUDA(cudaMemcpyAsync(master + mine + slice, local + mine + slice,
                                    (size_t)(off + len - mine - slice) * sizeof(float), cudaMemcpyDeviceToDevice, s));
        return mark(b, 6, s);
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
