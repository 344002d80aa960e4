// nd_peer.cu -- the outer sync as ONE kernel per GPU over NVLink peer memory.
//
// Every rank's working parameters live in a buffer that all ranks of the node map (CUDA IPC).
// Rank r owns the contiguous shard [r*n/W, (r+1)*n/W) of the flat parameter space: it keeps the
// last-sync master and the outer momentum for that shard only (1/W of each), and its launch of the
// fused kernel
//     reads  local_w[shard r] from every worker w   (W-1 of them over NVLink, TMA bulk loads)
//     sums   (master - local_w) in worker order, divides by W, applies the Nesterov update
//     writes the new parameters to local_w[shard r] of every worker (W-1 over NVLink, bulk stores)
// so compute and the exchange overlap tile by tile inside one kernel; there is no reduce-scatter,
// no all-gather, no staging buffer and no master replica.  Only rank r ever touches shard r of
// anybody's buffer, so the kernel itself needs no cross-GPU synchronisation; a flag barrier before
// it (every worker has finished its inner steps) and after it (every shard has been written) is
// enough.  Same wire bytes as reduce-scatter + all-gather: 8*(W-1)/W bytes per element per direction.
//
// Replaces nanodiloco/diloco/diloco.py:46-54 including the collective of :49.
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "nd_common.cuh"
#include "nd_error.h"
#include "nd_tma.cuh"

namespace nd {

// nd_kernels.cu: the fused outer kernel with the two cross-GPU rendezvous folded in
int outer_fused_rw_peer(float* master, float* mom, void* const* reads, int W, void* const* writes, int n_writes,
                        int local_dtype, int64_t n, float lr, float mu, float dampening, float wd, int nesterov,
                        int first_step, float* sumsq_out, void* scratch, cudaStream_t s, const tma::PeerSync& peer,
                        bool* launched);

constexpr size_t kFlagBytes = 4096;  // per-rank flag page appended to the shared buffer

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Barrier across the ranks of a node: thread w publishes `epoch` to slot [rank] of rank w's flag
// page, then waits until slot [w] of our own page reaches `epoch`.  System-scope release/acquire.
// The first barrier of an outer step absorbs the skew between workers (a rank that checkpoints,
// evaluates or stalls on its data loader arrives late), so the wait is bounded by `timeout_ns` of
// wall time -- ND_PEER_TIMEOUT_S, default 1800 s, 0 = wait for ever -- not by a few seconds.  When
// it does expire the waiting rank records WHO it was waiting for in `diag` (page-locked host memory
// the host can still read after the context is lost; nd_last_error / nd_peer_last_ms report it)
// and traps, rather than going on to average parameters that are not there.
__global__ void peer_barrier_kernel(uint32_t* const* flag_pages, int first, int count, int rank, uint32_t epoch,
                                    int channel, unsigned long long timeout_ns, int* diag) {
    if ((int)threadIdx.x >= count) return;
    const int w = first + threadIdx.x;   // the ranks [first, first + count) meet; the others are not involved
    __threadfence_system();
    // each channel has its own ND_MAX_WORKERS slots, so barriers issued on different streams (whose
    // relative order on the device is not fixed) can never satisfy one another
    uint32_t* remote = flag_pages[w] + channel * ND_MAX_WORKERS + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
    const uint32_t* mine = flag_pages[rank] + channel * ND_MAX_WORKERS + w;
    const unsigned long long t0 = global_ns();
    for (unsigned spins = 0;; ++spins) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if ((int32_t)(v - epoch) >= 0) break;
        if (timeout_ns && (spins & 1023u) == 1023u && global_ns() - t0 > timeout_ns) {
            if (diag) { diag[1] = w; diag[2] = channel; diag[3] = (int)epoch; __threadfence_system(); diag[0] = 1; }
            __threadfence_system();
            __trap();
        }
    }
    __threadfence_system();
}

// ---- push: every worker WRITES its slice of the working copy into the owner's staging area ------
// The pull form of the fused kernel reads W-1 remote copies; a remote read costs a request packet
// against the flow of the data, so with reads and writes both crossing the links each direction
// carries payload plus the other direction's requests.  In the push form nothing is ever read over
// NVLink: rank w bulk-stores local_w[shard r] into slot [worker(w)] of rank r's staging area for every
// owner r, and after a flag barrier rank r's fused kernel finds all W copies in its own HBM (which has
// bandwidth to spare) and only WRITES the new parameters out.  Wire bytes are the same 8(W-1)/W per
// element per direction; the links see pure write traffic.
//
// The kernel is a copy: no arithmetic, so ONE thread per CTA drives a ring of shared-memory stages
// with bulk loads (local HBM -> shared) and bulk stores (shared -> peer HBM).  Tile t of the launch
// goes to destination t % n_dst so that all links carry traffic at any moment.
struct PushArgs {
    const char* src;                    // this rank's working copy (byte 0 of shard 0)
    char* dst[ND_MAX_WORKERS];          // per destination: its staging slot for this worker
    int dst_shard[ND_MAX_WORKERS];      // which shard of the flat space that destination owns
    int n_dst;
    int64_t shard_bytes;
};

template <int TILE_BYTES, int STAGES, int STORES>
__global__ void __launch_bounds__(32, 2) push_kernel(PushArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint64_t full[STAGES];
    if (threadIdx.x != 0) return;
    const int64_t tiles_per_dst = (a.shard_bytes + TILE_BYTES - 1) / TILE_BYTES;
    const int64_t ntiles = tiles_per_dst * a.n_dst;
    for (int s = 0; s < STAGES; ++s) tma::mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    auto where = [&](int64_t t, int& d, int64_t& off, uint32_t& len) {
        d = (int)(t % a.n_dst);
        off = (t / a.n_dst) * (int64_t)TILE_BYTES;
        const int64_t left = a.shard_bytes - off;
        len = (uint32_t)(left < TILE_BYTES ? left : TILE_BYTES);
    };
    auto load = [&](int64_t t, int s) {
        int d; int64_t off; uint32_t len;
        where(t, d, off, len);
        tma::mbar_expect_tx(&full[s], len);
        tma::bulk_load(smem_raw + (size_t)s * TILE_BYTES, a.src + (int64_t)a.dst_shard[d] * a.shard_bytes + off, len,
                       &full[s]);
    };
    // CTA-local tile k is launch tile blockIdx.x + k * gridDim.x
    const int64_t mine = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    for (int k = 0; k < STAGES && k < mine; ++k) load(blockIdx.x + (int64_t)k * gridDim.x, k);
    for (int64_t k = 0; k < mine; ++k) {
        const int s = (int)(k % STAGES);
        tma::mbar_wait(&full[s], (uint32_t)((k / STAGES) & 1));
        int d; int64_t off; uint32_t len;
        where(blockIdx.x + k * gridDim.x, d, off, len);
        tma::bulk_store(a.dst[d] + off, smem_raw + (size_t)s * TILE_BYTES, len);
        tma::bulk_commit();
        // at most STORES-1 stores may still be reading shared memory: the stage of tile k-(STORES-1) is free
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STORES - 1) : "memory");
        const int64_t j = k - (STORES - 1);
        if (j >= 0 && j + STAGES < mine) load(blockIdx.x + (j + STAGES) * gridDim.x, (int)(j % STAGES));
    }
    tma::bulk_wait_all();   // every byte has left for its destination before the kernel ends
}

// ---- NVLS: the cross-worker sum is taken INSIDE the NVSwitch ---------------------------------
// multimem.ld_reduce on a multicast address returns the sum over every GPU's copy of that address
// (SASS: LDGMC.E.ADD.F32x4); multimem.st writes every GPU's copy at once.  Per direction and GPU
// the wire then carries ~4.5 B/element at W=8 instead of 7 for reduce-scatter + all-gather.
__device__ __forceinline__ float4 mc_ld_reduce_add(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(float* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// This rank's shard: g = (in-switch sum of every worker's pseudo-gradient) / W, Nesterov update of
// the master and momentum shards (local HBM), new parameters multicast into every worker's copy.
template <int U, int MINB>
__global__ void __launch_bounds__(kThreads, MINB) nvls_shard_kernel(float* __restrict__ master, float* __restrict__ mom,
                                                                const float* mc_pg, float* mc_local, int64_t n,
                                                                float world, OuterHyper h) {
    const int64_t nvec = n >> 2;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto one = [&](float& p, float& b, float s) { outer_update(__fdiv_rn(s, world), p, b, h); };
    for (; i < nvec; i += (int64_t)U * nthreads) {
        float4 s[U], p[U], b[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = i + u * nthreads;
            if (j < nvec) {
                s[u] = mc_ld_reduce_add(mc_pg + (j << 2));
                p[u] = ld_stream(master + (j << 2));
                b[u] = h.first_step ? make_float4(0.f, 0.f, 0.f, 0.f) : ld_stream(mom + (j << 2));
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = i + u * nthreads;
            if (j < nvec) {
                one(p[u].x, b[u].x, s[u].x); one(p[u].y, b[u].y, s[u].y);
                one(p[u].z, b[u].z, s[u].z); one(p[u].w, b[u].w, s[u].w);
                st_stream(master + (j << 2), p[u]);
                st_stream(mom + (j << 2), b[u]);
                mc_st(mc_local + (j << 2), p[u]);
            }
        }
    }
}

}  // namespace nd

using namespace nd;

struct nd_peer_ctx {
    int world = 1, rank = 0;
    size_t bytes = 0;                 // payload bytes of each rank's buffer (flags follow)
    std::vector<void*> base;          // mapped base of rank w's buffer (own allocation for w == rank)
    uint32_t** flag_pages_dev = nullptr;
    static constexpr int kChannels = 6;   // 0 outer sync, 1-3 in-switch pipeline lanes, 4 replicas of one worker
    uint32_t epoch[kChannels] = {0, 0, 0, 0, 0, 0};
    // nd_outer_sync_p2p with the rendezvous folded into the fused kernel: a block counter, the time block 0 waited
    // for the slowest worker, and whether the last sync ran that way
    unsigned* done_dev = nullptr;
    unsigned long long* wait_ns_dev = nullptr;
    int fold = 1;                         // ND_P2P_FOLD=0 / nd_peer_set_fold(ctx, 0): barrier kernels around the fused kernel
    bool last_folded = false;
    void* scratch = nullptr;              // norm scratch of nd_inner_step_replicas (nd_scratch_bytes), created on first use
    int replicas = 1;                 // consecutive ranks that hold identical copies of one worker
    size_t staging_off = 0, staging_bytes = 0;   // push mode: where in every rank's buffer the staging slots start
    int mode = 0;                     // 0 = pull (remote reads + remote writes), 1 = push (remote writes only)
    cudaEvent_t ta = nullptr, tb = nullptr;      // around the flag barrier that absorbs the skew between workers
    bool owns_mappings = true;        // false: the peer addresses were mapped by the caller (symmetric memory)
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    bool timed = false;
    bool split = false;               // the last sync recorded ta/tb (nd_outer_sync_p2p does)
    // side streams for the host-resident-master variant (created on first use)
    static constexpr int kStreams = 4;
    cudaStream_t streams[kStreams] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t done[kStreams] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t fork = nullptr;
    std::vector<cudaEvent_t> bucket_events;  // NVLS pipeline: two per bucket
    unsigned long long timeout_ns = 0;       // bound on a flag wait (0 = none); ND_PEER_TIMEOUT_S
    int* diag_host = nullptr;                // page-locked, device-mapped: {timed_out, waited_for_rank, channel, epoch}
    int* diag_dev = nullptr;
};

static unsigned long long peer_timeout_ns() {
    const char* e = getenv("ND_PEER_TIMEOUT_S");
    const double s = e ? atof(e) : 1800.0;
    return s <= 0.0 ? 0ull : (unsigned long long)(s * 1e9);
}

// what the two constructors share once `base` is filled in; on failure the caller destroys the context
static int peer_ctx_finish(nd_peer_ctx* c) {
    std::vector<uint32_t*> pages(c->world);
    for (int w = 0; w < c->world; ++w)
        pages[w] = reinterpret_cast<uint32_t*>(static_cast<char*>(c->base[w]) + c->bytes);
    ND_CUDA(cudaMalloc(&c->flag_pages_dev, sizeof(uint32_t*) * c->world));
    ND_CUDA(cudaMemcpy(c->flag_pages_dev, pages.data(), sizeof(uint32_t*) * c->world, cudaMemcpyHostToDevice));
    ND_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&c->diag_host), 4 * sizeof(int), cudaHostAllocMapped));
    memset(c->diag_host, 0, 4 * sizeof(int));
    ND_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->diag_dev), c->diag_host, 0));
    c->timeout_ns = peer_timeout_ns();
    ND_CUDA(cudaMalloc(&c->done_dev, 16));
    ND_CUDA(cudaMemset(c->done_dev, 0, 16));
    c->wait_ns_dev = reinterpret_cast<unsigned long long*>(c->done_dev) + 1;
    if (const char* e = getenv("ND_P2P_FOLD")) c->fold = atoi(e) != 0;
    ND_CUDA(cudaEventCreate(&c->t0));
    ND_CUDA(cudaEventCreate(&c->t1));
    ND_CUDA(cudaEventCreate(&c->ta));
    ND_CUDA(cudaEventCreate(&c->tb));
    return ND_OK;
}

// A flag wait that expired left a note in page-locked memory before it trapped: turn the opaque
// "unspecified launch failure" the next CUDA call reports into something a user can act on.
static int peer_diag(const nd_peer_ctx* c, const char* where) {
    if (c && c->diag_host && c->diag_host[0])
        return set_error(ND_ESTATE, "%s: rank %d waited longer than ND_PEER_TIMEOUT_S for rank %d (flag channel %d, "
                         "epoch %d); the CUDA context of this process is lost", where, c->rank, c->diag_host[1],
                         c->diag_host[2], c->diag_host[3]);
    return ND_OK;
}

extern "C" {

int nd_peer_ctx_destroy(nd_peer_ctx* c);

int nd_peer_alloc(size_t bytes, void** dev_ptr_out, char handle_out[ND_IPC_HANDLE_BYTES]) {
    if (!dev_ptr_out || !handle_out || bytes == 0) return set_error(ND_EINVAL, "nd_peer_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == ND_IPC_HANDLE_BYTES, "IPC handle size");
    const size_t payload = (bytes + 255) / 256 * 256;
    void* p = nullptr;
    ND_CUDA(cudaMalloc(&p, payload + kFlagBytes));
    ND_CUDA(cudaMemset(p, 0, payload + kFlagBytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return set_error(ND_ECUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    memcpy(handle_out, &h, ND_IPC_HANDLE_BYTES);
    *dev_ptr_out = p;
    return ND_OK;
}

int nd_peer_free(void* dev_ptr) {
    if (dev_ptr) ND_CUDA(cudaFree(dev_ptr));
    return ND_OK;
}

int nd_peer_ctx_create(int world, int rank, void* my_buffer, size_t bytes, const char* handles_host,
                       nd_peer_ctx** out) {
    if (!out) return set_error(ND_EINVAL, "nd_peer_ctx_create: null out");
    *out = nullptr;
    if (world < 1 || world > ND_MAX_WORKERS || rank < 0 || rank >= world)
        return set_error(ND_EINVAL, "nd_peer_ctx_create: bad world/rank (world <= ND_MAX_WORKERS)");
    if (!my_buffer || bytes == 0 || (world > 1 && !handles_host))
        return set_error(ND_EINVAL, "nd_peer_ctx_create: null buffer/handles");
    nd_peer_ctx* c = new nd_peer_ctx();
    c->world = world; c->rank = rank; c->bytes = (bytes + 255) / 256 * 256;
    c->base.assign(world, nullptr);
    c->base[rank] = my_buffer;
    for (int w = 0; w < world; ++w) {
        if (w == rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles_host + (size_t)w * ND_IPC_HANDLE_BYTES, ND_IPC_HANDLE_BYTES);
        cudaError_t e = cudaIpcOpenMemHandle(&c->base[w], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            c->base[w] = nullptr;
            nd_peer_ctx_destroy(c);      // closes the mappings opened so far
            return set_error(ND_ECUDA, "cudaIpcOpenMemHandle(rank %d) failed: %s (is peer access available?)", w,
                             cudaGetErrorString(e));
        }
    }
    if (int rc = peer_ctx_finish(c)) {   // nd_last_error() keeps the message of the call that failed
        nd_peer_ctx_destroy(c);
        return rc;
    }
    *out = c;
    return ND_OK;
}

int nd_peer_ctx_create_from_ptrs(int world, int rank, void* const* bases_host, size_t bytes, nd_peer_ctx** out) {
    if (!out) return set_error(ND_EINVAL, "nd_peer_ctx_create_from_ptrs: null out");
    *out = nullptr;
    if (world < 1 || world > ND_MAX_WORKERS || rank < 0 || rank >= world)
        return set_error(ND_EINVAL, "nd_peer_ctx_create_from_ptrs: bad world/rank (world <= ND_MAX_WORKERS)");
    if (!bases_host || bytes == 0) return set_error(ND_EINVAL, "nd_peer_ctx_create_from_ptrs: null bases");
    for (int w = 0; w < world; ++w)
        if (!bases_host[w]) return set_error(ND_EINVAL, "nd_peer_ctx_create_from_ptrs: null base for rank %d", w);
    nd_peer_ctx* c = new nd_peer_ctx();
    c->world = world; c->rank = rank; c->bytes = (bytes + 255) / 256 * 256;
    c->owns_mappings = false;
    c->base.assign(bases_host, bases_host + world);
    if (int rc = peer_ctx_finish(c)) {
        nd_peer_ctx_destroy(c);
        return rc;
    }
    *out = c;
    return ND_OK;
}

int nd_peer_ctx_destroy(nd_peer_ctx* c) {
    if (!c) return ND_OK;
    if (c->owns_mappings)
        for (int w = 0; w < c->world; ++w)
            if (w != c->rank && c->base[w]) cudaIpcCloseMemHandle(c->base[w]);
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) {
        if (c->streams[i]) cudaStreamDestroy(c->streams[i]);
        if (c->done[i]) cudaEventDestroy(c->done[i]);
    }
    if (c->fork) cudaEventDestroy(c->fork);
    for (auto e : c->bucket_events) cudaEventDestroy(e);
    if (c->flag_pages_dev) cudaFree(c->flag_pages_dev);
    if (c->diag_host) cudaFreeHost(c->diag_host);
    if (c->scratch) cudaFree(c->scratch);
    if (c->done_dev) cudaFree(c->done_dev);
    if (c->t0) cudaEventDestroy(c->t0);
    if (c->t1) cudaEventDestroy(c->t1);
    if (c->ta) cudaEventDestroy(c->ta);
    if (c->tb) cudaEventDestroy(c->tb);
    delete c;
    return ND_OK;
}

int nd_peer_buffer(const nd_peer_ctx* c, int w, void** out) {
    if (!c || !out || w < 0 || w >= c->world) return set_error(ND_EINVAL, "nd_peer_buffer: bad argument");
    *out = c->base[w];
    return ND_OK;
}

int nd_peer_set_replicas(nd_peer_ctx* c, int replicas) {
    if (!c) return set_error(ND_ESTATE, "nd_peer_set_replicas: null context");
    if (replicas < 1 || c->world % replicas != 0)
        return set_error(ND_EINVAL, "nd_peer_set_replicas: replicas must divide the world size");
    c->replicas = replicas;
    return ND_OK;
}

// Pointers into every copy at element offset `off`: one READ per worker (its first replica), a WRITE
// for every rank.
static int fill_rw(const nd_peer_ctx* c, size_t byte_off, void** reads, void** writes) {
    const int workers = c->world / c->replicas;
    for (int k = 0; k < workers; ++k) reads[k] = static_cast<char*>(c->base[k * c->replicas]) + byte_off;
    for (int w = 0; w < c->world; ++w) writes[w] = static_cast<char*>(c->base[w]) + byte_off;
    return workers;
}

static int peer_barrier_group(nd_peer_ctx* c, int channel, int first, int count, cudaStream_t stream) {
    c->epoch[channel] += 1;
    peer_barrier_kernel<<<1, 32, 0, stream>>>(c->flag_pages_dev, first, count, c->rank, c->epoch[channel], channel,
                                              c->timeout_ns, c->diag_dev);
    return check_launch("peer_barrier_kernel");
}

static int peer_barrier_ch(nd_peer_ctx* c, int channel, cudaStream_t stream) {
    return peer_barrier_group(c, channel, 0, c->world, stream);
}

int nd_peer_barrier(nd_peer_ctx* c, nd_stream_t stream) {
    if (!c) return set_error(ND_ESTATE, "nd_peer_barrier: null context");
    return peer_barrier_ch(c, 0, reinterpret_cast<cudaStream_t>(stream));
}

int nd_peer_set_staging(nd_peer_ctx* c, size_t byte_offset, size_t bytes) {
    if (!c) return set_error(ND_ESTATE, "nd_peer_set_staging: null context");
    if (byte_offset % 256 != 0 || byte_offset + bytes > c->bytes)
        return set_error(ND_EINVAL, "nd_peer_set_staging: the staging area must start on a 256-byte boundary inside "
                                    "the peer buffer");
    c->staging_off = byte_offset;
    c->staging_bytes = bytes;
    return ND_OK;
}

int nd_peer_set_fold(nd_peer_ctx* c, int on) {
    if (!c) return set_error(ND_ESTATE, "nd_peer_set_fold: null context");
    c->fold = on != 0;
    return ND_OK;
}

int nd_peer_set_mode(nd_peer_ctx* c, int mode) {
    if (!c) return set_error(ND_ESTATE, "nd_peer_set_mode: null context");
    if (mode != ND_PEER_PULL && mode != ND_PEER_PUSH) return set_error(ND_EINVAL, "nd_peer_set_mode: unknown mode");
    if (mode == ND_PEER_PUSH && c->staging_bytes == 0)
        return set_error(ND_ESTATE, "nd_peer_set_mode: push needs a staging area (nd_peer_set_staging)");
    c->mode = mode;
    return ND_OK;
}

// ND_PUSH_VARIANT (tuning knob): stage shape of the push kernel
static int push_launch(const PushArgs& a, cudaStream_t stream) {
    static int v = -1;
    if (v < 0) { const char* e = getenv("ND_PUSH_VARIANT"); v = e ? atoi(e) : 0; }
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
#define ND_PUSH(TB, ST, SO, CTAS)                                                                            \
    do {                                                                                                     \
        auto k = push_kernel<TB, ST, SO>;                                                                    \
        static std::atomic<bool> configured[64];                                                             \
        const int d = (dev >= 0 && dev < 64) ? dev : 0;                                                      \
        if (!configured[d].load(std::memory_order_acquire)) {                                                \
            ND_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (TB) * (ST)));       \
            configured[d].store(true, std::memory_order_release);                                            \
        }                                                                                                    \
        k<<<sms * (CTAS), 32, (TB) * (ST), stream>>>(a);                                                     \
    } while (0)
    switch (v) {
        case 1: ND_PUSH(32768, 6, 3, 1); break;
        case 2: ND_PUSH(8192, 8, 4, 2); break;
        case 3: ND_PUSH(16384, 4, 2, 3); break;
        default: ND_PUSH(16384, 6, 3, 2); break;
    }
#undef ND_PUSH
    return check_launch("push_kernel");
}

int nd_outer_sync_p2p(nd_peer_ctx* c, float* master_shard, float* mom_shard, int local_dtype, int64_t n_padded,
                      float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                      float* sumsq_out, void* scratch, nd_stream_t stream_) {
    if (!c) return set_error(ND_ESTATE, "nd_outer_sync_p2p: null context");
    if (int rc = peer_diag(c, "nd_outer_sync_p2p")) return rc;
    const int W = c->world;
    if (local_dtype != ND_F32 && local_dtype != ND_BF16) return set_error(ND_EINVAL, "nd_outer_sync_p2p: unknown dtype");
    const int gran = local_dtype == ND_BF16 ? 8 : 4;
    if (n_padded < 0 || n_padded % ((int64_t)gran * W) != 0)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p: n_padded must be a multiple of %d*world", gran);
    const size_t esz = local_dtype == ND_BF16 ? 2 : 4;
    const bool push = c->mode == ND_PEER_PUSH && W > 1;
    const size_t local_limit = c->staging_bytes ? c->staging_off : c->bytes;
    if ((size_t)n_padded * esz > local_limit) return set_error(ND_EINVAL, "nd_outer_sync_p2p: n_padded exceeds the peer buffers");
    if (n_padded && (!master_shard || !mom_shard)) return set_error(ND_EINVAL, "nd_outer_sync_p2p: null pointer");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const int64_t shard = n_padded / W;
    const size_t shard_bytes = (size_t)shard * esz;
    const int workers = W / c->replicas;
    if (push && (size_t)workers * shard_bytes > c->staging_bytes)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p: the staging area holds fewer than workers*shard bytes");
    NvtxRange range_all(push ? "nd_outer_sync_p2p (push)" : "nd_outer_sync_p2p (pull)");
    ND_CUDA(cudaEventRecord(c->t0, stream));
    int rc;
    void *reads[ND_MAX_WORKERS], *writes[ND_MAX_WORKERS];
    fill_rw(c, (size_t)c->rank * shard_bytes, reads, writes);
    if (push) {
        NvtxRange range("nd.p2p: push slices to the shard owners");
        // phase A: the first replica of every worker writes its copy of shard r into slot [worker] of rank r
        if (c->rank % c->replicas == 0 && shard_bytes > 0) {
            PushArgs a;
            a.src = static_cast<const char*>(c->base[c->rank]);
            a.shard_bytes = (int64_t)shard_bytes;
            a.n_dst = 0;
            const int me = c->rank / c->replicas;
            for (int r = 0; r < W; ++r) {
                if (r == c->rank) continue;      // our own shard is read in place
                a.dst[a.n_dst] = static_cast<char*>(c->base[r]) + c->staging_off + (size_t)me * shard_bytes;
                a.dst_shard[a.n_dst] = r;
                ++a.n_dst;
            }
            if ((rc = push_launch(a, stream))) return rc;
        }
        // every worker's copy of OUR shard now sits in our own staging slots (after the barrier below)
        for (int k = 0; k < workers; ++k)
            if (k * c->replicas != c->rank)
                reads[k] = static_cast<char*>(c->base[c->rank]) + c->staging_off + (size_t)k * shard_bytes;
    }
    if (W > 1 && c->fold && !push) {
        // ONE launch: arrival rendezvous in the kernel's prologue, completion rendezvous in its epilogue
        tma::PeerSync ps;
        ps.flag_pages = c->flag_pages_dev; ps.first = 0; ps.count = W; ps.rank = c->rank;
        ps.ch_in = 0; ps.ch_out = 5;
        ps.epoch_in = c->epoch[0] + 1; ps.epoch_out = c->epoch[5] + 1;
        ps.timeout_ns = c->timeout_ns; ps.diag = c->diag_dev; ps.done = c->done_dev; ps.wait_ns = c->wait_ns_dev;
        bool launched = false;
        NvtxRange range("nd.p2p: rendezvous + fused average + Nesterov update + copy-back + rendezvous");
        rc = outer_fused_rw_peer(master_shard, mom_shard, reads, workers, writes, W, local_dtype, shard, lr, mu,
                                 dampening, wd, nesterov, first_step, sumsq_out, scratch, stream, ps, &launched);
        if (rc) return rc;
        if (launched) {
            c->epoch[0] += 1; c->epoch[5] += 1;
            ND_CUDA(cudaEventRecord(c->t1, stream));
            c->timed = true;
            c->split = false;
            c->last_folded = true;
            return ND_OK;
        }
    }
    c->last_folded = false;
    ND_CUDA(cudaEventRecord(c->ta, stream));
    // pull: every worker's inner steps are done.  push: every worker's slices have landed here (which
    // also means its inner steps are done and it has finished reading its own copy)
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;
    ND_CUDA(cudaEventRecord(c->tb, stream));
    {
        NvtxRange range("nd.p2p: fused average + Nesterov update + copy-back");
        rc = nd_outer_fused_rw(master_shard, mom_shard, reads, workers, writes, W, local_dtype, shard, lr, mu, dampening,
                               wd, nesterov, first_step, sumsq_out, scratch, stream_);
    }
    if (rc) return rc;
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;  // every shard of our copy has been written
    ND_CUDA(cudaEventRecord(c->t1, stream));
    c->timed = true;
    c->split = true;
    return ND_OK;
}

static int ensure_side_streams(nd_peer_ctx* c) {
    if (c->fork) return ND_OK;  // first use: side streams and the events that join them to the caller's stream
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) {
        ND_CUDA(cudaStreamCreateWithFlags(&c->streams[i], cudaStreamNonBlocking));
        ND_CUDA(cudaEventCreateWithFlags(&c->done[i], cudaEventDisableTiming));
    }
    ND_CUDA(cudaEventCreateWithFlags(&c->fork, cudaEventDisableTiming));
    return ND_OK;
}

int nd_outer_sync_p2p_offloaded(nd_peer_ctx* c, float* master_host_shard, float* staging, int64_t chunk_elems,
                                float* mom_shard, int local_dtype, int64_t n_padded, float lr, float mu,
                                float dampening, float wd, int nesterov, int first_step, nd_stream_t stream_) {
    if (!c) return set_error(ND_ESTATE, "nd_outer_sync_p2p_offloaded: null context");
    const int W = c->world;
    if (local_dtype != ND_F32 && local_dtype != ND_BF16)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p_offloaded: unknown dtype");
    const int gran = local_dtype == ND_BF16 ? 8 : 4;
    if (n_padded < 0 || n_padded % ((int64_t)gran * W) != 0)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p_offloaded: n_padded must be a multiple of %d*world", gran);
    if (chunk_elems < gran || chunk_elems % gran != 0)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p_offloaded: chunk_elems must be a positive multiple of %d", gran);
    const size_t esz = local_dtype == ND_BF16 ? 2 : 4;
    if ((size_t)n_padded * esz > c->bytes)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p_offloaded: n_padded exceeds the peer buffers");
    if (n_padded && (!master_host_shard || !staging || !mom_shard))
        return set_error(ND_EINVAL, "nd_outer_sync_p2p_offloaded: null pointer");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    int rc;
    if ((rc = ensure_side_streams(c))) return rc;
    const int64_t shard = n_padded / W;
    ND_CUDA(cudaEventRecord(c->t0, stream));
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;
    ND_CUDA(cudaEventRecord(c->fork, stream));
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) ND_CUDA(cudaStreamWaitEvent(c->streams[i], c->fork, 0));
    int64_t k = 0;
    for (int64_t off = 0; off < shard; off += chunk_elems, ++k) {
        const int64_t len = (shard - off < chunk_elems) ? (shard - off) : chunk_elems;
        const int slot = (int)(k % 3);  // three staging slots (see the header), one stream each
        cudaStream_t s = c->streams[slot];  // stream order also keeps a slot from being reused too early
        float* m_dev = staging + (int64_t)slot * chunk_elems;
        ND_CUDA(cudaMemcpyAsync(m_dev, master_host_shard + off, (size_t)len * 4, cudaMemcpyHostToDevice, s));
        void *reads[ND_MAX_WORKERS], *writes[ND_MAX_WORKERS];
        const int workers = fill_rw(c, ((size_t)c->rank * shard + (size_t)off) * esz, reads, writes);
        rc = nd_outer_fused_rw(m_dev, mom_shard + off, reads, workers, writes, W, local_dtype, len, lr, mu, dampening,
                               wd, nesterov, first_step, nullptr, nullptr, reinterpret_cast<nd_stream_t>(s));
        if (rc) return rc;
        ND_CUDA(cudaMemcpyAsync(master_host_shard + off, m_dev, (size_t)len * 4, cudaMemcpyDeviceToHost, s));
    }
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) {
        ND_CUDA(cudaEventRecord(c->done[i], c->streams[i]));
        ND_CUDA(cudaStreamWaitEvent(stream, c->done[i], 0));
    }
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;
    ND_CUDA(cudaEventRecord(c->t1, stream));
    c->timed = true;
    c->split = false;
    return ND_OK;
}

// In-flight depth of the in-switch loads: U independent 16-byte multimem.ld_reduce per thread, MINB
// resident CTAs of 512 threads per SM.  ND_NVLS_VARIANT (0..3) selects among the compiled shapes.
static int nvls_variant() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("ND_NVLS_VARIANT");
        v = e ? atoi(e) : 0;
        if (v < 0 || v > 3) v = 0;
    }
    return v;
}

static int nvls_launch(nd_peer_ctx* c, float* master, float* mom, const float* mc_pg, float* mc_local, int64_t n,
                       const OuterHyper& h, cudaStream_t stream) {
    if (n <= 0) return ND_OK;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t need = ((n >> 2) + kThreads - 1) / kThreads;
    auto grid_for = [&](int minb) {
        return (int)(need < (int64_t)sms * minb ? (need > 0 ? need : 1) : (int64_t)sms * minb);
    };
    const float w = (float)c->world;
    switch (nvls_variant()) {
        case 1: nvls_shard_kernel<4, 2><<<grid_for(2), kThreads, 0, stream>>>(master, mom, mc_pg, mc_local, n, w, h); break;
        case 2: nvls_shard_kernel<2, 4><<<grid_for(4), kThreads, 0, stream>>>(master, mom, mc_pg, mc_local, n, w, h); break;
        case 3: nvls_shard_kernel<8, 1><<<grid_for(1), kThreads, 0, stream>>>(master, mom, mc_pg, mc_local, n, w, h); break;
        default: nvls_shard_kernel<2, 2><<<grid_for(2), kThreads, 0, stream>>>(master, mom, mc_pg, mc_local, n, w, h); break;
    }
    return check_launch("nvls_shard_kernel");
}

int nd_outer_sync_hybrid(nd_peer_ctx* c, float* local, float* pg, float* mc_local, float* mc_pg, float* master,
                         float* mom_shard, int64_t n_padded, int64_t n_nvls, int64_t bucket_elems, float lr, float mu,
                         float dampening, float wd, int nesterov, int first_step, nd_stream_t stream_) {
    if (!c) return set_error(ND_ESTATE, "nd_outer_sync_hybrid: null context");
    const int W = c->world;
    if (c->replicas != 1) return set_error(ND_ESTATE, "nd_outer_sync_hybrid: replicas are not supported");
    const int64_t q = 4 * (int64_t)W;
    if (n_padded < 0 || n_padded % q != 0)
        return set_error(ND_EINVAL, "nd_outer_sync_hybrid: n_padded must be a multiple of 4*world");
    if ((size_t)n_padded * 4 > c->bytes) return set_error(ND_EINVAL, "nd_outer_sync_hybrid: n_padded exceeds the buffers");
    if (n_padded && (!local || !pg || !mc_local || !mc_pg || !master || !mom_shard))
        return set_error(ND_EINVAL, "nd_outer_sync_hybrid: null pointer");
    if (bucket_elems < 1) return set_error(ND_EINVAL, "nd_outer_sync_hybrid: bucket_elems < 1");
    if (n_nvls < 0 || n_nvls > n_padded || n_nvls % q != 0)
        return set_error(ND_EINVAL, "nd_outer_sync_hybrid: n_nvls must be a multiple of 4*world in [0, n_padded]");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const int64_t bucket = (bucket_elems + q - 1) / q * q;
    const int64_t nb = (n_nvls + bucket - 1) / bucket;
    const OuterHyper h{-lr, mu, 1.0f - dampening, wd, nesterov, first_step};
    int rc;
    if ((rc = ensure_side_streams(c))) return rc;
    // Three lanes, bucket by bucket (bucket b = [b*B, min((b+1)*B, n)); rank r owns its r-th slice):
    //   lane A  K1: pg_b = master_b - local_b                                   (HBM)
    //   lane B  barrier(all ranks' pg_b are in memory) -> in-switch sum + update + multicast store   (NVLink)
    //   lane C  barrier(every copy of bucket b is written) -> master replica refresh of bucket b       (HBM)
    // so the HBM-only passes of buckets b+1 and b-1 hide under the NVLink pass of bucket b.
    cudaStream_t sa = c->streams[0], sb = c->streams[1], sc = c->streams[2], sd = c->streams[3];
    ND_CUDA(cudaEventRecord(c->t0, stream));
    ND_CUDA(cudaEventRecord(c->fork, stream));
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) ND_CUDA(cudaStreamWaitEvent(c->streams[i], c->fork, 0));
    if ((int64_t)c->bucket_events.size() < 2 * nb) {
        const size_t old = c->bucket_events.size();
        c->bucket_events.resize(2 * nb);
        for (size_t i = old; i < c->bucket_events.size(); ++i)
            ND_CUDA(cudaEventCreateWithFlags(&c->bucket_events[i], cudaEventDisableTiming));
    }
    // Region B = [n_nvls, n_padded): the peer-memory fused kernel on its own lane, concurrently.  The
    // in-switch path tops out below what the links carry (the switch's reduction rate), so handing
    // part of the space to plain peer loads/stores uses the bandwidth it leaves idle.
    if (n_padded > n_nvls) {
        const int64_t sB = (n_padded - n_nvls) / W, mineB = n_nvls + (int64_t)c->rank * sB;
        if ((rc = peer_barrier_ch(c, 3, sd))) return rc;  // every worker's inner steps are done
        void *reads[ND_MAX_WORKERS], *writes[ND_MAX_WORKERS];
        const int workers = fill_rw(c, (size_t)mineB * 4, reads, writes);
        if ((rc = nd_outer_fused_rw(master + mineB, mom_shard + n_nvls / W, reads, workers, writes, W, ND_F32, sB, lr,
                                    mu, dampening, wd, nesterov, first_step, nullptr, nullptr,
                                    reinterpret_cast<nd_stream_t>(sd))))
            return rc;
        if ((rc = peer_barrier_ch(c, 3, sd))) return rc;  // every shard of region B has been written
        if (mineB > n_nvls)
            ND_CUDA(cudaMemcpyAsync(master + n_nvls, local + n_nvls, (size_t)(mineB - n_nvls) * 4,
                                    cudaMemcpyDeviceToDevice, sd));
        if (mineB + sB < n_padded)
            ND_CUDA(cudaMemcpyAsync(master + mineB + sB, local + mineB + sB, (size_t)(n_padded - mineB - sB) * 4,
                                    cudaMemcpyDeviceToDevice, sd));
    }
    for (int64_t b = 0; b < nb; ++b) {
        const int64_t off = b * bucket;
        const int64_t len = (n_nvls - off < bucket) ? (n_nvls - off) : bucket;
        const int64_t slice = len / W, mine = off + (int64_t)c->rank * slice;
        const int64_t shard_off = b * (bucket / W);
        cudaEvent_t k1_done = c->bucket_events[2 * b], upd_done = c->bucket_events[2 * b + 1];
        if ((rc = nd_pseudo_grad(master + off, local + off, ND_F32, pg + off, len, reinterpret_cast<nd_stream_t>(sa))))
            return rc;
        ND_CUDA(cudaEventRecord(k1_done, sa));
        ND_CUDA(cudaStreamWaitEvent(sb, k1_done, 0));
        if ((rc = peer_barrier_ch(c, 1, sb))) return rc;
        if ((rc = nvls_launch(c, master + mine, mom_shard + shard_off, mc_pg + mine, mc_local + mine, slice, h, sb)))
            return rc;
        ND_CUDA(cudaEventRecord(upd_done, sb));
        ND_CUDA(cudaStreamWaitEvent(sc, upd_done, 0));
        if ((rc = peer_barrier_ch(c, 2, sc))) return rc;
        // theta_sync replica: the slices other ranks own are refreshed from the working copy (diloco.py:54)
        if (mine > off)
            ND_CUDA(cudaMemcpyAsync(master + off, local + off, (size_t)(mine - off) * 4, cudaMemcpyDeviceToDevice, sc));
        if (mine + slice < off + len)
            ND_CUDA(cudaMemcpyAsync(master + mine + slice, local + mine + slice,
                                    (size_t)(off + len - mine - slice) * 4, cudaMemcpyDeviceToDevice, sc));
    }
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) {
        ND_CUDA(cudaEventRecord(c->done[i], c->streams[i]));
        ND_CUDA(cudaStreamWaitEvent(stream, c->done[i], 0));
    }
    ND_CUDA(cudaEventRecord(c->t1, stream));
    c->timed = true;
    c->split = false;
    return ND_OK;
}

int nd_outer_sync_nvls(nd_peer_ctx* c, float* local, float* pg, float* mc_local, float* mc_pg, float* master,
                       float* mom_shard, int64_t n_padded, int64_t bucket_elems, float lr, float mu, float dampening,
                       float wd, int nesterov, int first_step, nd_stream_t stream_) {
    return nd_outer_sync_hybrid(c, local, pg, mc_local, mc_pg, master, mom_shard, n_padded, n_padded, bucket_elems, lr,
                                mu, dampening, wd, nesterov, first_step, stream_);
}

int nd_peer_last_ms(nd_peer_ctx* c, float* ms_out) {
    if (!c || !ms_out) return set_error(ND_EINVAL, "nd_peer_last_ms: null argument");
    if (!c->timed) return set_error(ND_ESTATE, "nd_peer_last_ms: no nd_outer_sync_p2p has run");
    cudaError_t es = cudaEventSynchronize(c->t1);
    if (es != cudaSuccess) {
        if (int rc = peer_diag(c, "nd_peer_last_ms")) return rc;
        return set_error(ND_ECUDA, "nd_peer_last_ms: cudaEventSynchronize failed: %s", cudaGetErrorString(es));
    }
    ND_CUDA(cudaEventElapsedTime(ms_out, c->t0, c->t1));
    return ND_OK;
}

// ---- inner step of a worker that spans several GPUs (BASELINE config 5; not in the reference) --------
// The R replicas of a worker hold identical parameters and each has its own gradient.  Replica j owns the
// j-th 1/R of the flat space for this step (ZeRO-1 style: it alone keeps exp_avg / exp_avg_sq there):
//   flag barrier among the R replicas (every backward pass is done);
//   K_a  reads all R gradients of its part (R-1 over NVLink), averages them in replica order, writes the
//        average over its own gradient and leaves the sum of squares of its part in its flag page;
//   flag barrier among the R replicas (gradients read, partial norms published);
//   K_b  clip coefficient from the R partial norms (same order on every replica), AdamW on its part, the
//        new parameters bulk-stored into every replica's copy, its gradient part zeroed; the parts of its
//        gradient buffer that the other replicas consumed are cleared with a memset;
//   flag barrier (every replica's parameters are complete).
// Against all_reduce(AVG) + sum of squares + AdamW on every replica this moves 22 instead of ~44 bytes per
// element through each GPU's HBM and 4(R-1)/R bytes per element per direction over NVLink.
constexpr size_t kPartialOffset = 1024;   // byte offset of the partial-norm slot inside a flag page

struct PartPtrs { const float* p[ND_MAX_REPLICAS]; };
__global__ void sum_parts_kernel_v(PartPtrs parts, int n, float* out) {
    float s = 0.0f;
    for (int k = 0; k < n; ++k) {
        float v;
        asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(parts.p[k]) : "memory");
        s = k == 0 ? v : __fadd_rn(s, v);
    }
    *out = s;
}

int nd_inner_step_replicas(nd_peer_ctx* pc, nd_peer_ctx* gc, float* m, float* v, int dtype, int64_t n_padded,
                           double lr, double beta1, double beta2, double eps, double weight_decay, int64_t step,
                           float max_norm, float* sumsq_total_out, nd_stream_t stream_) {
    if (!pc || !gc) return set_error(ND_ESTATE, "nd_inner_step_replicas: null context");
    if (int rc = peer_diag(gc, "nd_inner_step_replicas")) return rc;
    if (pc->world != gc->world || pc->rank != gc->rank || pc->replicas != gc->replicas)
        return set_error(ND_EINVAL, "nd_inner_step_replicas: the parameter and gradient contexts differ in world/rank/replicas");
    const int R = gc->replicas;
    if (R > ND_MAX_REPLICAS) return set_error(ND_EINVAL, "nd_inner_step_replicas: more than ND_MAX_REPLICAS replicas");
    if (dtype != ND_F32 && dtype != ND_BF16) return set_error(ND_EINVAL, "nd_inner_step_replicas: unknown dtype");
    const int gran = dtype == ND_BF16 ? 8 : 4;
    const size_t esz = dtype == ND_BF16 ? 2 : 4;
    if (n_padded < 0 || n_padded % ((int64_t)gran * R) != 0)
        return set_error(ND_EINVAL, "nd_inner_step_replicas: n_padded must be a multiple of %d*replicas", gran);
    const size_t plimit = pc->staging_bytes ? pc->staging_off : pc->bytes;
    if ((size_t)n_padded * esz > plimit || (size_t)n_padded * esz > gc->bytes)
        return set_error(ND_EINVAL, "nd_inner_step_replicas: n_padded exceeds the peer buffers");
    if (n_padded && (!m || !v)) return set_error(ND_EINVAL, "nd_inner_step_replicas: null pointer");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!gc->scratch) {
        ND_CUDA(cudaMalloc(&gc->scratch, nd_scratch_bytes()));
        ND_CUDA(cudaMemset(gc->scratch, 0, nd_scratch_bytes()));
    }
    const int j = gc->rank % R, first = gc->rank - j;
    const int64_t part = n_padded / R;
    const size_t off = (size_t)j * part * esz;
    void *grads[ND_MAX_REPLICAS], *pw[ND_MAX_REPLICAS];
    const float* parts[ND_MAX_REPLICAS];
    for (int k = 0; k < R; ++k) {
        grads[k] = static_cast<char*>(gc->base[first + k]) + off;
        pw[k] = static_cast<char*>(pc->base[first + k]) + off;
        parts[k] = reinterpret_cast<const float*>(static_cast<char*>(gc->base[first + k]) + gc->bytes + kPartialOffset);
    }
    char* mine_g = static_cast<char*>(gc->base[gc->rank]);
    float* my_part = reinterpret_cast<float*>(mine_g + gc->bytes + kPartialOffset);
    int rc;
    NvtxRange range_all("nd_inner_step_replicas");
    // every replica's backward pass has finished writing its gradient buffer
    if (R > 1 && (rc = peer_barrier_group(gc, 4, first, R, stream))) return rc;
    if ((rc = nd_grad_avg(grads, R, mine_g + off, dtype, part, my_part, gc->scratch, stream_))) return rc;
    if (R > 1 && (rc = peer_barrier_group(gc, 4, first, R, stream))) return rc;
    if (sumsq_total_out) {
        PartPtrs pp;
        for (int k = 0; k < ND_MAX_REPLICAS; ++k) pp.p[k] = k < R ? parts[k] : nullptr;
        sum_parts_kernel_v<<<1, 1, 0, stream>>>(pp, R, sumsq_total_out);
        if ((rc = check_launch("sum_parts_kernel"))) return rc;
    }
    // the rest of our gradient buffer has been consumed by the replicas that own those parts
    if (off) ND_CUDA(cudaMemsetAsync(mine_g, 0, off, stream));
    if ((size_t)(j + 1) * part < (size_t)n_padded)
        ND_CUDA(cudaMemsetAsync(mine_g + off + (size_t)part * esz, 0, ((size_t)n_padded - (size_t)(j + 1) * part) * esz, stream));
    rc = nd_adamw_rw(static_cast<char*>(pc->base[pc->rank]) + off, pw, R, mine_g + off, m + (int64_t)j * part,
                     v + (int64_t)j * part, dtype, part, lr, beta1, beta2, eps, weight_decay, step, parts, R, max_norm, 1,
                     stream_);
    if (rc) return rc;
    if (R > 1 && (rc = peer_barrier_group(gc, 4, first, R, stream))) return rc;
    return ND_OK;
}

int nd_peer_last_ms2(nd_peer_ctx* c, float* total_ms_out, float* wait_ms_out) {
    if (!wait_ms_out) return set_error(ND_EINVAL, "nd_peer_last_ms2: null argument");
    if (int rc = nd_peer_last_ms(c, total_ms_out)) return rc;
    *wait_ms_out = 0.0f;
    // only nd_outer_sync_p2p brackets its first barrier; the other variants report wait = 0
    if (c->split) ND_CUDA(cudaEventElapsedTime(wait_ms_out, c->ta, c->tb));
    if (c->last_folded) {   // the kernel's block 0 timed its own arrival wait
        unsigned long long ns = 0;
        ND_CUDA(cudaMemcpy(&ns, c->wait_ns_dev, sizeof(ns), cudaMemcpyDeviceToHost));
        *wait_ms_out = (float)((double)ns * 1e-6);
    }
    return ND_OK;
}

}  // extern "C"
