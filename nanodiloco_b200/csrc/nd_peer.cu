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
#include <cstring>
#include <vector>

#include "nd_common.cuh"
#include "nd_error.h"

namespace nd {

constexpr size_t kFlagBytes = 4096;  // per-rank flag page appended to the shared buffer

// Barrier across the ranks of a node: thread w publishes `epoch` to slot [rank] of rank w's flag
// page, then waits until slot [w] of our own page reaches `epoch`.  System-scope release/acquire;
// bounded wait (~10 s of SM clock) so that a dead peer traps this rank instead of hanging the box.
__global__ void peer_barrier_kernel(uint32_t* const* flag_pages, int world, int rank, uint32_t epoch) {
    const int w = threadIdx.x;
    if (w >= world) return;
    __threadfence_system();
    uint32_t* remote = flag_pages[w] + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
    const uint32_t* mine = flag_pages[rank] + w;
    const long long t0 = clock64();
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if ((int32_t)(v - epoch) >= 0) break;
        if (clock64() - t0 > 20000000000LL) __trap();
    }
    __threadfence_system();
}

}  // namespace nd

using namespace nd;

struct nd_peer_ctx {
    int world = 1, rank = 0;
    size_t bytes = 0;                 // payload bytes of each rank's buffer (flags follow)
    std::vector<void*> base;          // mapped base of rank w's buffer (own allocation for w == rank)
    uint32_t** flag_pages_dev = nullptr;
    uint32_t epoch = 0;
    int replicas = 1;                 // consecutive ranks that hold identical copies of one worker
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    bool timed = false;
    // side streams for the host-resident-master variant (created on first use)
    static constexpr int kStreams = 3;
    cudaStream_t streams[kStreams] = {nullptr, nullptr, nullptr};
    cudaEvent_t done[kStreams] = {nullptr, nullptr, nullptr};
    cudaEvent_t fork = nullptr;
};

extern "C" {

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
            for (int v = 0; v < w; ++v)
                if (v != rank && c->base[v]) cudaIpcCloseMemHandle(c->base[v]);
            delete c;
            return set_error(ND_ECUDA, "cudaIpcOpenMemHandle(rank %d) failed: %s (is peer access available?)", w,
                             cudaGetErrorString(e));
        }
    }
    std::vector<uint32_t*> pages(world);
    for (int w = 0; w < world; ++w) pages[w] = reinterpret_cast<uint32_t*>(static_cast<char*>(c->base[w]) + c->bytes);
    ND_CUDA(cudaMalloc(&c->flag_pages_dev, sizeof(uint32_t*) * world));
    ND_CUDA(cudaMemcpy(c->flag_pages_dev, pages.data(), sizeof(uint32_t*) * world, cudaMemcpyHostToDevice));
    ND_CUDA(cudaEventCreate(&c->t0));
    ND_CUDA(cudaEventCreate(&c->t1));
    *out = c;
    return ND_OK;
}

int nd_peer_ctx_destroy(nd_peer_ctx* c) {
    if (!c) return ND_OK;
    for (int w = 0; w < c->world; ++w)
        if (w != c->rank && c->base[w]) cudaIpcCloseMemHandle(c->base[w]);
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) {
        if (c->streams[i]) cudaStreamDestroy(c->streams[i]);
        if (c->done[i]) cudaEventDestroy(c->done[i]);
    }
    if (c->fork) cudaEventDestroy(c->fork);
    if (c->flag_pages_dev) cudaFree(c->flag_pages_dev);
    if (c->t0) cudaEventDestroy(c->t0);
    if (c->t1) cudaEventDestroy(c->t1);
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

int nd_peer_barrier(nd_peer_ctx* c, nd_stream_t stream) {
    if (!c) return set_error(ND_ESTATE, "nd_peer_barrier: null context");
    c->epoch += 1;
    peer_barrier_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(c->flag_pages_dev, c->world, c->rank,
                                                                              c->epoch);
    return check_launch("peer_barrier_kernel");
}

int nd_outer_sync_p2p(nd_peer_ctx* c, float* master_shard, float* mom_shard, int local_dtype, int64_t n_padded,
                      float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                      float* sumsq_out, void* scratch, nd_stream_t stream_) {
    if (!c) return set_error(ND_ESTATE, "nd_outer_sync_p2p: null context");
    const int W = c->world;
    if (local_dtype != ND_F32 && local_dtype != ND_BF16) return set_error(ND_EINVAL, "nd_outer_sync_p2p: unknown dtype");
    const int gran = local_dtype == ND_BF16 ? 8 : 4;
    if (n_padded < 0 || n_padded % ((int64_t)gran * W) != 0)
        return set_error(ND_EINVAL, "nd_outer_sync_p2p: n_padded must be a multiple of %d*world", gran);
    const size_t esz = local_dtype == ND_BF16 ? 2 : 4;
    if ((size_t)n_padded * esz > c->bytes) return set_error(ND_EINVAL, "nd_outer_sync_p2p: n_padded exceeds the peer buffers");
    if (n_padded && (!master_shard || !mom_shard)) return set_error(ND_EINVAL, "nd_outer_sync_p2p: null pointer");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const int64_t shard = n_padded / W;
    ND_CUDA(cudaEventRecord(c->t0, stream));
    int rc;
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;  // every worker's inner steps are done
    void *reads[ND_MAX_WORKERS], *writes[ND_MAX_WORKERS];
    const int workers = fill_rw(c, (size_t)c->rank * shard * esz, reads, writes);
    rc = nd_outer_fused_rw(master_shard, mom_shard, reads, workers, writes, W, local_dtype, shard, lr, mu, dampening,
                           wd, nesterov, first_step, sumsq_out, scratch, stream_);
    if (rc) return rc;
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;  // every shard of our copy has been written
    ND_CUDA(cudaEventRecord(c->t1, stream));
    c->timed = true;
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
    if (!c->fork) {  // first use: side streams and the events that join them to the caller's stream
        for (int i = 0; i < nd_peer_ctx::kStreams; ++i) {
            ND_CUDA(cudaStreamCreateWithFlags(&c->streams[i], cudaStreamNonBlocking));
            ND_CUDA(cudaEventCreateWithFlags(&c->done[i], cudaEventDisableTiming));
        }
        ND_CUDA(cudaEventCreateWithFlags(&c->fork, cudaEventDisableTiming));
    }
    const int64_t shard = n_padded / W;
    ND_CUDA(cudaEventRecord(c->t0, stream));
    int rc;
    if (W > 1 && (rc = nd_peer_barrier(c, stream_))) return rc;
    ND_CUDA(cudaEventRecord(c->fork, stream));
    for (int i = 0; i < nd_peer_ctx::kStreams; ++i) ND_CUDA(cudaStreamWaitEvent(c->streams[i], c->fork, 0));
    int64_t k = 0;
    for (int64_t off = 0; off < shard; off += chunk_elems, ++k) {
        const int64_t len = (shard - off < chunk_elems) ? (shard - off) : chunk_elems;
        const int slot = (int)(k % nd_peer_ctx::kStreams);
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
    return ND_OK;
}

int nd_peer_last_ms(nd_peer_ctx* c, float* ms_out) {
    if (!c || !ms_out) return set_error(ND_EINVAL, "nd_peer_last_ms: null argument");
    if (!c->timed) return set_error(ND_ESTATE, "nd_peer_last_ms: no nd_outer_sync_p2p has run");
    ND_CUDA(cudaEventSynchronize(c->t1));
    ND_CUDA(cudaEventElapsedTime(ms_out, c->t0, c->t1));
    return ND_OK;
}

}  // extern "C"
