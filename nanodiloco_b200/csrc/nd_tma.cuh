// nd_tma.cuh -- TMA-staged streaming driver (cp.async.bulk + mbarrier) for sm_100a.
//
// Why: these kernels have zero reuse, so shared memory is not a cache here -- it is a landing
// zone that lets ONE thread keep tens of KB of large, contiguous, 16-byte-aligned HBM bursts in
// flight per CTA (bulk async copies, tracked by an mbarrier's transaction count) while the other
// threads do the arithmetic, and that drains results with bulk stores instead of per-thread STG.
// Measured on B200 (tools/tune.cu): 6.29 TB/s on the fused outer kernel vs 5.5-6.1 TB/s for the
// best register-staged LDG.128 variants, against 6.57 TB/s for a device-to-device memcpy.
//
// Pipeline per CTA (tiles t = blockIdx.x, +gridDim.x, ...; STAGES shared-memory stages):
//   thread 0 : [wait until the bulk stores that last read stage s' are done] -> expect_tx + bulk
//              loads of tile t+STAGES-1 into s'
//   all      : wait full[s] (mbarrier parity) -> compute the tile in place in shared memory
//              -> fence.proxy.async -> __syncthreads
//   thread 0 : bulk stores of the results + commit_group
// A "TileOp" supplies: kBufs (buffers per stage), kGran (elements per 16 bytes of its narrowest
// array), loads(), compute(), stores(), tail().
#pragma once

#include "nd_common.cuh"

namespace nd {
namespace tma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
// Bounded wait (~2 s of SM clock): a mis-programmed pipeline traps -- reported as a launch
// failure at the next synchronisation -- instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    const long long t0 = clock64();
    for (;;) {
        uint32_t ok;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_u32(b)), "r"(parity)
            : "memory");
        if (ok) return;
        if (clock64() - t0 > 4000000000LL) __trap();
    }
}
// ND_TMA_L2_EVICT_FIRST (compile-time experiment, see tools/tune.cu): tag every bulk copy with an
// evict-first L2 policy.  Off by default: it measured no better than the default policy.
#ifdef ND_TMA_L2_EVICT_FIRST
__device__ __forceinline__ uint64_t evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
#endif
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
#ifdef ND_TMA_L2_EVICT_FIRST
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(evict_first_policy())
        : "memory");
#else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
#endif
}
__device__ __forceinline__ void bulk_store(void* dst_gmem, const void* src_smem, uint32_t bytes) {
#ifdef ND_TMA_L2_EVICT_FIRST
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "r"(bytes), "l"(evict_first_policy())
                 : "memory");
#else
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
#endif
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- flag barriers across GPUs folded into a tile kernel ------------------------------------------------
// The peer-memory outer sync needs two rendezvous of the ranks: before the kernel reads the peers' copies
// (every worker has finished its inner steps) and after it has written them (every shard of every copy is
// complete).  As kernels of their own they cost two launches and two drains of the GPU per sync, which is most
// of the step for small models.  Folded: block 0 publishes this rank's arrival to every peer's flag page and
// EVERY block polls the local page before its first load; after the last bulk store has completed, the block
// that finishes last publishes "done" and alone waits for the peers' "done" flags, so that the end of the
// kernel still means "my copy is complete".  Flags are monotonically increasing epochs in peer memory
// (system-scope release/acquire), one slot per (channel, sender).  A wait is bounded by wall time
// (`timeout_ns`, 0 = none): on expiry the rank records whom it waited for in `diag` (host-mapped) and traps.
struct PeerSync {
    uint32_t* const* flag_pages;   // device array: flag page of every rank (peer-mapped)
    int first, count, rank;        // the ranks [first, first + count) meet
    uint32_t epoch_in, epoch_out;  // value this sync publishes / awaits on the two channels
    int ch_in, ch_out;
    unsigned long long timeout_ns;
    int* diag;                     // {timed_out, waited_for_rank, channel, epoch}
    unsigned* done;                // device counter of finished blocks, zero at rest
    unsigned long long* wait_ns;   // out: nanoseconds block 0 spent waiting for the slowest rank to arrive
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// thread t < count of the calling block: tell rank first+t that this rank has reached `epoch` on `channel`
__device__ __forceinline__ void peer_publish(const PeerSync& ps, int channel, uint32_t epoch) {
    if ((int)threadIdx.x < ps.count) {
        __threadfence_system();
        uint32_t* remote = ps.flag_pages[ps.first + threadIdx.x] + channel * ND_MAX_WORKERS + ps.rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
    }
}

// thread t < count of the calling block: wait until rank first+t has reached `epoch` on `channel`
__device__ __forceinline__ void peer_await(const PeerSync& ps, int channel, uint32_t epoch) {
    if ((int)threadIdx.x < ps.count) {
        const int w = ps.first + threadIdx.x;
        const uint32_t* mine = ps.flag_pages[ps.rank] + channel * ND_MAX_WORKERS + w;
        const unsigned long long t0 = global_ns();
        for (unsigned spins = 0;; ++spins) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
            if ((int32_t)(v - epoch) >= 0) break;
            if (ps.timeout_ns && (spins & 1023u) == 1023u && global_ns() - t0 > ps.timeout_ns) {
                if (ps.diag) {
                    ps.diag[1] = w; ps.diag[2] = channel; ps.diag[3] = (int)epoch;
                    __threadfence_system();
                    ps.diag[0] = 1;
                }
                __threadfence_system();
                __trap();
            }
        }
        __threadfence_system();
    }
}

// One stage of shared memory: kBufs buffers of TILE fp32 slots each.
template <int TILE>
struct Stage {
    float* base;
    __device__ __forceinline__ float* buf(int k) const { return base + (size_t)k * TILE; }
};

template <int TILE, int STAGES, int THREADS, int MINB, class TileOp, bool PEER = false>
__global__ void __launch_bounds__(THREADS, MINB) tile_kernel(TileOp op, int64_t n, Scratch* sc, float* sum_out,
                                                             PeerSync ps) {
    static_assert(!PEER || THREADS >= ND_MAX_WORKERS, "one polling thread per rank");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    __shared__ uint64_t full[STAGES];
    constexpr int kStageFloats = TileOp::kBufs * TILE;
    // bulk copies move multiples of 16 bytes: kGran elements of the narrowest array (4 fp32, 8 bf16)
    const int64_t n4 = n & ~(int64_t)(TileOp::kGran - 1);
    const int64_t ntiles = (n4 + TILE - 1) / TILE;
    const int tid = threadIdx.x;
    op.prologue();
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if constexpr (PEER) {
        // arrival rendezvous: nothing of the peers' copies is read before every worker has published its arrival
        unsigned long long t0 = 0;
        if (blockIdx.x == 0) {
            peer_publish(ps, ps.ch_in, ps.epoch_in);
            if (tid == 0) t0 = global_ns();
        }
        peer_await(ps, ps.ch_in, ps.epoch_in);
        asm volatile("fence.proxy.async;" ::: "memory");   // the bulk-copy engine reads what the acquire made visible
        if (blockIdx.x == 0 && tid == 0 && ps.wait_ns) *ps.wait_ns = global_ns() - t0;
    }
    __syncthreads();

    auto issue = [&](int64_t tile, int s) {
        const int64_t off = tile * TILE;
        const int len = (int)(n4 - off < TILE ? n4 - off : TILE);
        Stage<TILE> st{sm + (size_t)s * kStageFloats};
        mbar_expect_tx(&full[s], op.load_bytes(len));
        op.loads(st, off, len, &full[s]);
    };

    if (tid == 0) {
        for (int k = 0; k < STAGES - 1; ++k) {
            const int64_t t = blockIdx.x + (int64_t)k * gridDim.x;
            if (t < ntiles) issue(t, k);
        }
    }
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int s = it % STAGES;
        if (tid == 0) {
            const int64_t tnext = tile + (int64_t)(STAGES - 1) * gridDim.x;
            if (tnext < ntiles) {
                bulk_wait_read_all();  // the stores that read the stage being refilled have drained
                issue(tnext, (it + STAGES - 1) % STAGES);
            }
        }
        mbar_wait(&full[s], (it / STAGES) & 1);
        const int64_t off = tile * TILE;
        const int len = (int)(n4 - off < TILE ? n4 - off : TILE);
        Stage<TILE> st{sm + (size_t)s * kStageFloats};
        op.compute(st, len, tid, THREADS);
        fence_async();  // generic-proxy writes to shared memory -> visible to the bulk-copy engine
        __syncthreads();
        if (tid == 0) {
            op.stores(st, off, len);
            bulk_commit();
        }
    }
    if (tid == 0) bulk_wait_all();
    // ragged tail: the last n % kGran elements, plain loads/stores
    if (blockIdx.x == 0 && tid < (int)(n - n4)) op.tail(n4 + tid);
    if constexpr (TileOp::kHasSum) grid_sum_to(op.acc(), sc, sum_out);
    if constexpr (PEER) {
        // completion rendezvous: the block that finishes last tells the peers and waits for theirs
        __shared__ bool last_to_finish;
        __syncthreads();                       // this block's tail stores are issued; thread 0's bulk stores are complete
        if (tid == 0) {
            __threadfence_system();            // ... and ordered before the count below for whoever observes it
            last_to_finish = atomicAdd(ps.done, 1u) == gridDim.x - 1;
        }
        __syncthreads();
        if (last_to_finish) {
            if (tid == 0) *ps.done = 0u;       // ready for the next launch (stream order)
            peer_publish(ps, ps.ch_out, ps.epoch_out);
            peer_await(ps, ps.ch_out, ps.epoch_out);
        }
    }
}

// ---- the same pipeline over a device TABLE of tensors -------------------------------------------
// For callers whose arrays are not one flat range: the inner AdamW step when the gradients are the
// tensors autograd allocated (reference semantics: `.grad` is None after every step, diloco.py:60),
// so only p, exp_avg and exp_avg_sq live in flat buffers.  A tensor is cut into ND_CHUNK-element
// chunks (the table's `first_chunk` numbering) and every chunk into ND_CHUNK / TILE tiles; tile
// g = blockIdx.x + k * gridDim.x belongs to chunk g / (ND_CHUNK / TILE).  The producer thread finds
// the row by binary search (the table is a few hundred rows, L1/L2-resident), publishes a TileDesc
// for the consumers through shared memory (ordered by the mbarrier it then arrives on) and issues
// the bulk loads.  Rows whose pointers are not 16-byte aligned, and the last numel % kGran
// elements of a tensor, take plain loads and stores straight from global memory.
struct TileDesc {
    void* ptr[4];  // the row's pointers, advanced to this tile's first element
    int len;       // elements of the tile (0: a chunk slot past the tensor's end)
    int bulk;      // leading elements that travel as bulk copies (multiple of kGran; 0 when unaligned)
};

__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}

template <int TILE, int STAGES, int THREADS, int MINB, class TileOp>
__global__ void __launch_bounds__(THREADS, MINB)
    tile_table_kernel(TileOp op, const nd_tensor_desc* __restrict__ tab, int T, int64_t total_chunks, Scratch* sc,
                      float* sum_out) {
    static_assert(ND_CHUNK % TILE == 0, "a chunk is a whole number of tiles");
    constexpr int kTilesPerChunk = ND_CHUNK / TILE;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    __shared__ uint64_t full[STAGES];
    __shared__ TileDesc desc[STAGES];
    constexpr int kStageFloats = TileOp::kBufs * TILE;
    const int64_t ntiles = total_chunks * kTilesPerChunk;
    const int tid = threadIdx.x;
    op.prologue();
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto issue = [&](int64_t g, int s) {   // producer thread only
        const int64_t chunk = g / kTilesPerChunk;
        const int sub = (int)(g % kTilesPerChunk);
        int lo = 0, hi = T - 1;
        while (lo < hi) {  // last row with first_chunk <= chunk
            const int mid = (lo + hi + 1) >> 1;
            if (__ldg(&tab[mid].first_chunk) <= chunk) lo = mid; else hi = mid - 1;
        }
        const nd_tensor_desc d = tab[lo];
        const int64_t off = (chunk - d.first_chunk) * (int64_t)ND_CHUNK + (int64_t)sub * TILE;
        int64_t len = d.numel - off;
        len = len < 0 ? 0 : (len > TILE ? TILE : len);
        TileDesc td;
        uintptr_t bits = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            td.ptr[k] = d.ptr[k] ? static_cast<char*>(d.ptr[k]) + off * TileOp::esz(k) : nullptr;
            bits |= reinterpret_cast<uintptr_t>(td.ptr[k]);
        }
        td.len = (int)len;
        td.bulk = (bits & 15) ? 0 : ((int)len & ~(TileOp::kGran - 1));
        desc[s] = td;
        if (td.bulk > 0) {
            TileOp o = op;
            o.rebase(td);
            Stage<TILE> st{sm + (size_t)s * kStageFloats};
            mbar_expect_tx(&full[s], o.load_bytes(td.bulk));
            o.loads(st, 0, td.bulk, &full[s]);
        } else {
            mbar_arrive(&full[s]);   // nothing in flight for this tile; the arrival publishes desc[s]
        }
    };

    if (tid == 0) {
        for (int k = 0; k < STAGES - 1; ++k) {
            const int64_t g = blockIdx.x + (int64_t)k * gridDim.x;
            if (g < ntiles) issue(g, k);
        }
    }
    int it = 0;
    for (int64_t g = blockIdx.x; g < ntiles; g += gridDim.x, ++it) {
        const int s = it % STAGES;
        if (tid == 0) {
            const int64_t gnext = g + (int64_t)(STAGES - 1) * gridDim.x;
            if (gnext < ntiles) {
                bulk_wait_read_all();  // the stores that read the stage being refilled have drained
                issue(gnext, (it + STAGES - 1) % STAGES);
            }
        }
        mbar_wait(&full[s], (it / STAGES) & 1);
        const TileDesc td = desc[s];   // into registers: the producer rewrites desc[s] STAGES-1 tiles from now
        TileOp o = op;
        o.rebase(td);
        Stage<TILE> st{sm + (size_t)s * kStageFloats};
        if (td.bulk > 0) o.compute(st, td.bulk, tid, THREADS);
        fence_async();
        __syncthreads();
        if (tid == 0 && td.bulk > 0) {
            o.stores(st, 0, td.bulk);
            bulk_commit();
        }
        for (int i = td.bulk + tid; i < td.len; i += THREADS) o.tail(i);
        if constexpr (TileOp::kHasSum) op.add_acc(o.acc());
    }
    if (tid == 0) bulk_wait_all();
    if constexpr (TileOp::kHasSum) grid_sum_to(op.acc(), sc, sum_out);
}

}  // namespace tma
}  // namespace nd
