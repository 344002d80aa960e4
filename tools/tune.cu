// tune.cu -- standalone tuning harness for the streaming kernels (not part of the product).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 tools/tune.cu -o tools/tune
//   ./tools/tune [elements]        (default 1,261,529,088: the 1B-class Llama parameter count)
//
// Times, with CUDA events, every (U = groups in flight per thread, B = resident CTAs per SM)
// variant of the fused outer kernel, the pseudo-gradient kernel, sumsq and AdamW, a plain device
// copy (the roofline denominator's own method) and an experimental TMA-bulk (cp.async.bulk +
// mbarrier) staging variant of the fused kernel.  Prints GB/s of ALGORITHMIC bytes.
#include "../nanodiloco_b200/csrc/nd_kernels.cu"

#include <vector>

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e = (x);                                                          \
        if (e != cudaSuccess) {                                                       \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                  \
        }                                                                             \
    } while (0)

// ---------------------------------------------------------------------------------------------
// TMA-bulk variant of the fused W=1 outer kernel.
// Each CTA owns tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...; a tile is TILE floats of each
// of the three arrays.  STAGES smem stages; thread 0 issues cp.async.bulk loads (mbarrier
// complete_tx) and bulk stores; all threads compute in place in shared memory.
// ---------------------------------------------------------------------------------------------
namespace tma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(b)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst_gmem, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int TILE, int STAGES, int THREADS>
__global__ void __launch_bounds__(THREADS) fused_w1_tma(float* __restrict__ master, float* __restrict__ mom,
                                                        float* __restrict__ local, int64_t n, nd::OuterHyper h) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);  // [STAGES][3][TILE]
    __shared__ uint64_t full[STAGES];
    const int64_t ntiles = (n + TILE - 1) / TILE;
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto issue_load = [&](int64_t tile, int s) {
        const int64_t off = tile * TILE;
        const uint32_t len = (uint32_t)((n - off < TILE ? n - off : TILE) * 4);
        float* base = sm + (size_t)s * 3 * TILE;
        mbar_expect_tx(&full[s], 3 * len);
        bulk_load(base, master + off, len, &full[s]);
        bulk_load(base + TILE, mom + off, len, &full[s]);
        bulk_load(base + 2 * TILE, local + off, len, &full[s]);
    };

    // prologue: fill STAGES-1 stages
    if (tid == 0) {
        for (int k = 0; k < STAGES - 1; ++k) {
            int64_t t = blockIdx.x + (int64_t)k * gridDim.x;
            if (t < ntiles) issue_load(t, k);
        }
    }
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int s = it % STAGES;
        const uint32_t parity = (it / STAGES) & 1;
        // refill: the stage used STAGES-1 iterations ahead is the one consumed last iteration
        if (tid == 0) {
            int64_t tnext = tile + (int64_t)(STAGES - 1) * gridDim.x;
            if (tnext < ntiles) {
                bulk_wait_read<0>();  // its stores have finished reading shared memory
                issue_load(tnext, (it + STAGES - 1) % STAGES);
            }
        }
        mbar_wait(&full[s], parity);
        const int64_t off = tile * TILE;
        const int len = (int)(n - off < TILE ? n - off : TILE);
        float* pm = sm + (size_t)s * 3 * TILE;
        float* pb = pm + TILE;
        float* pl = pm + 2 * TILE;
        for (int i = tid * 4; i < len; i += THREADS * 4) {
            float4 p = *reinterpret_cast<float4*>(pm + i);
            float4 b = *reinterpret_cast<float4*>(pb + i);
            float4 l = *reinterpret_cast<float4*>(pl + i);
            nd::outer_update(__fsub_rn(p.x, l.x), p.x, b.x, h);
            nd::outer_update(__fsub_rn(p.y, l.y), p.y, b.y, h);
            nd::outer_update(__fsub_rn(p.z, l.z), p.z, b.z, h);
            nd::outer_update(__fsub_rn(p.w, l.w), p.w, b.w, h);
            *reinterpret_cast<float4*>(pm + i) = p;
            *reinterpret_cast<float4*>(pb + i) = b;
        }
        fence_async();
        __syncthreads();
        if (tid == 0) {
            const uint32_t bytes = (uint32_t)len * 4;
            bulk_store(master + off, pm, bytes);
            bulk_store(local + off, pm, bytes);
            bulk_store(mom + off, pb, bytes);
            bulk_commit();
        }
    }
    if (tid == 0) bulk_wait_read<0>();
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace tma

__global__ void copy_kernel(const float4* __restrict__ a, float4* __restrict__ b, int64_t nvec) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i + stride < nvec; i += 2 * stride) {
        float4 x = __ldcs(a + i), y = __ldcs(a + i + stride);
        __stcs(b + i, x);
        __stcs(b + i + stride, y);
    }
    for (; i < nvec; i += stride) __stcs(b + i, __ldcs(a + i));
}

template <class F>
static float time_ms(F f, int warm = 3, int reps = 10) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < warm; ++i) f();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; ++i) f();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms / reps;
}

static void report(const char* name, float ms, double bytes) {
    printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms, bytes / (ms * 1e-3) / 1e9);
    fflush(stdout);
}

template <int U, int B>
static void fused_variant(float* master, float* mom, float* local, int64_t n, int sms) {
    nd::OuterFusedOp<1, false, false> op;
    op.master = master; op.mom = mom; op.w_rt = 1; op.world = 1.f; op.acc = 0.f;
    op.h = nd::make_outer_hyper(0.7f, 0.9f, 0.f, 0.f, 1, 0);
    for (int k = 0; k < ND_MAX_WORKERS; ++k) op.locals.p[k] = k == 0 ? local : nullptr;
    char name[64];
    snprintf(name, sizeof name, "fused_w1 LDG U=%d B=%d", U, B);
    float ms = time_ms([&] {
        nd::flat_kernel<U, B, decltype(op)><<<sms * B, nd::kThreads>>>(op, n, nullptr, nullptr);
    });
    report(name, ms, 24.0 * n);
}

template <int U, int B>
static void pg_variant(float* master, float* local, float* pg, int64_t n, int sms) {
    nd::PseudoGradOp<false> op{master, local, pg};
    char name[64];
    snprintf(name, sizeof name, "pseudo_grad LDG U=%d B=%d", U, B);
    float ms = time_ms([&] { nd::flat_kernel<U, B, decltype(op)><<<sms * B, nd::kThreads>>>(op, n, nullptr, nullptr); });
    report(name, ms, 12.0 * n);
}

template <int U, int B>
static void adam_variant(float* p, float* g, float* m, float* v, int64_t n, int sms) {
    nd::AdamWOp op;
    op.p = p; op.g = g; op.m = m; op.v = v; op.sumsq = nullptr; op.coef = 1.f;
    op.h = nd::make_adam_hyper(4e-4, 0.9, 0.999, 1e-8, 1e-2, 10, 1.0f, 1);
    char name[64];
    snprintf(name, sizeof name, "adamw(+zero_grad) LDG U=%d B=%d", U, B);
    float ms = time_ms([&] { nd::flat_kernel<U, B, decltype(op)><<<sms * B, nd::kThreads>>>(op, n, nullptr, nullptr); });
    report(name, ms, 32.0 * n);
}

template <int U, int B>
static void sumsq_variant(float* g, int64_t n, int sms, void* scratch, float* out) {
    nd::SumsqOp<false> op{g, 0.f};
    char name[64];
    snprintf(name, sizeof name, "sumsq LDG U=%d B=%d", U, B);
    float ms = time_ms([&] {
        nd::flat_kernel<U, B, decltype(op)><<<sms * B, nd::kThreads>>>(op, n, (nd::Scratch*)scratch, out);
    });
    report(name, ms, 4.0 * n);
}

template <int TILE, int STAGES, int THREADS>
static void tma_variant(float* master, float* mom, float* local, int64_t n, int sms, int ctas_per_sm) {
    auto k = tma::fused_w1_tma<TILE, STAGES, THREADS>;
    size_t smem = (size_t)STAGES * 3 * TILE * 4;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nd::OuterHyper h = nd::make_outer_hyper(0.7f, 0.9f, 0.f, 0.f, 1, 0);
    char name[96];
    snprintf(name, sizeof name, "fused_w1 TMA tile=%dKB st=%d thr=%d cta/sm=%d", TILE * 4 / 1024, STAGES, THREADS,
             ctas_per_sm);
    float ms = time_ms([&] { k<<<sms * ctas_per_sm, THREADS, smem>>>(master, mom, local, n, h); });
    report(name, ms, 24.0 * n);
}

int main(int argc, char** argv) {
    int64_t n = argc > 1 ? atoll(argv[1]) : 1261529088LL;
    int dev = 0, sms = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    printf("n = %lld elements, %d SMs\n", (long long)n, sms);
    float *a, *b, *c, *d;
    CK(cudaMalloc(&a, n * 4)); CK(cudaMalloc(&b, n * 4)); CK(cudaMalloc(&c, n * 4)); CK(cudaMalloc(&d, n * 4));
    CK(cudaMemset(a, 0, n * 4)); CK(cudaMemset(b, 0, n * 4)); CK(cudaMemset(c, 0, n * 4)); CK(cudaMemset(d, 0, n * 4));
    void* scratch; float* out;
    CK(cudaMalloc(&scratch, sizeof(nd::Scratch))); CK(cudaMemset(scratch, 0, sizeof(nd::Scratch)));
    CK(cudaMalloc(&out, 4));

    report("cudaMemcpyAsync D2D (8 B/elem)", time_ms([&] { CK(cudaMemcpyAsync(b, a, n * 4, cudaMemcpyDeviceToDevice)); }), 8.0 * n);
    report("copy_kernel float4 x2 (8 B/elem)", time_ms([&] { copy_kernel<<<sms * 4, 512>>>((float4*)a, (float4*)b, n / 4); }), 8.0 * n);

    // correctness of the TMA variant against the LDG kernel (small n, non-trivial values)
    {
        int64_t m = 1000003 / 4 * 4;
        std::vector<float> hm(m), hl(m), hb(m);
        for (int64_t i = 0; i < m; ++i) { hm[i] = 0.02f * sinf(i * 0.37f); hl[i] = hm[i] + 1e-3f * cosf(i * 0.11f); hb[i] = 1e-3f * sinf(i * 0.05f); }
        float *x[6];
        for (auto& p : x) CK(cudaMalloc(&p, m * 4));
        for (int r = 0; r < 2; ++r) {
            CK(cudaMemcpy(x[3 * r], hm.data(), m * 4, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(x[3 * r + 1], hb.data(), m * 4, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(x[3 * r + 2], hl.data(), m * 4, cudaMemcpyHostToDevice));
        }
        nd::OuterHyper h = nd::make_outer_hyper(0.7f, 0.9f, 0.f, 0.f, 1, 0);
        void* locals[1] = {x[2]};
        nd_outer_fused(x[0], x[1], locals, 1, ND_F32, m, 0.7f, 0.9f, 0.f, 0.f, 1, 0, nullptr, nullptr, nullptr);
        auto k = tma::fused_w1_tma<2048, 4, 256>;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 3 * 2048 * 4));
        k<<<sms * 2, 256, 4 * 3 * 2048 * 4>>>(x[3], x[4], x[5], m, h);
        CK(cudaDeviceSynchronize());
        std::vector<float> r0(m), r1(m);
        bool ok = true;
        for (int j = 0; j < 3; ++j) {
            CK(cudaMemcpy(r0.data(), x[j], m * 4, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(r1.data(), x[3 + j], m * 4, cudaMemcpyDeviceToHost));
            ok &= memcmp(r0.data(), r1.data(), m * 4) == 0;
        }
        printf("TMA variant bit-identical to LDG kernel: %s\n", ok ? "yes" : "NO");
        for (auto& p : x) CK(cudaFree(p));
    }

    fused_variant<1, 4>(a, b, c, n, sms);
    fused_variant<2, 4>(a, b, c, n, sms);
    fused_variant<1, 3>(a, b, c, n, sms);
    fused_variant<2, 3>(a, b, c, n, sms);
    fused_variant<1, 2>(a, b, c, n, sms);
    fused_variant<2, 2>(a, b, c, n, sms);
    fused_variant<3, 2>(a, b, c, n, sms);
    fused_variant<4, 2>(a, b, c, n, sms);
    fused_variant<2, 1>(a, b, c, n, sms);
    fused_variant<4, 1>(a, b, c, n, sms);
    fused_variant<8, 1>(a, b, c, n, sms);

    tma_variant<2048, 4, 256>(a, b, c, n, sms, 2);
    tma_variant<2048, 3, 256>(a, b, c, n, sms, 3);
    tma_variant<4096, 4, 256>(a, b, c, n, sms, 1);
    tma_variant<4096, 2, 512>(a, b, c, n, sms, 2);
    tma_variant<1024, 4, 256>(a, b, c, n, sms, 4);
    tma_variant<1024, 6, 128>(a, b, c, n, sms, 3);
    tma_variant<2048, 8, 256>(a, b, c, n, sms, 1);

    pg_variant<1, 4>(a, c, d, n, sms);
    pg_variant<2, 4>(a, c, d, n, sms);
    pg_variant<4, 2>(a, c, d, n, sms);
    pg_variant<4, 4>(a, c, d, n, sms);
    pg_variant<8, 2>(a, c, d, n, sms);

    sumsq_variant<2, 4>(d, n, sms, scratch, out);
    sumsq_variant<4, 4>(d, n, sms, scratch, out);
    sumsq_variant<8, 2>(d, n, sms, scratch, out);
    sumsq_variant<8, 4>(d, n, sms, scratch, out);

    adam_variant<1, 4>(a, d, b, c, n, sms);
    adam_variant<1, 3>(a, d, b, c, n, sms);
    adam_variant<1, 2>(a, d, b, c, n, sms);
    adam_variant<2, 2>(a, d, b, c, n, sms);
    adam_variant<2, 1>(a, d, b, c, n, sms);
    adam_variant<4, 1>(a, d, b, c, n, sms);
    return 0;
}
