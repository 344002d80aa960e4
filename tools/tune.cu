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
    for (int k = 0; k < ND_MAX_WORKERS; ++k) op.locals.p[k] = op.writes.p[k] = k == 0 ? local : nullptr;
    op.n_writes = 1;
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

template <int TILE, int STAGES, int THREADS, int MINB>
static void tma_fused(float* master, float* mom, float* local, int64_t n, int sms) {
    nd::OuterFusedOp<1, false, false> op;
    op.master = master; op.mom = mom; op.w_rt = 1; op.world = 1.f; op.acc = 0.f;
    op.h = nd::make_outer_hyper(0.7f, 0.9f, 0.f, 0.f, 1, 0);
    for (int k = 0; k < ND_MAX_WORKERS; ++k) op.locals.p[k] = op.writes.p[k] = k == 0 ? local : nullptr;
    op.n_writes = 1;
    nd::FusedTileOp<1, false, false> t{op};
    char name[96];
    snprintf(name, sizeof name, "fused_w1 TMA tile=%dKB st=%d thr=%d cta/sm=%d", TILE * 4 / 1024, STAGES, THREADS, MINB);
    float ms = time_ms([&] { nd::launch_tile<TILE, STAGES, THREADS, MINB>(t, n, nullptr, nullptr, 0); });
    report(name, ms, 24.0 * n);
}

template <int TILE, int STAGES, int THREADS, int MINB>
static void tma_adam(float* p, float* g, float* m, float* v, int64_t n, int sms) {
    nd::AdamWOp op;
    op.p = p; op.g = g; op.m = m; op.v = v; op.sumsq = nullptr; op.coef = 1.f;
    op.h = nd::make_adam_hyper(4e-4, 0.9, 0.999, 1e-8, 1e-2, 10, 1.0f, 1);
    nd::AdamTileOp t{op};
    char name[96];
    snprintf(name, sizeof name, "adamw(+zero_grad) TMA tile=%dKB st=%d thr=%d cta/sm=%d", TILE * 4 / 1024, STAGES, THREADS, MINB);
    float ms = time_ms([&] { nd::launch_tile<TILE, STAGES, THREADS, MINB>(t, n, nullptr, nullptr, 0); });
    report(name, ms, 32.0 * n);
}

template <int W, int TILE, int STAGES, int THREADS, int MINB>
static void tma_fused_w(float* master, float* mom, float** locals, int64_t n, int sms) {
    nd::OuterFusedOp<W, false, false> op;
    op.master = master; op.mom = mom; op.w_rt = W; op.world = (float)W; op.acc = 0.f;
    op.h = nd::make_outer_hyper(0.7f, 0.9f, 0.f, 0.f, 1, 0);
    for (int k = 0; k < ND_MAX_WORKERS; ++k) op.locals.p[k] = op.writes.p[k] = k < W ? locals[k] : nullptr;
    op.n_writes = W;
    nd::FusedTileOp<W, false, false> t{op};
    char name[96];
    snprintf(name, sizeof name, "fused W=%d TMA tile=%dKB st=%d thr=%d cta/sm=%d", W, TILE * 4 / 1024, STAGES, THREADS, MINB);
    float ms = time_ms([&] { nd::launch_tile<TILE, STAGES, THREADS, MINB>(t, n, nullptr, nullptr, 0); });
    report(name, ms, (16.0 + 8.0 * W) * n);
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

    const bool tma_only = argc > 2;
    if (tma_only) {
        tma_fused<4096, 2, 512, 2>(a, b, c, n, sms);
        tma_fused<4096, 2, 256, 2>(a, b, c, n, sms);
        tma_fused<4096, 2, 1024, 2>(a, b, c, n, sms);
        tma_fused<2048, 3, 256, 3>(a, b, c, n, sms);
        tma_fused<8192, 2, 1024, 1>(a, b, c, n, sms);
        tma_adam<2048, 2, 512, 2>(a, d, b, c, n, sms);
        tma_adam<2048, 2, 512, 3>(a, d, b, c, n, sms);
        tma_adam<2048, 2, 1024, 2>(a, d, b, c, n, sms);
        tma_adam<1024, 2, 512, 4>(a, d, b, c, n, sms);
        tma_adam<1024, 3, 512, 3>(a, d, b, c, n, sms);
        tma_adam<1024, 2, 256, 5>(a, d, b, c, n, sms);
        tma_adam<4096, 2, 1024, 1>(a, d, b, c, n, sms);
        tma_adam<512, 3, 256, 6>(a, d, b, c, n, sms);
        tma_adam<3072, 2, 768, 2>(a, d, b, c, n, sms);
        return 0;
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

    tma_fused<4096, 2, 512, 2>(a, b, c, n, sms);
    tma_fused<4096, 2, 256, 2>(a, b, c, n, sms);
    tma_fused<4096, 3, 512, 1>(a, b, c, n, sms);
    tma_fused<4096, 4, 512, 1>(a, b, c, n, sms);
    tma_fused<8192, 2, 512, 1>(a, b, c, n, sms);
    tma_fused<2048, 3, 256, 3>(a, b, c, n, sms);
    tma_fused<2048, 2, 256, 4>(a, b, c, n, sms);
    tma_fused<2048, 4, 256, 2>(a, b, c, n, sms);
    tma_fused<2048, 2, 128, 4>(a, b, c, n, sms);
    tma_fused<1024, 4, 128, 4>(a, b, c, n, sms);
    tma_fused<1024, 3, 128, 6>(a, b, c, n, sms);
    tma_fused<1024, 2, 128, 8>(a, b, c, n, sms);
    tma_fused<512, 4, 128, 8>(a, b, c, n, sms);
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
    adam_variant<1, 1>(a, d, b, c, n, sms);
    tma_adam<2048, 3, 256, 2>(a, d, b, c, n, sms);
    tma_adam<2048, 2, 256, 3>(a, d, b, c, n, sms);
    tma_adam<2048, 2, 512, 2>(a, d, b, c, n, sms);
    tma_adam<4096, 2, 512, 1>(a, d, b, c, n, sms);
    tma_adam<4096, 3, 512, 1>(a, d, b, c, n, sms);
    tma_adam<1024, 3, 128, 4>(a, d, b, c, n, sms);
    tma_adam<1024, 2, 256, 6>(a, d, b, c, n, sms);
    tma_adam<1024, 4, 256, 3>(a, d, b, c, n, sms);
    // several worker copies emulated on one GPU (each of n/4 elements so that everything fits)
    {
        int64_t m = n / 4 / 4 * 4;
        float* loc[8];
        for (int k = 0; k < 8; ++k) { CK(cudaMalloc(&loc[k], m * 4)); CK(cudaMemset(loc[k], 0, m * 4)); }
        tma_fused_w<2, 2048, 3, 256, 2>(a, b, loc, m, sms);
        tma_fused_w<2, 4096, 2, 512, 1>(a, b, loc, m, sms);
        tma_fused_w<4, 2048, 2, 256, 2>(a, b, loc, m, sms);
        tma_fused_w<4, 1024, 3, 256, 3>(a, b, loc, m, sms);
        tma_fused_w<8, 2048, 2, 256, 1>(a, b, loc, m, sms);
        tma_fused_w<8, 1024, 2, 256, 2>(a, b, loc, m, sms);
        tma_fused_w<8, 1024, 3, 256, 1>(a, b, loc, m, sms);
    }
    return 0;
}
