// nd_common.cuh -- shared device helpers for the nanodiloco_b200 kernels (sm_100a).
//
// Everything on this path is a streaming, HBM-bound elementwise pass (no data
// reuse, no dense contraction), so the design rules are: 128-bit coalesced
// accesses, enough independent loads in flight per SM to cover HBM latency
// (>= ~50 KB/SM at ~7 TB/s), streaming cache hints so single-use data does not
// linger in L1, and grids sized as a multiple of the SM count.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/nanodiloco_b200.h"

namespace nd {

constexpr int kThreads = 512;        // threads per CTA for the streaming kernels
constexpr int kCtasPerSm = 4;        // 4 x 512 = 2048 resident threads / SM (<= 32 regs each)
constexpr int kMaxPartials = 148 * kCtasPerSm * 2;  // upper bound on gridDim.x for norm kernels

// ---- 128-bit streaming accesses -------------------------------------------
// ld.global.cs / st.global.cs: "cache streaming" -- the line is marked
// evict-first; right for data touched exactly once per launch.
__device__ __forceinline__ float4 ld_stream(const float* p) {
    return __ldcs(reinterpret_cast<const float4*>(p));
}
__device__ __forceinline__ void st_stream(float* p, const float4& v) {
    __stcs(reinterpret_cast<float4*>(p), v);
}
__device__ __forceinline__ float ld_stream1(const float* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream1(float* p, float v) { __stcs(p, v); }

// four bf16 <-> float4 (64-bit access)
__device__ __forceinline__ float4 ld_stream_bf16x4(const __nv_bfloat16* p) {
    uint2 raw = __ldcs(reinterpret_cast<const uint2*>(p));
    __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&raw.x);
    __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
    float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st_stream_bf16x4(__nv_bfloat16* p, const float4& v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 raw;
    raw.x = *reinterpret_cast<unsigned*>(&lo);
    raw.y = *reinterpret_cast<unsigned*>(&hi);
    __stcs(reinterpret_cast<uint2*>(p), raw);
}

// ---- typed access to the workers' local copy (fp32 or bf16) -----------------
template <bool BF16>
struct Local;
template <>
struct Local<false> {
    using T = float;
    static __device__ __forceinline__ float4 ld4(const void* base, int64_t i) {
        return ld_stream(static_cast<const float*>(base) + i);
    }
    static __device__ __forceinline__ void st4(void* base, int64_t i, const float4& v) {
        st_stream(static_cast<float*>(base) + i, v);
    }
    static __device__ __forceinline__ float ld1(const void* base, int64_t i) {
        return __ldcs(static_cast<const float*>(base) + i);
    }
    static __device__ __forceinline__ void st1(void* base, int64_t i, float v) {
        __stcs(static_cast<float*>(base) + i, v);
    }
};
template <>
struct Local<true> {
    using T = __nv_bfloat16;
    static __device__ __forceinline__ float4 ld4(const void* base, int64_t i) {
        return ld_stream_bf16x4(static_cast<const __nv_bfloat16*>(base) + i);
    }
    static __device__ __forceinline__ void st4(void* base, int64_t i, const float4& v) {
        st_stream_bf16x4(static_cast<__nv_bfloat16*>(base) + i, v);
    }
    static __device__ __forceinline__ float ld1(const void* base, int64_t i) {
        return __bfloat162float(static_cast<const __nv_bfloat16*>(base)[i]);
    }
    static __device__ __forceinline__ void st1(void* base, int64_t i, float v) {
        static_cast<__nv_bfloat16*>(base)[i] = __float2bfloat16_rn(v);
    }
};

// ---- the outer Nesterov-SGD recipe, one element -----------------------------
// torch:optim/sgd.py:343-380.  Explicit intrinsics pin the rounding sequence:
// nvcc must neither fuse b*mu into the following add nor split the FMAs.
struct OuterHyper {
    float neg_lr, mu, one_minus_damp, wd;
    int nesterov, first_step;
};

__device__ __forceinline__ void outer_update(float g, float& p, float& b, const OuterHyper& h) {
    if (h.wd != 0.0f) g = __fmaf_rn(h.wd, p, g);
    float d = g;
    if (h.mu != 0.0f) {
        if (h.first_step) {
            b = g;
        } else {
            b = __fmul_rn(b, h.mu);
            b = __fmaf_rn(h.one_minus_damp, g, b);
        }
        d = h.nesterov ? __fmaf_rn(h.mu, b, g) : b;
    }
    p = __fmaf_rn(h.neg_lr, d, p);
}

// ---- deterministic grid-wide sum ---------------------------------------------
// scratch layout: float partials[kMaxPartials]; unsigned counter (zero at rest).
struct Scratch {
    float partials[kMaxPartials];
    unsigned counter;
    unsigned pad[3];
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Every thread of every block calls this once with its local fp32 accumulator.
// Block partials are combined in a fixed order, in double, by whichever block
// arrives last; the counter resets itself so the scratch is reusable.
__device__ __forceinline__ void grid_sum_to(float acc, Scratch* sc, float* out) {
    __shared__ float warp_part[kThreads / 32];
    __shared__ bool last_block;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    acc = warp_sum(acc);
    if (lane == 0) warp_part[wid] = acc;
    __syncthreads();
    if (wid == 0) {
        float v = (lane < (blockDim.x >> 5)) ? warp_part[lane] : 0.0f;
        v = warp_sum(v);
        if (lane == 0) {
            sc->partials[blockIdx.x] = v;
            __threadfence();
            unsigned t = atomicAdd(&sc->counter, 1u);
            last_block = (t == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (last_block) {
        __threadfence();
        double s = 0.0;
        for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x)
            s += (double)__ldcg(&sc->partials[i]);
        // block-reduce the doubles in a fixed order
        __shared__ double dpart[kThreads / 32];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) dpart[wid] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += dpart[w];
            *out = (float)t;
            sc->counter = 0u;
        }
    }
}

}  // namespace nd
