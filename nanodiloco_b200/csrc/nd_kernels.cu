// nd_kernels.cu -- sm_100a kernels of the DiLoCo outer-sync hot path + their C ABI.
//
// Kernel inventory (each is one HBM pass; algorithmic bytes/element in brackets):
//   pseudo_grad        pg = master - local                         [12 fp32 | 10 bf16 local]
//   outer_shard        AVG-scale + Nesterov + master/mom update + copy-back on a shard
//                                                                  [24 fp32 | 22 bf16]
//   outer_fused<W>     both of the above for W addressable worker copies, pseudo-gradient
//                      kept in registers                           [16 + 8W fp32]
//   sumsq              sum of squares (clip_grad_norm_ stage 1)     [4]
//   adamw              clip-scale + AdamW (+ optional zero_grad)    [28 (+4)]
// and table-driven multi-tensor forms of each for non-flat callers.
//
// Structure: an "Op" describes how to load / compute / store one group of four
// consecutive elements (or one element on the scalar path).  Register-staged
// drivers run an Op: `flat_kernel` over one contiguous range and `table_kernel`
// over a device table of tensors cut into ND_CHUNK-element chunks.  Both issue
// all loads of U independent groups before any dependent arithmetic or store,
// so each thread keeps U x (#inputs) 128-bit requests in flight.
// A "TileOp" wraps an Op for the TMA-staged driver (nd_tma.cuh: cp.async.bulk +
// mbarrier, whole tiles landing in shared memory and leaving with bulk stores),
// which is what the hot calls use whenever their buffers are 16-byte aligned:
// fused outer kernel, shard update and AdamW.  The register-staged path serves
// unaligned buffers, run-time worker counts, pseudo_grad, sumsq and the tables.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <atomic>
#include <type_traits>

#include "nd_common.cuh"
#include "nd_error.h"
#include "nd_tma.cuh"

namespace nd {

// ===========================================================================
// Ops
// ===========================================================================

template <bool BF16>
struct PseudoGradOp {
    static constexpr bool kHasSum = false;
    const float* master;
    const void* local;
    float* pg;
    struct Regs { float4 a, b; };

    __device__ __forceinline__ void rebase(const nd_tensor_desc& d, int64_t off) {
        master = static_cast<const float*>(d.ptr[0]) + off;
        local = static_cast<const typename Local<BF16>::T*>(d.ptr[1]) + off;
        pg = static_cast<float*>(d.ptr[2]) + off;
    }
    __device__ __forceinline__ bool aligned() const {
        return ((reinterpret_cast<uintptr_t>(master) | reinterpret_cast<uintptr_t>(pg)) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(local) & (BF16 ? 7 : 15)) == 0;
    }
    __device__ __forceinline__ void load4(int64_t i, Regs& r) const {
        r.a = ld_stream(master + i);
        r.b = Local<BF16>::ld4(local, i);
    }
    __device__ __forceinline__ void compute4(Regs& r) {
        r.a.x = __fsub_rn(r.a.x, r.b.x); r.a.y = __fsub_rn(r.a.y, r.b.y);
        r.a.z = __fsub_rn(r.a.z, r.b.z); r.a.w = __fsub_rn(r.a.w, r.b.w);
    }
    __device__ __forceinline__ void store4(int64_t i, const Regs& r) const { st_stream(pg + i, r.a); }
    __device__ __forceinline__ void scalar(int64_t i) {
        pg[i] = __fsub_rn(master[i], Local<BF16>::ld1(local, i));
    }
};

// Outer update of a shard whose pseudo-gradient SUM is already in memory.
template <bool BF16, bool SUM>
struct OuterShardOp {
    static constexpr bool kHasSum = SUM;
    float* master;
    float* mom;
    const float* pg_sum;
    void* local_out;  // nullable
    float world;
    OuterHyper h;
    float acc;
    struct Regs { float4 p, b, g; };

    __device__ __forceinline__ void rebase(const nd_tensor_desc&, int64_t) {}
    __device__ __forceinline__ bool aligned() const {
        return ((reinterpret_cast<uintptr_t>(master) | reinterpret_cast<uintptr_t>(mom) |
                 reinterpret_cast<uintptr_t>(pg_sum)) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(local_out) & (BF16 ? 7 : 15)) == 0;
    }
    __device__ __forceinline__ void load4(int64_t i, Regs& r) const {
        r.p = ld_stream(master + i);
        r.g = ld_stream(pg_sum + i);
        r.b = h.first_step ? make_float4(0.f, 0.f, 0.f, 0.f) : ld_stream(mom + i);
    }
    __device__ __forceinline__ void one(float& p, float& b, float s) {
        float g = __fdiv_rn(s, world);
        if (SUM) acc = __fmaf_rn(g, g, acc);
        outer_update(g, p, b, h);
    }
    __device__ __forceinline__ void compute4(Regs& r) {
        one(r.p.x, r.b.x, r.g.x); one(r.p.y, r.b.y, r.g.y);
        one(r.p.z, r.b.z, r.g.z); one(r.p.w, r.b.w, r.g.w);
    }
    __device__ __forceinline__ void store4(int64_t i, const Regs& r) const {
        st_stream(master + i, r.p);
        st_stream(mom + i, r.b);
        if (local_out) Local<BF16>::st4(local_out, i, r.p);
    }
    __device__ __forceinline__ void scalar(int64_t i) {
        float p = master[i], b = h.first_step ? 0.0f : mom[i];
        one(p, b, pg_sum[i]);
        master[i] = p; mom[i] = b;
        if (local_out) Local<BF16>::st1(local_out, i, p);
    }
};

struct LocalPtrs { void* p[ND_MAX_WORKERS]; };

// Kernels (1)+(2) fused over W worker copies.  W > 0: compile-time count, all W
// local loads issued up front; W == 0: run-time count (9..ND_MAX_WORKERS).
template <int W, bool BF16, bool SUM>
struct OuterFusedOp {
    static constexpr bool kHasSum = SUM;
    static constexpr int kW = W > 0 ? W : 1;
    float* master;
    float* mom;
    LocalPtrs locals;  // the worker copies that are READ (one per DiLoCo worker)
    LocalPtrs writes;  // the copies that receive the new parameters (every replica of every worker)
    int n_writes;
    int w_rt;  // run-time W
    float world;
    OuterHyper h;
    float acc;
    struct Regs { float4 p, b; float4 l[kW]; };

    __device__ __forceinline__ void rebase(const nd_tensor_desc& d, int64_t off) {
        master = static_cast<float*>(d.ptr[0]) + off;
        mom = static_cast<float*>(d.ptr[1]) + off;
        locals.p[0] = static_cast<typename Local<BF16>::T*>(d.ptr[2]) + off;
        writes.p[0] = locals.p[0];
        n_writes = 1;
    }
    __device__ __forceinline__ bool aligned() const {
        uintptr_t a = reinterpret_cast<uintptr_t>(master) | reinterpret_cast<uintptr_t>(mom);
        uintptr_t l = 0;
        const int w = W > 0 ? W : w_rt;
        for (int k = 0; k < w; ++k) l |= reinterpret_cast<uintptr_t>(locals.p[k]);
        for (int k = 0; k < n_writes; ++k) l |= reinterpret_cast<uintptr_t>(writes.p[k]);
        return (a & 15) == 0 && (l & (BF16 ? 7 : 15)) == 0;
    }
    __device__ __forceinline__ void load4(int64_t i, Regs& r) const {
        r.p = ld_stream(master + i);
        r.b = h.first_step ? make_float4(0.f, 0.f, 0.f, 0.f) : ld_stream(mom + i);
        if (W > 0) {
#pragma unroll
            for (int k = 0; k < kW; ++k) r.l[k] = Local<BF16>::ld4(locals.p[k], i);
        } else {
            // run-time W: accumulate as we go (sum of (master - local_w) in worker order)
            float4 s;
            for (int k = 0; k < w_rt; ++k) {
                float4 l = Local<BF16>::ld4(locals.p[k], i);
                float4 d = make_float4(__fsub_rn(r.p.x, l.x), __fsub_rn(r.p.y, l.y),
                                       __fsub_rn(r.p.z, l.z), __fsub_rn(r.p.w, l.w));
                if (k == 0) s = d;
                else s = make_float4(__fadd_rn(s.x, d.x), __fadd_rn(s.y, d.y),
                                     __fadd_rn(s.z, d.z), __fadd_rn(s.w, d.w));
            }
            r.l[0] = s;
        }
    }
    __device__ __forceinline__ void one(float& p, float& b, float s) {
        float g = (W == 1) ? s : __fdiv_rn(s, world);
        if (SUM) acc = __fmaf_rn(g, g, acc);
        outer_update(g, p, b, h);
    }
    __device__ __forceinline__ void compute4(Regs& r) {
        float4 s;
        if (W > 0) {
            s = make_float4(__fsub_rn(r.p.x, r.l[0].x), __fsub_rn(r.p.y, r.l[0].y),
                            __fsub_rn(r.p.z, r.l[0].z), __fsub_rn(r.p.w, r.l[0].w));
#pragma unroll
            for (int k = 1; k < kW; ++k) {
                s.x = __fadd_rn(s.x, __fsub_rn(r.p.x, r.l[k].x));
                s.y = __fadd_rn(s.y, __fsub_rn(r.p.y, r.l[k].y));
                s.z = __fadd_rn(s.z, __fsub_rn(r.p.z, r.l[k].z));
                s.w = __fadd_rn(s.w, __fsub_rn(r.p.w, r.l[k].w));
            }
        } else {
            s = r.l[0];
        }
        one(r.p.x, r.b.x, s.x); one(r.p.y, r.b.y, s.y);
        one(r.p.z, r.b.z, s.z); one(r.p.w, r.b.w, s.w);
    }
    __device__ __forceinline__ void store4(int64_t i, const Regs& r) const {
        st_stream(master + i, r.p);
        st_stream(mom + i, r.b);
        for (int k = 0; k < n_writes; ++k) Local<BF16>::st4(writes.p[k], i, r.p);
    }
    __device__ __forceinline__ void scalar(int64_t i) {
        const int w = W > 0 ? W : w_rt;
        float p = master[i], b = h.first_step ? 0.0f : mom[i];
        float s = __fsub_rn(p, Local<BF16>::ld1(locals.p[0], i));
        for (int k = 1; k < w; ++k) s = __fadd_rn(s, __fsub_rn(p, Local<BF16>::ld1(locals.p[k], i)));
        float g = (w == 1) ? s : __fdiv_rn(s, world);
        if (SUM) acc = __fmaf_rn(g, g, acc);
        outer_update(g, p, b, h);
        master[i] = p; mom[i] = b;
        for (int k = 0; k < n_writes; ++k) Local<BF16>::st1(writes.p[k], i, p);
    }
};

template <bool BF16>
struct SumsqOp {
    static constexpr bool kHasSum = true;
    const void* g;
    float acc;
    struct Regs { float4 a; };
    __device__ __forceinline__ void rebase(const nd_tensor_desc& d, int64_t off) {
        g = static_cast<const typename Local<BF16>::T*>(d.ptr[0]) + off;
    }
    __device__ __forceinline__ bool aligned() const {
        return (reinterpret_cast<uintptr_t>(g) & (BF16 ? 7 : 15)) == 0;
    }
    __device__ __forceinline__ void load4(int64_t i, Regs& r) const { r.a = Local<BF16>::ld4(g, i); }
    __device__ __forceinline__ void compute4(Regs& r) {
        acc = __fmaf_rn(r.a.x, r.a.x, acc); acc = __fmaf_rn(r.a.y, r.a.y, acc);
        acc = __fmaf_rn(r.a.z, r.a.z, acc); acc = __fmaf_rn(r.a.w, r.a.w, acc);
    }
    __device__ __forceinline__ void store4(int64_t, const Regs&) const {}
    __device__ __forceinline__ void scalar(int64_t i) {
        float x = Local<BF16>::ld1(g, i);
        acc = __fmaf_rn(x, x, acc);
    }
};

struct AdamHyper {
    float decay;      // 1 - lr*wd              (adam.py:692)
    float w1;         // 1 - beta1              (adam.py:705-707)
    float b2;         // beta2                  (adam.py:709)
    float w2;         // 1 - beta2              (adam.py:722-724)
    float step_size;  // -lr / bc1              (adam.py:779)
    float bc2_sqrt;   // sqrt(1 - beta2^step)   (adam.py:781)
    float eps;
    float max_norm;
    int has_decay, zero_grad;
};

// BF16: parameters and gradients are bf16 (the "bf16 inner / fp32 outer" working copy of BASELINE
// config 2 -- an extension, the reference is fp32 end to end); moments stay fp32 and all arithmetic
// is fp32 on the exactly-widened values, with one rounding of the new parameter back to bf16.
template <bool BF16>
struct AdamWOpT {
    static constexpr bool kHasSum = false;
    using PG = typename Local<BF16>::T;
    PG* p; PG* g; float* m; float* v;
    const float* sumsq;  // nullable
    AdamHyper h;
    float coef;
    // replicas of one worker (nd_adamw_rw): the squared norm arrives as one partial sum per replica (each
    // replica reduced its own part of the gradient) and the new parameters go to every replica's copy
    const float* sumsq_parts[ND_MAX_REPLICAS];
    int n_parts = 0;
    PG* p_extra[ND_MAX_REPLICAS];
    int n_extra = 0;
    struct Regs { float4 p, g, m, v; };

    __device__ __forceinline__ void rebase(const nd_tensor_desc& d, int64_t off) {
        p = static_cast<PG*>(d.ptr[0]) + off; g = static_cast<PG*>(d.ptr[1]) + off;
        m = static_cast<float*>(d.ptr[2]) + off; v = static_cast<float*>(d.ptr[3]) + off;
    }
    // clip_grad.py:165-168, evaluated on the device so the host never waits for the norm
    __device__ __forceinline__ void prologue() {
        coef = 1.0f;
        if (sumsq || n_parts > 0) {
            float ss = sumsq ? __ldg(sumsq) : 0.0f;
            for (int k = 0; k < n_parts; ++k) {   // fixed order: every replica forms the same total
                float part;
                asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(part) : "l"(sumsq_parts[k]) : "memory");
                ss = (k == 0 && !sumsq) ? part : __fadd_rn(ss, part);
            }
            float total = sqrtf(ss);
            coef = fminf(__fdiv_rn(h.max_norm, total + 1e-6f), 1.0f);
        }
    }
    __device__ __forceinline__ bool aligned() const {
        return ((reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0 &&
               ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g)) & (BF16 ? 7 : 15)) == 0;
    }
    __device__ __forceinline__ void load4(int64_t i, Regs& r) const {
        r.p = Local<BF16>::ld4(p, i); r.g = Local<BF16>::ld4(g, i);
        r.m = ld_stream(m + i); r.v = ld_stream(v + i);
    }
    __device__ __forceinline__ void one(float& pi, float gi, float& mi, float& vi) const {
        gi = __fmul_rn(gi, coef);
        if (h.has_decay) pi = __fmul_rn(pi, h.decay);
        mi = __fmaf_rn(h.w1, __fsub_rn(gi, mi), mi);
        vi = __fmul_rn(vi, h.b2);
        vi = __fmaf_rn(h.w2, __fmul_rn(gi, gi), vi);
        float den = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), h.bc2_sqrt), h.eps);
        pi = __fmaf_rn(h.step_size, __fdiv_rn(mi, den), pi);
    }
    __device__ __forceinline__ void compute4(Regs& r) const {
        one(r.p.x, r.g.x, r.m.x, r.v.x); one(r.p.y, r.g.y, r.m.y, r.v.y);
        one(r.p.z, r.g.z, r.m.z, r.v.z); one(r.p.w, r.g.w, r.m.w, r.v.w);
    }
    __device__ __forceinline__ void store4(int64_t i, const Regs& r) const {
        Local<BF16>::st4(p, i, r.p); st_stream(m + i, r.m); st_stream(v + i, r.v);
        for (int k = 0; k < n_extra; ++k) Local<BF16>::st4(p_extra[k], i, r.p);
        if (h.zero_grad) Local<BF16>::st4(g, i, make_float4(0.f, 0.f, 0.f, 0.f));
    }
    __device__ __forceinline__ void scalar(int64_t i) const {
        float pi = Local<BF16>::ld1(p, i), mi = m[i], vi = v[i];
        one(pi, Local<BF16>::ld1(g, i), mi, vi);
        Local<BF16>::st1(p, i, pi); m[i] = mi; v[i] = vi;
        for (int k = 0; k < n_extra; ++k) Local<BF16>::st1(p_extra[k], i, pi);
        if (h.zero_grad) Local<BF16>::st1(g, i, 0.0f);
    }
};
using AdamWOp = AdamWOpT<false>;

template <class Op, class = void>
struct HasPrologue { static constexpr bool value = false; };
template <class Op>
struct HasPrologue<Op, decltype(std::declval<Op&>().prologue(), void())> {
    static constexpr bool value = true;
};


// ===========================================================================
// TileOps: the same arithmetic on shared-memory tiles for the TMA-staged driver (nd_tma.cuh).
// Each wraps the register-path Op above (pointers, hyper-parameters, scalar tail).
// ===========================================================================

__device__ __forceinline__ float4 lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void sts4(float* p, const float4& v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 lds_bf16x4(const float* buf, int i) {
    uint2 raw = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(buf) + i);
    float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&raw.x));
    float2 b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&raw.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void sts_bf16x4(float* buf, int i, const float4& v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 raw;
    raw.x = *reinterpret_cast<unsigned*>(&lo);
    raw.y = *reinterpret_cast<unsigned*>(&hi);
    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(buf) + i) = raw;
}

// buffers: 0 = master -> new params, 1 = momentum, 2.. = the W local copies (buffer 2 is reused
// for the bf16 image of the new params when the working copy is bf16)
template <int W, bool BF16, bool SUM>
struct FusedTileOp {
    static_assert(W >= 1, "compile-time worker count");
    static constexpr int kBufs = 2 + W;
    static constexpr bool kHasSum = SUM;
    static constexpr int kEsz = BF16 ? 2 : 4;
    static constexpr int kGran = 16 / kEsz;
    OuterFusedOp<W, BF16, SUM> o;
    __device__ __forceinline__ void prologue() { o.acc = 0.0f; }
    __device__ __forceinline__ float acc() const { return o.acc; }
    __device__ __forceinline__ uint32_t load_bytes(int len) const {
        return (uint32_t)len * ((o.h.first_step ? 4u : 8u) + (uint32_t)(W * kEsz));
    }
    template <int TILE>
    __device__ __forceinline__ void loads(const tma::Stage<TILE>& st, int64_t off, int len, uint64_t* bar) const {
        tma::bulk_load(st.buf(0), o.master + off, (uint32_t)len * 4, bar);
        if (!o.h.first_step) tma::bulk_load(st.buf(1), o.mom + off, (uint32_t)len * 4, bar);
#pragma unroll
        for (int k = 0; k < W; ++k)
            tma::bulk_load(st.buf(2 + k), static_cast<const char*>(o.locals.p[k]) + off * kEsz, (uint32_t)len * kEsz, bar);
    }
    template <int TILE>
    __device__ __forceinline__ void compute(const tma::Stage<TILE>& st, int len, int tid, int nthreads) {
        for (int i = tid * 4; i < len; i += nthreads * 4) {
            float4 p = lds4(st.buf(0) + i);
            float4 b = o.h.first_step ? make_float4(0.f, 0.f, 0.f, 0.f) : lds4(st.buf(1) + i);
            float4 s;
#pragma unroll
            for (int k = 0; k < W; ++k) {
                float4 l = BF16 ? lds_bf16x4(st.buf(2 + k), i) : lds4(st.buf(2 + k) + i);
                float4 d = make_float4(__fsub_rn(p.x, l.x), __fsub_rn(p.y, l.y), __fsub_rn(p.z, l.z), __fsub_rn(p.w, l.w));
                if (k == 0) s = d;
                else s = make_float4(__fadd_rn(s.x, d.x), __fadd_rn(s.y, d.y), __fadd_rn(s.z, d.z), __fadd_rn(s.w, d.w));
            }
            o.one(p.x, b.x, s.x); o.one(p.y, b.y, s.y); o.one(p.z, b.z, s.z); o.one(p.w, b.w, s.w);
            sts4(st.buf(0) + i, p);
            sts4(st.buf(1) + i, b);
            if (BF16) sts_bf16x4(st.buf(2), i, p);
        }
    }
    template <int TILE>
    __device__ __forceinline__ void stores(const tma::Stage<TILE>& st, int64_t off, int len) const {
        tma::bulk_store(o.master + off, st.buf(0), (uint32_t)len * 4);
        tma::bulk_store(o.mom + off, st.buf(1), (uint32_t)len * 4);
        for (int k = 0; k < o.n_writes; ++k)
            tma::bulk_store(static_cast<char*>(o.writes.p[k]) + off * kEsz, st.buf(BF16 ? 2 : 0), (uint32_t)len * kEsz);
    }
    __device__ __forceinline__ void tail(int64_t i) { o.scalar(i); }
};

// buffers: 0 = master shard, 1 = momentum shard, 2 = reduced pseudo-gradient, 3 = bf16 image of the
// new parameters (bf16 working copy only).  The image cannot share buffer 2: a thread's bf16 store
// at bytes [2i, 2i+8) would land on fp32 pseudo-gradient elements that another warp has yet to read.
template <bool BF16, bool SUM>
struct ShardTileOp {
    static constexpr int kBufs = BF16 ? 4 : 3;
    static constexpr int kOutBuf = BF16 ? 3 : 0;
    static constexpr bool kHasSum = SUM;
    static constexpr int kEsz = BF16 ? 2 : 4;
    static constexpr int kGran = 16 / kEsz;
    OuterShardOp<BF16, SUM> o;
    __device__ __forceinline__ void prologue() { o.acc = 0.0f; }
    __device__ __forceinline__ float acc() const { return o.acc; }
    __device__ __forceinline__ uint32_t load_bytes(int len) const { return (uint32_t)len * (o.h.first_step ? 8u : 12u); }
    template <int TILE>
    __device__ __forceinline__ void loads(const tma::Stage<TILE>& st, int64_t off, int len, uint64_t* bar) const {
        tma::bulk_load(st.buf(0), o.master + off, (uint32_t)len * 4, bar);
        if (!o.h.first_step) tma::bulk_load(st.buf(1), o.mom + off, (uint32_t)len * 4, bar);
        tma::bulk_load(st.buf(2), o.pg_sum + off, (uint32_t)len * 4, bar);
    }
    template <int TILE>
    __device__ __forceinline__ void compute(const tma::Stage<TILE>& st, int len, int tid, int nthreads) {
        for (int i = tid * 4; i < len; i += nthreads * 4) {
            float4 p = lds4(st.buf(0) + i);
            float4 b = o.h.first_step ? make_float4(0.f, 0.f, 0.f, 0.f) : lds4(st.buf(1) + i);
            float4 g = lds4(st.buf(2) + i);
            o.one(p.x, b.x, g.x); o.one(p.y, b.y, g.y); o.one(p.z, b.z, g.z); o.one(p.w, b.w, g.w);
            sts4(st.buf(0) + i, p);
            sts4(st.buf(1) + i, b);
            if (BF16 && o.local_out) sts_bf16x4(st.buf(kOutBuf), i, p);
        }
    }
    template <int TILE>
    __device__ __forceinline__ void stores(const tma::Stage<TILE>& st, int64_t off, int len) const {
        tma::bulk_store(o.master + off, st.buf(0), (uint32_t)len * 4);
        tma::bulk_store(o.mom + off, st.buf(1), (uint32_t)len * 4);
        if (o.local_out)
            tma::bulk_store(static_cast<char*>(o.local_out) + off * kEsz, st.buf(kOutBuf), (uint32_t)len * kEsz);
    }
    __device__ __forceinline__ void tail(int64_t i) { o.scalar(i); }
};

// buffers: 0 = p, 1 = g (zeroed in place when zero_grad), 2 = exp_avg, 3 = exp_avg_sq
template <bool BF16>
struct AdamTileOpT {
    static constexpr int kBufs = 4;
    static constexpr bool kHasSum = false;
    static constexpr int kEsz = BF16 ? 2 : 4;
    static constexpr int kGran = 16 / kEsz;
    AdamWOpT<BF16> o;
    __device__ __forceinline__ void prologue() { o.prologue(); }
    __device__ __forceinline__ float acc() const { return 0.0f; }
    // table-driven launches: bytes per element of table column k (p, g, exp_avg, exp_avg_sq)
    static __device__ __forceinline__ int esz(int k) { return k < 2 ? kEsz : 4; }
    __device__ __forceinline__ void rebase(const tma::TileDesc& td) {
        using PG = typename AdamWOpT<BF16>::PG;
        o.p = static_cast<PG*>(td.ptr[0]); o.g = static_cast<PG*>(td.ptr[1]);
        o.m = static_cast<float*>(td.ptr[2]); o.v = static_cast<float*>(td.ptr[3]);
    }
    __device__ __forceinline__ uint32_t load_bytes(int len) const { return (uint32_t)len * (8u + 2u * kEsz); }
    template <int TILE>
    __device__ __forceinline__ void loads(const tma::Stage<TILE>& st, int64_t off, int len, uint64_t* bar) const {
        tma::bulk_load(st.buf(0), o.p + off, (uint32_t)len * kEsz, bar);
        tma::bulk_load(st.buf(1), o.g + off, (uint32_t)len * kEsz, bar);
        tma::bulk_load(st.buf(2), o.m + off, (uint32_t)len * 4, bar);
        tma::bulk_load(st.buf(3), o.v + off, (uint32_t)len * 4, bar);
    }
    template <int TILE>
    __device__ __forceinline__ void compute(const tma::Stage<TILE>& st, int len, int tid, int nthreads) {
        for (int i = tid * 4; i < len; i += nthreads * 4) {
            float4 p = BF16 ? lds_bf16x4(st.buf(0), i) : lds4(st.buf(0) + i);
            float4 g = BF16 ? lds_bf16x4(st.buf(1), i) : lds4(st.buf(1) + i);
            float4 m = lds4(st.buf(2) + i), v = lds4(st.buf(3) + i);
            o.one(p.x, g.x, m.x, v.x); o.one(p.y, g.y, m.y, v.y);
            o.one(p.z, g.z, m.z, v.z); o.one(p.w, g.w, m.w, v.w);
            if (BF16) sts_bf16x4(st.buf(0), i, p); else sts4(st.buf(0) + i, p);
            sts4(st.buf(2) + i, m); sts4(st.buf(3) + i, v);
            if (o.h.zero_grad) {
                if (BF16) sts_bf16x4(st.buf(1), i, make_float4(0.f, 0.f, 0.f, 0.f));
                else sts4(st.buf(1) + i, make_float4(0.f, 0.f, 0.f, 0.f));
            }
        }
    }
    template <int TILE>
    __device__ __forceinline__ void stores(const tma::Stage<TILE>& st, int64_t off, int len) const {
        tma::bulk_store(o.p + off, st.buf(0), (uint32_t)len * kEsz);
        for (int k = 0; k < o.n_extra; ++k) tma::bulk_store(o.p_extra[k] + off, st.buf(0), (uint32_t)len * kEsz);
        tma::bulk_store(o.m + off, st.buf(2), (uint32_t)len * 4);
        tma::bulk_store(o.v + off, st.buf(3), (uint32_t)len * 4);
        if (o.h.zero_grad) tma::bulk_store(o.g + off, st.buf(1), (uint32_t)len * kEsz);
    }
    __device__ __forceinline__ void tail(int64_t i) { o.scalar(i); }
};
using AdamTileOp = AdamTileOpT<false>;

// Gradient average over the R replicas of one worker (BASELINE config 5; an extension -- in the reference
// every rank is a worker, so there is no gradient exchange): out = (g_0 + ... + g_{R-1}) / R in replica
// order, plus the sum of squares of the result.  buffers: one per replica; buffer 0 is reused for the result.
template <int R, bool BF16>
struct GradAvgTileOp {
    static constexpr int kBufs = R;
    static constexpr bool kHasSum = true;
    static constexpr int kEsz = BF16 ? 2 : 4;
    static constexpr int kGran = 16 / kEsz;
    const char* g[R];
    char* out;
    float inv_r_num;   // R as a float (division, like dist.all_reduce(AVG) after a sum)
    float acc_;
    __device__ __forceinline__ void prologue() { acc_ = 0.0f; }
    __device__ __forceinline__ float acc() const { return acc_; }
    __device__ __forceinline__ uint32_t load_bytes(int len) const { return (uint32_t)len * (uint32_t)(R * kEsz); }
    template <int TILE>
    __device__ __forceinline__ void loads(const tma::Stage<TILE>& st, int64_t off, int len, uint64_t* bar) const {
#pragma unroll
        for (int k = 0; k < R; ++k) tma::bulk_load(st.buf(k), g[k] + off * kEsz, (uint32_t)len * kEsz, bar);
    }
    __device__ __forceinline__ float avg1(float s) {
        float a = __fdiv_rn(s, inv_r_num);
        if (BF16) a = __bfloat162float(__float2bfloat16_rn(a));   // the stored value is what gets squared
        acc_ = __fmaf_rn(a, a, acc_);
        return a;
    }
    template <int TILE>
    __device__ __forceinline__ void compute(const tma::Stage<TILE>& st, int len, int tid, int nthreads) {
        for (int i = tid * 4; i < len; i += nthreads * 4) {
            float4 s = BF16 ? lds_bf16x4(st.buf(0), i) : lds4(st.buf(0) + i);
#pragma unroll
            for (int k = 1; k < R; ++k) {
                float4 x = BF16 ? lds_bf16x4(st.buf(k), i) : lds4(st.buf(k) + i);
                s = make_float4(__fadd_rn(s.x, x.x), __fadd_rn(s.y, x.y), __fadd_rn(s.z, x.z), __fadd_rn(s.w, x.w));
            }
            s = make_float4(avg1(s.x), avg1(s.y), avg1(s.z), avg1(s.w));
            if (BF16) sts_bf16x4(st.buf(0), i, s); else sts4(st.buf(0) + i, s);
        }
    }
    template <int TILE>
    __device__ __forceinline__ void stores(const tma::Stage<TILE>& st, int64_t off, int len) const {
        tma::bulk_store(out + off * kEsz, st.buf(0), (uint32_t)len * kEsz);
    }
    __device__ __forceinline__ void tail(int64_t i) {
        float s = Local<BF16>::ld1(g[0], i);
        for (int k = 1; k < R; ++k) s = __fadd_rn(s, Local<BF16>::ld1(g[k], i));
        Local<BF16>::st1(out, i, avg1(s));
    }
};

// ===========================================================================
// Drivers
// ===========================================================================

// Runs `op` over elements [0, count) of its (already rebased) pointers using
// `nthreads` cooperating threads of which this one is `tid`.
template <int U, class Op>
__device__ __forceinline__ void run_range(Op& op, int64_t count, int64_t tid, int64_t nthreads) {
    if (op.aligned()) {
        const int64_t nvec = count >> 2;
        int64_t i = tid;
        // main loop: U independent groups per thread, all loads first
        for (; i + (U - 1) * nthreads < nvec; i += U * nthreads) {
            typename Op::Regs r[U];
#pragma unroll
            for (int u = 0; u < U; ++u) op.load4((i + u * nthreads) << 2, r[u]);
#pragma unroll
            for (int u = 0; u < U; ++u) op.compute4(r[u]);
#pragma unroll
            for (int u = 0; u < U; ++u) op.store4((i + u * nthreads) << 2, r[u]);
        }
        for (; i < nvec; i += nthreads) {
            typename Op::Regs r;
            op.load4(i << 2, r);
            op.compute4(r);
            op.store4(i << 2, r);
        }
        // ragged tail: the last count % 4 elements
        const int64_t t = (nvec << 2) + tid;
        if (t < count) op.scalar(t);
    } else {
        for (int64_t i = tid; i < count; i += nthreads) op.scalar(i);
    }
}

template <int U, int MINB, class Op>
__global__ void __launch_bounds__(kThreads, MINB) flat_kernel(Op op, int64_t n, Scratch* sc, float* sum_out) {
    if constexpr (HasPrologue<Op>::value) op.prologue();
    if constexpr (Op::kHasSum) op.acc = 0.0f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    run_range<U>(op, n, tid, nthreads);
    if constexpr (Op::kHasSum) grid_sum_to(op.acc, sc, sum_out);
}

// One CTA per chunk (grid-stride).  The table is small (hundreds of rows) and
// L2/L1-resident; a binary search per chunk is noise next to 16K elements of
// HBM traffic.
template <int U, int MINB, class Op>
__global__ void __launch_bounds__(kThreads, MINB) table_kernel(Op op, const nd_tensor_desc* __restrict__ tab, int T,
                                                               int64_t total_chunks, Scratch* sc, float* sum_out) {
    if constexpr (HasPrologue<Op>::value) op.prologue();
    if constexpr (Op::kHasSum) op.acc = 0.0f;
    for (int64_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
        int lo = 0, hi = T - 1;
        while (lo < hi) {  // last row with first_chunk <= c
            int mid = (lo + hi + 1) >> 1;
            if (__ldg(&tab[mid].first_chunk) <= c) lo = mid; else hi = mid - 1;
        }
        nd_tensor_desc d = tab[lo];
        const int64_t off = (c - d.first_chunk) * (int64_t)ND_CHUNK;
        int64_t count = d.numel - off;
        if (count > ND_CHUNK) count = ND_CHUNK;
        if (count <= 0) continue;
        Op local_op = op;
        local_op.rebase(d, off);
        run_range<U>(local_op, count, threadIdx.x, blockDim.x);
        if constexpr (Op::kHasSum) op.acc = local_op.acc;
    }
    if constexpr (Op::kHasSum) grid_sum_to(op.acc, sc, sum_out);
}

// ===========================================================================
// Host side
// ===========================================================================

static std::atomic<int> g_sm_count[64];

static int sm_count() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = g_sm_count[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        g_sm_count[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

// grid: every SM gets MINB resident CTAs (one full wave); never more CTAs than work.
static int grid_for(int64_t work_items_per_thread_group, int minb) {
    int64_t full = (int64_t)sm_count() * minb;
    int64_t need = (work_items_per_thread_group + kThreads - 1) / kThreads;
    if (need < 1) need = 1;
    int64_t g = need < full ? need : full;
    if (g > kMaxPartials) g = kMaxPartials;
    return (int)g;
}

template <int U, int MINB, class Op>
static int launch_flat(const Op& op, int64_t n, void* scratch, float* sum_out, cudaStream_t s) {
    if (n == 0 && !Op::kHasSum) return ND_OK;
    int grid = grid_for((n + 3) / 4, MINB);
    flat_kernel<U, MINB, Op><<<grid, kThreads, 0, s>>>(op, n, static_cast<Scratch*>(scratch), sum_out);
    return check_launch("flat_kernel");
}

template <int U, int MINB, class Op>
static int launch_table(const Op& op, const nd_tensor_desc* tab, int T, int64_t chunks, void* scratch,
                        float* sum_out, cudaStream_t s) {
    if ((T == 0 || chunks == 0) && !Op::kHasSum) return ND_OK;
    int64_t full = (int64_t)sm_count() * MINB;
    int grid = (int)(chunks < full ? (chunks > 0 ? chunks : 1) : full);
    if (grid > kMaxPartials) grid = kMaxPartials;
    table_kernel<U, MINB, Op><<<grid, kThreads, 0, s>>>(op, tab, T, chunks, static_cast<Scratch*>(scratch), sum_out);
    return check_launch("table_kernel");
}


// TMA-staged launch: TILE fp32 slots per buffer, STAGES stages, THREADS threads, MINB CTAs per SM.
template <int TILE, int STAGES, int THREADS, int MINB, class TileOp, bool PEER = false>
static int launch_tile(const TileOp& op, int64_t n, void* scratch, float* sum_out, cudaStream_t s,
                       const tma::PeerSync* peer = nullptr) {
    if (n == 0 && !TileOp::kHasSum && !PEER) return ND_OK;   // a rendezvous is kept even when there is no data
    auto kernel = tma::tile_kernel<TILE, STAGES, THREADS, MINB, TileOp, PEER>;
    constexpr size_t smem = (size_t)STAGES * TileOp::kBufs * TILE * sizeof(float);
    static_assert(smem <= 227 * 1024, "stage set exceeds the 227 KB of shared memory a CTA can have");
    static std::atomic<bool> configured[64];   // per device; two threads racing here both set the same attribute
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (!configured[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return set_error(ND_ECUDA, "cudaFuncSetAttribute(smem=%zu): %s", smem, cudaGetErrorString(e));
        configured[dev].store(true, std::memory_order_release);
    }
    const int64_t ntiles = ((n & ~(int64_t)(TileOp::kGran - 1)) + TILE - 1) / TILE;
    int64_t full = (int64_t)sm_count() * MINB;
    int64_t g = ntiles < full ? (ntiles > 0 ? ntiles : 1) : full;
    if (g > kMaxPartials) g = kMaxPartials;
    kernel<<<(int)g, THREADS, smem, s>>>(op, n, static_cast<Scratch*>(scratch), sum_out,
                                         peer ? *peer : tma::PeerSync{});
    return check_launch("tile_kernel");
}

// Table-driven TMA-staged launch (tma::tile_table_kernel): one tile = TILE elements of one tensor.
template <int TILE, int STAGES, int THREADS, int MINB, class TileOp>
static int launch_tile_table(const TileOp& op, const nd_tensor_desc* tab, int T, int64_t chunks, void* scratch,
                             float* sum_out, cudaStream_t s) {
    if ((T == 0 || chunks == 0) && !TileOp::kHasSum) return ND_OK;
    auto kernel = tma::tile_table_kernel<TILE, STAGES, THREADS, MINB, TileOp>;
    constexpr size_t smem = (size_t)STAGES * TileOp::kBufs * TILE * sizeof(float);
    static_assert(smem <= 200 * 1024, "stage set too large");
    static std::atomic<bool> configured[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (!configured[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return set_error(ND_ECUDA, "cudaFuncSetAttribute(smem=%zu): %s", smem, cudaGetErrorString(e));
        configured[dev].store(true, std::memory_order_release);
    }
    const int64_t ntiles = chunks * (ND_CHUNK / TILE);
    const int64_t full = (int64_t)sm_count() * MINB;
    int64_t g = ntiles < full ? (ntiles > 0 ? ntiles : 1) : full;
    if (g > kMaxPartials) g = kMaxPartials;
    kernel<<<(int)g, THREADS, smem, s>>>(op, tab, T, chunks, static_cast<Scratch*>(scratch), sum_out);
    return check_launch("tile_table_kernel");
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static OuterHyper make_outer_hyper(float lr, float mu, float dampening, float wd, int nesterov, int first) {
    OuterHyper h;
    h.neg_lr = -lr; h.mu = mu; h.one_minus_damp = 1.0f - dampening; h.wd = wd;
    h.nesterov = nesterov; h.first_step = first;
    return h;
}

static AdamHyper make_adam_hyper(double lr, double b1, double b2, double eps, double wd, int64_t step,
                                 float max_norm, int zero_grad) {
    // mirrors the host arithmetic of torch:optim/adam.py:692,773-781 (double, then cast)
    AdamHyper h;
    h.decay = (float)(1.0 - lr * wd);
    h.w1 = (float)(1.0 - b1);
    h.b2 = (float)b2;
    h.w2 = (float)(1.0 - b2);
    const double bc1 = 1.0 - std::pow(b1, (double)step);
    const double bc2 = 1.0 - std::pow(b2, (double)step);
    h.step_size = (float)((lr / bc1) * -1.0);
    h.bc2_sqrt = (float)std::sqrt(bc2);
    h.eps = (float)eps;
    h.max_norm = max_norm;
    h.has_decay = wd != 0.0;
    h.zero_grad = zero_grad;
    return h;
}

// Tuning knobs (tools/tune.cu sweeps them on the box; profiles/ holds the sweep).
// Register-staged path: U groups in flight per thread / B resident CTAs per SM.  Lower occupancy
// with fewer, fatter threads streams better here than 2048 threads/SM.
constexpr int U_PG = 1, B_PG = 4;
constexpr int U_SHARD = 1, B_SHARD = 2;
constexpr int U_SUMSQ = 4, B_SUMSQ = 4;
constexpr int U_ADAM = 1, B_ADAM = 2;
// TMA-staged path: fp32 slots per buffer per stage / stages / threads / CTAs per SM.
#define ND_TILE_FUSED1 4096, 2, 512, 2
#define ND_TILE_FUSED2 4096, 2, 512, 1
#define ND_TILE_FUSED4 1024, 3, 256, 3
#define ND_TILE_FUSED8 1024, 3, 256, 1
#define ND_TILE_SHARD 4096, 2, 512, 2
#define ND_TILE_SHARD_BF16 2048, 2, 512, 2   // four buffers per stage: 64 KB per CTA
#define ND_TILE_ADAM 2048, 2, 512, 2

template <int W, bool BF16, bool SUM>
static int launch_fused(float* master, float* mom, void* const* locals, int w_rt, void* const* writes, int n_writes,
                        int64_t n, const OuterHyper& h, float* sumsq_out, void* scratch, cudaStream_t s,
                        const tma::PeerSync* peer = nullptr, bool* peer_done = nullptr) {
    OuterFusedOp<W, BF16, SUM> op;
    op.master = master; op.mom = mom; op.w_rt = w_rt; op.world = (float)w_rt; op.h = h; op.acc = 0.f;
    op.n_writes = n_writes;
    bool al = aligned16(master) && aligned16(mom);
    for (int k = 0; k < ND_MAX_WORKERS; ++k) {
        op.locals.p[k] = k < w_rt ? locals[k] : nullptr;
        op.writes.p[k] = k < n_writes ? writes[k] : nullptr;
        al = al && aligned16(op.locals.p[k]) && aligned16(op.writes.p[k]);
    }
    if constexpr (W > 0) {
        if (al && peer) {   // the same kernels with the two cross-GPU rendezvous folded in (nd_outer_sync_p2p)
            FusedTileOp<W, BF16, SUM> t{op};
            *peer_done = true;
            if constexpr (W == 1) return launch_tile<ND_TILE_FUSED1, FusedTileOp<W, BF16, SUM>, true>(t, n, scratch, sumsq_out, s, peer);
            else if constexpr (W == 2) return launch_tile<ND_TILE_FUSED2, FusedTileOp<W, BF16, SUM>, true>(t, n, scratch, sumsq_out, s, peer);
            else if constexpr (W <= 4) return launch_tile<ND_TILE_FUSED4, FusedTileOp<W, BF16, SUM>, true>(t, n, scratch, sumsq_out, s, peer);
            else return launch_tile<ND_TILE_FUSED8, FusedTileOp<W, BF16, SUM>, true>(t, n, scratch, sumsq_out, s, peer);
        }
        if (al) {  // TMA-staged path: every bulk copy needs 16-byte aligned addresses
            FusedTileOp<W, BF16, SUM> t{op};
            if constexpr (W == 1) return launch_tile<ND_TILE_FUSED1>(t, n, scratch, sumsq_out, s);
            else if constexpr (W == 2) return launch_tile<ND_TILE_FUSED2>(t, n, scratch, sumsq_out, s);
            else if constexpr (W <= 4) return launch_tile<ND_TILE_FUSED4>(t, n, scratch, sumsq_out, s);
            else {
                // ND_P2P8_VARIANT: alternative stage shapes for 5..8 worker copies (tuning knob)
                static int v = -1;
                if (v < 0) { const char* e = getenv("ND_P2P8_VARIANT"); v = e ? atoi(e) : 0; }
                if (v == 1) return launch_tile<2048, 2, 256, 1>(t, n, scratch, sumsq_out, s);
                if (v == 2) return launch_tile<1024, 4, 512, 1>(t, n, scratch, sumsq_out, s);
                if (v == 3) return launch_tile<512, 3, 128, 3>(t, n, scratch, sumsq_out, s);
                return launch_tile<ND_TILE_FUSED8>(t, n, scratch, sumsq_out, s);
            }
        }
    }
    // register-staged path: unaligned buffers (scalar accesses) and run-time worker counts
    if (peer) return ND_OK;   // not launched: *peer_done stays false and the caller uses separate barrier launches
    return launch_flat<1, 2>(op, n, scratch, sumsq_out, s);
}

template <bool BF16, bool SUM>
static int dispatch_fused(int W, float* master, float* mom, void* const* locals, void* const* writes, int n_writes,
                          int64_t n, const OuterHyper& h, float* sumsq_out, void* scratch, cudaStream_t s,
                          const tma::PeerSync* peer = nullptr, bool* peer_done = nullptr) {
#define ND_CASE(w) \
    case w: return launch_fused<w, BF16, SUM>(master, mom, locals, W, writes, n_writes, n, h, sumsq_out, scratch, s, \
                                              peer, peer_done)
    switch (W) {
        ND_CASE(1); ND_CASE(2); ND_CASE(3); ND_CASE(4); ND_CASE(8);
        default: return launch_fused<0, BF16, SUM>(master, mom, locals, W, writes, n_writes, n, h, sumsq_out, scratch, s,
                                                   peer, peer_done);
    }
#undef ND_CASE
}

// nd_outer_fused_rw with the two cross-GPU rendezvous of nd_outer_sync_p2p folded into the kernel (internal to the
// library: nd_peer.cu).  *launched = false when the buffers do not qualify for the TMA-staged kernel; nothing was
// enqueued then and the caller falls back to barrier launches around nd_outer_fused_rw.
int outer_fused_rw_peer(float* master, float* mom, void* const* reads, int W, void* const* writes, int n_writes,
                        int local_dtype, int64_t n, float lr, float mu, float dampening, float wd, int nesterov,
                        int first_step, float* sumsq_out, void* scratch, cudaStream_t s, const tma::PeerSync& peer,
                        bool* launched) {
    *launched = false;
    OuterHyper h = make_outer_hyper(lr, mu, dampening, wd, nesterov, first_step);
    if (local_dtype == ND_BF16)
        return sumsq_out ? dispatch_fused<true, true>(W, master, mom, reads, writes, n_writes, n, h, sumsq_out, scratch, s, &peer, launched)
                         : dispatch_fused<true, false>(W, master, mom, reads, writes, n_writes, n, h, nullptr, nullptr, s, &peer, launched);
    return sumsq_out ? dispatch_fused<false, true>(W, master, mom, reads, writes, n_writes, n, h, sumsq_out, scratch, s, &peer, launched)
                     : dispatch_fused<false, false>(W, master, mom, reads, writes, n_writes, n, h, nullptr, nullptr, s, &peer, launched);
}

}  // namespace nd

using namespace nd;

// ===========================================================================
// C ABI
// ===========================================================================

#define ND_REQUIRE(cond, msg)                       \
    do {                                            \
        if (!(cond)) return set_error(ND_EINVAL, "%s: %s", __func__, msg); \
    } while (0)

extern "C" {

int nd_version(void) { return 200; }

const char* nd_last_error(void) { return nd::last_error(); }

size_t nd_scratch_bytes(void) { return sizeof(Scratch); }

int nd_pseudo_grad(const float* master, const void* local, int local_dtype, float* pg, int64_t n,
                   nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(n == 0 || (master && local && pg), "null pointer");
    ND_REQUIRE(local_dtype == ND_F32 || local_dtype == ND_BF16, "unknown local_dtype");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (local_dtype == ND_BF16) {
        PseudoGradOp<true> op{master, local, pg};
        return launch_flat<U_PG, B_PG>(op, n, nullptr, nullptr, s);
    }
    PseudoGradOp<false> op{master, local, pg};
    return launch_flat<U_PG, B_PG>(op, n, nullptr, nullptr, s);
}

int nd_pseudo_grad_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks, int local_dtype,
                      nd_stream_t stream) {
    ND_REQUIRE(n_tensors >= 0 && total_chunks >= 0, "negative count");
    ND_REQUIRE(n_tensors == 0 || table_dev, "null table");
    ND_REQUIRE(local_dtype == ND_F32 || local_dtype == ND_BF16, "unknown local_dtype");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (local_dtype == ND_BF16) {
        PseudoGradOp<true> op{nullptr, nullptr, nullptr};
        return launch_table<U_PG, B_PG>(op, table_dev, n_tensors, total_chunks, nullptr, nullptr, s);
    }
    PseudoGradOp<false> op{nullptr, nullptr, nullptr};
    return launch_table<U_PG, B_PG>(op, table_dev, n_tensors, total_chunks, nullptr, nullptr, s);
}

int nd_outer_nesterov(float* master, float* mom, const float* pg_sum, void* local_out, int local_dtype, int64_t n,
                      float world, float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                      float* sumsq_out, void* scratch, nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(n == 0 || (master && mom && pg_sum), "null pointer");
    ND_REQUIRE(local_dtype == ND_F32 || local_dtype == ND_BF16, "unknown local_dtype");
    ND_REQUIRE(world >= 1.0f, "world < 1");
    ND_REQUIRE(!sumsq_out || scratch, "sumsq_out needs scratch");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    OuterHyper h = make_outer_hyper(lr, mu, dampening, wd, nesterov, first_step);
#define ND_SHARD(BF, SM)                                                                     \
    do {                                                                                     \
        OuterShardOp<BF, SM> op;                                                             \
        op.master = master; op.mom = mom; op.pg_sum = pg_sum; op.local_out = local_out;      \
        op.world = world; op.h = h; op.acc = 0.f;                                            \
        if (aligned16(master) && aligned16(mom) && aligned16(pg_sum) && aligned16(local_out)) { \
            ShardTileOp<BF, SM> t{op};                                                       \
            if constexpr (BF) return launch_tile<ND_TILE_SHARD_BF16>(t, n, scratch, sumsq_out, s); \
            else return launch_tile<ND_TILE_SHARD>(t, n, scratch, sumsq_out, s);             \
        }                                                                                    \
        return launch_flat<U_SHARD, B_SHARD>(op, n, scratch, sumsq_out, s);                  \
    } while (0)
    if (local_dtype == ND_BF16) { if (sumsq_out) ND_SHARD(true, true); else ND_SHARD(true, false); }
    if (sumsq_out) ND_SHARD(false, true); else ND_SHARD(false, false);
#undef ND_SHARD
}

int nd_outer_fused_rw(float* master, float* mom, void* const* reads_host, int W, void* const* writes_host,
                      int n_writes, int local_dtype, int64_t n, float lr, float mu, float dampening, float wd,
                      int nesterov, int first_step, float* sumsq_out, void* scratch, nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(W >= 1 && W <= ND_MAX_WORKERS, "W out of range [1, ND_MAX_WORKERS]");
    ND_REQUIRE(n_writes >= 0 && n_writes <= ND_MAX_WORKERS, "n_writes out of range [0, ND_MAX_WORKERS]");
    ND_REQUIRE(reads_host && (n_writes == 0 || writes_host), "null pointer array");
    ND_REQUIRE(n == 0 || (master && mom), "null pointer");
    for (int k = 0; k < W; ++k) ND_REQUIRE(n == 0 || reads_host[k], "null read pointer");
    for (int k = 0; k < n_writes; ++k) ND_REQUIRE(n == 0 || writes_host[k], "null write pointer");
    ND_REQUIRE(local_dtype == ND_F32 || local_dtype == ND_BF16, "unknown local_dtype");
    ND_REQUIRE(!sumsq_out || scratch, "sumsq_out needs scratch");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    OuterHyper h = make_outer_hyper(lr, mu, dampening, wd, nesterov, first_step);
    if (local_dtype == ND_BF16) {
        return sumsq_out ? dispatch_fused<true, true>(W, master, mom, reads_host, writes_host, n_writes, n, h,
                                                      sumsq_out, scratch, s)
                         : dispatch_fused<true, false>(W, master, mom, reads_host, writes_host, n_writes, n, h,
                                                       nullptr, nullptr, s);
    }
    return sumsq_out ? dispatch_fused<false, true>(W, master, mom, reads_host, writes_host, n_writes, n, h,
                                                   sumsq_out, scratch, s)
                     : dispatch_fused<false, false>(W, master, mom, reads_host, writes_host, n_writes, n, h,
                                                    nullptr, nullptr, s);
}

int nd_outer_fused(float* master, float* mom, void* const* locals_host, int W, int local_dtype, int64_t n, float lr,
                   float mu, float dampening, float wd, int nesterov, int first_step, float* sumsq_out,
                   void* scratch, nd_stream_t stream) {
    // every copy that is read is also rewound (diloco.py:50)
    return nd_outer_fused_rw(master, mom, locals_host, W, locals_host, W, local_dtype, n, lr, mu, dampening, wd,
                             nesterov, first_step, sumsq_out, scratch, stream);
}

int nd_outer_fused_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks, int local_dtype,
                      float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                      nd_stream_t stream) {
    ND_REQUIRE(n_tensors >= 0 && total_chunks >= 0, "negative count");
    ND_REQUIRE(n_tensors == 0 || table_dev, "null table");
    ND_REQUIRE(local_dtype == ND_F32 || local_dtype == ND_BF16, "unknown local_dtype");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    OuterHyper h = make_outer_hyper(lr, mu, dampening, wd, nesterov, first_step);
    if (local_dtype == ND_BF16) {
        OuterFusedOp<1, true, false> op;
        op.master = nullptr; op.mom = nullptr; op.w_rt = 1; op.world = 1.f; op.h = h; op.acc = 0.f; op.n_writes = 1;
        return launch_table<2, 3>(op, table_dev, n_tensors, total_chunks, nullptr, nullptr, s);
    }
    OuterFusedOp<1, false, false> op;
    op.master = nullptr; op.mom = nullptr; op.w_rt = 1; op.world = 1.f; op.h = h; op.acc = 0.f; op.n_writes = 1;
    return launch_table<2, 3>(op, table_dev, n_tensors, total_chunks, nullptr, nullptr, s);
}

int nd_sumsq(const void* g, int dtype, int64_t n, float* out, void* scratch, nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE((n == 0 || g) && out && scratch, "null pointer");
    ND_REQUIRE(dtype == ND_F32 || dtype == ND_BF16, "unknown dtype");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == ND_BF16) {
        SumsqOp<true> op{g, 0.f};
        return launch_flat<U_SUMSQ, B_SUMSQ>(op, n, scratch, out, s);
    }
    SumsqOp<false> op{g, 0.f};
    return launch_flat<U_SUMSQ, B_SUMSQ>(op, n, scratch, out, s);
}

// ---- nd_sumsq_static: the sum of squares with no caller scratch and no allocation ------------
//
// Some users of this library can hand us their inputs and outputs but cannot give us scratch:
//   * a trainer or server that has already given all of HBM to its own pool (torch's caching
//     allocator under a hard memory cap, an inference server's fixed arena): a stray cudaMalloc
//     from a helper library either fails or breaks their memory plan;
//   * a latency-critical loop, where the device-wide synchronisation hidden inside
//     cudaMalloc/cudaFree would stall every other stream of the process;
//   * a thin FFI caller (ctypes, cgo, a DLPack consumer) that owns no allocator at all and only
//     forwards tensors somebody else created.
// For them the per-block partial sums and the arrival counter live in the array below, which is
// part of the library image: every GPU that loads the library carries these bytes until the
// process exits.  All calls on a device share it, so two calls that overlap on the GPU (different
// streams without an event between them, or two host threads) race on it -- order them.
static __device__ __align__(256) unsigned char g_sumsq_static_scratch[ND_SUMSQ_STATIC_SCRATCH_BYTES];
static_assert(sizeof(Scratch) <= ND_SUMSQ_STATIC_SCRATCH_BYTES, "static scratch too small for nd::Scratch");

// In host code `g_sumsq_static_scratch` is only a handle for the runtime, not a device address:
// the address a kernel can use comes from cudaGetSymbolAddress, looked up once per device with
// that device current.  The cache is atomic so that two host threads making their first call on a
// device at the same time is not a data race (both would store the same address).  It does not
// survive cudaDeviceReset: a process that resets a device must not call this entry point afterwards.
static std::atomic<void*> g_sumsq_static_ptr[64];

int nd_sumsq_static(const void* g, int dtype, int64_t n, float* out, nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    if (n > (int64_t)ND_SUMSQ_STATIC_MAX_ELEMS)
        return set_error(ND_EINVAL, "nd_sumsq_static: n = %lld exceeds ND_SUMSQ_STATIC_MAX_ELEMS = %lld; "
                                    "use nd_sumsq with caller-provided scratch",
                         (long long)n, (long long)ND_SUMSQ_STATIC_MAX_ELEMS);
    ND_REQUIRE((n == 0 || g) && out, "null pointer");
    ND_REQUIRE(dtype == ND_F32 || dtype == ND_BF16, "unknown dtype");
    int dev = 0;
    ND_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return set_error(ND_EINVAL, "nd_sumsq_static: device index %d out of range", dev);
    void* sptr = g_sumsq_static_ptr[dev].load(std::memory_order_acquire);
    if (!sptr) {  // set-up: first call on this device
        cudaError_t e = cudaGetSymbolAddress(&sptr, g_sumsq_static_scratch);
        if (e != cudaSuccess || !sptr)
            return set_error(ND_ECUDA, "nd_sumsq_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        g_sumsq_static_ptr[dev].store(sptr, std::memory_order_release);
    }
    Scratch* sc = static_cast<Scratch*>(sptr);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    // the array keeps whatever the previous call (or a call that died half-way) left behind: the
    // arrival counter is the one field that must be zero when the kernel starts
    ND_CUDA(cudaMemsetAsync(&sc->counter, 0, sizeof(unsigned), s));
    if (dtype == ND_BF16) {
        SumsqOp<true> op{g, 0.f};
        return launch_flat<U_SUMSQ, B_SUMSQ>(op, n, sc, out, s);
    }
    SumsqOp<false> op{g, 0.f};
    return launch_flat<U_SUMSQ, B_SUMSQ>(op, n, sc, out, s);
}

int nd_sumsq_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks, int dtype, float* out,
                void* scratch, nd_stream_t stream) {
    ND_REQUIRE(n_tensors >= 0 && total_chunks >= 0, "negative count");
    ND_REQUIRE((n_tensors == 0 || table_dev) && out && scratch, "null pointer");
    ND_REQUIRE(dtype == ND_F32 || dtype == ND_BF16, "unknown dtype");
    if (dtype == ND_BF16) {
        SumsqOp<true> op{nullptr, 0.f};
        return launch_table<U_SUMSQ, B_SUMSQ>(op, table_dev, n_tensors, total_chunks, scratch, out,
                                              reinterpret_cast<cudaStream_t>(stream));
    }
    SumsqOp<false> op{nullptr, 0.f};
    return launch_table<U_SUMSQ, B_SUMSQ>(op, table_dev, n_tensors, total_chunks, scratch, out,
                                          reinterpret_cast<cudaStream_t>(stream));
}

int nd_adamw(float* p, float* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2, double eps,
             double weight_decay, int64_t step, const float* sumsq_dev, float max_norm, int zero_grad,
             nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(n == 0 || (p && g && m && v), "null pointer");
    ND_REQUIRE(step >= 1, "step must be >= 1 (1-based)");
    AdamWOp op;
    op.p = p; op.g = g; op.m = m; op.v = v; op.sumsq = sumsq_dev; op.coef = 1.f;
    op.h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, step, max_norm, zero_grad);
    if (aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v)) {
        AdamTileOp t{op};
        return launch_tile<ND_TILE_ADAM>(t, n, nullptr, nullptr, reinterpret_cast<cudaStream_t>(stream));
    }
    return launch_flat<U_ADAM, B_ADAM>(op, n, nullptr, nullptr, reinterpret_cast<cudaStream_t>(stream));
}

int nd_adamw_bf16(void* p, void* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2, double eps,
                  double weight_decay, int64_t step, const float* sumsq_dev, float max_norm, int zero_grad,
                  nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(n == 0 || (p && g && m && v), "null pointer");
    ND_REQUIRE(step >= 1, "step must be >= 1 (1-based)");
    AdamWOpT<true> op;
    op.p = static_cast<__nv_bfloat16*>(p); op.g = static_cast<__nv_bfloat16*>(g); op.m = m; op.v = v;
    op.sumsq = sumsq_dev; op.coef = 1.f;
    op.h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, step, max_norm, zero_grad);
    if (aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v)) {
        AdamTileOpT<true> t{op};
        return launch_tile<ND_TILE_ADAM>(t, n, nullptr, nullptr, reinterpret_cast<cudaStream_t>(stream));
    }
    return launch_flat<U_ADAM, B_ADAM>(op, n, nullptr, nullptr, reinterpret_cast<cudaStream_t>(stream));
}

int nd_adamw_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks, int dtype, double lr,
                double beta1, double beta2, double eps, double weight_decay, int64_t step, const float* sumsq_dev,
                float max_norm, int zero_grad, nd_stream_t stream) {
    ND_REQUIRE(n_tensors >= 0 && total_chunks >= 0, "negative count");
    ND_REQUIRE(n_tensors == 0 || table_dev, "null table");
    ND_REQUIRE(dtype == ND_F32 || dtype == ND_BF16, "unknown dtype");
    ND_REQUIRE(step >= 1, "step must be >= 1 (1-based)");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    // TMA-staged, table-driven: rows with unaligned pointers fall back to plain accesses inside the kernel
    if (dtype == ND_BF16) {
        AdamWOpT<true> op;
        op.p = op.g = nullptr; op.m = op.v = nullptr; op.sumsq = sumsq_dev; op.coef = 1.f;
        op.h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, step, max_norm, zero_grad);
        AdamTileOpT<true> t{op};
        return launch_tile_table<ND_TILE_ADAM>(t, table_dev, n_tensors, total_chunks, nullptr, nullptr, s);
    }
    AdamWOp op;
    op.p = op.g = op.m = op.v = nullptr; op.sumsq = sumsq_dev; op.coef = 1.f;
    op.h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, step, max_norm, zero_grad);
    AdamTileOp t{op};
    return launch_tile_table<ND_TILE_ADAM>(t, table_dev, n_tensors, total_chunks, nullptr, nullptr, s);
}

// ---- workers that span several GPUs: the per-inner-step gradient exchange (BASELINE config 5) ----------
#define ND_TILE_GAVG 2048, 3, 512, 2

}  // extern "C"

template <int R, bool BF16>
static int launch_grad_avg(void* const* grads, void* out, int64_t n, float* sumsq_out, void* scratch, cudaStream_t s) {
    GradAvgTileOp<R, BF16> t;
    for (int k = 0; k < R; ++k) t.g[k] = static_cast<const char*>(grads[k]);
    t.out = static_cast<char*>(out);
    t.inv_r_num = (float)R;
    t.acc_ = 0.f;
    return launch_tile<ND_TILE_GAVG>(t, n, scratch, sumsq_out, s);
}

extern "C" {

int nd_grad_avg(void* const* grads_host, int R, void* g_out, int dtype, int64_t n, float* sumsq_out, void* scratch,
                nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(R >= 1 && R <= ND_MAX_REPLICAS, "R out of range [1, ND_MAX_REPLICAS]");
    ND_REQUIRE(grads_host && sumsq_out && scratch && (n == 0 || g_out), "null pointer");
    ND_REQUIRE(dtype == ND_F32 || dtype == ND_BF16, "unknown dtype");
    bool al = aligned16(g_out);
    for (int k = 0; k < R; ++k) {
        ND_REQUIRE(n == 0 || grads_host[k], "null gradient pointer");
        al = al && aligned16(grads_host[k]);
    }
    ND_REQUIRE(al, "gradient buffers must be 16-byte aligned (bulk copies)");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define ND_GA(r) case r: return dtype == ND_BF16 ? launch_grad_avg<r, true>(grads_host, g_out, n, sumsq_out, scratch, s) \
                                                   : launch_grad_avg<r, false>(grads_host, g_out, n, sumsq_out, scratch, s)
    switch (R) {
        ND_GA(1); ND_GA(2); ND_GA(3); ND_GA(4); ND_GA(8);
        default: return set_error(ND_EINVAL, "nd_grad_avg: R must be 1, 2, 3, 4 or 8");
    }
#undef ND_GA
}

int nd_adamw_rw(void* p, void* const* p_writes_host, int n_writes, void* g, float* m, float* v, int dtype, int64_t n,
                double lr, double beta1, double beta2, double eps, double weight_decay, int64_t step,
                const float* const* sumsq_parts_host, int n_parts, float max_norm, int zero_grad, nd_stream_t stream) {
    ND_REQUIRE(n >= 0, "n < 0");
    ND_REQUIRE(n == 0 || (p && g && m && v), "null pointer");
    ND_REQUIRE(dtype == ND_F32 || dtype == ND_BF16, "unknown dtype");
    ND_REQUIRE(step >= 1, "step must be >= 1 (1-based)");
    ND_REQUIRE(n_writes >= 0 && n_writes <= ND_MAX_REPLICAS && (n_writes == 0 || p_writes_host), "bad write set");
    ND_REQUIRE(n_parts >= 0 && n_parts <= ND_MAX_REPLICAS && (n_parts == 0 || sumsq_parts_host), "bad partial sums");
    bool al = aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v);
    for (int k = 0; k < n_writes; ++k) {
        ND_REQUIRE(p_writes_host[k], "null write pointer");
        al = al && aligned16(p_writes_host[k]);
    }
    ND_REQUIRE(al, "buffers must be 16-byte aligned (bulk copies)");
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    auto fill = [&](auto& op) {
        using PG = typename std::remove_reference<decltype(op)>::type::PG;
        op.p = static_cast<PG*>(p); op.g = static_cast<PG*>(g); op.m = m; op.v = v;
        op.sumsq = nullptr; op.coef = 1.f;
        op.h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, step, max_norm, zero_grad);
        op.n_parts = n_parts;
        for (int k = 0; k < n_parts; ++k) op.sumsq_parts[k] = sumsq_parts_host[k];
        op.n_extra = 0;
        for (int k = 0; k < n_writes; ++k)      // the copy that is read is written by the regular store
            if (p_writes_host[k] != p) op.p_extra[op.n_extra++] = static_cast<PG*>(p_writes_host[k]);
    };
    if (dtype == ND_BF16) {
        AdamWOpT<true> op;
        fill(op);
        AdamTileOpT<true> t{op};
        return launch_tile<ND_TILE_ADAM>(t, n, nullptr, nullptr, s);
    }
    AdamWOp op;
    fill(op);
    AdamTileOp t{op};
    return launch_tile<ND_TILE_ADAM>(t, n, nullptr, nullptr, s);
}

}  // extern "C"

