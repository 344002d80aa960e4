/*
 * nd_oracle.c -- CPU restatement of the reference's DiLoCo outer-sync hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in nanodiloco_b200/ may import, link or
 * execute this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or the
 * timed CPU baseline -- never as the product path.
 *
 * Parity status: the reference (SJCaldwell/NanoDiloco) has no tests, fixtures
 * or golden vectors, so nothing of its own pins this oracle.  It is pinned
 * instead against outputs of the reference's own `Diloco` class run in the
 * build container on gloo/CPU (tests/golden/make_golden.py, fixtures committed
 * under tests/golden/), and against torch's CPU optimizers, which is where the
 * reference's arithmetic actually lives (torch is a third-party dependency of
 * the reference: pyproject.toml:12 `torch>=2.8.0`, uv.lock pins 2.8.0; this
 * image has 2.11.0).
 *
 * Each function cites the reference line it restates.  Paths are relative to
 * /root/reference unless prefixed torch: (the installed torch).
 *
 * Rounding recipe (what makes it bit-exact against torch CPU and CUDA):
 * torch's `a.add_(b, alpha=s)` evaluates a + s*b as ONE fused multiply-add in
 * fp32, while `buf.mul_(mu)` is its own rounding.  fmaf() below is therefore
 * load-bearing; compile with -ffp-contract=off so that gcc neither adds nor
 * removes contractions.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>

#ifdef _OPENMP
#include <omp.h>
#define ND_PAR _Pragma("omp parallel for schedule(static)")
#else
#define ND_PAR
#endif

int ndo_version(void) { return 1; }

int ndo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ndo_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* nanodiloco/diloco/diloco.py:48
 *   replica_param.grad = last_sync_param_on_device - replica_param.data
 * One fp32 subtraction per element, exact specification. */
void ndo_pseudo_grad(const float* master, const float* local, float* pg, int64_t n) {
    ND_PAR
    for (int64_t i = 0; i < n; ++i) pg[i] = master[i] - local[i];
}

/* nanodiloco/diloco/diloco.py:49
 *   dist.all_reduce(tensor=replica_param.grad, op=dist.ReduceOp.AVG)
 * gloo: sum over workers, then divide by W.  The sum is taken here in worker
 * order 0..W-1; for W<=2 any order gives the same bits, for W>=4 gloo's ring
 * order (and NCCL's) may differ by an ulp of the sum, which is why W>=4 parity
 * is stated with a tolerance.  pg_w[w] points at worker w's pseudo-gradient. */
void ndo_average(const float* const* pg_w, int W, float* out, int64_t n) {
    const float fw = (float)W;
    ND_PAR
    for (int64_t i = 0; i < n; ++i) {
        float s = pg_w[0][i];
        for (int w = 1; w < W; ++w) s = s + pg_w[w][i];
        out[i] = s / fw;
    }
}

/* nanodiloco/diloco/diloco.py:52  self.outer_optimizer.step()
 * with the optimizer built at nanodiloco/main.py:101
 *   torch.optim.SGD(model.parameters(), lr=outer_lr, momentum=0.9, nesterov=True)
 * Arithmetic: torch:optim/sgd.py:343-380 (_single_tensor_sgd; the CUDA foreach
 * form :425-473 computes the same sequence):
 *   g   = grad (+ wd*p when weight_decay != 0)         sgd.py:345-356
 *   buf = clone(g)                      on first step  sgd.py:361-363
 *   buf = buf*mu ; buf = buf + (1-damp)*g  otherwise   sgd.py:365
 *   g   = g + mu*buf          (nesterov) sgd.py:367-368 ; else g = buf
 *   p   = p + (-lr)*g                                  sgd.py:380
 * Every `alpha=` form is one FMA.  g is NOT written back (diloco.py:53 drops
 * the grads right afterwards). */
void ndo_sgd_step(float* p, float* buf, const float* grad, int64_t n,
                  float lr, float mu, float dampening, float wd,
                  int nesterov, int first_step) {
    const float neg_lr = -lr;
    const float one_minus_damp = 1.0f - dampening;
    ND_PAR
    for (int64_t i = 0; i < n; ++i) {
        float g = grad[i];
        if (wd != 0.0f) g = fmaf(wd, p[i], g);
        float d = g;
        if (mu != 0.0f) {
            float b;
            if (first_step) {
                b = g;
            } else {
                b = buf[i] * mu;
                b = fmaf(one_minus_damp, g, b);
            }
            buf[i] = b;
            d = nesterov ? fmaf(mu, b, g) : b;
        }
        p[i] = fmaf(neg_lr, d, p[i]);
    }
}

/* The whole outer step for one worker's view, diloco.py:46-54, W workers held
 * in one address space (local_w[w] = worker w's params after H inner steps).
 * master is theta_sync (offloaded_last_sync_parameters, diloco.py:25,54);
 * on return master holds the new theta_sync and every local_w[w] equals it
 * (diloco.py:50 rebinds param.data to the last-sync copy before the step, so
 * the optimizer updates theta_sync and all workers end identical). */
void ndo_outer_step(float* master, float* const* local_w, int W, float* buf,
                    float* scratch_pg /* n floats */, int64_t n,
                    float lr, float mu, float dampening, float wd,
                    int nesterov, int first_step) {
    const float fw = (float)W;
    ND_PAR
    for (int64_t i = 0; i < n; ++i) {
        float s = master[i] - local_w[0][i];
        for (int w = 1; w < W; ++w) s = s + (master[i] - local_w[w][i]);
        scratch_pg[i] = s / fw;
    }
    ndo_sgd_step(master, buf, scratch_pg, n, lr, mu, dampening, wd, nesterov, first_step);
    for (int w = 0; w < W; ++w) {
        float* dst = local_w[w];
        ND_PAR
        for (int64_t i = 0; i < n; ++i) dst[i] = master[i];
    }
}

/* nanodiloco/diloco/diloco.py:57
 *   torch.nn.utils.clip_grad_norm_(self.model.parameters(), max_norm=1.0)
 * torch:nn/utils/clip_grad.py:93-108: total_norm = || (||g_t||_2)_t ||_2 .
 * Mathematically sqrt(sum of all squares); torch's per-tensor then
 * across-tensor two-level order is not reproduced (vectorised CPU reductions
 * are order-dependent), so this accumulates in double and rounds once: it is
 * the *more* accurate answer and parity on it is stated with a tolerance. */
double ndo_sumsq(const float* g, int64_t n) {
    double acc = 0.0;
#ifdef _OPENMP
#pragma omp parallel for reduction(+ : acc) schedule(static)
#endif
    for (int64_t i = 0; i < n; ++i) acc += (double)g[i] * (double)g[i];
    return acc;
}

/* torch:nn/utils/clip_grad.py:165-168
 *   clip_coef = max_norm / (total_norm + 1e-6); clamp(max=1.0)
 * evaluated in fp32 as torch does on a 0-d fp32 tensor. */
float ndo_clip_coef(float total_norm, float max_norm) {
    float c = max_norm / (total_norm + 1e-6f);
    return c > 1.0f ? 1.0f : c;
}

/* nanodiloco/diloco/diloco.py:57-58 : clip (g *= coef, clip_grad.py:173) then
 * inner_optimizer.step() with torch.optim.AdamW(lr) (main.py:100; defaults
 * betas (0.9,0.999), eps 1e-8, weight_decay 1e-2).
 * Arithmetic: torch:optim/adam.py:692-800 (_multi_tensor_adam, the CUDA
 * default) / :419-547 (_single_tensor_adam):
 *   p  *= 1 - lr*wd                     (scalar formed in double, cast to fp32)
 *   m   = m + (1-b1)*(g-m)              lerp, weight<0.5 branch (Lerp.h:22-35), one FMA
 *   v   = v*b2 ; v += (1-b2)*g*g        addcmul
 *   den = sqrt(v)/sqrt(bc2) + eps       true division by the fp32 scalar
 *   p  += (-lr/bc1) * m/den             addcdiv
 * torch's CPU and CUDA kernels round addcmul/addcdiv differently, so there are
 * two recipes (both measured, see tests/test_oracle_golden.py):
 *   recipe 0 = torch CPU   : v = fma(w2*g, g, v)  ; p = p + (s*m)/den
 *              (pinned bit-exact by tests/golden/inner.npz, generated from the
 *               reference's inner_step on CPU)
 *   recipe 1 = torch CUDA  : v = fma(w2, g*g, v)  ; p = fma(s, m/den, p)
 *              (torch:include/ATen/native/cuda/DeviceAddCmulCdiv.cuh:10-32 --
 *               what the reference executes on a GPU, and what the product
 *               kernel follows; pinned on the GPU box against torch's CUDA AdamW)
 * `step` is the 1-based step count; bias corrections are formed in double on
 * the host (adam.py:773-781) and cast to fp32 when they enter the kernels.
 * clip_coef == 1.0f is the "no clipping" case (g*1 is exact). */
void ndo_adamw_step(float* p, const float* grad, float* m, float* v, int64_t n,
                    double lr, double beta1, double beta2, double eps, double wd,
                    int64_t step, float clip_coef, int recipe) {
    const float decay = (float)(1.0 - lr * wd);
    const float w1 = (float)(1.0 - beta1);
    const float b2 = (float)beta2;
    const float w2 = (float)(1.0 - beta2);
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    const float step_size = (float)((lr / bc1) * -1.0);
    const float bc2_sqrt = (float)sqrt(bc2);
    const float feps = (float)eps;
    ND_PAR
    for (int64_t i = 0; i < n; ++i) {
        float g = grad[i] * clip_coef;
        float pi = p[i];
        if (wd != 0.0) pi = pi * decay;
        float mi = m[i];
        mi = fmaf(w1, g - mi, mi);
        float vi = v[i] * b2;
        vi = recipe ? fmaf(w2, g * g, vi) : fmaf(w2 * g, g, vi);
        float den = sqrtf(vi) / bc2_sqrt + feps;
        if (recipe) {
            pi = fmaf(step_size, mi / den, pi);
        } else {
            float q = (step_size * mi) / den;
            pi = pi + q;
        }
        p[i] = pi; m[i] = mi; v[i] = vi;
    }
}
