/*
 * nanodiloco_b200.h -- C ABI of libnanodiloco_b200.so
 *
 * Hand-written sm_100a CUDA kernels for the one hot path of SJCaldwell/NanoDiloco:
 * the DiLoCo outer synchronisation (pseudo-gradient -> cross-worker average ->
 * Nesterov outer SGD -> copy-back) and the inner clip + AdamW step.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross this boundary.
 *   - every pointer is DEVICE memory owned by the caller unless the name says
 *     `host`; the library never allocates or frees caller-visible memory.
 *   - all work is enqueued on the caller's stream (`nd_stream_t` is a
 *     `cudaStream_t`); no call synchronises the device unless documented.
 *   - every function returns 0 on success, a non-zero ND_E* code otherwise;
 *     `nd_last_error()` then holds a human-readable message (thread-local).
 *   - `n` counts ELEMENTS.  Buffers may have any alignment; the 128-bit
 *     vectorised path is taken when every pointer of a call is 16-byte aligned
 *     (always true for the flat buffers nanodiloco_b200.Diloco builds).
 *
 * Citations `diloco.py:NN`, `main.py:NN` are into the reference repository
 * (nanodiloco/diloco/diloco.py, nanodiloco/main.py); `torch:` into the torch
 * sources the reference calls.
 */
#ifndef NANODILOCO_B200_H
#define NANODILOCO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* nd_stream_t; /* == cudaStream_t */

/* element types of the workers' local ("working") parameter copy */
enum { ND_F32 = 0, ND_BF16 = 1 };

/* error codes (negative: library; the message carries the CUDA/NCCL string) */
enum {
    ND_OK = 0,
    ND_EINVAL = 1,   /* bad argument (null pointer, n < 0, unknown dtype, W out of range) */
    ND_ECUDA = 2,    /* a CUDA runtime call or kernel launch failed */
    ND_ENCCL = 3,    /* an NCCL call failed or libnccl could not be loaded */
    ND_ESTATE = 4    /* call made in the wrong state (e.g. context not initialised) */
};

#define ND_MAX_WORKERS 16 /* most worker copies one fused launch can read */
#define ND_MAX_REPLICAS 8 /* most GPUs one DiLoCo worker can span (nd_grad_avg, nd_adamw_rw) */

/* One row of a multi-tensor table (lives in DEVICE memory, built once by the
 * caller because parameter storage never moves -- unlike diloco.py:50, which
 * rebinds .data each outer step).  `ptr[k]` meaning depends on the call. */
typedef struct {
    void* ptr[4];
    int64_t numel;
    int64_t first_chunk; /* index of this tensor's first ND_CHUNK-sized chunk in the launch */
} nd_tensor_desc;

#define ND_CHUNK 16384 /* elements per multi-tensor chunk */

/* ---- library ---------------------------------------------------------- */

int nd_version(void);
const char* nd_last_error(void);

/* Bytes of zero-initialised device scratch the norm-producing calls need
 * (per-block partial sums + a self-resetting arrival counter).  Zero it once
 * after allocation; calls leave it ready for the next call.  Calls that share
 * one scratch buffer must be ordered on the GPU (same stream or events). */
size_t nd_scratch_bytes(void);

/* ---- (1) pseudo-gradient ---------------------------------------------- */

/* pg[i] = master[i] - (float)local[i]
 * Replaces diloco.py:47-48 (`last_sync.to(device)`; `last_sync - replica.data`)
 * -- the master copy is already in HBM, so the per-tensor H2D copy is gone. */
int nd_pseudo_grad(const float* master, const void* local, int local_dtype,
                   float* pg, int64_t n, nd_stream_t stream);

/* Multi-tensor form over a device table: ptr[0]=master, ptr[1]=local, ptr[2]=pg.
 * Replaces the per-tensor Python loop diloco.py:46-48 with one launch. */
int nd_pseudo_grad_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks,
                      int local_dtype, nd_stream_t stream);

/* ---- (2) fused outer update ------------------------------------------- */

/* Outer Nesterov-SGD on one shard after the cross-worker SUM:
 *   g = pg_sum / world                         (the AVG of diloco.py:49)
 *   g = g + wd*p        if wd != 0             torch:optim/sgd.py:345-356
 *   b = g               if first_step          torch:optim/sgd.py:361-363
 *   b = b*mu ; b = b + (1-dampening)*g  else   torch:optim/sgd.py:365
 *   d = g + mu*b        if nesterov else b     torch:optim/sgd.py:367-370
 *   p = p + (-lr)*d                            torch:optim/sgd.py:380
 * master <- p, mom <- b, local_out <- p (cast to local_dtype): the rewind of
 * diloco.py:50, the step of diloco.py:52 and the snapshot of diloco.py:54 in a
 * single HBM pass.  Rounding follows torch exactly (each `alpha=` form is one
 * FMA, `b*mu` is its own rounding), so W<=2 results are bit-identical.
 * local_out may be NULL (shard-only update; the all-gather writes the copy).
 * sumsq_out (nullable): receives sum(g*g) over the n elements (the averaged
 * pseudo-gradient's squared norm); needs `scratch` (nd_scratch_bytes()). */
int nd_outer_nesterov(float* master, float* mom, const float* pg_sum,
                      void* local_out, int local_dtype, int64_t n,
                      float world, float lr, float mu, float dampening, float wd,
                      int nesterov, int first_step,
                      float* sumsq_out, void* scratch, nd_stream_t stream);

/* Kernels (1)+(2) in one pass for W worker copies addressable from this GPU
 * (W=1: the single-GPU outer step; W>1: peer-mapped copies over NVLink, or
 * several workers emulated on one GPU):
 *   g = ( sum_{w=0..W-1} (master - local_w) ) / W     diloco.py:48-49, summed in worker order
 *   ...update as nd_outer_nesterov...
 *   master <- p, mom <- b, local_w <- p for every w    diloco.py:50-54
 * `locals_host` is a HOST array of W device pointers (copied into the launch
 * arguments).  24 B/element of HBM traffic at W=1; the pseudo-gradient never
 * touches memory. */
int nd_outer_fused(float* master, float* mom, void* const* locals_host, int W,
                   int local_dtype, int64_t n,
                   float lr, float mu, float dampening, float wd,
                   int nesterov, int first_step,
                   float* sumsq_out, void* scratch, nd_stream_t stream);

/* nd_outer_fused with separate READ and WRITE sets: the average runs over the W copies in
 * `reads_host` (one per DiLoCo worker) and the new parameters are written to the `n_writes` copies
 * in `writes_host` (every replica of every worker).  For workers that span several GPUs
 * (BASELINE config 5: 4 workers x 2 GPUs -- the reference has no intra-worker parallelism, so
 * this is an extension): replicas of one worker hold identical parameters, so reading one of
 * them and writing all of them is the same outer step.  n_writes may be 0 (update master and
 * momentum only). */
int nd_outer_fused_rw(float* master, float* mom, void* const* reads_host, int W,
                      void* const* writes_host, int n_writes, int local_dtype, int64_t n,
                      float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                      float* sumsq_out, void* scratch, nd_stream_t stream);

/* Multi-tensor form of nd_outer_fused for W=1:
 * ptr[0]=master, ptr[1]=mom, ptr[2]=local.  Replaces diloco.py:46-54. */
int nd_outer_fused_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks,
                      int local_dtype, float lr, float mu, float dampening, float wd,
                      int nesterov, int first_step, nd_stream_t stream);

/* ---- (3) inner step: clip_grad_norm_ + AdamW --------------------------- */

/* out[0] = sum(g[i]^2), deterministic (fixed reduction tree, fp32 partials
 * combined in double).  First half of torch.nn.utils.clip_grad_norm_
 * (diloco.py:57; torch:nn/utils/clip_grad.py:93-108).  Needs `scratch`. */
int nd_sumsq(const void* g, int dtype, int64_t n, float* out, void* scratch,
             nd_stream_t stream);

/* nd_sumsq without caller scratch and without any allocation (no counterpart in the
 * reference: it exists for how this library gets embedded, not for the algorithm).
 * For callers that can pass inputs and outputs but cannot provide scratch or tolerate a
 * cudaMalloc: a process whose HBM is entirely owned by another pool (a capped caching
 * allocator, a server's fixed arena), a latency-critical loop that cannot afford the device-wide
 * synchronisation hidden in cudaMalloc/cudaFree, or a thin FFI caller with no allocator of its
 * own.  The partial sums live in a fixed ND_SUMSQ_STATIC_SCRATCH_BYTES array inside the library
 * image (held by every GPU that loads the library until the process exits).
 *   - out[0] = sum(g[i]^2) exactly as nd_sumsq computes it.
 *   - n > ND_SUMSQ_STATIC_MAX_ELEMS is rejected with ND_EINVAL; use nd_sumsq beyond it.
 *   - the first call on each device looks up the array's device address (set-up); make that
 *     first call outside any stream capture.  No call allocates or frees memory.
 *   - all calls on a device share the array: calls that overlap on the GPU race; order them
 *     (same stream, or an event). */
#define ND_SUMSQ_STATIC_SCRATCH_BYTES 8192
#define ND_SUMSQ_STATIC_MAX_ELEMS (1LL << 31)
int nd_sumsq_static(const void* g, int dtype, int64_t n, float* out, nd_stream_t stream);

/* Multi-tensor form: ptr[0]=grad (`dtype` of the gradients: ND_F32 or ND_BF16).  One launch over the
 * gradient tensors autograd allocated -- what clip_grad_norm_ (diloco.py:57) does in ~3 foreach
 * launches over per-tensor norms. */
int nd_sumsq_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks, int dtype,
                float* out, void* scratch, nd_stream_t stream);

/* Fused clip + AdamW (diloco.py:57-58; torch:nn/utils/clip_grad.py:165-174,
 * torch:optim/adam.py:692-800 with decoupled weight decay):
 *   coef = min(1, max_norm / (sqrt(*sumsq_dev) + 1e-6))   (1 if sumsq_dev is NULL)
 *   g  = g*coef
 *   p  = p*(1 - lr*wd)
 *   m  = m + (1-beta1)*(g - m)
 *   v  = v*beta2 ; v = v + (1-beta2)*(g*g)
 *   p  = p + (-lr/bc1) * ( m / (sqrt(v)/sqrt(bc2) + eps) )
 * with bc1 = 1-beta1^step, bc2 = 1-beta2^step formed in double on the host like
 * torch does, `step` 1-based.  Rounding follows torch's CUDA foreach kernels
 * (torch:include/ATen/native/cuda/DeviceAddCmulCdiv.cuh:10-32).  The clip
 * coefficient is read on the device: no host synchronisation anywhere.
 * zero_grad != 0 also writes g <- 0 (the zero_grad of diloco.py:60, fused). */
int nd_adamw(float* p, float* g, float* m, float* v, int64_t n,
             double lr, double beta1, double beta2, double eps, double weight_decay,
             int64_t step, const float* sumsq_dev, float max_norm, int zero_grad,
             nd_stream_t stream);

/* nd_adamw for a bf16 working copy (parameters and gradients bf16, moments fp32): the
 * "bf16 inner / fp32 outer" mode of BASELINE config 2.  An extension -- the reference is fp32
 * end to end (main.py:97-99).  Arithmetic as nd_adamw on the exactly-widened values, one rounding
 * of the new parameter to bf16. */
int nd_adamw_bf16(void* p_bf16, void* g_bf16, float* m, float* v, int64_t n,
                  double lr, double beta1, double beta2, double eps, double weight_decay,
                  int64_t step, const float* sumsq_dev, float max_norm, int zero_grad,
                  nd_stream_t stream);

/* Gradient average over the R copies of one worker's gradient (a DiLoCo worker that spans R GPUs:
 * BASELINE config 5 -- an extension, in the reference every rank is a worker and nothing averages
 * gradients): g_out = (g_0 + ... + g_{R-1}) / R summed in the order given, sumsq_out[0] = sum(g_out^2).
 * g_out may alias one of the inputs.  What `dist.all_reduce(grad, AVG)` + the first half of
 * clip_grad_norm_ (diloco.py:57) would do in two library passes.  R in {1,2,3,4,8}; buffers 16-byte aligned. */
int nd_grad_avg(void* const* grads_host, int R, void* g_out, int dtype, int64_t n,
                float* sumsq_out, void* scratch, nd_stream_t stream);

/* nd_adamw / nd_adamw_bf16 for one replica's part of the flat space: the squared gradient norm is the
 * sum, in the order given, of `n_parts` partial sums (device pointers, possibly in peer memory) and the new
 * parameters are written to every copy in `p_writes_host` (peer-mapped copies of the other replicas; `p`
 * itself may be among them).  n_parts = 0: no clipping.  Buffers 16-byte aligned. */
int nd_adamw_rw(void* p, void* const* p_writes_host, int n_writes, void* g, float* m, float* v, int dtype,
                int64_t n, double lr, double beta1, double beta2, double eps, double weight_decay,
                int64_t step, const float* const* sumsq_parts_host, int n_parts, float max_norm,
                int zero_grad, nd_stream_t stream);

/* Multi-tensor form: ptr[0]=p, ptr[1]=g, ptr[2]=m, ptr[3]=v; `dtype` is the type of p and g (ND_F32,
 * or ND_BF16 with fp32 moments as in nd_adamw_bf16).  This is the form the drop-in wrapper uses by
 * default: the gradients stay the tensors autograd allocated, so the step can end with
 * `.grad = None` exactly like diloco.py:60 (`zero_grad()`, set_to_none) and pays no zeroing store
 * (zero_grad = 0: 28 B/element).  TMA-staged like nd_adamw; a row whose pointers are not 16-byte
 * aligned is processed with plain loads and stores.  The table is read when the kernel RUNS: keep it
 * (and the tensors it names) alive and unchanged until then. */
int nd_adamw_mt(const nd_tensor_desc* table_dev, int n_tensors, int64_t total_chunks, int dtype,
                double lr, double beta1, double beta2, double eps, double weight_decay,
                int64_t step, const float* sumsq_dev, float max_norm, int zero_grad,
                nd_stream_t stream);

/* ---- sharded multi-GPU outer sync (one process per GPU) ----------------- */

/* 128-byte NCCL unique id, produced on rank 0 and handed to every rank by the
 * host (torch.distributed broadcast / store). */
#define ND_UNIQUE_ID_BYTES 128
int nd_nccl_unique_id(char out_host[ND_UNIQUE_ID_BYTES]);

typedef struct nd_sync_ctx nd_sync_ctx;

/* Creates this rank's context: an NCCL communicator over `world` ranks (the
 * replacement of utils.py:41-43's process group for this path), `n_streams`
 * side streams and the events that join them to the caller's stream.  Must be
 * called by all ranks together, with the calling thread's current CUDA device
 * set.  bucket_elems: elements per pipeline bucket (rounded up to a multiple of
 * 4*world). */
int nd_sync_ctx_create(const char unique_id_host[ND_UNIQUE_ID_BYTES], int world, int rank,
                       int64_t bucket_elems, int n_streams, nd_sync_ctx** out);
int nd_sync_ctx_destroy(nd_sync_ctx* ctx);

/* Optional: a SECOND communicator used for the all-gathers only.  NCCL runs one communicator's
 * collectives in issue order, so with a single one AG(b) queues behind RS(b+1); with two, the
 * all-gather of bucket b and the reduce-scatter of bucket b+1 are independent NCCL kernels.
 * Collective over the ranks; `unique_id_host` is a second id from nd_nccl_unique_id on rank 0. */
int nd_sync_ctx_add_ag_comm(nd_sync_ctx* ctx, const char unique_id_host[128]);

/* Buffers from NCCL's own allocator (ncclMemAlloc: CUDA VMM memory NCCL can bind to the NVSwitch
 * multicast object) and their registration with the context's communicators (ncclCommRegister), so
 * that reduce-scatter / all-gather may run in-switch (NVLS) straight on the user buffers instead of
 * staging through NCCL's internal ones.  These replace nothing in the reference -- its buffers are
 * ordinary torch tensors (diloco.py:48-49) -- they are how the pipeline reaches the fabric's peak.
 * Register `local` and `pg` after nd_sync_ctx_create (and nd_sync_ctx_add_ag_comm). */
int nd_nccl_mem_alloc(size_t bytes, void** dev_ptr_out);
int nd_nccl_mem_free(void* dev_ptr);
int nd_sync_register(nd_sync_ctx* ctx, void* buf, size_t bytes);

/* Per-bucket stage timeline of the LAST nd_outer_sync (row f4 of the scope table: per-bucket
 * counters).  With profiling on, every bucket records 7 CUDA events; nd_sync_timeline returns, per
 * bucket, milliseconds since the call started for: K1 start, K1 end, reduce-scatter end, K2 start,
 * K2 end, all-gather end, master-refresh end (rows_out: n_rows x 7 floats; waits for the sync). */
int nd_sync_set_profiling(nd_sync_ctx* ctx, int on);
int nd_sync_timeline(nd_sync_ctx* ctx, float* rows_out, int64_t cap_rows, int64_t* n_rows_out);

/* Elements of the padded flat parameter space for n real elements (multiple of
 * bucket size boundaries so that every bucket splits into `world` equal,
 * 16-byte-aligned shards), and the number of elements of sharded state
 * (master shard / momentum shard) this rank must hold. */
int64_t nd_sync_padded_elems(const nd_sync_ctx* ctx, int64_t n);
int64_t nd_sync_shard_elems(const nd_sync_ctx* ctx, int64_t n);
/* Elements per pipeline bucket after rounding.  Bucket b covers padded elements
 * [b*B, min((b+1)*B, n_padded)); rank r owns the r-th of its `world` equal
 * slices, and the rank's shard is the concatenation of its slices in bucket order. */
int64_t nd_sync_bucket_elems(const nd_sync_ctx* ctx);

/* The whole outer step of diloco.py:34-54 across `world` GPUs.  Per bucket b,
 * pipelined over the side streams:
 *   K1   pg_b = master_b - local_b                       (all ranks, full bucket)
 *   RS   ncclReduceScatter(pg_b, sum) -> this rank's 1/world slice
 *   K2   nd_outer_nesterov on the slice (master shard, momentum shard), writing
 *        the new parameters into this rank's slice of local_b
 *   AG   ncclAllGather(local_b) in place; master_b <- local_b rides with it
 * local:      padded flat working parameters (fp32), n_padded elements
 * master:     padded flat last-sync copy (fp32), n_padded elements (replicated)
 * mom_shard:  this rank's momentum slices, nd_sync_shard_elems() elements
 * pg:         staging for the pseudo-gradient, n_padded elements
 * On return the work is ordered on `stream`; nothing is synchronised. */
int nd_outer_sync(nd_sync_ctx* ctx, float* local, float* master, float* mom_shard, float* pg,
                  int64_t n_padded, float lr, float mu, float dampening, float wd,
                  int nesterov, int first_step, nd_stream_t stream);

/* The outer step with the last-sync copy resident in HOST memory, where the reference
 * keeps it (diloco.py:25-32 offloads it, :47 copies it back per tensor, :54 offloads
 * again).  World size 1.  The flat space is cut into chunks; chunk k rides side stream
 * k % n_streams:  H2D(master chunk, pinned host -> staging slot)  ->  nd_outer_fused on
 * the slot  ->  D2H(new master chunk), so the two PCIe directions and the kernel overlap.
 * master_host: n floats of page-locked host memory (updated in place)
 * staging:     device buffer of n_streams * chunk_elems floats
 * Ordered on `stream` on return; nothing is synchronised. */
int nd_outer_step_offloaded(nd_sync_ctx* ctx, float* master_host, float* staging, int64_t chunk_elems,
                            float* mom, void* local, int local_dtype, int64_t n,
                            float lr, float mu, float dampening, float wd,
                            int nesterov, int first_step, nd_stream_t stream);

/* Device milliseconds of the last nd_outer_sync / nd_outer_step_offloaded on this rank (blocks until it
 * has finished) -- what the reference's dead `avg_sync_time` (diloco.py:62-64)
 * promised. */
int nd_sync_last_ms(nd_sync_ctx* ctx, float* ms_out);

/* ---- peer-memory outer sync: one kernel per GPU over NVLink --------------------- */

#define ND_IPC_HANDLE_BYTES 64 /* sizeof(cudaIpcMemHandle_t) */

/* Allocates this rank's SHARED working-parameter buffer (device memory every rank of the node
 * can map; zero-filled; `bytes` of payload followed by a flag page used by the barrier) and
 * returns its CUDA IPC handle for the host to pass to the other ranks (torch.distributed
 * all_gather / a store).  This is the one place the library allocates memory the caller sees:
 * IPC needs cudaMalloc'ed memory.  Free with nd_peer_free after nd_peer_ctx_destroy. */
int nd_peer_alloc(size_t bytes, void** dev_ptr_out, char handle_out[ND_IPC_HANDLE_BYTES]);
int nd_peer_free(void* dev_ptr);

typedef struct nd_peer_ctx nd_peer_ctx;

/* handles_host: world * ND_IPC_HANDLE_BYTES bytes, the handle of every rank in rank order (this
 * rank's own entry is ignored).  Maps the peers' buffers (cudaIpcOpenMemHandle with lazy peer
 * access).  world <= ND_MAX_WORKERS; all ranks on one NVLink/NVSwitch domain. */
int nd_peer_ctx_create(int world, int rank, void* my_buffer, size_t bytes, const char* handles_host,
                       nd_peer_ctx** out);
int nd_peer_ctx_destroy(nd_peer_ctx* ctx);
/* Address, in this process, of rank w's buffer. */
int nd_peer_buffer(const nd_peer_ctx* ctx, int w, void** out);
/* Workers that span several GPUs: ranks [k*replicas, (k+1)*replicas) hold identical copies of
 * worker k (they all-reduce their gradients every inner step).  The outer sync then averages over
 * world/replicas workers (reading each worker's first replica) and writes every rank's copy.
 * Default 1.  `replicas` must divide the world size.  (Extension: BASELINE config 5.) */
int nd_peer_set_replicas(nd_peer_ctx* ctx, int replicas);
/* Device-side barrier over the ranks (flag exchange in peer memory), enqueued on `stream`.
 * Every rank must enqueue the same sequence of barriers. */
int nd_peer_barrier(nd_peer_ctx* ctx, nd_stream_t stream);

/* The whole outer step of diloco.py:34-54 across `world` GPUs as one fused kernel per GPU:
 *   barrier -> nd_outer_fused over this rank's shard [rank*n/W, (rank+1)*n/W) with the W worker
 *   copies addressed through peer memory (reads AND the copy-back go over NVLink inside the
 *   kernel) -> barrier.
 * The working parameters of every rank are the first n_padded elements of its shared buffer.
 * master_shard / mom_shard: this rank's n_padded/W elements of last-sync copy and outer momentum
 * (ordinary device memory; there is no replica of either).  n_padded: multiple of 4*world
 * (8*world for a bf16 working copy).  The average is summed in worker order 0..W-1 on every
 * shard, so all ranks end bit-identical and results do not depend on timing.
 * sumsq_out/scratch as in nd_outer_fused (this rank's shard only).  Replaces the collective of
 * diloco.py:49 together with :46-54. */
int nd_outer_sync_p2p(nd_peer_ctx* ctx, float* master_shard, float* mom_shard, int local_dtype,
                      int64_t n_padded, float lr, float mu, float dampening, float wd,
                      int nesterov, int first_step, float* sumsq_out, void* scratch,
                      nd_stream_t stream);
/* nd_outer_sync_p2p with this rank's master shard resident in page-locked HOST memory, where the
 * reference keeps the last-sync copy (diloco.py:25-32,47,54): the shard is streamed in chunks
 * (H2D -> fused peer kernel on the chunk -> D2H) over three internal side streams, between the two
 * barriers.  staging: device buffer of 3 * chunk_elems floats.  Any world size >= 1. */
int nd_outer_sync_p2p_offloaded(nd_peer_ctx* ctx, float* master_host_shard, float* staging,
                                int64_t chunk_elems, float* mom_shard, int local_dtype,
                                int64_t n_padded, float lr, float mu, float dampening, float wd,
                                int nesterov, int first_step, nd_stream_t stream);
/* A peer context over buffers the CALLER has already mapped into every rank (e.g. torch
 * symmetric memory): bases_host[w] = address, in this process, of rank w's buffer of
 * `bytes` payload + a 4096-byte flag page (zero-filled).  Nothing is mapped or unmapped here. */
int nd_peer_ctx_create_from_ptrs(int world, int rank, void* const* bases_host, size_t bytes,
                                 nd_peer_ctx** out);

/* The outer step with the cross-worker sum taken INSIDE the NVSwitch (NVLS multicast):
 *   K1  pg = master - local on every rank (whole space)            diloco.py:48
 *   barrier
 *   this rank's shard: g = multimem.ld_reduce.add(pg)/W (the switch sums every rank's pg),
 *   Nesterov update of master/momentum shard, multimem.st of the new parameters into every
 *   rank's working copy                                            diloco.py:49-52
 *   barrier; the master replica's other shards are refreshed from the working copy (diloco.py:54)
 * run bucket by bucket (bucket_elems, rounded up to a multiple of 4*world; rank r owns the r-th
 * slice of every bucket, mom_shard is the concatenation of its slices) on three internal streams so
 * that the two HBM-only passes hide under the NVLink pass of the neighbouring buckets.
 * local, pg: this rank's two symmetric buffers (plain addresses); mc_local, mc_pg: the multicast
 * addresses of the same two buffers; master: full fp32 replica (n_padded); mom_shard: n_padded/W.
 * fp32 only.  The summation order inside the switch is the hardware's: results equal the
 * reference bit for bit at W = 2 and within the stated tolerance beyond (like NCCL's or gloo's
 * order).  Per direction and GPU ~4*(1+1/W) B/element cross NVLink instead of 8*(W-1)/W. */
int nd_outer_sync_nvls(nd_peer_ctx* ctx, float* local, float* pg, float* mc_local, float* mc_pg,
                       float* master, float* mom_shard, int64_t n_padded, int64_t bucket_elems,
                       float lr, float mu, float dampening, float wd, int nesterov, int first_step,
                       nd_stream_t stream);
/* nd_outer_sync_nvls on the first n_nvls elements and, concurrently on a fourth internal stream,
 * the peer-memory fused kernel (nd_outer_fused_rw over the mapped copies) on the rest
 * [n_nvls, n_padded), cut into `world` contiguous shards.  The in-switch reduction sustains fewer
 * link bytes per second than plain peer traffic; splitting the space lets the two mechanisms share
 * the links.  n_nvls: multiple of 4*world in [0, n_padded] (n_padded = pure NVLS, 0 = pure peer
 * kernel with a full master replica).  mom_shard: the in-switch part's n_nvls/world elements (bucket
 * order) followed by the peer part's (n_padded - n_nvls)/world. */
int nd_outer_sync_hybrid(nd_peer_ctx* ctx, float* local, float* pg, float* mc_local, float* mc_pg,
                         float* master, float* mom_shard, int64_t n_padded, int64_t n_nvls,
                         int64_t bucket_elems, float lr, float mu, float dampening, float wd,
                         int nesterov, int first_step, nd_stream_t stream);
int nd_peer_last_ms(nd_peer_ctx* ctx, float* ms_out);

/* How the W worker copies reach the owner of a shard in nd_outer_sync_p2p (replaces the collective
 * of diloco.py:49 either way; same result bit for bit, same 8(W-1)/W wire bytes per element):
 *   ND_PEER_PULL  the owner's fused kernel reads the remote copies (bulk loads over NVLink).
 *   ND_PEER_PUSH  every worker first writes its slice into the owner's staging slots (a copy kernel of
 *                 bulk stores), a flag barrier, then the owner's fused kernel reads local HBM only and
 *                 writes the new parameters out: the links carry write traffic only.
 * Push needs a staging area INSIDE every rank's peer buffer (the same offset on every rank), of at
 * least (world / replicas) * (n_padded / world) * element bytes: nd_peer_set_staging, then
 * nd_peer_set_mode.  The working copy then ends where the staging area starts. */
enum { ND_PEER_PULL = 0, ND_PEER_PUSH = 1 };
int nd_peer_set_staging(nd_peer_ctx* ctx, size_t byte_offset, size_t bytes);
int nd_peer_set_mode(nd_peer_ctx* ctx, int mode);

/* The whole inner step (diloco.py:56-60 minus the scheduler) of a worker that spans `replicas` GPUs, as
 * two kernels per GPU over peer memory: nd_grad_avg on this replica's 1/R of the flat space reading every
 * replica's gradient, a flag barrier among the replicas, nd_adamw_rw on that part writing every replica's
 * parameters (exp_avg / exp_avg_sq are kept for the owned part only), a flag barrier.  `params_ctx` maps the
 * working copies (the context of nd_outer_sync_p2p), `grads_ctx` the gradient buffers (a second
 * nd_peer_alloc / nd_peer_ctx_create with the same replicas setting); m and v point at element 0 of the full
 * flat moment buffers.  On return (stream order) the gradient buffer is zero and sumsq_total_out (nullable)
 * holds the squared norm of the averaged gradient.  The first call on a gradient context allocates that
 * context's norm scratch (a few KB, freed by nd_peer_ctx_destroy): make it outside any stream capture.
 * An extension: the reference has no such mode. */
int nd_inner_step_replicas(nd_peer_ctx* params_ctx, nd_peer_ctx* grads_ctx, float* m, float* v, int dtype,
                           int64_t n_padded, double lr, double beta1, double beta2, double eps,
                           double weight_decay, int64_t step, float max_norm, float* sumsq_total_out,
                           nd_stream_t stream);

/* nd_outer_sync_p2p (pull) normally runs as ONE launch per GPU: the arrival rendezvous of the ranks sits in
 * the fused kernel's prologue (no block reads a peer's copy before every worker has published its arrival) and
 * the completion rendezvous in its epilogue (the block that finishes last publishes and awaits the "done"
 * flags).  on = 0 brackets the kernel with two barrier launches instead (ND_P2P_FOLD=0 does the same); the
 * results are the same bits.  Calls on all ranks must agree. */
int nd_peer_set_fold(nd_peer_ctx* ctx, int on);

/* nd_peer_last_ms plus the part of it spent in the first flag barrier of nd_outer_sync_p2p, i.e.
 * waiting for the slowest worker to arrive (the skew the reference's blocking all_reduce, diloco.py:49,
 * hides inside the collective).  work = total - wait. */
int nd_peer_last_ms2(nd_peer_ctx* ctx, float* total_ms_out, float* wait_ms_out);

#ifdef __cplusplus
}
#endif
#endif /* NANODILOCO_B200_H */
