"""ctypes binding of libnanodiloco_b200.so (the C ABI in include/nanodiloco_b200.h).

There is NO fallback: if the CUDA library is missing or a call fails, this raises.
Nothing here (or anywhere in nanodiloco_b200/) touches oracle/.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

from . import build as _build

ND_F32, ND_BF16 = 0, 1
ND_CHUNK = 16384
ND_MAX_WORKERS = 16
ND_MAX_REPLICAS = 8
ND_UNIQUE_ID_BYTES = 128
ND_IPC_HANDLE_BYTES = 64
ND_PEER_PULL, ND_PEER_PUSH = 0, 1

_c_i64 = ctypes.c_int64
_c_f = ctypes.c_float
_c_d = ctypes.c_double
_c_p = ctypes.c_void_p
_c_i = ctypes.c_int

# name -> (restype, argtypes); one entry per declaration in include/nanodiloco_b200.h
SIGNATURES = {
    "nd_version": (_c_i, []),
    "nd_last_error": (ctypes.c_char_p, []),
    "nd_scratch_bytes": (ctypes.c_size_t, []),
    "nd_pseudo_grad": (_c_i, [_c_p, _c_p, _c_i, _c_p, _c_i64, _c_p]),
    "nd_pseudo_grad_mt": (_c_i, [_c_p, _c_i, _c_i64, _c_i, _c_p]),
    "nd_outer_nesterov": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_i, _c_i64, _c_f, _c_f, _c_f, _c_f, _c_f,
                                 _c_i, _c_i, _c_p, _c_p, _c_p]),
    "nd_outer_fused": (_c_i, [_c_p, _c_p, ctypes.POINTER(_c_p), _c_i, _c_i, _c_i64, _c_f, _c_f, _c_f, _c_f,
                              _c_i, _c_i, _c_p, _c_p, _c_p]),
    "nd_outer_fused_rw": (_c_i, [_c_p, _c_p, ctypes.POINTER(_c_p), _c_i, ctypes.POINTER(_c_p), _c_i, _c_i, _c_i64,
                                 _c_f, _c_f, _c_f, _c_f, _c_i, _c_i, _c_p, _c_p, _c_p]),
    "nd_outer_fused_mt": (_c_i, [_c_p, _c_i, _c_i64, _c_i, _c_f, _c_f, _c_f, _c_f, _c_i, _c_i, _c_p]),
    "nd_sumsq": (_c_i, [_c_p, _c_i, _c_i64, _c_p, _c_p, _c_p]),
    "nd_sumsq_static": (_c_i, [_c_p, _c_i, _c_i64, _c_p, _c_p]),
    "nd_sumsq_mt": (_c_i, [_c_p, _c_i, _c_i64, _c_i, _c_p, _c_p, _c_p]),
    "nd_adamw": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_i64, _c_d, _c_d, _c_d, _c_d, _c_d, _c_i64, _c_p, _c_f,
                        _c_i, _c_p]),
    "nd_adamw_bf16": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_i64, _c_d, _c_d, _c_d, _c_d, _c_d, _c_i64, _c_p, _c_f,
                             _c_i, _c_p]),
    "nd_adamw_mt": (_c_i, [_c_p, _c_i, _c_i64, _c_i, _c_d, _c_d, _c_d, _c_d, _c_d, _c_i64, _c_p, _c_f, _c_i, _c_p]),
    "nd_grad_avg": (_c_i, [ctypes.POINTER(_c_p), _c_i, _c_p, _c_i, _c_i64, _c_p, _c_p, _c_p]),
    "nd_adamw_rw": (_c_i, [_c_p, ctypes.POINTER(_c_p), _c_i, _c_p, _c_p, _c_p, _c_i, _c_i64, _c_d, _c_d, _c_d, _c_d, _c_d,
                           _c_i64, ctypes.POINTER(_c_p), _c_i, _c_f, _c_i, _c_p]),
    "nd_inner_step_replicas": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_i, _c_i64, _c_d, _c_d, _c_d, _c_d, _c_d, _c_i64, _c_f,
                                      _c_p, _c_p]),
    "nd_nccl_unique_id": (_c_i, [ctypes.c_char_p]),
    "nd_sync_ctx_create": (_c_i, [ctypes.c_char_p, _c_i, _c_i, _c_i64, _c_i, ctypes.POINTER(_c_p)]),
    "nd_sync_ctx_destroy": (_c_i, [_c_p]),
    "nd_sync_ctx_add_ag_comm": (_c_i, [_c_p, ctypes.c_char_p]),
    "nd_nccl_mem_alloc": (_c_i, [ctypes.c_size_t, ctypes.POINTER(_c_p)]),
    "nd_nccl_mem_free": (_c_i, [_c_p]),
    "nd_sync_register": (_c_i, [_c_p, _c_p, ctypes.c_size_t]),
    "nd_sync_set_profiling": (_c_i, [_c_p, _c_i]),
    "nd_sync_timeline": (_c_i, [_c_p, ctypes.POINTER(_c_f), _c_i64, ctypes.POINTER(_c_i64)]),
    "nd_sync_padded_elems": (_c_i64, [_c_p, _c_i64]),
    "nd_sync_shard_elems": (_c_i64, [_c_p, _c_i64]),
    "nd_sync_bucket_elems": (_c_i64, [_c_p]),
    "nd_outer_sync": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_p, _c_i64, _c_f, _c_f, _c_f, _c_f, _c_i, _c_i, _c_p]),
    "nd_outer_step_offloaded": (_c_i, [_c_p, _c_p, _c_p, _c_i64, _c_p, _c_p, _c_i, _c_i64, _c_f, _c_f, _c_f, _c_f,
                                       _c_i, _c_i, _c_p]),
    "nd_sync_last_ms": (_c_i, [_c_p, ctypes.POINTER(_c_f)]),
    "nd_peer_alloc": (_c_i, [ctypes.c_size_t, ctypes.POINTER(_c_p), ctypes.c_char_p]),
    "nd_peer_free": (_c_i, [_c_p]),
    "nd_peer_ctx_create": (_c_i, [_c_i, _c_i, _c_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.POINTER(_c_p)]),
    "nd_peer_ctx_destroy": (_c_i, [_c_p]),
    "nd_peer_buffer": (_c_i, [_c_p, _c_i, ctypes.POINTER(_c_p)]),
    "nd_peer_set_replicas": (_c_i, [_c_p, _c_i]),
    "nd_peer_barrier": (_c_i, [_c_p, _c_p]),
    "nd_outer_sync_p2p": (_c_i, [_c_p, _c_p, _c_p, _c_i, _c_i64, _c_f, _c_f, _c_f, _c_f, _c_i, _c_i, _c_p, _c_p,
                                 _c_p]),
    "nd_outer_sync_p2p_offloaded": (_c_i, [_c_p, _c_p, _c_p, _c_i64, _c_p, _c_i, _c_i64, _c_f, _c_f, _c_f, _c_f,
                                           _c_i, _c_i, _c_p]),
    "nd_peer_ctx_create_from_ptrs": (_c_i, [_c_i, _c_i, ctypes.POINTER(_c_p), ctypes.c_size_t, ctypes.POINTER(_c_p)]),
    "nd_outer_sync_nvls": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_p, _c_p, _c_p, _c_i64, _c_i64, _c_f, _c_f, _c_f, _c_f, _c_i, _c_i,
                                  _c_p]),
    "nd_outer_sync_hybrid": (_c_i, [_c_p, _c_p, _c_p, _c_p, _c_p, _c_p, _c_p, _c_i64, _c_i64, _c_i64, _c_f, _c_f, _c_f,
                                    _c_f, _c_i, _c_i, _c_p]),
    "nd_peer_last_ms": (_c_i, [_c_p, ctypes.POINTER(_c_f)]),
    "nd_peer_set_staging": (_c_i, [_c_p, ctypes.c_size_t, ctypes.c_size_t]),
    "nd_peer_set_mode": (_c_i, [_c_p, _c_i]),
    "nd_peer_set_fold": (_c_i, [_c_p, _c_i]),
    "nd_peer_last_ms2": (_c_i, [_c_p, ctypes.POINTER(_c_f), ctypes.POINTER(_c_f)]),
}

_lib: Optional[ctypes.CDLL] = None


class NDError(RuntimeError):
    """A call into libnanodiloco_b200.so returned a non-zero code."""


def library_path() -> str:
    return _build.SO


def load() -> ctypes.CDLL:
    """Loads the CUDA library, building it first (under a lock, atomically) only when it is missing
    and nvcc exists.  An existing library is trusted; `python -m nanodiloco_b200.build --force` or
    `__graft_entry__.build()` rebuild it.

    Fails loudly otherwise -- there is no CPU or PyTorch fallback for this path."""
    global _lib
    if _lib is not None:
        return _lib
    so = _build.SO
    if _build.needs_build():
        try:
            _build.build()
        except Exception as e:  # noqa: BLE001
            if not os.path.exists(so):
                raise NDError(
                    f"libnanodiloco_b200.so is not built and could not be built ({e}); "
                    "run `python -m nanodiloco_b200.build` where nvcc is available") from e
    # libcudart.so.12 / libnccl.so.2 are resolved from the process (torch loads both)
    try:
        import torch  # noqa: F401  (pulls in libcudart / libnccl so the sonames resolve)
    except Exception:  # pragma: no cover
        pass
    L = ctypes.CDLL(so, mode=ctypes.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError here == header and library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().nd_last_error()
        raise NDError(f"{what or 'nanodiloco_b200'} failed (code {rc}): {msg.decode() if msg else ''}")


def _ptr(t) -> int:
    """Device pointer of a torch tensor (or None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def _stream(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def _dtype_code(t) -> int:
    import torch
    if t.dtype == torch.float32:
        return ND_F32
    if t.dtype == torch.bfloat16:
        return ND_BF16
    raise NDError(f"unsupported dtype {t.dtype}; the working copy must be float32 or bfloat16")


def _require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NDError("nanodiloco_b200 kernels need CUDA tensors; there is no CPU path")
        if t is not None and not t.is_contiguous():
            raise NDError("nanodiloco_b200 kernels need contiguous tensors")


# ---- thin tensor-level wrappers (all enqueue on the current stream) ----------------------------

def scratch(device):
    """Zeroed scratch for the norm-producing calls (see nd_scratch_bytes)."""
    import torch
    n = load().nd_scratch_bytes()
    return torch.zeros((n + 3) // 4, dtype=torch.int32, device=device)


def pseudo_grad(master, local, out, stream=None) -> None:
    _require_cuda(master, local, out)
    check(load().nd_pseudo_grad(_ptr(master), _ptr(local), _dtype_code(local), _ptr(out), master.numel(),
                                _stream(stream)), "nd_pseudo_grad")


def outer_nesterov(master, mom, pg_sum, local_out, *, world: float, lr: float, momentum: float,
                   dampening: float = 0.0, weight_decay: float = 0.0, nesterov: bool = True,
                   first_step: bool = False, sumsq_out=None, scratch_buf=None, stream=None) -> None:
    _require_cuda(master, mom, pg_sum, local_out, sumsq_out, scratch_buf)
    dt = ND_F32 if local_out is None else _dtype_code(local_out)
    check(load().nd_outer_nesterov(_ptr(master), _ptr(mom), _ptr(pg_sum), _ptr(local_out), dt, master.numel(),
                                   world, lr, momentum, dampening, weight_decay, int(nesterov), int(first_step),
                                   _ptr(sumsq_out), _ptr(scratch_buf), _stream(stream)), "nd_outer_nesterov")


def outer_fused(master, mom, locals_: Sequence, *, lr: float, momentum: float, dampening: float = 0.0,
                weight_decay: float = 0.0, nesterov: bool = True, first_step: bool = False,
                sumsq_out=None, scratch_buf=None, stream=None, n: Optional[int] = None) -> None:
    _require_cuda(master, mom, *locals_, sumsq_out, scratch_buf)
    W = len(locals_)
    arr = (_c_p * W)(*[_ptr(t) for t in locals_])
    check(load().nd_outer_fused(_ptr(master), _ptr(mom), arr, W, _dtype_code(locals_[0]),
                                master.numel() if n is None else n, lr, momentum, dampening, weight_decay,
                                int(nesterov), int(first_step), _ptr(sumsq_out), _ptr(scratch_buf),
                                _stream(stream)), "nd_outer_fused")


def outer_fused_rw(master, mom, reads: Sequence, writes: Sequence, *, lr: float, momentum: float,
                   dampening: float = 0.0, weight_decay: float = 0.0, nesterov: bool = True,
                   first_step: bool = False, stream=None) -> None:
    """nd_outer_fused_rw: average over `reads` (one copy per worker), write the result to `writes`."""
    _require_cuda(master, mom, *reads, *writes)
    ra = (_c_p * len(reads))(*[_ptr(t) for t in reads])
    wa = (_c_p * max(1, len(writes)))(*[_ptr(t) for t in writes])
    check(load().nd_outer_fused_rw(_ptr(master), _ptr(mom), ra, len(reads), wa, len(writes), _dtype_code(reads[0]),
                                   master.numel(), lr, momentum, dampening, weight_decay, int(nesterov),
                                   int(first_step), None, None, _stream(stream)), "nd_outer_fused_rw")


def sumsq(g, out, scratch_buf, stream=None) -> None:
    _require_cuda(g, out, scratch_buf)
    check(load().nd_sumsq(_ptr(g), _dtype_code(g), g.numel(), _ptr(out), _ptr(scratch_buf), _stream(stream)),
          "nd_sumsq")


def sumsq_static(g, out, stream=None) -> None:
    """nd_sumsq_static: same result as `sumsq`, scratch inside the library, nothing allocated."""
    _require_cuda(g, out)
    check(load().nd_sumsq_static(_ptr(g), _dtype_code(g), g.numel(), _ptr(out), _stream(stream)),
          "nd_sumsq_static")


def adamw(p, g, m, v, *, lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
          weight_decay: float = 1e-2, step: int, sumsq_dev=None, max_norm: float = 1.0,
          zero_grad: bool = False, stream=None) -> None:
    _require_cuda(p, g, m, v, sumsq_dev)
    if _dtype_code(p) == ND_BF16:
        if _dtype_code(g) != ND_BF16:
            raise NDError("a bf16 working copy needs bf16 gradients")
        check(load().nd_adamw_bf16(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), lr, beta1, beta2, eps,
                                   weight_decay, int(step), _ptr(sumsq_dev), max_norm, int(zero_grad),
                                   _stream(stream)), "nd_adamw_bf16")
        return
    check(load().nd_adamw(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), lr, beta1, beta2, eps, weight_decay,
                          int(step), _ptr(sumsq_dev), max_norm, int(zero_grad), _stream(stream)), "nd_adamw")


def make_table(rows, device):
    """Device multi-tensor table (nd_tensor_desc[]) from rows of (tensors..., ) with equal numel per row.

    Returns (table_tensor, n_tensors, total_chunks).  Each row lists up to four tensors."""
    import torch
    flat = []
    chunk = 0
    for row in rows:
        n = row[0].numel()
        ptrs = [t.data_ptr() for t in row] + [0] * (4 - len(row))
        flat.extend(ptrs + [n, chunk])
        chunk += (n + ND_CHUNK - 1) // ND_CHUNK
    t = torch.tensor(flat, dtype=torch.int64).reshape(-1, 6) if flat else torch.zeros((0, 6), dtype=torch.int64)
    return t.to(device), len(rows), chunk
