"""CPU: argument validation at the C ABI.  Every entry point checks its arguments before it touches CUDA, returns
an ND_E* code (never throws across the boundary) and leaves a message in nd_last_error() -- the error behaviour
SURVEY.md section 8(b) asks of the boundary.  No kernel is launched here, so this runs without a GPU."""
import ctypes

import pytest

from nanodiloco_b200 import _lib

ND_EINVAL, ND_ESTATE = 1, 4


@pytest.fixture(scope="module")
def L():
    return _lib.load()


def _err(L):
    return (L.nd_last_error() or b"").decode()


def test_static_scratch_entry_point_refuses_sizes_beyond_its_limit(L):
    rc = L.nd_sumsq_static(None, _lib.ND_F32, (1 << 31) + 1, None, None)
    assert rc == ND_EINVAL and "ND_SUMSQ_STATIC_MAX_ELEMS" in _err(L)
    rc = L.nd_sumsq_static(None, _lib.ND_F32, -1, None, None)
    assert rc == ND_EINVAL and "n < 0" in _err(L)


def test_kernel_entry_points_reject_bad_arguments(L):
    one = (ctypes.c_void_p * 1)(0)
    assert L.nd_pseudo_grad(None, None, 7, None, 0, None) == ND_EINVAL and "local_dtype" in _err(L)
    assert L.nd_pseudo_grad(None, None, _lib.ND_F32, None, 16, None) == ND_EINVAL and "null pointer" in _err(L)
    assert L.nd_outer_fused(None, None, one, 0, _lib.ND_F32, 0, 0.7, 0.9, 0.0, 0.0, 1, 0, None, None, None) == ND_EINVAL
    assert "W out of range" in _err(L)
    assert L.nd_outer_fused(None, None, one, 17, _lib.ND_F32, 0, 0.7, 0.9, 0.0, 0.0, 1, 0, None, None, None) == ND_EINVAL
    assert L.nd_outer_nesterov(None, None, None, None, _lib.ND_F32, 0, 0.5, 0.7, 0.9, 0.0, 0.0, 1, 0, None, None,
                               None) == ND_EINVAL and "world < 1" in _err(L)
    assert L.nd_adamw(None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, 0.01, 0, None, 1.0, 0, None) == ND_EINVAL
    assert "step must be >= 1" in _err(L)
    assert L.nd_adamw_mt(None, 3, 1, _lib.ND_F32, 1e-3, 0.9, 0.999, 1e-8, 0.01, 1, None, 1.0, 0, None) == ND_EINVAL
    assert "null table" in _err(L)
    assert L.nd_adamw_mt(None, 0, 0, 5, 1e-3, 0.9, 0.999, 1e-8, 0.01, 1, None, 1.0, 0, None) == ND_EINVAL
    assert L.nd_sumsq_mt(None, 0, 0, _lib.ND_F32, None, None, None) == ND_EINVAL      # out and scratch are required


def test_replica_entry_points_reject_bad_arguments(L):
    ptrs = (ctypes.c_void_p * 9)(*([0] * 9))
    assert L.nd_grad_avg(ptrs, 9, None, _lib.ND_F32, 0, None, None, None) == ND_EINVAL and "R out of range" in _err(L)
    assert L.nd_grad_avg(ptrs, 0, None, _lib.ND_F32, 0, None, None, None) == ND_EINVAL
    assert L.nd_adamw_rw(None, ptrs, 9, None, None, None, _lib.ND_F32, 0, 1e-3, 0.9, 0.999, 1e-8, 0.01, 1, ptrs, 0, 1.0,
                         0, None) == ND_EINVAL and "bad write set" in _err(L)
    assert L.nd_adamw_rw(None, ptrs, 0, None, None, None, _lib.ND_F32, 0, 1e-3, 0.9, 0.999, 1e-8, 0.01, 0, ptrs, 0, 1.0,
                         0, None) == ND_EINVAL and "step" in _err(L)
    assert L.nd_inner_step_replicas(None, None, None, None, _lib.ND_F32, 0, 1e-3, 0.9, 0.999, 1e-8, 0.01, 1, 1.0, None,
                                    None) == ND_ESTATE and "null context" in _err(L)


def test_context_entry_points_reject_null_contexts_and_bad_shapes(L):
    h = ctypes.c_void_p()
    assert L.nd_peer_ctx_create(0, 0, None, 0, None, ctypes.byref(h)) == ND_EINVAL and h.value is None
    assert L.nd_peer_ctx_create(17, 0, None, 0, None, ctypes.byref(h)) == ND_EINVAL and "ND_MAX_WORKERS" in _err(L)
    assert L.nd_peer_ctx_create(2, 2, None, 0, None, ctypes.byref(h)) == ND_EINVAL
    assert L.nd_sync_ctx_create(None, 0, 0, 1024, 2, ctypes.byref(h)) == ND_EINVAL and h.value is None
    assert L.nd_sync_ctx_create(None, 2, 0, 1024, 2, ctypes.byref(h)) == ND_EINVAL and "unique id" in _err(L)
    assert L.nd_sync_ctx_create(None, 1, 0, 0, 2, ctypes.byref(h)) == ND_EINVAL
    for fn, args in ((L.nd_peer_set_mode, (None, 1)), (L.nd_peer_set_fold, (None, 1)), (L.nd_peer_set_staging, (None, 0, 0)),
                     (L.nd_peer_set_replicas, (None, 2)), (L.nd_peer_barrier, (None, None)),
                     (L.nd_sync_set_profiling, (None, 1)), (L.nd_sync_ctx_add_ag_comm, (None, None))):
        assert fn(*args) == ND_ESTATE, fn
        assert "null context" in _err(L)
    n = ctypes.c_int64(-1)
    assert L.nd_sync_timeline(None, None, 0, ctypes.byref(n)) == ND_EINVAL
    ms = ctypes.c_float()
    assert L.nd_peer_last_ms(None, ctypes.byref(ms)) == ND_EINVAL
    assert L.nd_outer_sync_p2p(None, None, None, _lib.ND_F32, 0, 0.7, 0.9, 0.0, 0.0, 1, 0, None, None, None) == ND_ESTATE
    assert L.nd_outer_sync(None, None, None, None, None, 0, 0.7, 0.9, 0.0, 0.0, 1, 0, None) == ND_ESTATE
    assert L.nd_peer_ctx_destroy(None) == 0 and L.nd_sync_ctx_destroy(None) == 0      # destroying nothing is fine
    assert L.nd_nccl_mem_free(None) == 0
