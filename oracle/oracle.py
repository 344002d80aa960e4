"""ctypes front-end of oracle/libnd_oracle.so (nd_oracle.c) over numpy arrays.

TEST INFRASTRUCTURE ONLY -- see nd_oracle.c.  Parity status: the reference has
no tests or golden vectors of its own; this oracle is pinned against fixtures
generated from the reference's `Diloco` class (tests/golden/make_golden.py) and
against torch's CPU optimizers (tests/test_oracle_golden.py).

Every function works in place on C-contiguous float32 numpy arrays.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libnd_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile nd_oracle.c with gcc (seconds) when the library is missing (or `force`).  Built into a
    temporary file and renamed into place under a lock.  Building the checker is not using it."""
    import fcntl
    if not force and os.path.exists(_SO):
        return _SO
    with open(_SO + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if force or not os.path.exists(_SO):
            tmp = f"{_SO}.tmp.{os.getpid()}"
            subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fopenmp", "-Wall", "-Wextra",
                                   "-o", tmp, os.path.join(_HERE, "nd_oracle.c"), "-lm"])
            os.replace(tmp, _SO)
        fcntl.flock(lock, fcntl.LOCK_UN)
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        f32p = ctypes.POINTER(ctypes.c_float)
        i64 = ctypes.c_int64
        L.ndo_version.restype = ctypes.c_int
        L.ndo_max_threads.restype = ctypes.c_int
        L.ndo_set_threads.argtypes = [ctypes.c_int]
        L.ndo_pseudo_grad.argtypes = [f32p, f32p, f32p, i64]
        L.ndo_average.argtypes = [ctypes.POINTER(f32p), ctypes.c_int, f32p, i64]
        L.ndo_sgd_step.argtypes = [f32p, f32p, f32p, i64] + [ctypes.c_float] * 4 + [ctypes.c_int] * 2
        L.ndo_outer_step.argtypes = ([f32p, ctypes.POINTER(f32p), ctypes.c_int, f32p, f32p, i64]
                                     + [ctypes.c_float] * 4 + [ctypes.c_int] * 2)
        L.ndo_sumsq.argtypes = [f32p, i64]
        L.ndo_sumsq.restype = ctypes.c_double
        L.ndo_clip_coef.argtypes = [ctypes.c_float, ctypes.c_float]
        L.ndo_clip_coef.restype = ctypes.c_float
        L.ndo_adamw_step.argtypes = ([f32p, f32p, f32p, f32p, i64] + [ctypes.c_double] * 5
                                     + [i64, ctypes.c_float, ctypes.c_int])
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags.c_contiguous, (a.dtype, a.flags)
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _pp(arrs: Sequence[np.ndarray]):
    T = ctypes.POINTER(ctypes.c_float) * len(arrs)
    return T(*[_p(a) for a in arrs])


def set_threads(n: int) -> None:
    lib().ndo_set_threads(int(n))


def max_threads() -> int:
    return int(lib().ndo_max_threads())


def pseudo_grad(master: np.ndarray, local: np.ndarray) -> np.ndarray:
    """diloco.py:48"""
    out = np.empty_like(master)
    lib().ndo_pseudo_grad(_p(master), _p(local), _p(out), master.size)
    return out


def average(pgs: Sequence[np.ndarray]) -> np.ndarray:
    """diloco.py:49 (sum in worker order, / W)"""
    out = np.empty_like(pgs[0])
    lib().ndo_average(_pp(pgs), len(pgs), _p(out), out.size)
    return out


def sgd_step(p, buf, grad, *, lr, momentum=0.9, dampening=0.0, weight_decay=0.0,
             nesterov=True, first_step=False) -> None:
    """diloco.py:52 / torch:optim/sgd.py:343-380, in place on p and buf."""
    lib().ndo_sgd_step(_p(p), _p(buf), _p(grad), p.size, lr, momentum, dampening,
                       weight_decay, int(nesterov), int(first_step))


def outer_step(master, locals_, buf, *, lr, momentum=0.9, dampening=0.0, weight_decay=0.0,
               nesterov=True, first_step=False) -> None:
    """diloco.py:46-54 for W workers in one address space, in place."""
    scratch = np.empty_like(master)
    lib().ndo_outer_step(_p(master), _pp(locals_), len(locals_), _p(buf), _p(scratch),
                         master.size, lr, momentum, dampening, weight_decay,
                         int(nesterov), int(first_step))


def sumsq(g: np.ndarray) -> float:
    return float(lib().ndo_sumsq(_p(g), g.size))


def clip_coef(total_norm: float, max_norm: float = 1.0) -> float:
    """clip_grad.py:165-168"""
    return float(lib().ndo_clip_coef(total_norm, max_norm))


def adamw_step(p, grad, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2,
               step: int, clip: float = 1.0, recipe: str = "cuda") -> None:
    """diloco.py:57-58 / torch:optim/adam.py:692-800, in place on p, m, v.

    recipe "cpu" follows torch's CPU rounding (pinned by tests/golden/inner.npz),
    "cuda" torch's CUDA foreach rounding (what the product kernel follows)."""
    lib().ndo_adamw_step(_p(p), _p(grad), _p(m), _p(v), p.size, lr, beta1, beta2, eps,
                         weight_decay, int(step), clip, {"cpu": 0, "cuda": 1}[recipe])


def cosine_lr(base_lr: float, step: int, warmup: int, total: int) -> float:
    """transformers.get_cosine_schedule_with_warmup (diloco.py:20), num_cycles=0.5.

    `step` is the number of scheduler.step() calls made so far."""
    import math
    if step < warmup:
        return base_lr * float(step) / float(max(1, warmup))
    progress = float(step - warmup) / float(max(1, total - warmup))
    return base_lr * max(0.0, 0.5 * (1.0 + math.cos(math.pi * 0.5 * 2.0 * progress)))
