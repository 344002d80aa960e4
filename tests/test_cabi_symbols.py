"""CPU: libnanodiloco_b200.so loads and exports every symbol include/nanodiloco_b200.h declares.
No compute call is made (there is no GPU here); argument validation happens before any launch."""
import ctypes
import os
import re

import pytest

from nanodiloco_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "nanodiloco_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nd_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    names = declared_functions()
    for must in ("nd_pseudo_grad", "nd_outer_nesterov", "nd_outer_fused", "nd_sumsq", "nd_adamw",
                 "nd_outer_sync", "nd_sync_ctx_create", "nd_last_error", "nd_version"):
        assert must in names


def test_library_builds_and_exports_every_declared_symbol():
    so = build.build()
    assert os.path.exists(so)
    L = ctypes.CDLL(so)
    missing = [n for n in declared_functions() if not hasattr(L, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_binding_covers_every_declared_symbol():
    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_version_and_scratch_size():
    L = _lib.load()
    assert L.nd_version() >= 200
    assert 0 < L.nd_scratch_bytes() < (1 << 20)


def test_argument_errors_are_reported_not_thrown():
    L = _lib.load()
    rc = L.nd_pseudo_grad(None, None, 0, None, -1, None)
    assert rc == 1  # ND_EINVAL
    assert b"n < 0" in L.nd_last_error()
    rc = L.nd_outer_fused(None, None, None, 99, 0, 0, 0.7, 0.9, 0.0, 0.0, 1, 0, None, None, None)
    assert rc == 1 and b"ND_MAX_WORKERS" in L.nd_last_error()
    rc = L.nd_adamw(None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, 0.01, 0, None, 1.0, 0, None)
    assert rc == 1 and b"step" in L.nd_last_error()
    with pytest.raises(_lib.NDError):
        _lib.check(rc, "nd_adamw")


def test_sync_ctx_rejects_bad_geometry():
    L = _lib.load()
    h = ctypes.c_void_p()
    assert L.nd_sync_ctx_create(None, 2, 5, 1024, 2, ctypes.byref(h)) == 1
    assert L.nd_sync_ctx_create(None, 0, 0, 1024, 2, ctypes.byref(h)) == 1
    assert L.nd_sync_padded_elems(None, 10) == -1
