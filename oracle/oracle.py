"""ctypes front end of oracle/liboracle.so (store_oracle.c) -- TEST INFRASTRUCTURE ONLY.

`OracleStore` mirrors the reference's `Store` methods
(/root/reference/backend/go/stores/store.go:106,200,266,477) so that parity tests
can drive the oracle and the CUDA-backed store with the same calls.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")


def build(force: bool = False) -> str:
    """Compile store_oracle.c with the recipe in oracle/Makefile."""
    src = os.path.join(_HERE, "store_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        fp = C.POINTER(C.c_float)
        i64p = C.POINTER(C.c_int64)
        L.orc_new.restype = C.c_void_p
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_last_error.restype = C.c_char_p
        L.orc_last_error.argtypes = [C.c_void_p]
        L.orc_len.restype = C.c_int64
        L.orc_len.argtypes = [C.c_void_p]
        L.orc_key_len.argtypes = [C.c_void_p]
        L.orc_keys_are_normalized.argtypes = [C.c_void_p]
        L.orc_row.argtypes = [C.c_void_p, C.c_int64, fp, i64p]
        L.orc_is_normalized.argtypes = [fp, C.c_int]
        L.orc_dot_f32.restype = C.c_float
        L.orc_dot_f32.argtypes = [fp, fp, C.c_int]
        L.orc_query_mag.restype = C.c_double
        L.orc_query_mag.argtypes = [fp, C.c_int]
        L.orc_cosine.restype = C.c_float
        L.orc_cosine.argtypes = [fp, fp, C.c_int, C.c_double]
        L.orc_set.argtypes = [C.c_void_p, fp, C.c_int64, C.c_int, i64p, C.c_int64]
        L.orc_delete.argtypes = [C.c_void_p, fp, C.c_int64, C.c_int]
        L.orc_get.argtypes = [C.c_void_p, fp, C.c_int64, C.c_int, fp, i64p, i64p]
        L.orc_find.argtypes = [C.c_void_p, fp, C.c_int, C.c_int, C.c_int, i64p, fp, i64p, C.POINTER(C.c_int)]
        L.orc_scan_block.argtypes = [fp, C.c_int64, C.c_int, fp, C.c_int, C.c_int, i64p, fp, i64p]
        L.orc_scores_block.argtypes = [fp, C.c_int64, C.c_int, fp, C.c_int, fp]
        _lib = L
    return _lib


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(C.POINTER(C.c_int64))


class OracleError(Exception):
    pass


class OraclePanic(Exception):
    """The reference would crash here (e.g. StoresGet with no keys, store.go:276)."""


def _as_rows(keys):
    """Rows of equal length -> (2-D fp32 array, dim).  dim of an empty list is 0."""
    if isinstance(keys, np.ndarray) and keys.ndim == 2:
        return np.ascontiguousarray(keys, dtype=np.float32), int(keys.shape[1])
    keys = [np.asarray(k, dtype=np.float32).reshape(-1) for k in keys]
    if not keys:
        return np.zeros((0, 0), np.float32), 0
    d = len(keys[0])
    if any(len(k) != d for k in keys):
        raise ValueError("the oracle restates store.go for equal-length keys only")
    return np.ascontiguousarray(np.stack(keys), dtype=np.float32), d


class OracleStore:
    """store.go `Store`, restated.  Values are bytes; the C side keeps int64 handles."""

    def __init__(self):
        self._L = lib()
        self._h = C.c_void_p(self._L.orc_new())
        self._vals = {}
        self._next = 0

    def __del__(self):
        try:
            self._L.orc_free(self._h)
        except Exception:
            pass

    # -- introspection --------------------------------------------------
    def __len__(self):
        return int(self._L.orc_len(self._h))

    @property
    def key_len(self):
        return int(self._L.orc_key_len(self._h))

    @property
    def keys_are_normalized(self):
        return bool(self._L.orc_keys_are_normalized(self._h))

    def row(self, i):
        k = np.empty(self.key_len, np.float32)
        v = C.c_int64()
        self._L.orc_row(self._h, i, k.ctypes.data_as(C.POINTER(C.c_float)), C.byref(v))
        return k, self._vals[v.value]

    def _err(self):
        return self._L.orc_last_error(self._h).decode()

    # -- the four store RPCs --------------------------------------------
    def StoresSet(self, keys, values):
        ks, d = _as_rows(keys)
        handles = []
        for v in values:
            self._vals[self._next] = bytes(v)
            handles.append(self._next)
            self._next += 1
        ha, hp = _i64(handles)
        _, kp = _f32(ks)
        rc = self._L.orc_set(self._h, kp, len(ks), d, hp, len(handles))
        if rc:
            raise OracleError(self._err())

    def StoresDelete(self, keys):
        ks, d = _as_rows(keys)
        _, kp = _f32(ks)
        rc = self._L.orc_delete(self._h, kp, len(ks), d)
        if rc:
            raise OracleError(self._err())

    def StoresGet(self, keys):
        ks, d = _as_rows(keys)
        _, kp = _f32(ks)
        ok = np.empty((max(len(ks), 1), max(d, 1)), np.float32)
        ov = np.empty(max(len(ks), 1), np.int64)
        n = C.c_int64()
        rc = self._L.orc_get(self._h, kp, len(ks), d, ok.ctypes.data_as(C.POINTER(C.c_float)),
                             ov.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(n))
        if rc == 2:
            raise OraclePanic(self._err())
        if rc:
            raise OracleError(self._err())
        m = n.value
        return [ok[i, :d].copy() for i in range(m)], [self._vals[int(ov[i])] for i in range(m)]

    def StoresFind(self, key, topk, faithful=False):
        q, qp = _f32(np.asarray(key, np.float32).reshape(-1))
        cap = max(1, min(int(topk) if topk > 0 else 1, len(self)))
        rows = np.empty(cap, np.int64)
        sims = np.empty(cap, np.float32)
        n = C.c_int64()
        path = C.c_int()
        rc = self._L.orc_find(self._h, qp, len(q), int(topk), int(faithful),
                              rows.ctypes.data_as(C.POINTER(C.c_int64)),
                              sims.ctypes.data_as(C.POINTER(C.c_float)), C.byref(n), C.byref(path))
        if rc:
            raise OracleError(self._err())
        m = n.value
        self.last_path = "normalized" if path.value else "fallback"
        ks, vs = [], []
        for i in range(m):
            k, v = self.row(int(rows[i]))
            ks.append(k)
            vs.append(v)
        return ks, vs, sims[:m].copy()


# -- block helpers (bulk parity and the timed CPU baseline) ----------------

def is_normalized(k):
    a, p = _f32(np.asarray(k, np.float32).reshape(-1))
    return bool(lib().orc_is_normalized(p, len(a)))


def scan_block(keys, q, topk, normalized):
    """Top-k over a contiguous [n, d] block in the reference arithmetic, scan order = row order."""
    ks, kp = _f32(keys)
    qa, qp = _f32(np.asarray(q, np.float32).reshape(-1))
    n, d = ks.shape
    cap = max(1, min(topk, n))
    rows = np.empty(cap, np.int64)
    sims = np.empty(cap, np.float32)
    m = C.c_int64()
    lib().orc_scan_block(kp, n, d, qp, int(topk), int(bool(normalized)),
                         rows.ctypes.data_as(C.POINTER(C.c_int64)),
                         sims.ctypes.data_as(C.POINTER(C.c_float)), C.byref(m))
    return rows[:m.value].copy(), sims[:m.value].copy()


def scores_block(keys, q, normalized):
    ks, kp = _f32(keys)
    qa, qp = _f32(np.asarray(q, np.float32).reshape(-1))
    n, d = ks.shape
    out = np.empty(n, np.float32)
    lib().orc_scores_block(kp, n, d, qp, int(bool(normalized)), out.ctypes.data_as(C.POINTER(C.c_float)))
    return out
