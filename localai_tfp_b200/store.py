"""Host-side mirror of the reference `Store` (backend/go/stores/store.go) over the C ABI.

This is what the Go host (go/local-store/store_gpu.go) does, written in Python
because the build image has no Go toolchain: it keeps `id -> value bytes`, calls
libtfpstore.so for everything that touches keys, and re-raises the library's
validation errors with the reference's exact texts.

Method names and argument meaning follow store.go:
    StoresSet(keys, values)           store.go:106-198
    StoresDelete(keys)                store.go:200-264
    StoresGet(keys) -> (keys, values) store.go:266-311  (sorted-key order)
    StoresFind(key, topk) -> (keys, values, similarities)  store.go:477-511
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import List, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import TfpCfg, TfpStats, fp, i32p, i64p, u8p


def nccl_unique_id() -> bytes:
    """A fresh 128-byte NCCL unique id (rank 0 makes it, every rank passes it to comm_init)."""
    L = _lib.load()
    buf = C.create_string_buffer(128)
    if L.tfp_nccl_unique_id(buf, 128) != 0:
        raise RuntimeError(L.tfp_last_error().decode())
    return buf.raw


class StoreError(Exception):
    """An error the reference would also return (same message text)."""

    def __init__(self, msg: str, code: int = 1):
        super().__init__(msg)
        self.code = code


def _rows(keys) -> Tuple[np.ndarray, int]:
    """Pack keys into one contiguous [n, dim] fp32 block (the Go host does the same gather copy).

    Deviation from the reference, documented in DESIGN.md: a ragged batch is rejected here; the
    reference only checks the first key's length (store.go:115-121) and would corrupt its order."""
    if isinstance(keys, np.ndarray) and keys.ndim == 2:
        return np.ascontiguousarray(keys, dtype=np.float32), int(keys.shape[1])
    keys = [np.asarray(k, dtype=np.float32).reshape(-1) for k in keys]
    if not keys:
        return np.zeros((0, 1), np.float32), 0
    d = len(keys[0])
    for k in keys:
        if len(k) != d:
            raise StoreError(f"Try to add key with length {len(k)} when existing length is {d}")
    return np.ascontiguousarray(np.stack(keys), dtype=np.float32), d


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(t)


def _addr(a: np.ndarray) -> int:
    """Address of a contiguous array (several times cheaper than a.ctypes on the per-query path)."""
    return a.__array_interface__["data"][0]


class Store:
    """One store resident on one B200 -- or, with `devices=[...]`, ONE logical store over several B200s of
    this box inside this process (tfp_store_open_devices): same methods, same results."""

    def __init__(self, device: int = 0, capacity_rows: int = 0, id_base: int = 0, devices: Sequence[int] | None = None):
        self._L = _lib.load()
        h = C.c_void_p()
        if devices is not None:
            devs = [int(d) for d in devices]
            if not devs:
                raise ValueError("devices must name at least one GPU")
            device = devs[0]
            cfg = TfpCfg(device=device, capacity_rows=capacity_rows, flags=0, reserved=0, id_base=id_base)
            arr = (C.c_int * len(devs))(*devs)
            rc = self._L.tfp_store_open_devices(C.byref(cfg), arr, len(devs), C.byref(h))
        else:
            cfg = TfpCfg(device=device, capacity_rows=capacity_rows, flags=0, reserved=0, id_base=id_base)
            rc = self._L.tfp_store_open(C.byref(cfg), C.byref(h))
        if rc:
            raise RuntimeError(f"tfp_store_open failed ({rc}): {self._err()}")
        self._h = h
        self.device = device
        self.devices = list(devices) if devices is not None else [device]
        self.values = {}  # id -> bytes; values never cross the C ABI (store.go:23)
        self._set_mu = threading.Lock()   # held by a StoresSet from its library call until its values are in the map
        self._every_id_has_a_value = True  # False once rows were loaded without values (set_device(values=None))
        self.last_path = None

    # -- plumbing ----------------------------------------------------------
    def _err(self) -> str:
        return self._L.tfp_last_error().decode(errors="replace")

    def _check(self, rc: int):
        if rc > 0:
            raise StoreError(self._err(), rc)
        if rc < 0:
            raise RuntimeError(f"libtfpstore failure ({rc}): {self._err()}")

    def close(self):
        if getattr(self, "_h", None):
            self._L.tfp_store_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stats(self) -> TfpStats:
        st = TfpStats()
        self._check(self._L.tfp_store_stats(self._h, C.byref(st)))
        return st

    @property
    def shard_count(self) -> int:
        return int(self._L.tfp_store_shard_count(self._h))

    def shard_stats(self, shard: int) -> TfpStats:
        st = TfpStats()
        self._check(self._L.tfp_store_shard_stats(self._h, shard, C.byref(st)))
        return st

    def last_error(self) -> str:
        """Text of the last failing call on this handle, whichever thread made it."""
        buf = C.create_string_buffer(512)
        self._L.tfp_store_last_error(self._h, buf, 512)
        return buf.value.decode(errors="replace")

    def __len__(self):
        return int(self.stats().rows)

    @property
    def key_len(self) -> int:
        return int(self.stats().dim)

    @property
    def keys_are_normalized(self) -> bool:
        return bool(self.stats().keys_are_normalized)

    # -- StoresSet -----------------------------------------------------------
    def StoresSet(self, keys, values: Sequence[bytes]):
        ks, d = _rows(keys)
        n = len(ks)
        ids = np.empty(max(n, 1), np.int64)
        new = np.empty(max(n, 1), np.uint8)
        with self._set_mu:      # finds are not held up by this lock; see _values_of
            self._check(self._L.tfp_store_set(self._h, _p(ks, fp), n, d, len(values), _p(ids, i64p), _p(new, u8p)))
            for i in range(n):
                self.values[int(ids[i])] = bytes(values[i])
        return ids[:n], new[:n]

    def _values_of(self, ids):
        """Values of the ids a find returned.  An append publishes its rows to the finds before this host has put the
        values into the map (finds never wait for an append), so a find can come back with an id whose value is a few
        microseconds away: only then does the find wait for the StoresSet in progress to finish."""
        vs = [self.values.get(i) for i in ids]
        if self._every_id_has_a_value and any(v is None for v in vs):
            with self._set_mu:
                pass
            vs = [self.values.get(i) for i in ids]
        return vs

    def set_device(self, d_ptr: int, n: int, dim: int, values: Sequence[bytes] | None = None):
        """Bulk load rows that already sit in this GPU's HBM (d_ptr = device address of [n, dim] fp32)."""
        ids = np.empty(max(n, 1), np.int64)
        new = np.empty(max(n, 1), np.uint8)
        self._check(self._L.tfp_store_set_device(self._h, C.c_void_p(d_ptr), n, dim, _p(ids, i64p), _p(new, u8p)))
        if values is not None:
            for i in range(n):
                self.values[int(ids[i])] = bytes(values[i])
        else:
            self._every_id_has_a_value = False
        return ids[:n], new[:n]

    # -- StoresDelete --------------------------------------------------------
    def StoresDelete(self, keys):
        ks, d = _rows(keys)
        n = len(ks)
        ids = np.empty(max(n, 1), np.int64)
        with self._set_mu:
            self._check(self._L.tfp_store_delete(self._h, _p(ks, fp), n, d, _p(ids, i64p)))
            for i in range(n):
                if ids[i] >= 0:
                    self.values.pop(int(ids[i]), None)
        return ids[:n]

    # -- StoresGet -----------------------------------------------------------
    def StoresGet(self, keys) -> Tuple[List[np.ndarray], List[bytes]]:
        ks, d = _rows(keys)
        n = len(ks)
        ids = np.empty(max(n, 1), np.int64)
        order = np.empty(max(n, 1), np.int64)
        nf = C.c_int64()
        self._check(self._L.tfp_store_get(self._h, _p(ks, fp), n, d, _p(ids, i64p), _p(order, i64p), C.byref(nf)))
        hit = [int(order[j]) for j in range(nf.value)]
        out_k = [ks[i].copy() for i in hit]  # the reference echoes the request's key (store.go:289-291)
        return out_k, self._values_of([int(ids[i]) for i in hit])

    # -- StoresFind ----------------------------------------------------------
    def StoresFind(self, key, topk: int):
        q = np.ascontiguousarray(np.asarray(key, dtype=np.float32).reshape(-1))
        k = int(topk)
        kk = max(k, 1)
        if kk > 1024:               # (a small top-k needs no look at the row count: the buffers are cheap either way)
            kk = min(kk, max(len(self), 1))  # the result holds min(topk, rows) entries (store.go:386-388)
        n = C.c_int()
        path = C.c_int(-1)
        ids = np.empty(kk, np.int64)
        sims = np.empty(kk, np.float32)
        outk = np.empty((kk, max(len(q), 1)), np.float32)
        for attempt in range(4):
            self._check(self._L.tfp_store_find(self._h, _addr(q), len(q), k if k < 1 else kk, C.byref(n), _addr(ids),
                                               _addr(sims), _addr(outk), C.byref(path)))
            m = n.value
            values = self._values_of(ids[:m].tolist())
            # A row deleted between the find and the look-up has no value any more: answer from the state after the
            # delete instead of handing out a row without its value.
            if not self._every_id_has_a_value or all(v is not None for v in values):
                break
        self.last_path = "normalized" if path.value == 1 else "fallback"
        # outk / sims are this call's own buffers: hand out views, not copies
        return (list(outk[:m, :len(q)]), values, sims[:m])

    def find_ids(self, key, topk: int):
        """StoresFind without the key echo: (ids, similarities)."""
        q = np.ascontiguousarray(np.asarray(key, dtype=np.float32).reshape(-1))
        kk = min(max(int(topk), 1), max(len(self), 1))
        n = C.c_int()
        ids = np.empty(kk, np.int64)
        sims = np.empty(kk, np.float32)
        self._check(self._L.tfp_store_find(self._h, _addr(q), len(q), int(topk) if topk < 1 else kk, C.byref(n),
                                           _addr(ids), _addr(sims), None, None))
        return ids[:n.value], sims[:n.value]

    def find_batch(self, queries, topk: int, with_keys: bool = False):
        qs = np.ascontiguousarray(np.asarray(queries, dtype=np.float32))
        nq, d = qs.shape
        kk = min(max(int(topk), 1), max(len(self), 1))
        n = np.empty(nq, np.int32)
        ids = np.empty((nq, kk), np.int64)
        sims = np.empty((nq, kk), np.float32)
        outk = np.empty((nq, kk, d), np.float32) if with_keys else None
        self._check(self._L.tfp_store_find_batch(self._h, _p(qs, fp), nq, d, int(topk) if topk < 1 else kk,
                                                 _p(n, i32p), _p(ids, i64p), _p(sims, fp),
                                                 _p(outk, fp) if with_keys else None))
        res = [(ids[i, :n[i]].copy(), sims[i, :n[i]].copy()) for i in range(nq)]
        return (res, outk) if with_keys else res

    # -- snapshot / restore ---------------------------------------------------------
    # The library writes keys, ids and flags to `path` (atomically: tmp + fsync + rename).  The values belong to
    # the host; they go to `path.values.<tag>`, a flat little-endian file a Go host reads just as easily:
    #     b"TFPVALS1" | u64 tag | u64 count | count x (i64 id | u32 len | bytes)
    # A save writes the values under a NEW tag first, then the keys file carrying that tag (its rename is the
    # commit point), then removes older values files -- so a kill at any moment leaves a matching pair.
    def save(self, path: str):
        import glob
        import struct
        tag = int.from_bytes(os.urandom(8), "little") | 1
        vpath = f"{path}.values.{tag:016x}"
        tmp = vpath + ".tmp"
        with open(tmp, "wb") as f:
            f.write(b"TFPVALS1" + struct.pack("<QQ", tag, len(self.values)))
            for i, v in self.values.items():
                v = b"" if v is None else bytes(v)
                f.write(struct.pack("<qI", int(i), len(v)))
                f.write(v)
            f.flush()
            os.fsync(f.fileno())
        os.replace(tmp, vpath)
        self._check(self._L.tfp_store_save_tagged(self._h, os.fsencode(path), tag))
        for old in glob.glob(glob.escape(path) + ".values.*"):
            if old != vpath:
                try:
                    os.remove(old)
                except OSError:
                    pass

    def load(self, path: str):
        """Fill this (empty) store from a snapshot written by save().  Raises (and leaves the values empty) when
        the keys file and the values file do not belong together."""
        import struct
        tag = C.c_uint64()
        self._check(self._L.tfp_snapshot_tag(os.fsencode(path), C.byref(tag)))
        vpath = f"{path}.values.{tag.value:016x}"
        with open(vpath, "rb") as f:
            head = f.read(24)
            if len(head) != 24 or head[:8] != b"TFPVALS1":
                raise StoreError(f"{vpath} is not a values file", 2)
            vtag, count = struct.unpack("<QQ", head[8:])
            if vtag != tag.value:
                raise StoreError(f"{vpath} carries tag {vtag:x}, the snapshot {tag.value:x}", 2)
            values = {}
            for _ in range(count):
                rec = f.read(12)
                if len(rec) != 12:
                    raise StoreError(f"{vpath} is truncated", 2)
                i, n = struct.unpack("<qI", rec)
                v = f.read(n)
                if len(v) != n:
                    raise StoreError(f"{vpath} is truncated", 2)
                values[i] = v
        self._check(self._L.tfp_store_load(self._h, os.fsencode(path)))
        self.values = values
        self._every_id_has_a_value = len(values) == len(self)

    def mark_not_normalized(self):
        self._check(self._L.tfp_store_mark_not_normalized(self._h))

    def find_context(self) -> "FindContext":
        return FindContext(self)

    def mma_counters(self):
        """(tensor-core passes, queries escalated to the 3xTF32 filter, queries re-run exactly) of the built-in
        context that serves find_batch."""
        a, b = C.c_int64(), C.c_int64()
        ctx = C.c_void_p(self._L.tfp_store_default_ctx(self._h))
        self._check(self._L.tfp_find_ctx_mma_counters(ctx, C.byref(a), C.byref(b)))
        return a.value, int(self._L.tfp_find_ctx_mma_escalated(ctx)), b.value


class FindContext:
    """Asynchronous, device-resident finds (include/tfp_store.h `tfp_find_ctx_*`)."""

    def __init__(self, store: Store):
        self.store = store
        self._L = store._L
        h = C.c_void_p()
        store._check(self._L.tfp_find_ctx_create(store._h, C.byref(h)))
        self._h = h
        self.dim = 0
        self.nq = 0
        self._count = 0
        self._topk = 0

    def close(self):
        if getattr(self, "_h", None) and getattr(self.store, "_h", None):
            self._L.tfp_find_ctx_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, queries):
        qs = np.ascontiguousarray(np.asarray(queries, dtype=np.float32))
        if qs.ndim == 1:
            qs = qs[None, :]
        self.nq, self.dim = qs.shape
        self.store._check(self._L.tfp_find_ctx_upload(self._h, _addr(qs), self.nq, self.dim))

    def launch(self, first: int, count: int, topk: int):
        self.store._check(self._L.tfp_find_ctx_launch(self._h, first, count, topk))
        self._count, self._topk = count, topk

    def download(self, with_keys: bool = False):
        cnt, k = self._count, self._topk
        n = np.empty(cnt, np.int32)
        ids = np.empty((cnt, k), np.int64)
        sims = np.empty((cnt, k), np.float32)
        outk = np.empty((cnt, k, self.dim), np.float32) if with_keys else None
        self.store._check(self._L.tfp_find_ctx_download(self._h, _p(n, i32p), _p(ids, i64p), _p(sims, fp),
                                                        _p(outk, fp) if with_keys else None))
        res = [(ids[i, :n[i]].copy(), sims[i, :n[i]].copy()) for i in range(cnt)]
        return (res, outk) if with_keys else res

    def sync(self):
        self.store._check(self._L.tfp_find_ctx_sync(self._h))

    @property
    def stream(self) -> int:
        return int(self._L.tfp_find_ctx_stream(self._h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self._L.tfp_find_ctx_launch_count(self._h))

    def gather_keys(self) -> int:
        """Enqueue the gather of the winning rows; returns the device address of [count, stride, dim] fp32."""
        p = C.c_void_p()
        self.store._check(self._L.tfp_find_ctx_gather_keys(self._h, C.byref(p)))
        return p.value

    # -- sharded finds in one call (native NCCL) ------------------------------------------------
    def comm_init(self, unique_id: bytes, rank: int, world: int):
        buf = C.create_string_buffer(unique_id, 128)
        self.store._check(self._L.tfp_find_ctx_comm_init(self._h, buf, rank, world))

    def comm_share(self, owner: "FindContext"):
        self.store._check(self._L.tfp_find_ctx_comm_share(self._h, owner._h))

    # -- sharded finds over peer memory (records pushed by the scan kernel itself) ---------------------
    def p2p_export(self, world: int, slot_bytes: int) -> bytes:
        """Allocate this rank's gather block; returns its 64-byte cudaIpc handle for the other ranks."""
        buf = C.create_string_buffer(64)
        self.store._check(self._L.tfp_find_ctx_p2p_export(self._h, world, slot_bytes, buf, 64))
        return buf.raw

    def p2p_attach(self, handles: bytes, rank: int, world: int):
        """Map every rank's gather block (`handles`: world x 64 bytes, rank order)."""
        assert len(handles) == 64 * world
        buf = C.create_string_buffer(handles, len(handles))
        self.store._check(self._L.tfp_find_ctx_p2p_attach(self._h, buf, rank, world))

    def p2p_detach(self):
        if self._h:
            self.store._check(self._L.tfp_find_ctx_p2p_detach(self._h))

    def launch_sharded_keys(self, qi: int, topk: int, count: int = 1):
        """Sharded find whose merged answer carries the winners' key rows (peer exchange only)."""
        self.store._check(self._L.tfp_find_ctx_launch_sharded_keys(self._h, qi, count, topk))
        self._topk = topk
        self._count = count

    def download_merged_keys(self, dim: int):
        """(ids, sims, keys) of a single-query launch_sharded_keys, or a list of them for a batch."""
        k, cnt = self._topk, self._count
        n = np.empty(cnt, np.int32)
        ids = np.empty((cnt, k), np.int64); sims = np.empty((cnt, k), np.float32)
        keys = np.empty((cnt, k, dim), np.float32)
        self.store._check(self._L.tfp_find_ctx_download_merged_keys(self._h, _addr(n), _addr(ids), _addr(sims), _addr(keys)))
        out = [(ids[i, :n[i]], sims[i, :n[i]], keys[i, :n[i]]) for i in range(cnt)]   # views of this call's own buffers
        return out[0] if cnt == 1 else out

    def launch_sharded(self, qi: int, topk: int, count: int = 1):
        self.store._check(self._L.tfp_find_ctx_launch_sharded_batch(self._h, qi, count, topk))
        self._topk = topk
        self._count = count

    def merged_results(self):
        """Device addresses (sims, ids, shard, index, n) of the last sharded launch's merged answer."""
        p = [C.c_void_p() for _ in range(5)]
        self.store._check(self._L.tfp_find_ctx_merged_results(self._h, *[C.byref(x) for x in p]))
        return tuple(x.value for x in p)

    def download_merged(self):
        k, cnt = self._topk, self._count
        n = np.empty(cnt, np.int32)
        ids = np.empty((cnt, k), np.int64); sims = np.empty((cnt, k), np.float32)
        shard = np.empty((cnt, k), np.int32); index = np.empty((cnt, k), np.int32)
        self.store._check(self._L.tfp_find_ctx_download_merged(self._h, _p(n, i32p), _p(ids, i64p), _p(sims, fp),
                                                               _p(shard, i32p), _p(index, i32p)))
        if cnt == 1:
            m = int(n[0])
            return ids[0, :m].copy(), sims[0, :m].copy(), shard[0, :m].copy(), index[0, :m].copy()
        return [(ids[i, :n[i]].copy(), sims[i, :n[i]].copy(), shard[i, :n[i]].copy(), index[i, :n[i]].copy())
                for i in range(cnt)]

    def mma_counters(self):
        """(tensor-core passes, queries re-run on the exact kernel) so far on this context."""
        a, b = C.c_int64(), C.c_int64()
        self.store._check(self._L.tfp_find_ctx_mma_counters(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    @property
    def mma_escalated(self) -> int:
        """Queries that needed a second tensor-core pass with the 3xTF32 filter so far on this context."""
        return int(self._L.tfp_find_ctx_mma_escalated(self._h))

    def device_records(self):
        """(device address, record_bytes) of the last launch's packed per-query records."""
        p, nb = C.c_void_p(), C.c_size_t()
        self.store._check(self._L.tfp_find_ctx_device_records(self._h, C.byref(p), C.byref(nb)))
        return p.value, nb.value

    def device_results(self):
        """(stride, cand_ptr, ids_ptr, sims_ptr, n_ptr): device addresses of the last launch's results."""
        a, b, c, d = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        stride = self._L.tfp_find_ctx_device_results(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d))
        if stride < 0:
            self.store._check(stride)
        return stride, a.value, b.value, c.value, d.value
