"""Key-set sharding across the GPUs of one box: one process per GPU, each
holding one shard of the store; a find scans every shard, all-gathers the per-shard
(similarity, id) candidates with torch.distributed (NCCL over NVLink on GPUs) and merges them
with `tfp_merge_candidates` on the device.

The reference has no counterpart (backend/go/stores/store.go is single-process, single-thread);
the observable behaviour is that of ONE `Store` holding the union of the shards:
StoresSet / StoresDelete / StoresGet / StoresFind with the reference's semantics.

Every rank calls every method with the same arguments (SPMD).  Keys are routed to their owner
rank by a hash of their canonical bits, so exact-match operations touch one shard only.

The device work is behind a small `ShardOps` interface so that the routing / gather / assembly
logic can be exercised on CPU with the gloo backend (tests/ supplies a CPU double of ShardOps);
the product implementation below is CUDA-only.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np
import torch
import torch.distributed as dist

from .store import Store, StoreError, nccl_unique_id

ID_SHIFT = 40  # shard r hands out ids in [r << 40, (r + 1) << 40)


def canonical_bits(keys: np.ndarray) -> np.ndarray:
    """uint32 view of fp32 keys with Go's cmp.Compare equality folded in: -0 -> +0, any NaN -> one NaN."""
    k = np.ascontiguousarray(keys, dtype=np.float32)
    bits = k.view(np.uint32).copy()
    bits[(bits & 0x7FFFFFFF) > 0x7F800000] = 0x7FC00000
    bits[bits == 0x80000000] = 0
    return bits


def owner_of(keys: np.ndarray, world: int) -> np.ndarray:
    """Owner rank of each key row: a multiplicative hash of the canonical bits, mod world."""
    bits = canonical_bits(keys).astype(np.uint64)
    d = bits.shape[1]
    mult = (np.arange(1, d + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)) | np.uint64(1)
    with np.errstate(over="ignore"):
        h = (bits * mult[None, :]).sum(axis=1, dtype=np.uint64)
        h ^= h >> np.uint64(29)
        h *= np.uint64(0xBF58476D1CE4E5B9)
        h ^= h >> np.uint64(32)
    return (h % np.uint64(world)).astype(np.int64)


def record_bytes(topk: int) -> int:
    """Size of one packed per-query record (include/tfp_store.h): n, ids[topk], sims[topk], padded to 16."""
    return (8 + 12 * topk + 15) & ~15


def pack_record(n: int, ids: np.ndarray, sims: np.ndarray, topk: int) -> np.ndarray:
    buf = np.zeros(record_bytes(topk), np.uint8)
    buf[:4] = np.array([n], np.int32).view(np.uint8)
    buf[8:8 + 8 * topk] = np.ascontiguousarray(ids, np.int64).view(np.uint8)
    buf[8 + 8 * topk:8 + 12 * topk] = np.ascontiguousarray(sims, np.float32).view(np.uint8)
    return buf


def unpack_record(buf: np.ndarray, topk: int):
    n = int(buf[:4].view(np.int32)[0])
    ids = buf[8:8 + 8 * topk].view(np.int64)
    sims = buf[8 + 8 * topk:8 + 12 * topk].view(np.float32)
    return n, ids, sims


class _DevArray:
    """A raw device pointer dressed as __cuda_array_interface__ so torch can alias it (zero copy)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


def _alias(ptr: int, shape, typestr: str, device) -> torch.Tensor:
    return torch.as_tensor(_DevArray(ptr, shape, typestr), device=device)


class _Lane:
    """One find in flight on a shard: a find context, its stream, and cached device tensors."""

    def __init__(self, store: Store, device):
        self.ctx = store.find_context()
        self.stream = torch.cuda.ExternalStream(self.ctx.stream, device=device)
        self.alias = {}       # (ptr, shape, typestr) -> tensor aliasing library memory
        self.bufs = {}        # (tag, shape) -> reusable tensor


class CudaShardOps:
    """Device work of one shard: the local store, its find contexts ("lanes": finds that may be in
    flight at once, each with its own stream and scratch), and the merge kernel."""

    def __init__(self, rank: int, device: int, capacity_rows: int = 0, lanes: int = 1):
        self.store = Store(device=device, capacity_rows=capacity_rows, id_base=rank << ID_SHIFT)
        self.device = torch.device("cuda", device)
        self._L = self.store._L
        self.lanes = [_Lane(self.store, self.device) for _ in range(max(1, lanes))]
        self.ctx = self.lanes[0].ctx
        self.native = False
        self.p2p = False
        self._p2p_group = None

    def stream(self, lane: int = 0):
        """Everything a find enqueues (scan, collective, merge) is ordered on its lane's stream."""
        return torch.cuda.stream(self.lanes[lane].stream)

    def _alias(self, lane: _Lane, ptr: int, shape, typestr: str) -> torch.Tensor:
        key = (ptr, tuple(shape), typestr)
        t = lane.alias.get(key)
        if t is None:
            if len(lane.alias) > 64:
                lane.alias.clear()
            t = lane.alias[key] = _alias(ptr, shape, typestr, self.device)
        return t

    def _buf(self, lane: _Lane, tag: str, shape, dtype) -> torch.Tensor:
        key = (tag, tuple(shape))
        t = lane.bufs.get(key)
        if t is None:
            t = lane.bufs[key] = torch.empty(shape, dtype=dtype, device=self.device)
        return t

    # exact-match operations
    def set(self, keys, values):
        return self.store.StoresSet(keys, values)

    def delete(self, keys):
        return self.store.StoresDelete(keys)

    def get(self, keys):
        return self.store.StoresGet(keys)

    def rows(self):
        return len(self.store)

    def key_len(self):
        return self.store.key_len

    def keys_are_normalized(self):
        return self.store.keys_are_normalized

    def mark_not_normalized(self):
        self.store.mark_not_normalized()

    def latch(self, dim: int):
        """Fix the key length of a shard that may own no key yet.  A delete on an empty store
        latches keyLen and removes nothing (store.go:209-210), which is exactly what is needed."""
        if self.store.key_len == -1:
            self.store.StoresDelete(np.zeros((1, dim), np.float32))

    def stage_queries(self, qs: np.ndarray):
        for lane in self.lanes:
            lane.ctx.upload(qs)

    def init_native_comm(self, rank: int, world: int, group=None, shared: bool = True):
        """NCCL communicator(s) created inside libtfpstore (the id travels over torch.distributed): one
        shared by all lanes (NCCL orders their all-gathers) or one per lane.  After this, find_merged()
        does scan + all-gather + merge in ONE library call."""
        for i, lane in enumerate(self.lanes):
            if shared and i > 0:
                lane.ctx.comm_share(self.lanes[0].ctx)
                continue
            box = [nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0, group=group)
            lane.ctx.comm_init(box[0], rank, world)
        self.native = True

    def init_p2p(self, rank: int, world: int, group=None, slot_bytes: int = 1 << 20):
        """Peer exchange: every lane's gather block is mapped into every rank (cudaIpc handles travel
        over torch.distributed); after this find_merged() is ONE library call whose scan kernel pushes
        the records to the peers over NVLink -- no collective call, no extra kernel per find."""
        # Every step is agreed on by all ranks before the next one (a rank that fails alone would leave the
        # others waiting in a collective): either every rank ends up attached on every lane, or none does.
        def agree(step, err):
            errs = [None] * world
            dist.all_gather_object(errs, err, group=group)
            bad = [f"rank {r}: {e}" for r, e in enumerate(errs) if e]
            if bad:
                for ln in self.lanes:
                    try:
                        ln.ctx.p2p_detach()
                    except Exception:
                        pass
                raise StoreError(f"peer exchange unavailable ({step}): " + "; ".join(bad))

        for lane in self.lanes:
            mine, err = None, None
            try:
                mine = lane.ctx.p2p_export(world, slot_bytes)
            except Exception as e:
                err = str(e)
            agree("export", err)
            handles = [None] * world
            dist.all_gather_object(handles, mine, group=group)
            err = None
            try:
                lane.ctx.p2p_attach(b"".join(handles), rank, world)
            except Exception as e:
                err = str(e)
            agree("attach", err)
        dist.barrier(group=group)      # nobody launches before everybody has attached
        self.native = True
        self.p2p = True
        self._p2p_group = group
        self._p2p_slot = (slot_bytes + 255) & ~255

    def find_host(self, q: np.ndarray, topk: int):
        """One StoresFind end to end over the peer exchange: host query in, (ids, sims, key rows) out.
        Returns None when the answer does not fit the exchange slot (the caller takes the general path)."""
        if not self.p2p or self.store.key_len < 1:
            return None
        if self._L.tfp_p2p_slot_bytes_with_keys(1, topk, self.store.key_len) > self._p2p_slot:
            return None
        ctx = self.lanes[0].ctx
        ctx.upload(q[None, :])
        ctx.launch_sharded_keys(0, topk)
        return ctx.download_merged_keys(self.store.key_len)

    def find_merged(self, qi: int, topk: int, lane: int = 0):
        """Native path: returns device tensors (sims, ids, shard, index, n) aliasing the lane's merged buffers."""
        ln = self.lanes[lane]
        ln.ctx.launch_sharded(qi, topk)
        key = ("merged", topk)
        t = ln.bufs.get(key)
        if t is None:
            ps, pi, psh, pix, pn = ln.ctx.merged_results()
            t = ln.bufs[key] = (_alias(ps, (topk,), "<f4", self.device), _alias(pi, (topk,), "<i8", self.device),
                                _alias(psh, (topk,), "<i4", self.device), _alias(pix, (topk,), "<i4", self.device),
                                _alias(pn, (1,), "<i4", self.device))
        return t

    def find_local(self, qi: int, topk: int, with_keys: bool, lane: int = 0):
        """Scan this shard for staged query qi.  Returns (record, keys): the packed result record as a
        uint8 device tensor that aliases the library's buffer, and, if asked, the matching key rows
        [topk, dim] (zero padded)."""
        ln = self.lanes[lane]
        ln.ctx.launch(qi, 1, topk)
        ptr, nbytes = ln.ctx.device_records()
        rec = self._alias(ln, ptr, (nbytes,), "|u1")
        keys = None
        if with_keys:
            d = max(self.key_len(), 1)
            keys = torch.zeros((topk, d), dtype=torch.float32, device=self.device)
            rows = self.rows()
            if rows > 0:
                stride = min(topk, rows)
                keys[:stride] = self._alias(ln, ln.ctx.gather_keys(), (stride, d), "<f4")
        return rec, keys

    def gather_buffer(self, nbytes: int, lane: int = 0) -> torch.Tensor:
        return self._buf(self.lanes[lane], "g_rec", (nbytes,), torch.uint8)

    def merge(self, g_rec, world: int, topk: int, lane: int = 0):
        ln = self.lanes[lane]
        o_sims = self._buf(ln, "o_sims", (topk,), torch.float32)
        o_ids = self._buf(ln, "o_ids", (topk,), torch.int64)
        o_shard = self._buf(ln, "o_shard", (topk,), torch.int32)
        o_index = self._buf(ln, "o_index", (topk,), torch.int32)
        o_n = self._buf(ln, "o_n", (1,), torch.int32)
        rc = self._L.tfp_merge_records(g_rec.data_ptr(), world, 1, topk, o_sims.data_ptr(), o_ids.data_ptr(),
                                       o_shard.data_ptr(), o_index.data_ptr(), o_n.data_ptr(),
                                       C.c_void_p(ln.stream.cuda_stream))
        self.store._check(rc)
        return o_sims, o_ids, o_shard, o_index, o_n

    def close(self):
        if self.p2p:
            # unmap the peers' blocks everywhere, THEN free: an exported block must outlive its mappings
            for ln in self.lanes:
                ln.ctx.p2p_detach()
            if dist.is_initialized():
                dist.barrier(group=self._p2p_group)
            self.p2p = False
        for ln in self.lanes:
            ln.alias.clear(); ln.bufs.clear()
            ln.ctx.close()
        self.store.close()


class ShardedStore:
    """One logical store over `world` shards.  SPMD: every rank makes the same calls."""

    def __init__(self, ops, rank: int, world: int, group=None):
        self.ops = ops
        self.rank = rank
        self.world = world
        self.group = group
        self.values = {}           # id -> bytes, replicated on every rank (values are tiny; store.go:23)
        self.key_len = -1          # store.go:29,44
        self.keys_are_normalized = True
        self.last_path = None

    # -- helpers -------------------------------------------------------------
    def _pack(self, keys):
        if isinstance(keys, np.ndarray) and keys.ndim == 2:
            return np.ascontiguousarray(keys, dtype=np.float32)
        ks = [np.asarray(k, dtype=np.float32).reshape(-1) for k in keys]
        if not ks:
            return np.zeros((0, 1), np.float32)
        d = len(ks[0])
        for k in ks:
            if len(k) != d:
                raise StoreError(f"Try to add key with length {len(k)} when existing length is {d}")
        return np.ascontiguousarray(np.stack(ks), dtype=np.float32)

    def _latch(self, d: int, fmt: str):
        if self.key_len == -1:
            self.key_len = d
            self.ops.latch(d)
        elif d != self.key_len:
            raise StoreError(fmt % (d, self.key_len))

    def _allgather_obj(self, obj):
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def __len__(self):
        return int(sum(self._allgather_obj(self.ops.rows())))

    # -- StoresSet (store.go:106-198) ------------------------------------------
    def StoresSet(self, keys, values: Sequence[bytes]):
        ks = self._pack(keys)
        if len(ks) == 0:
            raise StoreError("no keys to add")
        if len(ks) != len(values):
            raise StoreError(f"len(keys) = {len(ks)}, len(values) = {len(values)}")
        self._latch(ks.shape[1], "Try to add key with length %d when existing length is %d")
        own = owner_of(ks, self.world)
        mine = np.nonzero(own == self.rank)[0]
        local = []
        if len(mine):
            ids, _ = self.ops.set(ks[mine], [values[i] for i in mine])
            local = [(int(i), int(ids[j])) for j, i in enumerate(mine)]
        for part in self._allgather_obj(local):   # replicate id -> value everywhere
            for i, gid in part:
                self.values[gid] = bytes(values[i])
        flags = self._allgather_obj(bool(self.ops.keys_are_normalized()))
        if not all(flags):
            # the sticky flag (store.go:126-127) belongs to the whole store: every shard leaves
            # the normalised path together, or scores from different shards would not compare
            self.keys_are_normalized = False
            self.ops.mark_not_normalized()

    # -- StoresDelete (store.go:200-264) ----------------------------------------
    def StoresDelete(self, keys):
        ks = self._pack(keys)
        if len(ks) == 0:
            raise StoreError("no keys to delete")
        self._latch(ks.shape[1], "Trying to delete key with length %d when existing length is %d")
        own = owner_of(ks, self.world)
        mine = np.nonzero(own == self.rank)[0]
        gone = []
        if len(mine):
            ids = self.ops.delete(ks[mine])
            gone = [int(g) for g in ids if g >= 0]
        for part in self._allgather_obj(gone):
            for gid in part:
                self.values.pop(gid, None)

    # -- StoresGet (store.go:266-311) --------------------------------------------
    def StoresGet(self, keys):
        ks = self._pack(keys)
        if len(ks) == 0:
            raise StoreError("no keys to get")
        self._latch(ks.shape[1], "Try to get a key with length %d when existing length is %d")
        own = owner_of(ks, self.world)
        mine = np.nonzero(own == self.rank)[0]
        found = []
        if len(mine) and self.ops.rows() > 0:
            fk, fv = self.ops.get(ks[mine])
            found = [(k.tobytes(), v) for k, v in zip(fk, fv)]
        allf = [item for part in self._allgather_obj(found) for item in part]
        rows = [np.frombuffer(kb, dtype=np.float32) for kb, _ in allf]
        order = sorted(range(len(rows)), key=lambda i: _go_sort_key(rows[i]))   # sorted-key order, store.go:269
        return [rows[i].copy() for i in order], [allf[i][1] for i in order]

    # -- StoresFind (store.go:477-511) ---------------------------------------------
    def StoresFind(self, key, topk: int, with_keys: bool = True):
        q = np.ascontiguousarray(np.asarray(key, dtype=np.float32).reshape(-1))
        if len(q) != self.key_len:
            raise StoreError(f"Try to find key with length {len(q)} when existing length is {self.key_len}")
        if topk < 1:
            raise StoreError(f"opts.TopK = {topk}, must be >= 1")
        if with_keys and self.world * int(topk) <= 16384 and hasattr(self.ops, "find_host"):
            got = self.ops.find_host(q, int(topk))     # peer exchange: three library calls, one D2H copy
            if got is not None:
                ids_h, sims_h, keys_h = got
                return list(keys_h), [self.values.get(i) for i in ids_h.tolist()], sims_h
        self.ops.stage_queries(q[None, :])
        sims, ids, shard, index, n, g_keys = self.find_device(0, int(topk), with_keys)
        with self.ops.stream():
            m = int(n.item())
        sims_h = sims[:m].cpu().numpy()
        ids_h = ids[:m].cpu().numpy()
        out_keys = []
        if with_keys and m:
            sh = shard[:m].long()
            ix = index[:m].long()
            out_keys = [r.copy() for r in g_keys[sh, ix].cpu().numpy()]
        return out_keys, [self.values.get(int(i)) for i in ids_h], sims_h

    def find_device(self, qi: int, topk: int, with_keys: bool = False, lane: int = 0):
        """Scan + all-gather + merge for staged query qi on one lane; results stay on the device (what
        bench.py times) in buffers that the lane reuses for its next find."""
        if not with_keys and getattr(self.ops, "native", False):
            return self.ops.find_merged(qi, topk, lane) + (None,)     # one library call: scan + all-gather + merge
        with self.ops.stream(lane):
            rec, keys = self.ops.find_local(qi, topk, with_keys, lane)
            # ONE collective per query: every shard's whole answer is one fixed-size record
            g_rec = self.ops.gather_buffer(self.world * rec.numel(), lane)
            dist.all_gather_into_tensor(g_rec, rec, group=self.group)
            g_keys = None
            if with_keys:
                g_keys = torch.empty(self.world * keys.numel(), dtype=keys.dtype, device=keys.device)
                dist.all_gather_into_tensor(g_keys, keys.contiguous().view(-1), group=self.group)
                g_keys = g_keys.view((self.world,) + tuple(keys.shape))
            o_sims, o_ids, o_shard, o_index, o_n = self.ops.merge(g_rec, self.world, topk, lane)
        return o_sims, o_ids, o_shard, o_index, o_n, g_keys

    def close(self):
        self.ops.close()


def _go_sort_key(row: np.ndarray):
    """Sort key reproducing Go's cmp.Compare order on float32 rows: NaN lowest, -0 == +0."""
    out = []
    for x in row:
        x = float(x)
        out.append((0, 0.0) if x != x else (1, x + 0.0))
    return tuple(out)
