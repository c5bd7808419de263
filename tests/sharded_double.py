"""CPU stand-in for localai_tfp_b200.sharded.CudaShardOps, used only by the gloo tests: same
interface, the oracle does the per-shard work and numpy does the merge (same tie rule as
merge_candidates_kernel: higher score, then lower shard, then the shard's own order)."""
import contextlib

import numpy as np
import torch

from oracle import oracle as orc


class CpuShardOps:
    def __init__(self, rank):
        self.rank = rank
        self.o = orc.OracleStore()
        self.forced_fallback = False
        self.next = rank << 40
        self.qs = None
        self.vals = {}

    def stream(self, lane=0):
        return contextlib.nullcontext()

    def latch(self, dim):
        if self.o.key_len == -1:
            self.o.StoresDelete(np.zeros((1, dim), np.float32))

    def set(self, keys, values):
        ids = np.arange(self.next, self.next + len(keys), dtype=np.int64)
        self.next += len(keys)
        self.o.StoresSet(keys, [int(i).to_bytes(8, "little") for i in ids])
        for i, v in zip(ids, values):
            self.vals[int(i)] = bytes(v)
        return ids, np.ones(len(keys), np.uint8)

    def delete(self, keys):
        _, vals = self.o.StoresGet(keys)
        self.o.StoresDelete(keys)
        return np.array([int.from_bytes(v, "little") for v in vals], np.int64)

    def get(self, keys):
        ks, vals = self.o.StoresGet(keys)
        return ks, [self.vals[int.from_bytes(v, "little")] for v in vals]

    def rows(self):
        return len(self.o)

    def key_len(self):
        return self.o.key_len

    def keys_are_normalized(self):
        return self.o.keys_are_normalized and not self.forced_fallback

    def mark_not_normalized(self):
        self.forced_fallback = True

    def stage_queries(self, qs):
        self.qs = np.array(qs, np.float32)

    def find_local(self, qi, topk, with_keys, lane=0):
        from localai_tfp_b200.sharded import pack_record
        q = self.qs[qi]
        n = len(self.o)
        d = max(self.o.key_len, 1)
        sims = np.zeros(topk, np.float32)
        ids = np.full(topk, -1, np.int64)
        keys = torch.zeros((topk, d), dtype=torch.float32) if with_keys else None
        m = 0
        if n:
            rows = [self.o.row(i) for i in range(n)]
            K = np.stack([r[0] for r in rows])
            normalized = self.keys_are_normalized() and orc.is_normalized(q)
            scores = orc.scores_block(K, q, normalized)
            order = sorted(range(n), key=lambda i: (-float(scores[i]), int.from_bytes(rows[i][1], "little")))[:topk]
            m = len(order)
            sims[:m] = scores[order]
            ids[:m] = [int.from_bytes(rows[i][1], "little") for i in order]
            if with_keys:
                keys[:m] = torch.from_numpy(K[order])
        return torch.from_numpy(pack_record(m, ids, sims, topk)), keys

    def gather_buffer(self, nbytes, lane=0):
        return torch.empty(nbytes, dtype=torch.uint8)

    def merge(self, g_rec, world, topk, lane=0):
        from localai_tfp_b200.sharded import record_bytes, unpack_record
        return merge_records_numpy(g_rec.numpy(), world, topk)

    def close(self):
        pass


def merge_records_numpy(g, world, topk):
    """Reference for merge_records_kernel: higher score, then lower shard, then the shard's own order."""
    from localai_tfp_b200.sharded import record_bytes, unpack_record
    rb = record_bytes(topk)
    recs = [unpack_record(g[sh * rb:(sh + 1) * rb], topk) for sh in range(world)]
    cand = []
    for sh, (n, ids, sims) in enumerate(recs):
        for j in range(min(n, topk)):
            cand.append((-float(sims[j]), sh, j))
    cand.sort()
    cand = cand[:topk]
    m = len(cand)
    o_sims = torch.zeros(topk); o_ids = torch.full((topk,), -1, dtype=torch.int64)
    o_shard = torch.full((topk,), -1, dtype=torch.int32); o_index = torch.full((topk,), -1, dtype=torch.int32)
    for i, (_, sh, j) in enumerate(cand):
        o_sims[i] = float(recs[sh][2][j]); o_ids[i] = int(recs[sh][1][j]); o_shard[i] = sh; o_index[i] = j
    return o_sims, o_ids, o_shard, o_index, torch.tensor([m], dtype=torch.int32)
