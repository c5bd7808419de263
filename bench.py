#!/usr/bin/env python
"""bench.py -- StoresFind throughput of the B200 local-store (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (all N): BASELINE.json configs[1] (C2) -- 1M keys x 768-d fp32, unit-length rows,
single-query top-10 on the normalised-key fast path.  A "step" is one StoresFind over the whole store.
N > 1 shards the SAME store by rows across the ranks (strong scaling): every step each rank scans its
shard, the per-shard top-k records travel to every rank over NVLink and are merged on the device.

Prints ONE JSON line (rank 0).
  value      queries/s with keys and queries resident in HBM: the median over `repeats` timed regions
             of exactly --steps finds each (CUDA events, max over ranks), `finds_in_flight` finds
             overlapping; `single_find_ms` is the latency of one find with nothing else in flight.
  e2e        the same metric through the public call with host buffers (H2D + find + D2H every step).
  roofline   the scan kernel alone against measured HBM bandwidth.
  parity     the answers of the timed configuration checked against the CPU oracle after the timed
             regions, at every N: returned similarities bit for bit against the oracle's score of the
             returned rows, plus an oracle scan of a >= 100k-row sample (and, on one GPU, of the whole
             C2 store) that must not hold a better row than the k-th returned one.  A mismatch fails the run.
  also       the same measurement (ms, roofline, parity) on the north star's target shape (10M x 1024,
             top-10, batch 1), on C3 (10M x 1024, 256 queries, top-100, tensor cores) and, on 8 GPUs,
             on C4 (100M x 384); on one GPU C1 (10k x 384 through the gRPC backend process, microseconds
             per StoresFind); on several GPUs `inproc`: the same store as ONE process over all N GPUs
             (tfp_store_open_devices), which is also where `e2e` comes from there.  `--also none` skips them.
  cpu_baseline  the CPU oracle (a port of the reference's Go code) timed on this box (N = 1).
`--impl reference` times that CPU port as its own arm.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from dataclasses import dataclass

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_QUERIES = 64
MMA_TERMS = 3 if os.environ.get("TFP_MMA_TERMS") == "3" else 1   # tensor-core filter: TF32 MMAs per K-step
BLOCK_ROWS = 62_500            # generation granularity: shards of 1/2/4/8 ranks align with blocks
ID_SHIFT = 40                  # shard r hands out ids from r << 40 (sharded.py)
REPEATS = 7                    # timed regions per measurement; the median is reported
SAMPLE_ROWS = 100_000          # the oracle scans at least this many rows of the store per checked query


@dataclass
class Workload:
    name: str
    n_keys: int
    dim: int
    topk: int
    queries_per_step: int
    seed: int
    label: str
    metric: str = "StoresFind queries/sec (top-10)"

    @property
    def batched(self):
        return self.queries_per_step > 1


WORKLOADS = {
    "c2": Workload("c2", 1_000_000, 768, 10, 1, 0x5EED0002,
                   "C2 (BASELINE.json configs[1]): 1M keys x 768-d fp32, single-query top-10, normalised-key fast path"),
    "target": Workload("target", 10_000_000, 1024, 10, 1, 0x5EED0003,
                       "north-star target: 10M keys x 1024-d fp32, single-query top-10, normalised-key fast path"),
    "c3": Workload("c3", 10_000_000, 1024, 100, 256, 0x5EED0003,
                   "C3 (BASELINE.json configs[2]): 10M keys x 1024-d fp32, batch-256 queries top-100 on the "
                   "tensor-core path (tcgen05 %dxTF32 filter + exact re-score)" % MMA_TERMS,
                   "StoresFind queries/sec (top-100, batch 256)"),
    "c4": Workload("c4", 100_000_000, 384, 10, 1, 0x5EED0004,
                   "C4 (BASELINE.json configs[3]): 100M keys x 384-d fp32 (153.6 GB), single-query top-10, "
                   "sharded by rows, candidate merge over NVLink"),
}


# ----------------------------------------------------------------------------- inputs
def gen_block_torch(wl: Workload, block: int, device):
    import torch
    g = torch.Generator(device=device).manual_seed(wl.seed + block)
    k = torch.randn((BLOCK_ROWS, wl.dim), generator=g, device=device, dtype=torch.float32)
    k /= k.norm(dim=1, keepdim=True)
    return k


def gen_queries(wl: Workload) -> np.ndarray:
    rng = np.random.default_rng(wl.seed + 1)
    q = rng.standard_normal((max(N_QUERIES, wl.queries_per_step), wl.dim)).astype(np.float32)
    q /= np.linalg.norm(q.astype(np.float64), axis=1, keepdims=True).astype(np.float32)
    return np.ascontiguousarray(q.astype(np.float32))


def gen_keys_numpy(wl: Workload, rows: int) -> np.ndarray:
    rng = np.random.default_rng(wl.seed)
    k = rng.standard_normal((rows, wl.dim), dtype=np.float32)
    k /= np.sqrt((k.astype(np.float64) ** 2).sum(axis=1, keepdims=True)).astype(np.float32)
    return k


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons during the timed region (NVML; nvidia-smi is the fallback)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self.bits = 0
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _sample(self):
        try:
            self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            self.bits |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
        except Exception:
            pass

    def _loop(self):
        while not self._stop.is_set():
            self._sample()
            self._stop.wait(0.002)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self) -> dict:
        if self._thr:
            self._stop.set()
            self._thr.join()
        reasons = [name for bit, name in self.REASONS.items() if self.bits & bit]
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(self.samples)}


# ----------------------------------------------------------------------------- CPU baseline (the oracle)
def cpu_reference_qps(wl: Workload, budget_s: float, steps: int, warmup: int):
    """Times the CPU port of the reference (oracle/, `faithful` flavour: separately allocated rows in
    sorted order, one heap item allocated per key, container/heap mechanics) on a bounded row sample
    of the store and scales to queries/s over the full store (the scan is linear in rows)."""
    from oracle import oracle as orc
    orc.build()
    q = gen_queries(wl)
    probe_rows = 20_000
    probe = gen_keys_numpy(wl, probe_rows)
    st = orc.OracleStore()
    st.StoresSet(probe, [b""] * probe_rows)
    t0 = time.perf_counter()
    st.StoresFind(q[0], wl.topk, faithful=True)
    per_row = (time.perf_counter() - t0) / probe_rows
    rows = int(min(wl.n_keys, 1_000_000, max(probe_rows, budget_s / max(steps + warmup, 1) / per_row)))
    if rows > probe_rows:
        keys = gen_keys_numpy(wl, rows)
        st = orc.OracleStore()
        st.StoresSet(keys, [b""] * rows)
    assert st.keys_are_normalized
    for i in range(warmup):
        st.StoresFind(q[i % N_QUERIES], wl.topk, faithful=True)
    t0 = time.perf_counter()
    for i in range(steps):
        st.StoresFind(q[i % N_QUERIES], wl.topk, faithful=True)
    dt = (time.perf_counter() - t0) / steps
    assert st.last_path == "normalized"
    qps_full = 1.0 / (dt * wl.n_keys / rows)
    sample = (f"{steps} single-query top-{wl.topk} finds over a {rows}-row x {wl.dim}-d sample of the store, "
              f"1 thread (the reference scan is single-threaded by design, store.go:18), scaled by rows to {wl.n_keys}")
    return qps_full, dt * 1e3, rows, sample


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    qps, ms, rows, sample = cpu_reference_qps(wl, budget_s=120.0, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": wl.metric, "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        # same keys as our arm's `config` (same workload, shape and top-k), with this arm's own values where they differ
        "config": {"workload": wl.label, "keys": wl.n_keys, "dim": wl.dim, "topk": wl.topk, "queries_per_step": 1,
                   "path": "normalized", "rows_per_gpu": None, "finds_in_flight": 1, "warmup_finds": args.warmup,
                   "exchange": None, "l2": "host memory; %d-row sample per step, scaled by rows" % rows,
                   "parallelism": "CPU, one thread (the reference serialises every store call, store.go:18)",
                   "sample_rows_per_step": rows},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": 1, "kind": "port", "sample": sample,
                         "host_cores": os.cpu_count()},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- peaks
def hbm_peak():
    try:
        return (float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]),
                "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)")
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


_tf32_peak_cache = {}


def measure_tf32_peak(torch, device):
    """Dense TF32 throughput of this GPU: a library fp32 GEMM with TF32 tensor cores allowed (8192^3), best of
    10 launches (burst) and back to back for 2 s (sustained).  A denominator only: nothing here is on the hot path."""
    if "v" in _tf32_peak_cache:
        return _tf32_peak_cache["v"]
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn((n, n), device=device, dtype=torch.float32)
        b = torch.randn((n, n), device=device, dtype=torch.float32)
        c = torch.empty((n, n), device=device, dtype=torch.float32)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize(device)
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=c)
            e1.record()
            torch.cuda.synchronize(device)
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(2000.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize(device)
        sustained_ms = e0.elapsed_time(e1) / reps
        flops = 2.0 * n ** 3
        res = {"burst_tflops": flops / (best * 1e-3) / 1e12, "sustained_tflops": flops / (sustained_ms * 1e-3) / 1e12,
               "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS TF32 tensor cores): best of 10, and back to back for 2 s"}
        del a, b, c
        torch.cuda.empty_cache()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    _tf32_peak_cache["v"] = res
    return res


# ----------------------------------------------------------------------------- one measured workload
class Shard:
    """This rank's shard of one workload's store, with its find lanes and (N > 1) the exchange set up."""

    def __init__(self, env, wl: Workload, lanes: int, exchange: str):
        import torch
        from localai_tfp_b200.sharded import CudaShardOps, ShardedStore
        self.env, self.wl = env, wl
        dist = env["dist"]
        rank, world, device = env["rank"], env["world"], env["device"]
        n_blocks = wl.n_keys // BLOCK_ROWS
        self.b0, self.b1 = rank * n_blocks // world, (rank + 1) * n_blocks // world
        self.rows_local = (self.b1 - self.b0) * BLOCK_ROWS
        self.ops = CudaShardOps(rank, env["local_rank"], capacity_rows=self.rows_local, lanes=lanes)
        self.store = self.ops.store
        for b in range(self.b0, self.b1):
            blk = gen_block_torch(wl, b, device)
            ids, new = self.store.set_device(blk.data_ptr(), BLOCK_ROWS, wl.dim)
            assert bool(new.all()) and int(ids[0]) == (rank << ID_SHIFT) + (b - self.b0) * BLOCK_ROWS
            del blk
        torch.cuda.empty_cache()
        assert len(self.store) == self.rows_local and self.store.keys_are_normalized
        self.exchange = exchange if world > 1 else None
        self.sharded = ShardedStore(self.ops, rank, world) if world > 1 else None
        if self.sharded is not None:
            self.sharded.key_len = wl.dim
            if self.exchange == "p2p":              # the scan kernel pushes its records to the peers over NVLink
                try:
                    self.ops.init_p2p(rank, world)  # all ranks succeed or all ranks raise (it agrees step by step)
                except Exception as e:
                    if rank == 0:
                        print(f"bench: {e}; falling back to ncclAllGather", file=sys.stderr)
                    self.exchange = "native"
            if self.exchange in ("native", "native-per-lane"):   # scan + ncclAllGather + merge in one library call
                try:
                    self.ops.init_native_comm(rank, world, shared=(self.exchange == "native"))
                except Exception as e:              # NCCL not resolvable by dlopen: same exchange through torch.distributed
                    if rank == 0:
                        print(f"bench: native exchange unavailable ({e}); using torch.distributed", file=sys.stderr)
                    self.exchange = "torch"
        _ = dist

    def local_rows(self, rows: np.ndarray) -> np.ndarray:
        """Key rows of this shard by shard-local row number, REGENERATED from the seed (not read back through
        the store), so the parity check does not lean on the code it checks."""
        rows = np.asarray(rows, dtype=np.int64)
        out = np.empty((len(rows), self.wl.dim), np.float32)
        import torch
        blk_of = rows // BLOCK_ROWS
        for b in np.unique(blk_of):
            blk = gen_block_torch(self.wl, self.b0 + int(b), self.env["device"])
            sel = np.nonzero(blk_of == b)[0]
            idx = torch.as_tensor(rows[sel] - int(b) * BLOCK_ROWS, device=self.env["device"])
            out[sel] = blk[idx].cpu().numpy()
            del blk
        return out

    def close(self):
        self.ops.close()          # (with the peer exchange this is collective: unmap everywhere, barrier, free)


def check_parity(env, sh: Shard, Q: np.ndarray, answers, full_n: int):
    """answers: [(query index, ids, sims)] -- the merged answer every rank holds.  Each rank checks the part it
    can see: the returned rows that live in its shard, and a sample of its shard.  Returns a dict; `ok` is the
    AND over ranks."""
    import torch
    from oracle import oracle as orc
    wl, world, rank, dist = sh.wl, env["world"], env["rank"], env["dist"]
    per_rank = max(1, -(-SAMPLE_ROWS // world // BLOCK_ROWS))            # sample blocks per rank
    per_rank = min(per_rank, sh.b1 - sh.b0)
    sample = np.concatenate([gen_block_torch(wl, sh.b0 + b, env["device"]).cpu().numpy() for b in range(per_rank)])
    full = None
    if full_n and world == 1:
        full = np.concatenate([gen_block_torch(wl, b, env["device"]).cpu().numpy() for b in range(sh.b0, sh.b1)])
    problems = []
    full_done = 0
    want_n = min(wl.topk, wl.n_keys)
    for (qi, ids, sims) in answers:
        q = Q[qi]
        ids = np.asarray(ids, dtype=np.int64); sims = np.asarray(sims, dtype=np.float32)
        if len(ids) != want_n or len(sims) != want_n:
            problems.append(f"q{qi}: {len(ids)} results, expected {want_n}")
            continue
        if not (np.diff(sims.astype(np.float64)) <= 0).all():
            problems.append(f"q{qi}: similarities are not descending")
        if len(set(ids.tolist())) != len(ids):
            problems.append(f"q{qi}: a row was returned twice")
        mine = np.nonzero((ids >> ID_SHIFT) == rank)[0]
        if len(mine):
            rows = sh.local_rows(ids[mine] - (rank << ID_SHIFT))
            sc = orc.scores_block(rows, q, True)
            if not np.array_equal(sc, sims[mine]):                       # bit for bit
                problems.append(f"q{qi}: returned similarity differs from the oracle's score of the returned row")
        kth = sims[-1]
        sc = orc.scores_block(sample, q, True)
        better = np.nonzero(sc > kth)[0]
        missing = set(((rank << ID_SHIFT) + better).tolist()) - set(ids.tolist())
        if missing:
            problems.append(f"q{qi}: {len(missing)} sample rows beat the k-th returned similarity but were not returned")
        if full is not None and full_done < full_n:
            orows, osims = orc.scan_block(full, q, wl.topk, True)         # the whole store, reference arithmetic
            full_done += 1
            if not np.array_equal(osims, sims):
                problems.append(f"q{qi}: similarities differ from the oracle's scan of the whole store")
            else:
                for i in range(len(ids)):       # ids position by position; exactly equal scores are exempt (tie order)
                    tied = (i > 0 and osims[i - 1] == osims[i]) or (i + 1 < len(osims) and osims[i + 1] == osims[i])
                    if not tied and int(ids[i]) != int(orows[i]):
                        problems.append(f"q{qi}: id at rank {i} differs from the oracle")
                        break
    ok = torch.tensor([0 if problems else 1], dtype=torch.int32, device=env["device"])
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    res = {"ok": bool(ok.item()), "checked_queries": len(answers),
           "returned_rows_rescored_by_oracle": True,
           "sample_rows_scanned_by_oracle": int(len(sample)) * world,
           "full_store_queries": full_done,
           "rule": "similarities bit-identical to the oracle; ids in the oracle's order except among exactly equal scores"}
    if problems:
        res["problems"] = problems[:8]
    return res


def measure(env, wl: Workload, args, shard_cache: dict, with_cpu: bool, full_parity: int):
    """Build (or reuse) this rank's shard of `wl`, time it, check it.  Returns the pieces of a bench line."""
    import torch
    torch_mod = torch
    dist, rank, world, device = env["dist"], env["rank"], env["world"], env["device"]
    lanes = max(2, args.lanes)
    key = (wl.n_keys, wl.dim, wl.seed)
    sh = shard_cache.get(key)
    if sh is None:
        for old in list(shard_cache):           # one store at a time in HBM
            shard_cache.pop(old).close()
        torch.cuda.empty_cache()
        sh = shard_cache[key] = Shard(env, wl, lanes, args.exchange)
    sh.wl = wl
    ops, store, sharded = sh.ops, sh.store, sh.sharded
    TOPK, QPS, batched = wl.topk, wl.queries_per_step, wl.batched

    Q = gen_queries(wl)
    ctx = ops.ctx
    # batches: tfp_find_ctx_launch returns once a batch's first tensor-core pass is enqueued, so on one GPU two contexts
    # take the batches in turn and the next batch is queued while the host looks at the previous one's status words
    in_flight = lanes if not batched else (min(2, lanes, max(1, args.batch_lanes)) if world == 1 else 1)
    for ln in ops.lanes:
        ln.ctx.upload(Q)       # queries resident in HBM before the timed region
        ln.ctx.sync()
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    lane_ctx = [ln.ctx for ln in ops.lanes]
    side_streams = [ln.stream for ln in ops.lanes[1:in_flight]]

    def step_device(i: int):
        if batched:          # one batch of QPS queries per step (tensor-core path)
            if world == 1:
                lane_ctx[i % in_flight].launch(0, QPS, TOPK)
            else:
                ctx.launch_sharded(0, TOPK, count=QPS)
        elif world == 1:
            lane_ctx[i % in_flight].launch(i % N_QUERIES, 1, TOPK)
        else:
            sharded.find_device(i % N_QUERIES, TOPK, with_keys=False, lane=i % in_flight)

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def finish_batches():
        # a batch's later passes (taken only when its first pass flags queries) are enqueued by the context's next call:
        # make that call inside the timed region
        if batched and world == 1:
            for c in lane_ctx[:in_flight]:
                c.sync()

    def launches_now():
        return sum(c.launch_count for c in lane_ctx)

    # ---- device-resident throughput (`value`): REPEATS timed regions of exactly --steps finds each.
    # Warm-up: at least 4 finds per lane whatever --warmup says, so that every lane has been through both
    # parities of its exchange block (first touch of a peer mapping costs far more than a find).
    warm = max(args.warmup, 4 * in_flight if not batched else 3)
    rep_ms, launches = [], 0
    clocks = ClockSampler(env["local_rank"])
    with torch.cuda.stream(stream):
        for i in range(warm):
            step_device(i)
        barrier()
        clocks.start()
        for rep in range(REPEATS):
            launches0 = launches_now()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for sb in side_streams:
                sb.wait_event(e0)                # every lane's stream starts after e0 ...
            for i in range(args.steps):
                step_device(i)
            finish_batches()
            for sb in side_streams:
                eb = torch.cuda.Event()
                eb.record(sb)
                stream.wait_event(eb)            # ... and e1 follows the last kernel of every lane
            e1.record(stream)
            barrier()
            rep_ms.append(e0.elapsed_time(e1))
            launches = launches_now() - launches0 + (args.steps if world > 1 and sh.exchange == "torch" else 0)
        clk = clocks.stop()
    t = torch.tensor(rep_ms, dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)           # per timed region: the slowest rank
    rep_ms = [float(x) for x in t.tolist()]
    ms_total = float(np.median(rep_ms))
    ms_per_step = ms_total / args.steps
    qps = 1e3 / ms_per_step * QPS

    # ---- the scan kernel alone for the roofline: back to back on ONE stream, nothing overlapping it
    with torch.cuda.stream(stream):
        k = max(5, min(args.steps, 200)) if not batched else max(3, min(args.steps, 20))
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        s0.record(stream)
        for i in range(k):
            if batched:
                ctx.launch(0, QPS, TOPK)
            else:
                ctx.launch(i % N_QUERIES, 1, TOPK)
        if batched:
            ctx.sync()
        s1.record(stream)
        torch.cuda.synchronize(device)
        scan_ms = s0.elapsed_time(s1) / k
        # ---- one find at a time, exchange and merge included: the latency a lone caller sees on the device
        barrier()
        s0.record(stream)
        for i in range(k):
            if world == 1:
                ctx.launch(0 if batched else i % N_QUERIES, QPS, TOPK)
            elif batched:
                ctx.launch_sharded(0, TOPK, count=QPS)
            else:
                sharded.find_device(i % N_QUERIES, TOPK, with_keys=False, lane=0)
            if batched and world == 1:
                ctx.sync()           # one batch at a time: the whole answer, status words included
        s1.record(stream)
        torch.cuda.synchronize(device)
        single_ms = s0.elapsed_time(s1) / k
    t = torch.tensor([scan_ms, single_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    scan_ms, single_ms = float(t[0].item()), float(t[1].item())
    rows_local = sh.rows_local
    algo_bytes = rows_local * wl.dim * 4
    peak, peak_src = hbm_peak()
    achieved = algo_bytes / (scan_ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "peak_source": peak_src, "kernel": "scan_tma_kernel<1,fp32,fused>",
            "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": scan_ms,
            "note": "achieved = bytes / duration of one scan launched alone back to back on one stream; `value` "
                    "overlaps the tail of one scan with the start of the next (finds_in_flight)",
            "effective_gbs_in_timed_region": algo_bytes / (ms_per_step * 1e-3) / 1e9,
            "frac_of_nominal_8TBs": achieved / 8000.0}
    if batched:
        # credited flops = 2 * Q * rows * D per batch (the filter issues MMA_TERMS times that)
        tf = measure_tf32_peak(torch_mod, device)
        achieved = 2.0 * QPS * rows_local * wl.dim / (scan_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": achieved, "peak": tf["burst_tflops"], "unit": "TFLOP/s",
                "frac": achieved / tf["burst_tflops"], "traffic": None,
                "peak_source": "measured in this run: " + tf["how"] + " (burst; sustained %.1f)" % tf["sustained_tflops"],
                "peak_sustained": tf["sustained_tflops"], "frac_of_sustained": achieved / tf["sustained_tflops"],
                "kernel": "mma_scan_kernel (tcgen05.mma kind::tf32) + exact re-score",
                "credited_flops_per_batch": 2.0 * QPS * rows_local * wl.dim, "batch_ms": scan_ms,
                "hbm_gbs": algo_bytes / (scan_ms * 1e-3) / 1e9, "tf32_terms": MMA_TERMS,
                "issued_tf32_tflops": MMA_TERMS * achieved}
    elif wl.name == "c2" and world == 1:
        try:                                       # the ncu capture was taken on exactly this launch shape
            roof["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "scan_traffic.json")))["dram_bytes_per_launch"]
        except Exception:
            pass

    # ---- end to end through the public call: host query in, host (ids, sims, keys) out, every step
    e2e_steps = args.steps if not batched else min(args.steps, 20)
    if batched:
        qb = Q[:QPS]

        def e2e_once(i):
            if world == 1:
                return store.find_batch(qb, TOPK)[0][1]
            ctx.upload(qb)
            ctx.launch_sharded(0, TOPK, count=QPS)
            return ctx.download_merged()[0][1]
    elif world == 1:
        def e2e_once(i):
            return store.StoresFind(Q[i % N_QUERIES], TOPK)[2]
    else:
        def e2e_once(i):
            return sharded.StoresFind(Q[i % N_QUERIES], TOPK)[2]
    for i in range(3):
        e2e_once(i)
    e2e_rep = []
    for rep in range(3):
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            sims = e2e_once(i)
        torch.cuda.synchronize(device)
        e2e_rep.append(time.perf_counter() - t0)
    t = torch.tensor(e2e_rep, dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_dt = float(np.median(t.tolist()))
    e2e_qps = e2e_steps * QPS / e2e_dt
    assert len(sims) == TOPK and list(sims) == sorted(sims, reverse=True)
    h2d = QPS * (wl.dim * 4 + 16)              # queries + their QInfo records
    d2h = QPS * (4 + TOPK * (8 + 4)) + (TOPK * wl.dim * 4 if not batched else 0)   # n, ids, sims (+ key rows)
    for ln in ops.lanes:
        ln.ctx.upload(Q)
        ln.ctx.sync()

    # ---- parity of the timed configuration against the oracle (outside every timed region)
    answers = []
    if batched:
        if world == 1:
            ctx.launch(0, QPS, TOPK)
            res = ctx.download()
        else:
            ctx.launch_sharded(0, TOPK, count=QPS)
            res = ctx.download_merged()
        for qi in range(0, QPS, max(1, QPS // 8)):      # 8 of the batch's queries
            answers.append((qi, res[qi][0], res[qi][1]))
    else:
        for j, qi in enumerate((0, 1, 2, 3)):
            lane = j % in_flight
            if world == 1:
                lane_ctx[lane].launch(qi, 1, TOPK)
                ids, sims_ = lane_ctx[lane].download()[0]
            else:
                sharded.find_device(qi, TOPK, with_keys=False, lane=lane)
                ids, sims_ = lane_ctx[lane].download_merged()[:2]
            answers.append((qi, ids, sims_))
    parity = check_parity(env, sh, Q, answers, full_parity)

    cpu = None
    if with_cpu and world == 1:
        cqps, cms, crows, csample = cpu_reference_qps(wl, budget_s=15.0, steps=6, warmup=1)
        cpu = {"value": cqps, "unit": "queries/s", "cores": 1, "kind": "port", "sample": csample,
               "host_cores": os.cpu_count()}

    out = {
        "metric": wl.metric, "value": qps, "unit": "queries/s", "ms_per_step": ms_per_step,
        "repeats": REPEATS, "repeat_ms": [round(x, 4) for x in rep_ms],
        "timing": "median of %d timed regions of %d steps each (CUDA events on the launching stream, max over ranks per "
                  "region); throughput with %d find(s) in flight per GPU" % (REPEATS, args.steps, in_flight),
        "single_find_ms": single_ms,
        "config": {"workload": wl.label, "keys": wl.n_keys, "dim": wl.dim, "topk": TOPK, "queries_per_step": QPS,
                   "path": "normalized", "rows_per_gpu": rows_local, "finds_in_flight": in_flight,
                   "warmup_finds": warm, "exchange": sh.exchange,
                   "l2": "inputs larger than L2 (keys %.2f GB per GPU vs 126 MB); no flush" % (algo_bytes / 1e9)
                         if algo_bytes > 4 * 126e6 else "keys per GPU %.0f MB: partly L2-resident" % (algo_bytes / 1e6),
                   "parallelism": ("row-sharded x%d, " % world) +
                                  ("top-k records pushed to every peer over NVLink by the scan kernel, one waiting merge kernel per rank"
                                   if sh.exchange == "p2p" else "one all-gather of top-k records + device merge per query")
                                  if world > 1 else "single GPU"},
        "clocks": clk,
        "tensor_core_path": dict(zip(("passes", "exact_reruns", "escalated_to_3xtf32"), ctx.mma_counters() + (ctx.mma_escalated,)))
                            if batched else None,
        "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "roofline": roof,
        "parity": parity,
    }
    if cpu:
        out["cpu_baseline"] = cpu
    return out


def measure_c1(env):
    """C1 (BASELINE.json configs[0]): 10k keys x 384-d, single-query top-10 THROUGH gRPC -- the reference's own
    CPU-runnable case.  The backend process is spawned the way LocalAI's loader spawns `local-store` (binary + --addr),
    loaded over the wire, and timed from the client: microseconds per StoresFind in the reference's wire format, with the
    NoKeys extension, and -- in this process -- through the C ABI and on the device alone.  Every answer is checked
    against the oracle's scan of the whole store."""
    import socket
    import subprocess
    import grpc
    from oracle import oracle as orc
    from localai_tfp_b200 import Store
    from localai_tfp_b200.server import Client
    n, d, k, reps = 10_000, 384, 10, 300
    rng = np.random.default_rng(0x5EED0001)
    K = rng.standard_normal((n, d)).astype(np.float32)
    K /= np.sqrt((K.astype(np.float64) ** 2).sum(axis=1, keepdims=True)).astype(np.float32)
    Q = rng.standard_normal((16, d)).astype(np.float32)
    Q /= np.sqrt((Q.astype(np.float64) ** 2).sum(axis=1, keepdims=True)).astype(np.float32)
    vals = [b"%d" % i for i in range(n)]
    want = [orc.scan_block(K, q, k, True) for q in Q]

    def per_call(fn):
        for i in range(20):
            fn(i)
        t0 = time.perf_counter()
        for i in range(reps):
            fn(i)
        return (time.perf_counter() - t0) / reps * 1e6

    problems = []

    def check(tag, i, vs, sims):
        rows, osims = want[i % 16]
        if not np.array_equal(np.asarray(sims, np.float32), osims) or [int(v) for v in vs] != [int(r) for r in rows]:
            problems.append(f"{tag}: query {i % 16} differs from the oracle")

    st = Store(device=env["local_rank"])
    st.StoresSet(K, vals)
    c = st.find_context(); c.upload(Q); c.sync()

    def dev_only(i):
        c.launch(i % 16, 1, k); c.sync()
    us_dev = per_call(dev_only)
    us_abi = per_call(lambda i: st.StoresFind(Q[i % 16], k))
    for i in range(16):
        ks, vs, sims = st.StoresFind(Q[i], k)
        check("C ABI", i, vs, sims)
    c.close(); st.close()

    sock = socket.socket(); sock.bind(("127.0.0.1", 0)); port = sock.getsockname()[1]; sock.close()
    addr = f"127.0.0.1:{port}"
    proc = subprocess.Popen([sys.executable, "-m", "localai_tfp_b200.server", "--addr", addr, "--device", str(env["local_rank"])],
                            cwd=ROOT)
    cl = Client(addr)
    try:
        for _ in range(120):                      # the loader polls Health (pkg/model/initializers.go:375-387)
            try:
                if cl.health() == b"OK":
                    break
            except grpc.RpcError:
                time.sleep(0.25)
        else:
            raise RuntimeError("the backend process did not become healthy")
        for lo in range(0, n, 2500):
            cl.set(K[lo:lo + 2500], vals[lo:lo + 2500])
        us_grpc = per_call(lambda i: cl.find(Q[i % 16], k))
        us_lean = per_call(lambda i: cl.find(Q[i % 16], k, no_keys=True))
        us_health = per_call(lambda i: cl.health())
        for i in range(16):
            ks, vs, sims = cl.find(Q[i], k)
            check("gRPC", i, vs, sims)
            if not all(np.array_equal(a, K[int(v)]) for a, v in zip(ks, vs)):
                problems.append(f"gRPC: query {i} echoed a key that is not the stored row")
    finally:
        cl.close()
        proc.terminate()
        try:
            proc.wait(timeout=20)
        except Exception:
            proc.kill()
    t0 = time.perf_counter()
    for i in range(4):
        orc.scan_block(K, Q[i], k, True)
    us_cpu = (time.perf_counter() - t0) / 4 * 1e6
    out = {"workload": "C1 (BASELINE.json configs[0]): 10k keys x 384-d fp32, single query top-10 via gRPC",
           "unit": "microseconds per StoresFind, one caller",
           "grpc_reference_wire_format": us_grpc, "grpc_nokeys_extension": us_lean, "grpc_empty_health_rpc": us_health,
           "c_abi_host_buffers_with_key_rows": us_abi, "device_launch_plus_sync": us_dev,
           "cpu_port_scan_only_no_rpc": us_cpu, "value": 1e6 / us_grpc, "value_unit": "queries/s through gRPC",
           "parity": {"ok": not problems, "checked_queries": 32, "rule": "similarities and ids equal to the oracle's scan of the whole store"}}
    if problems:
        out["parity"]["problems"] = problems[:6]
    return out


def measure_inproc(env, wl: Workload, args):
    """The same store as ONE logical store over all N GPUs inside ONE process (tfp_store_open_devices) -- the shape
    in which LocalAI's loader runs a backend (one process, pkg/model/process.go:94-124).  Rank 0 drives it through
    the plain tfp_store_find entry point with host buffers from several host threads (each call takes one of the
    store's find lanes); the other ranks wait.  Wall-clock: this is an end-to-end figure by construction."""
    import ctypes as C
    import torch
    from oracle import oracle as orc
    from localai_tfp_b200.store import Store
    world, device = env["world"], env["device"]
    st = Store(devices=list(range(world)), capacity_rows=wl.n_keys)
    L = st._L
    fp, i64p, u8p = C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_uint8)
    n_blocks = wl.n_keys // BLOCK_ROWS
    all_ids = np.empty(wl.n_keys, np.int64)
    pin = torch.empty((BLOCK_ROWS, wl.dim), dtype=torch.float32).pin_memory()
    for b in range(n_blocks):
        pin.copy_(gen_block_torch(wl, b, device))
        new = np.empty(BLOCK_ROWS, np.uint8)
        ids = all_ids[b * BLOCK_ROWS:(b + 1) * BLOCK_ROWS]
        st._check(L.tfp_store_set(st._h, C.cast(pin.data_ptr(), fp), BLOCK_ROWS, wl.dim, BLOCK_ROWS,
                                  ids.ctypes.data_as(i64p), new.ctypes.data_as(u8p)))
        assert new.all()
    shard_rows = [int(st.shard_stats(g).rows) for g in range(st.shard_count)]
    Q = gen_queries(wl)
    k = wl.topk

    def finder(nfinds, first, out, with_keys):
        n = C.c_int()
        ids = np.empty(k, np.int64); sims = np.empty(k, np.float32)
        rows = np.empty((k, wl.dim), np.float32) if with_keys else None
        for i in range(nfinds):
            q = Q[(first + i) % N_QUERIES]
            st._check(L.tfp_store_find(st._h, q.ctypes.data, wl.dim, k, C.byref(n), ids.ctypes.data, sims.ctypes.data,
                                       rows.ctypes.data if with_keys else None, None))
        out.append((ids.copy(), sims.copy(), (first + nfinds - 1) % N_QUERIES))

    def run(threads, nfinds, with_keys=False):
        outs = []
        ts = [threading.Thread(target=finder, args=(nfinds, 7 * t, outs, with_keys)) for t in range(threads)]
        t0 = time.perf_counter()
        [t.start() for t in ts]
        [t.join() for t in ts]
        return time.perf_counter() - t0, outs

    lanes = 4
    run(lanes, 8)
    run(1, 8, True)
    per_thread = max(args.steps, 50)
    qps = lanes * per_thread / float(np.median([run(lanes, per_thread)[0] for _ in range(5)]))
    qps_keys = lanes * per_thread / float(np.median([run(lanes, per_thread, True)[0] for _ in range(5)]))
    single_ms = float(np.median([run(1, per_thread)[0] for _ in range(3)])) / per_thread * 1e3
    single_keys_ms = float(np.median([run(1, per_thread, True)[0] for _ in range(3)])) / per_thread * 1e3
    # parity: returned rows re-scored by the oracle, and a sample of the store scanned by it
    order = np.argsort(all_ids)
    problems = []
    _, outs = run(4, 1)
    sample = np.concatenate([gen_block_torch(wl, b, device).cpu().numpy() for b in range(2)])
    sample_ids = all_ids[:len(sample)]
    for ids, sims, qi in outs:
        rows = order[np.searchsorted(all_ids, ids, sorter=order)]
        keys = np.empty((len(rows), wl.dim), np.float32)
        for j, r in enumerate(rows):
            keys[j] = gen_block_torch(wl, int(r) // BLOCK_ROWS, device)[int(r) % BLOCK_ROWS].cpu().numpy()
        if not np.array_equal(orc.scores_block(keys, Q[qi], True), sims):
            problems.append(f"q{qi}: returned similarity differs from the oracle's score of the returned row")
        if not (np.diff(sims.astype(np.float64)) <= 0).all() or len(ids) != k:
            problems.append(f"q{qi}: bad ordering or count")
        sc = orc.scores_block(sample, Q[qi], True)
        if set(sample_ids[sc > sims[-1]].tolist()) - set(ids.tolist()):
            problems.append(f"q{qi}: a sample row beats the k-th returned similarity but was not returned")
    st.close()
    out = {"what": "C2 as ONE logical store over %d GPUs in ONE process (tfp_store_open_devices), driven through tfp_store_find "
                   "with host buffers from %d host threads" % (world, lanes),
           "value": qps, "unit": "queries/s", "timing": "wall clock, median of 5 runs of %d finds per thread" % per_thread,
           "value_with_key_rows": qps_keys,
           "single_find_ms": single_ms, "single_find_with_key_rows_ms": single_keys_ms, "shard_rows": shard_rows,
           "h2d_bytes_per_find": world * (wl.dim * 4 + 16), "d2h_bytes_per_find_with_key_rows": 16 + k * 20 + k * wl.dim * 4,
           "parity": {"ok": not problems, "checked_queries": len(outs), "sample_rows_scanned_by_oracle": int(len(sample))}}
    if problems:
        out["parity"]["problems"] = problems[:6]
    return out


# ----------------------------------------------------------------------------- our arm
def run_ours(args, rank: int, local_rank: int, world: int):
    import torch
    import torch.distributed as dist
    from localai_tfp_b200 import build as tfp_build

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the store has no CPU path (use --impl reference for the CPU arm)")
    tfp_build.build()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
        cpu_group = dist.new_group(backend="gloo")     # a barrier that keeps no kernel spinning on the waiting ranks' GPUs
    env = {"torch": torch, "dist": dist, "rank": rank, "local_rank": local_rank, "world": world, "device": device}
    cache = {}
    wl = WORKLOADS[args.workload]
    head = measure(env, wl, args, cache, with_cpu=True, full_parity=2 if wl.name == "c2" else 0)

    also = {}
    names = []
    if args.also == "auto" and args.workload == "c2":
        names = ["target", "c3"] + (["c4"] if world == 8 else [])
    elif args.also not in ("auto", "none"):
        names = [n for n in args.also.split(",") if n in WORKLOADS and n != args.workload]
    for name in names:
        try:
            r = measure(env, WORKLOADS[name], args, cache, with_cpu=False, full_parity=0)
            also[name] = {k: r[k] for k in ("metric", "value", "unit", "ms_per_step", "repeats", "repeat_ms", "single_find_ms",
                                            "config", "clocks", "tensor_core_path", "e2e", "gpu_launches", "roofline", "parity")}
        except Exception as e:                      # the headline stands on its own; say what happened
            also[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            if world > 1:
                break                               # ranks may have diverged: do not enter another collective
    for sh in cache.values():
        sh.close()
    cache.clear()
    torch.cuda.empty_cache()
    if world == 1 and args.also == "auto" and args.workload == "c2":
        try:
            also["c1"] = measure_c1(env)
        except Exception as e:
            also["c1"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    if world > 1 and args.inproc and args.workload == "c2" and not any("error" in v for v in also.values()):
        # The other ranks wait on the CPU (gloo): an NCCL barrier would leave a kernel spinning on their GPUs, and a GPU
        # time-slices between processes -- rank 0's kernels there would get every other millisecond.
        torch.cuda.synchronize(device)
        dist.barrier(group=cpu_group)
        if rank == 0:
            try:
                also["inproc"] = measure_inproc(env, wl, args)
            except Exception as e:
                also["inproc"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        dist.barrier(group=cpu_group)

    if rank == 0 and "inproc" in also and "error" not in also["inproc"]:
        # The call a user of the drop-in makes at N GPUs is tfp_store_find on ONE store handle in ONE process (LocalAI runs
        # one backend process): that is the end-to-end figure -- one caller, one find after the other, host query in, host
        # ids + similarities + key rows out, like the one-GPU e2e.  The SPMD host's figure stays beside it.
        ip = also["inproc"]
        head["e2e_spmd_host"] = dict(head["e2e"], path="ShardedStore.StoresFind, one process per GPU (torch.distributed bootstrap)")
        head["e2e"] = {"value": 1e3 / ip["single_find_with_key_rows_ms"], "unit": "queries/s",
                       "h2d_bytes_per_step": ip["h2d_bytes_per_find"], "d2h_bytes_per_step": ip["d2h_bytes_per_find_with_key_rows"],
                       "path": "tfp_store_find on one logical store over %d GPUs in one process, one caller thread, "
                               "wall clock; %.0f queries/s with 4 caller threads" % (world, ip["value_with_key_rows"])}

    ok = head["parity"]["ok"] and all(v.get("parity", {}).get("ok", True) for v in also.values())
    if rank == 0:
        line = {
            "metric": head["metric"], "value": head["value"], "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        }
        for k in ("repeats", "repeat_ms", "timing", "single_find_ms", "config", "clocks", "tensor_core_path", "e2e",
                  "e2e_spmd_host", "gpu_launches", "roofline", "parity", "cpu_baseline"):
            if k in head:
                line[k] = head[k]
        if also:
            line["also"] = also
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        raise SystemExit("bench.py: PARITY FAILURE against the oracle (see the `parity` blocks of the line above)")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=list(WORKLOADS), default="c2")
    ap.add_argument("--also", default="auto",
                    help="extra workloads measured after the headline in the same run: auto (target + c3, and c4 on 8 GPUs, "
                         "when the headline is c2), none, or a comma-separated list")
    ap.add_argument("--inproc", type=int, default=1,
                    help="N>1: also measure the store as one process over all N GPUs (rank 0 only, after the ranks' own runs)")
    ap.add_argument("--lanes", type=int, default=0,
                    help="single-query finds kept in flight per GPU (find contexts); default 2 on one GPU, 4 on several "
                         "(a sharded find waits for its slowest peer: more finds in flight absorb the skew between ranks)")
    ap.add_argument("--batch-lanes", type=int, default=2,
                    help="one GPU, batched workloads: contexts that take the batches in turn (1 or 2)")
    ap.add_argument("--exchange", choices=["p2p", "torch", "native", "native-per-lane"], default="p2p",
                    help="N>1: records pushed to the peers by the scan kernel (p2p), ncclAllGather inside libtfpstore "
                         "(native), or torch.distributed (torch)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.lanes <= 0:
        args.lanes = 4 if int(os.environ.get("WORLD_SIZE", "1")) > 1 else 2
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
