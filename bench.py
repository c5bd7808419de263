#!/usr/bin/env python
"""bench.py -- StoresFind throughput of the B200 local-store (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (all N): BASELINE.json configs[1] -- 1M keys x 768-d fp32, unit-length rows, single-query
top-10 on the normalised-key fast path.  A "step" is one StoresFind over the whole store.
N > 1 shards the SAME store by rows across the ranks (strong scaling): every step each rank scans
its shard, the per-shard top-k records are all-gathered over NCCL and merged on the device.

Prints ONE JSON line (rank 0).  `value` = queries/s with keys and queries resident in HBM;
`e2e` = the same metric through the public call with host buffers; `roofline` = the scan kernel
against measured HBM bandwidth; `cpu_baseline` = the CPU oracle (a port of the reference's Go
code) timed on this box.  `--impl reference` times that CPU port as its own arm.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_KEYS = 1_000_000
DIM = 768
TOPK = 10
N_QUERIES = 64
QUERIES_PER_STEP = 1
MMA_TERMS = 3 if os.environ.get("TFP_MMA_TERMS") == "3" else 1   # tensor-core filter: TF32 MMAs per K-step
BLOCK_ROWS = 62_500            # generation granularity: 16 blocks, so 1/2/4/8 shards align with blocks
SEED = 0x5EED0002
WORKLOAD = "C2 (BASELINE.json configs[1]): 1M keys x 768-d fp32, single-query top-10, normalised-key fast path"
METRIC = "StoresFind queries/sec (top-10)"


def select_workload(name: str):
    """`c2` is the bench contract's workload.  `target` is the north star's stated target shape
    (10M keys x 1024-d, top-10, batch 1: 41 GB, strong-scaled over the GPUs) for extra evidence."""
    global N_KEYS, DIM, WORKLOAD, SEED
    global TOPK, QUERIES_PER_STEP, METRIC
    if name == "target":
        N_KEYS, DIM, SEED = 10_000_000, 1024, 0x5EED0003
        WORKLOAD = "north-star target: 10M keys x 1024-d fp32, single-query top-10, normalised-key fast path"
    if name == "c4":
        N_KEYS, DIM, SEED = 100_000_000, 384, 0x5EED0004
        WORKLOAD = ("C4 (BASELINE.json configs[3]): 100M keys x 384-d fp32 (153.6 GB), single-query top-10, "
                    "sharded by rows, NCCL candidate merge")
    if name == "c3":
        N_KEYS, DIM, SEED, TOPK, QUERIES_PER_STEP = 10_000_000, 1024, 0x5EED0003, 100, 256
        METRIC = "StoresFind queries/sec (top-100, batch 256)"
        WORKLOAD = ("C3 (BASELINE.json configs[2]): 10M keys x 1024-d fp32, batch-256 queries top-100 on the "
                    "tensor-core path (tcgen05 %dxTF32 filter + exact re-score)" % MMA_TERMS)


# ----------------------------------------------------------------------------- inputs
def gen_block_torch(block: int, device):
    import torch
    g = torch.Generator(device=device).manual_seed(SEED + block)
    k = torch.randn((BLOCK_ROWS, DIM), generator=g, device=device, dtype=torch.float32)
    k /= k.norm(dim=1, keepdim=True)
    return k


def gen_queries() -> np.ndarray:
    rng = np.random.default_rng(SEED + 1)
    q = rng.standard_normal((max(N_QUERIES, QUERIES_PER_STEP), DIM)).astype(np.float32)
    q /= np.linalg.norm(q.astype(np.float64), axis=1, keepdims=True).astype(np.float32)
    return np.ascontiguousarray(q.astype(np.float32))


def gen_keys_numpy(rows: int) -> np.ndarray:
    rng = np.random.default_rng(SEED)
    k = rng.standard_normal((rows, DIM), dtype=np.float32)
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

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                self.bits |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            self._stop.wait(0.01)

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
def cpu_reference_qps(budget_s: float, steps: int, warmup: int, full: bool):
    """Times the CPU port of the reference (oracle/, `faithful` flavour: separately allocated rows in
    sorted order, one heap item allocated per key, container/heap mechanics) on a bounded row sample
    of the C2 store and scales to queries/s over the full 1M-row store (the scan is linear in rows)."""
    from oracle import oracle as orc
    orc.build()
    q = gen_queries()
    probe_rows = 20_000
    probe = gen_keys_numpy(probe_rows)
    st = orc.OracleStore()
    st.StoresSet(probe, [b""] * probe_rows)
    t0 = time.perf_counter()
    st.StoresFind(q[0], TOPK, faithful=True)
    per_row = (time.perf_counter() - t0) / probe_rows
    rows = int(min(N_KEYS, 1_000_000, max(probe_rows, budget_s / max(steps + warmup, 1) / per_row)))
    if rows > probe_rows:
        keys = gen_keys_numpy(rows)
        st = orc.OracleStore()
        st.StoresSet(keys, [b""] * rows)
    assert st.keys_are_normalized
    for i in range(warmup):
        st.StoresFind(q[i % N_QUERIES], TOPK, faithful=True)
    t0 = time.perf_counter()
    for i in range(steps):
        st.StoresFind(q[i % N_QUERIES], TOPK, faithful=True)
    dt = (time.perf_counter() - t0) / steps
    assert st.last_path == "normalized"
    qps_full = 1.0 / (dt * N_KEYS / rows)
    sample = (f"{steps} single-query top-{TOPK} finds over a {rows}-row x {DIM}-d sample of the store, "
              f"1 thread (the reference scan is single-threaded by design, store.go:18), scaled by rows to {N_KEYS}")
    return qps_full, dt * 1e3, rows, sample


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    qps, ms, rows, sample = cpu_reference_qps(budget_s=120.0, steps=args.steps, warmup=args.warmup, full=True)
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "keys": N_KEYS, "dim": DIM, "topk": TOPK, "queries_per_step": 1,
                   "sample_rows_per_step": rows},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": 1, "kind": "port", "sample": sample,
                         "host_cores": os.cpu_count()},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- our arm
def run_ours(args, rank: int, local_rank: int, world: int):
    import torch
    import torch.distributed as dist
    import localai_tfp_b200 as tfp
    from localai_tfp_b200 import build as tfp_build
    from localai_tfp_b200.sharded import CudaShardOps, ShardedStore

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the store has no CPU path (use --impl reference for the CPU arm)")
    tfp_build.build()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)

    # ---- this rank's shard: blocks [b0, b1) of the same 1M x 768 matrix at every N
    n_blocks = N_KEYS // BLOCK_ROWS
    b0, b1 = rank * n_blocks // world, (rank + 1) * n_blocks // world
    rows_local = (b1 - b0) * BLOCK_ROWS
    ops = CudaShardOps(rank, local_rank, capacity_rows=rows_local, lanes=max(2, args.lanes))
    store = ops.store
    for b in range(b0, b1):
        blk = gen_block_torch(b, device)
        ids, new = store.set_device(blk.data_ptr(), BLOCK_ROWS, DIM)
        assert bool(new.all())
        del blk
    torch.cuda.empty_cache()
    assert len(store) == rows_local and store.keys_are_normalized
    sharded = ShardedStore(ops, rank, world) if world > 1 else None
    if sharded is not None:
        sharded.key_len = DIM
        if args.exchange == "p2p":              # the scan kernel pushes its records to the peers over NVLink
            try:
                ops.init_p2p(rank, world)       # all ranks succeed or all ranks raise (it agrees step by step)
            except Exception as e:
                if rank == 0:
                    print(f"bench: {e}; falling back to ncclAllGather", file=sys.stderr)
                args.exchange = "native"
        if args.exchange in ("native", "native-per-lane"):   # scan + ncclAllGather + merge in one library call
            try:
                ops.init_native_comm(rank, world, shared=(args.exchange == "native"))
            except Exception as e:              # NCCL not resolvable by dlopen: same exchange through torch.distributed
                if rank == 0:
                    print(f"bench: native exchange unavailable ({e}); using torch.distributed", file=sys.stderr)
                args.exchange = "torch"

    Q = gen_queries()
    ctx = ops.ctx
    ctx.upload(Q)              # queries resident in HBM before the timed region
    ctx.sync()
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    # Two finds are kept in flight (two find-contexts, alternating): the next query's CTAs fill the
    # SMs that the previous query's tail and merge leave idle and, at N > 1, its scan overlaps the
    # previous query's all-gather.  Every query is still its own scan (batch 1).
    in_flight = max(2, args.lanes) if QUERIES_PER_STEP == 1 else 1
    for ln in ops.lanes[1:]:
        ln.ctx.upload(Q)
        ln.ctx.sync()
    lane_ctx = [ln.ctx for ln in ops.lanes]
    side_streams = [ln.stream for ln in ops.lanes[1:]]

    batched = QUERIES_PER_STEP > 1

    def step_device(i: int):
        if batched:          # one batch of QUERIES_PER_STEP queries per step (tensor-core path)
            if world == 1:
                ctx.launch(0, QUERIES_PER_STEP, TOPK)
            else:
                ctx.launch_sharded(0, TOPK, count=QUERIES_PER_STEP)
        elif world == 1:
            lane_ctx[i % in_flight].launch(i % N_QUERIES, 1, TOPK)
        else:
            sharded.find_device(i % N_QUERIES, TOPK, with_keys=False, lane=i % in_flight)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def launches_now():
        return sum(c.launch_count for c in lane_ctx)

    # ---- device-resident throughput (`value`)
    with torch.cuda.stream(stream):
        for i in range(args.warmup):
            step_device(i)
        barrier()
        launches0 = launches_now()
        clocks = ClockSampler(local_rank)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clocks.start()
        e0.record(stream)
        for sb in side_streams:
            sb.wait_event(e0)                # every lane's stream starts after e0 ...
        for i in range(args.steps):
            step_device(i)
        for sb in side_streams:
            eb = torch.cuda.Event()
            eb.record(sb)
            stream.wait_event(eb)            # ... and e1 follows the last kernel of every lane
        e1.record(stream)
        barrier()
        clk = clocks.stop()
        ms_total = e0.elapsed_time(e1)
        launches = launches_now() - launches0 + (args.steps if world > 1 and args.exchange == "torch" else 0)   # torch path: + the merge kernel
    t = torch.tensor([ms_total], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    qps = 1e3 / ms_per_step * QUERIES_PER_STEP

    # ---- the scan kernel alone for the roofline: back to back on ONE stream, nothing overlapping it
    with torch.cuda.stream(stream):
        k = min(args.steps, 200)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(device)
        s0.record(stream)
        for i in range(k):
            if batched:
                ctx.launch(0, QUERIES_PER_STEP, TOPK)
            else:
                ctx.launch(i % N_QUERIES, 1, TOPK)
        s1.record(stream)
        torch.cuda.synchronize(device)
        scan_ms = s0.elapsed_time(s1) / k
    algo_bytes = rows_local * DIM * 4
    achieved = algo_bytes / (scan_ms * 1e-3) / 1e9
    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        pass
    roof_bound, roof_unit, roof_kernel = "hbm", "GB/s", "scan_tma_kernel<1,fp32,fused>"
    if batched:
        # credited flops = 2 * Q * rows * D per batch (the filter issues MMA_TERMS times that)
        roof_bound, roof_unit, roof_kernel = "tensor", "TFLOP/s", "mma_scan_kernel (tcgen05.mma kind::tf32)"
        achieved = 2.0 * QUERIES_PER_STEP * rows_local * DIM / (scan_ms * 1e-3) / 1e12
        peak, peak_src = 1590.0 / 2, "fallback (B200_PROFILING.md bf16 1.59 PFLOP/s, halved for TF32)"
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]) / 2
            peak_src = "derived: half of the measured bf16 burst peak (MEASURED_PEAKS.json bf16_tflops); TF32 dense runs at half the bf16 rate"
        except Exception:
            pass
    traffic = None
    try:
        if args.workload == "c2" and world == 1:     # the ncu capture was taken on exactly this launch shape
            traffic = json.load(open(os.path.join(ROOT, "profiles", "scan_traffic.json")))["dram_bytes_per_launch"]
    except Exception:
        pass

    # ---- end to end through the public call: host query in, host (ids, sims, keys) out, every step
    e2e_steps = args.steps
    barrier()
    if batched:
        e2e_steps = min(args.steps, 20)
        qb = Q[:QUERIES_PER_STEP]

        def e2e_once():
            if world == 1:
                return store.find_batch(qb, TOPK)[0][1]
            ctx.upload(qb)
            ctx.launch_sharded(0, TOPK, count=QUERIES_PER_STEP)
            return ctx.download_merged()[0][1]
        e2e_once()
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            sims = e2e_once()
        torch.cuda.synchronize(device)
        e2e_dt = time.perf_counter() - t0
        ctx.upload(Q)
    elif world == 1:
        for i in range(min(args.warmup, 5)):
            store.StoresFind(Q[i % N_QUERIES], TOPK)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            ks, vs, sims = store.StoresFind(Q[i % N_QUERIES], TOPK)
        torch.cuda.synchronize(device)
        e2e_dt = time.perf_counter() - t0
    else:
        for i in range(min(args.warmup, 5)):
            sharded.StoresFind(Q[i % N_QUERIES], TOPK)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            ks, vs, sims = sharded.StoresFind(Q[i % N_QUERIES], TOPK)
        torch.cuda.synchronize(device)
        e2e_dt = time.perf_counter() - t0
    t = torch.tensor([e2e_dt], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_qps = e2e_steps * QUERIES_PER_STEP / float(t.item())
    assert len(sims) == TOPK and list(sims) == sorted(sims, reverse=True)
    h2d = QUERIES_PER_STEP * (DIM * 4 + 16)              # queries + their QInfo records
    d2h = QUERIES_PER_STEP * (4 + TOPK * (8 + 4)) + (TOPK * DIM * 4 if not batched else 0)   # n, ids, sims (+ key rows)

    # ---- CPU baseline beside it (rank 0, N=1 only)
    cpu = None
    if world == 1:
        cqps, cms, crows, csample = cpu_reference_qps(budget_s=15.0, steps=6, warmup=1, full=False)
        cpu = {"value": cqps, "unit": "queries/s", "cores": 1, "kind": "port", "sample": csample,
               "host_cores": os.cpu_count()}

    if rank == 0:
        line = {
            "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "keys": N_KEYS, "dim": DIM, "topk": TOPK, "queries_per_step": QUERIES_PER_STEP,
                       "path": "normalized", "rows_per_gpu": rows_local, "finds_in_flight": in_flight,
                       "exchange": args.exchange if world > 1 else None,
                       "l2": "inputs larger than L2 (keys %.2f GB per GPU vs 126 MB); no flush" % (algo_bytes / 1e9)
                             if algo_bytes > 4 * 126e6 else "keys per GPU %.0f MB: partly L2-resident" % (algo_bytes / 1e6),
                       "parallelism": ("row-sharded x%d, " % world) +
                                      ("top-k records pushed to every peer over NVLink by the scan kernel, stream wait, device merge"
                                       if args.exchange == "p2p" else "one all-gather of top-k records + device merge per query")
                                      if world > 1 else "single GPU"},
            "clocks": clk,
            "tensor_core_path": dict(zip(("passes", "exact_reruns"), ctx.mma_counters())) if batched else None,
            "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            "roofline": {"bound": roof_bound, "achieved": achieved, "peak": peak, "unit": roof_unit, "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": roof_kernel,
                         "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": scan_ms,
                         "note": "achieved = bytes / duration of one scan launched alone back to back on one stream; "
                                 "`value` overlaps the tail of one scan with the start of the next (finds_in_flight)",
                         "effective_gbs_in_timed_region": algo_bytes / (ms_per_step * 1e-3) / 1e9 if world == 1 else None,
                         "frac_of_nominal_8TBs": achieved / 8000.0 if not batched else None,
                         "tf32_terms": MMA_TERMS if batched else None,
                         "issued_tf32_tflops": MMA_TERMS * achieved if batched else None,
                         "frac_of_nominal_tf32_1100": MMA_TERMS * achieved / 1100.0 if batched else None},
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    ops.close()          # (with the peer exchange this is collective: unmap everywhere, barrier, free)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=["c2", "target", "c3", "c4"], default="c2")
    ap.add_argument("--lanes", type=int, default=0,
                    help="single-query finds kept in flight per GPU (find contexts); default 2 on one GPU, 4 on several "
                         "(a sharded find waits for its slowest peer: more finds in flight absorb the skew between ranks)")
    ap.add_argument("--exchange", choices=["p2p", "torch", "native", "native-per-lane"], default="p2p",
                    help="N>1: records pushed to the peers by the scan kernel (p2p), ncclAllGather inside libtfpstore "
                         "(native), or torch.distributed (torch)")
    args = ap.parse_args()
    select_workload(args.workload)
    args.warmup = max(args.warmup, 3)
    if args.lanes <= 0:
        args.lanes = 4 if int(os.environ.get("WORLD_SIZE", "1")) > 1 else 2
    os.environ.setdefault("NCCL_DEBUG", "WARN")     # keep NCCL's version banner off stdout: ONE JSON line is the contract
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
