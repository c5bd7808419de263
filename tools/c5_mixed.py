"""C5 (BASELINE.json configs[4]): mixed stream on N GPUs -- StoresSet ingest of 1536-d embeddings interleaved
with StoresDelete and StoresFind.  One process per GPU (torchrun); every rank ingests its own share of the
stream into its shard (the key -> shard routing of sharded.py, applied at the producer), deletes 10 % of what
it wrote earlier and takes part in sharded finds.  Prints aggregate set/s, delete/s, find/s and find latency."""
import os, sys, time
import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from localai_tfp_b200.sharded import CudaShardOps, ShardedStore

D, B, SECONDS = 1536, 8192, float(os.environ.get("C5_SECONDS", "8"))
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
os.environ.setdefault("NCCL_DEBUG", "WARN")
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
ops = CudaShardOps(rank, local, capacity_rows=2_000_000)
sh = ShardedStore(ops, rank, world) if world > 1 else None
store = ops.store
if sh is not None:
    sh.key_len = D
    ops.latch(D)
    ops.init_native_comm(rank, world)
rng = np.random.default_rng(1000 + rank)
pool = []
for i in range(6):                                   # pinned-size host batches, reused with a per-use twist
    k = rng.standard_normal((B, D), dtype=np.float32)
    k /= np.linalg.norm(k, axis=1, keepdims=True)
    pool.append(np.ascontiguousarray(k))
Q = np.ascontiguousarray(pool[0][:16].copy())
vals = [b""] * B
written = []
nset = ndel = nfind = 0
lat = []
store.StoresSet(pool[0], vals); written.append(pool[0][::10].copy())          # warm-up, not timed
ops.stage_queries(Q)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
step = 0
while time.perf_counter() - t0 < SECONDS:
    step += 1
    k = pool[step % len(pool)]
    k[:, 0] += np.float32(1e-3)                      # new keys every step, in place (no longer unit length: cosine path)
    store.StoresSet(k, vals); nset += B
    written.append(k[::10].copy())                   # the 10 % that will be deleted three steps later
    if len(written) > 3:
        old = written.pop(0)
        store.StoresDelete(old); ndel += len(old)
    for j in range(16):
        tq = time.perf_counter()
        if sh is not None:
            ops.lanes[0].ctx.launch_sharded(j, 10)
            ops.lanes[0].ctx.download_merged()
        else:
            store.find_ids(Q[j], 10)
        lat.append(time.perf_counter() - tq); nfind += 1
    if world > 1 and step % 4 == 0:                  # keep the ranks within a few steps of each other
        dist.barrier()
dt = time.perf_counter() - t0
tot = torch.tensor([nset, ndel, float(len(store))], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(tot)
if rank == 0:
    lat = np.array(lat) * 1e3
    print(f"C5 mixed stream, {world} GPU(s), {dt:.1f} s: {tot[0].item()/dt/1e6:.2f} M set/s, {tot[1].item()/dt/1e3:.0f} k delete/s, "
          f"{nfind/dt:.0f} find/s (sharded over all GPUs), find latency p50 {np.percentile(lat,50):.2f} ms p99 {np.percentile(lat,99):.2f} ms, "
          f"{int(tot[2].item())} rows resident at the end", flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
ops.close()
