"""Development probe for the mutation path (C5-like): StoresSet / StoresDelete / StoresFind interleaved."""
import os, sys, time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import localai_tfp_b200 as tfp
from localai_tfp_b200 import build as b

b.build()
d, B = 1536, 8192
rng = np.random.default_rng(0)
s = tfp.Store(capacity_rows=600_000)
batches = []
for i in range(24):
    k = rng.standard_normal((B, d), dtype=np.float32)
    k /= np.linalg.norm(k, axis=1, keepdims=True)
    batches.append(np.ascontiguousarray(k.astype(np.float32)))
vals = [b""] * B
s.StoresSet(batches[0], vals)          # warm-up (allocations, first launches)
t0 = time.perf_counter()
for k in batches[1:17]:
    s.StoresSet(k, vals)
dt = time.perf_counter() - t0
print(f"StoresSet (host keys, {B} x {d} per call): {16*B/dt/1e6:.3f} M keys/s, {dt/16*1e3:.2f} ms per call, {16*B*d*4/dt/1e9:.2f} GB/s", flush=True)
t0 = time.perf_counter()
s.StoresSet(batches[3], vals)
print(f"StoresSet overwrite of {B} existing keys: {(time.perf_counter()-t0)*1e3:.2f} ms", flush=True)
# mixed: set one batch, delete 10% of an older batch, 16 finds
q = batches[5][:16]
t0 = time.perf_counter(); nset = ndel = nfind = 0
for i, k in enumerate(batches[17:24]):
    s.StoresSet(k, vals); nset += B
    victims = batches[1 + i][:: 10]
    s.StoresDelete(victims); ndel += len(victims)
    for j in range(16):
        s.find_ids(q[j], 10); nfind += 1
dt = time.perf_counter() - t0
print(f"mixed: {nset/dt/1e6:.3f} M set/s, {ndel/dt/1e3:.1f} k delete/s, {nfind/dt:.0f} find/s over {len(s)} rows", flush=True)
# device-resident bulk ingest
g = torch.Generator(device="cuda").manual_seed(1)
K = torch.randn((200_000, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
s.set_device(K.data_ptr(), 200_000, d)
dt = time.perf_counter() - t0
print(f"set_device (keys already in HBM): {200_000/dt/1e6:.3f} M keys/s ({200_000*d*4/dt/1e9:.1f} GB/s)", flush=True)
s.close()
