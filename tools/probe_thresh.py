"""Where does the tensor-core path start to beat the exact multi-query kernel?  (C2 shape, small batches)
Run with TFP_MMA_MIN_QUERIES=1000 (exact only) and =2 (tensor cores from 2 queries up)."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import localai_tfp_b200 as tfp

n, d = 1_000_000, 768
g = torch.Generator(device="cuda").manual_seed(3)
K = torch.randn((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
s = tfp.Store(capacity_rows=n); s.set_device(K.data_ptr(), n, d)
Q = K[:16].cpu().numpy().copy(); del K
c = s.find_context(); c.upload(Q)
for nq in (2, 3, 4, 5, 6, 8, 12, 16):
    for k in (10,):
        c.launch(0, nq, k); c.sync()
        t0 = time.perf_counter()
        for _ in range(20):
            c.launch(0, nq, k)
        c.sync()
        us = (time.perf_counter() - t0) / 20 * 1e6
        print(f"min_queries={os.environ.get('TFP_MMA_MIN_QUERIES')} nq={nq} k={k}: {us:.0f} us/batch  {us/nq:.0f} us/query  mma={c.mma_counters()}", flush=True)
