"""Development probe for the tensor-core batched path (C3-like shapes)."""
import os, sys, time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import localai_tfp_b200 as tfp
from localai_tfp_b200 import build as b

b.build()


def run(n, d, nq, k, iters=5, label=""):
    g = torch.Generator(device="cuda").manual_seed(n + d)
    s = tfp.Store(capacity_rows=n)
    blk = 250_000
    for lo in range(0, n, blk):
        m = min(blk, n - lo)
        K = torch.randn((m, d), generator=g, device="cuda", dtype=torch.float32)
        K /= K.norm(dim=1, keepdim=True)
        s.set_device(K.data_ptr(), m, d)
        if lo == 0:
            Q = K[:nq].cpu().numpy().copy()
            Q += np.random.default_rng(1).standard_normal(Q.shape).astype(np.float32) * 0.05
            Q /= np.linalg.norm(Q, axis=1, keepdims=True)
            Q = Q.astype(np.float32)
        del K
    torch.cuda.empty_cache()
    c = s.find_context()
    c.upload(Q)
    c.launch(0, nq, k); c.sync()
    t0 = time.perf_counter()
    for _ in range(iters):
        c.launch(0, nq, k)
    c.sync()
    ms = (time.perf_counter() - t0) / iters * 1e3
    passes, reruns = c.mma_counters()
    tf = 2.0 * nq * n * d / (ms * 1e-3) / 1e12
    print(f"{label} n={n} d={d} nq={nq} k={k}: {ms:.2f} ms/batch  {nq/ms*1e3:.0f} qps  {tf:.1f} TFLOP/s credited "
          f"({3*tf:.1f} issued)  HBM {n*d*4/ms/1e6:.0f} GB/s  mma_passes={passes} reruns={reruns}", flush=True)
    c.close(); s.close()


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), flush=True)
    run(1_000_000, 768, 256, 10, label="C2xQ256")
    run(2_000_000, 1024, 256, 100, label="C3/5")
    run(2_000_000, 1024, 64, 100, label="C3/5 q64")
    run(2_000_000, 1024, 32, 100, label="C3/5 q32")
    run(4_000_000, 384, 256, 10, label="d384")
