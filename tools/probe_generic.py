"""Key lengths the TMA scan does not take (D % 4 != 0, or D < 32) go through scan_generic_kernel: µs per find and the
rate at which the keys are read, one find at a time and eight queries per pass."""
import os, sys, time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import localai_tfp_b200 as tfp


def run(n, d, k=10, iters=30):
    g = torch.Generator(device="cuda").manual_seed(n + d)
    s = tfp.Store(capacity_rows=n)
    blk = 500_000
    for lo in range(0, n, blk):
        m = min(blk, n - lo)
        K = torch.randn((m, d), generator=g, device="cuda", dtype=torch.float32)
        K /= K.norm(dim=1, keepdim=True)
        s.set_device(K.data_ptr(), m, d)
        del K
    Q = np.random.default_rng(1).standard_normal((8, d)).astype(np.float32)
    Q /= np.linalg.norm(Q, axis=1, keepdims=True)
    c = s.find_context()
    c.upload(Q)
    out = []
    for nq in (1, 2):
        for _ in range(3):
            c.launch(0, nq, k)
        c.sync()
        t0 = time.perf_counter()
        for i in range(iters):
            c.launch(0, nq, k)
        c.sync()
        us = (time.perf_counter() - t0) / iters * 1e6
        out.append("%d quer%s per pass: %.1f us, %.0f GB/s" % (nq, "y" if nq == 1 else "ies", us, n * d * 4 / us / 1e3))
    print("%d x %d: " % (n, d) + "; ".join(out), flush=True)
    c.close(); s.close()


if __name__ == "__main__":
    run(4_000_000, 50)
    run(1_000_000, 301)
    run(2_000_000, 25)
    run(500_000, 1023)
    run(1_000_000, 300)     # for comparison: a length the TMA scan takes
