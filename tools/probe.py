"""Quick GPU probe used during development: scan bandwidth at a few shapes, end-to-end latency.
Not part of the test suite or the bench contract."""
import os, sys, time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import localai_tfp_b200 as tfp
from localai_tfp_b200 import build as b

b.build()


def run(n, d, k, normalized, nq_per_launch=1, iters=50, label=""):
    g = torch.Generator(device="cuda").manual_seed(n + d)
    K = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float32)
    K /= K.norm(dim=1, keepdim=True)
    if not normalized:
        K *= (0.5 + 1.5 * torch.rand((n, 1), generator=g, device="cuda"))
    s = tfp.Store(capacity_rows=n)
    t0 = time.time()
    s.set_device(K.data_ptr(), n, d)
    t_ingest = time.time() - t0
    Q = K[:64].cpu().numpy().copy()
    del K
    torch.cuda.empty_cache()
    c = s.find_context()
    c.upload(Q)
    st = torch.cuda.ExternalStream(c.stream)
    for i in range(5):
        c.launch(i % 8, nq_per_launch, k)
    c.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(iters):
        c.launch((i * nq_per_launch) % (64 - nq_per_launch + 1), nq_per_launch, k)
    e1.record(st)
    c.sync(); e1.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gbs = n * d * 4 / ms / 1e6
    # two contexts alternating: the next query's CTAs fill the SMs the previous query's tail leaves idle
    c2 = s.find_context()
    c2.upload(Q)
    for i in range(4):
        (c if i % 2 == 0 else c2).launch(i % 8, nq_per_launch, k)
    c.sync(); c2.sync()
    t0 = time.perf_counter()
    for i in range(iters):
        (c if i % 2 == 0 else c2).launch((i * nq_per_launch) % (64 - nq_per_launch + 1), nq_per_launch, k)
    c.sync(); c2.sync()
    ms2 = (time.perf_counter() - t0) / iters * 1e3
    c2.close()
    # end-to-end single find through the synchronous ABI
    t0 = time.perf_counter()
    for i in range(20):
        s.find_ids(Q[i % 64], k)
    e2e = (time.perf_counter() - t0) / 20 * 1e3
    print(f"{label} n={n} d={d} k={k} norm={normalized} nq={nq_per_launch}: {ms*1e3:.1f} us/launch "
          f"{gbs:.0f} GB/s  qps={nq_per_launch/ms*1e3:.0f}  2ctx={ms2*1e3:.1f} us  e2e={e2e*1e3:.0f} us  ingest={t_ingest:.2f}s", flush=True)
    c.close(); s.close()


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), flush=True)
    if len(sys.argv) > 1 and sys.argv[1] == "quick":
        run(1_000_000, 768, 10, True, iters=200, label="C2")
        run(1_000_000, 768, 10, True, nq_per_launch=4, iters=100, label="C2-nq4")
        run(1_000_000, 768, 100, True, iters=200, label="C2-k100")
        run(2_000_000, 1024, 10, True, iters=100, label="d1024")
        sys.exit(0)
    run(10_000, 384, 10, True, label="C1")
    run(1_000_000, 768, 10, True, label="C2")
    run(1_000_000, 768, 10, False, label="C2-fallback")
    run(1_000_000, 768, 10, True, nq_per_launch=4, label="C2-nq4")
    run(1_000_000, 768, 10, True, nq_per_launch=8, label="C2-nq8")
    run(1_000_000, 768, 100, True, label="C2-k100")
    run(2_000_000, 1024, 10, True, label="d1024")
    run(4_000_000, 384, 10, True, label="d384")
