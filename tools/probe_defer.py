"""How long tfp_find_ctx_launch keeps the host for one tensor-core batch, against the batch's time on the GPU, and
what two contexts taking batches in turn gain from it.  Run once as is and once with TFP_MMA_DEFER=0 (the launch then
reads the batch's status words before it returns, as every launch did before):

    python tools/probe_defer.py            TFP_MMA_DEFER=0 python tools/probe_defer.py
"""
import json, os, sys, time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import localai_tfp_b200 as tfp


def run(n, d, nq, k, iters=40):
    g = torch.Generator(device="cuda").manual_seed(n + d)
    s = tfp.Store(capacity_rows=n)
    blk = 250_000
    for lo in range(0, n, blk):
        m = min(blk, n - lo)
        K = torch.randn((m, d), generator=g, device="cuda", dtype=torch.float32)
        K /= K.norm(dim=1, keepdim=True)
        s.set_device(K.data_ptr(), m, d)
        del K
    Q = np.random.default_rng(1).standard_normal((nq, d)).astype(np.float32)
    Q /= np.linalg.norm(Q, axis=1, keepdims=True)
    ctxs = [s.find_context(), s.find_context()]
    for c in ctxs:
        c.upload(Q)
        for _ in range(3):
            c.launch(0, nq, k)
        c.sync()
    out = {"rows": n, "dim": d, "queries": nq, "topk": k, "defer": os.environ.get("TFP_MMA_DEFER", "1")}
    # host time of the launch call itself, the GPU idle when it is made
    a = ctxs[0]
    held = []
    for _ in range(iters):
        t0 = time.perf_counter()
        a.launch(0, nq, k)
        held.append(time.perf_counter() - t0)
        a.sync()
    out["launch_call_us_median"] = float(np.median(held) * 1e6)
    # one context, batches back to back; then two contexts in turn
    for lanes in (1, 2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(iters):
            ctxs[i % lanes].launch(0, nq, k)
        for c in ctxs[:lanes]:
            c.sync()
        dt = time.perf_counter() - t0
        out["batch_ms_%d_context%s" % (lanes, "s" if lanes > 1 else "")] = dt / iters * 1e3
    out["mma_passes"], out["exact_reruns"] = a.mma_counters()
    print(json.dumps(out), flush=True)
    for c in ctxs:
        c.close()
    s.close()


if __name__ == "__main__":
    run(1_000_000, 768, 256, 10)
    run(250_000, 768, 64, 10)
