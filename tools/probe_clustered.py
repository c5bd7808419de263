"""Tensor-core batch path on data whose similarities crowd together: time per 256-query batch
(top-100, 1M x 1024 keys) and how many queries needed the 3xTF32 pass / an exact re-run, for three key distributions."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import localai_tfp_b200 as tfp

n, d, nq, k = 1_000_000, 1024, 256, 100
g = torch.Generator(device="cuda").manual_seed(5)


def gaussian():
    K = torch.randn((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
    Q = torch.randn((nq, d), generator=g, device="cuda"); Q /= Q.norm(dim=1, keepdim=True)
    return K, Q


def uniform01():        # the reference's own test distribution (stores_test.go:326-333), unit length
    K = torch.rand((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
    Q = torch.rand((nq, d), generator=g, device="cuda"); Q /= Q.norm(dim=1, keepdim=True)
    return K, Q


def cluster():          # every key within ~1e-3 of every other in similarity
    c = torch.randn((1, d), generator=g, device="cuda"); c /= c.norm()
    K = c + (0.05 / d ** 0.5) * torch.randn((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
    Q = c + (0.05 / d ** 0.5) * torch.randn((nq, d), generator=g, device="cuda"); Q /= Q.norm(dim=1, keepdim=True)
    return K, Q


for name, make in (("gaussian", gaussian), ("uniform[0,1)", uniform01), ("tight cluster", cluster)):
    K, Q = make()
    s = tfp.Store(capacity_rows=n); s.set_device(K.data_ptr(), n, d)
    Qh = Q.cpu().numpy(); del K, Q
    c = s.find_context(); c.upload(Qh)
    c.launch(0, nq, k); c.sync()
    p0, r0, e0 = c.mma_counters() + (c.mma_escalated,)
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        c.launch(0, nq, k)
    c.sync()
    ms = (time.perf_counter() - t0) / reps * 1e3
    p1, r1, e1 = c.mma_counters() + (c.mma_escalated,)
    res = c.download()
    ids, sims = s.find_ids(Qh[7], k)          # spot check against the exact single-query kernel
    ok = np.array_equal(res[7][1], sims)
    print(f"{name}: {ms:.2f} ms per {nq}-query batch ({nq / ms * 1e3:.0f} queries/s); per batch: passes {(p1 - p0) / reps:.1f}, "
          f"queries through the 3xTF32 pass {(e1 - e0) / reps:.0f}, exact re-runs {(r1 - r0) / reps:.0f}; spot check {'ok' if ok else 'MISMATCH'}",
          flush=True)
    c.close(); s.close()
