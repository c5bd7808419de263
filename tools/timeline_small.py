"""Per-CTA timeline of the exact scan on an 8-GPU-sized shard of C2 (125k x 768) and on C1 (10k x 384)."""
import os, sys, time
os.environ["TFP_SCAN_TIMES"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import localai_tfp_b200 as tfp
for n, d in ((125_000, 768), (10_000, 384)):
    g = torch.Generator(device="cuda").manual_seed(1)
    K = torch.randn((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
    s = tfp.Store(capacity_rows=n); s.set_device(K.data_ptr(), n, d)
    Q = K[:8].cpu().numpy(); del K
    c = s.find_context(); c.upload(Q)
    print(f"== n={n} d={d}", file=sys.stderr, flush=True)
    for i in range(6):
        c.launch(i, 1, 10); c.sync()
    t0 = time.perf_counter()
    for i in range(200):
        c.launch(i % 8, 1, 10)
    c.sync()
    print(f"back to back: {(time.perf_counter() - t0) / 200 * 1e6:.1f} us per scan", file=sys.stderr, flush=True)
    c.close(); s.close()
