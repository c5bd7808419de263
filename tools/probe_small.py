"""Pipelined throughput of the exact scan on small shards (what one rank of an N-GPU C2 run holds):
`lanes` find contexts alternating, CUDA events around 400 finds.  Run once per value of TFP_SCAN_STAGES."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import localai_tfp_b200 as tfp

out = []
for n, d in ((125_000, 768), (250_000, 768), (500_000, 768), (1_000_000, 768)):
    g = torch.Generator(device="cuda").manual_seed(1)
    K = torch.randn((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
    s = tfp.Store(capacity_rows=n); s.set_device(K.data_ptr(), n, d)
    Q = K[:8].cpu().numpy(); del K
    for lanes in (1, 2, 4):
        ctxs = [s.find_context() for _ in range(lanes)]
        for c in ctxs:
            c.upload(Q); c.sync()
        streams = [torch.cuda.ExternalStream(c.stream) for c in ctxs]
        for i in range(4 * lanes):
            ctxs[i % lanes].launch(i % 8, 1, 10)
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(streams[0])
            for st in streams[1:]:
                st.wait_event(e0)
            for i in range(400):
                ctxs[i % lanes].launch(i % 8, 1, 10)
            for st in streams[1:]:
                eb = torch.cuda.Event(); eb.record(st); streams[0].wait_event(eb)
            e1.record(streams[0])
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 400 * 1e3)
        out.append(f"{n}x{d} lanes={lanes}: {best:.1f} us/find ({n * d * 4 / best / 1e3:.0f} GB/s)")
        for c in ctxs:
            c.close()
    s.close()
print("TFP_SCAN_STAGES=%s | " % os.environ.get("TFP_SCAN_STAGES", "-") + " | ".join(out))
