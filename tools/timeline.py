import os, sys
os.environ["TFP_SCAN_TIMES"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import localai_tfp_b200 as tfp
n, d = 1_000_000, 768
g = torch.Generator(device="cuda").manual_seed(1)
K = torch.randn((n, d), generator=g, device="cuda"); K /= K.norm(dim=1, keepdim=True)
s = tfp.Store(capacity_rows=n); s.set_device(K.data_ptr(), n, d)
Q = K[:8].cpu().numpy(); del K
c = s.find_context(); c.upload(Q)
for i in range(6):
    c.launch(i, 1, 10); c.sync()
for k in (100,):
    c.launch(0, 1, k); c.sync()
