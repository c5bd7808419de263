"""Small end-to-end run for compute-sanitizer (memcheck): every kernel family once, tiny sizes."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TFP_MMA_MIN_ROWS", "4096")
import localai_tfp_b200 as tfp

rng = np.random.default_rng(0)
def unit(n, d):
    a = rng.standard_normal((n, d)).astype(np.float32)
    return (a / np.linalg.norm(a, axis=1, keepdims=True)).astype(np.float32)

s = tfp.Store()
K = unit(20000, 96)
s.StoresSet(K, [b"%d" % i for i in range(len(K))])          # row_stats, scatter
print("find", s.find_ids(K[5], 10)[0][:3])                   # TMA scan, fused top-k, merge
print("find k=1500", len(s.find_ids(K[5], 1500)[0]))         # spill + radix sort + decode
print("batch", len(s.find_batch(K[:7], 10)))                 # NQ = 4, 2, 1 groups
s.StoresSet((K[:3] * 2).astype(np.float32), [b"x"] * 3)      # fallback path from here on
print("fallback", s.StoresFind(K[9], 5)[1][:2])              # fp64 path + gather
print("mma", len(s.find_batch(unit(40, 96), 10)))            # tensor-core path (40 queries) + finalize
s.StoresDelete(K[100:4000])                                  # verify_pairs, move_rows
print("get", len(s.StoresGet(K[4000:4010])[0]))
s.close()
g = tfp.Store()
G = unit(3000, 37)                                           # generic kernel (dim % 4 != 0)
g.StoresSet(G, [b""] * len(G))
print("generic", g.find_ids(G[1], 7)[0][:3])
g.close()
print("sanitize target done")
