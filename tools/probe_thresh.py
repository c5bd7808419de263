import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import probe_mma
for nq in (9, 12, 16, 24, 31):
    probe_mma.run(1_000_000, 768, nq, 10, iters=5, label=f"thr q{nq}")
