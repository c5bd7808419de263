"""One C3-like batch on the tensor-core path (ncu target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import probe_mma
probe_mma.run(1_000_000, 1024, 256, 100, iters=2, label="ncu-target")
