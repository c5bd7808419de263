"""CPU oracle for the local-store path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this package.  The product
(localai_tfp_b200/) never does.
"""
