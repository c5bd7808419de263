import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (oracle/store_oracle.c) -- the checker, never the product."""
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def tfp():
    """The product package with its CUDA library built in-tree."""
    import localai_tfp_b200
    from localai_tfp_b200 import build as b
    b.build()
    localai_tfp_b200.load()
    return localai_tfp_b200


def two_devices():
    """Device list for the in-process multi-GPU store tests: two GPUs when the box has them, else two shards on GPU 0."""
    import torch
    return [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]


def normalize_rows(a):
    """Row-normalise in fp32 the way the reference's test helper does
    (tests/integration/stores_test.go:26-37): fp64 norm of fp32-rounded squares, fp32 divide."""
    a = np.array(a, dtype=np.float32, copy=True)
    sq = (a * a).astype(np.float32).astype(np.float64)
    norm = np.sqrt(sq.sum(axis=1))
    return (a / norm.astype(np.float32)[:, None]).astype(np.float32)


def make_keys(n, d, seed, normalized=True, scale_lo=0.5, scale_hi=2.0):
    rng = np.random.default_rng(seed)
    a = rng.standard_normal((n, d), dtype=np.float32)
    if normalized:
        return normalize_rows(a)
    s = rng.uniform(scale_lo, scale_hi, size=(n, 1)).astype(np.float32)
    return (normalize_rows(a) * s).astype(np.float32)


def check_find_against_scores(ids, sims, all_scores, k):
    """Parity check that is exact up to ties: `sims` must be bit-identical to the k best oracle
    scores in descending order, every returned row must really carry its score, rows are distinct."""
    n = len(all_scores)
    m = min(k, n)
    assert len(ids) == m and len(sims) == m
    order = np.argsort(-all_scores.astype(np.float64), kind="stable")
    want = all_scores[order[:m]]
    assert np.array_equal(sims, want), f"scores differ:\n{sims}\n{want}"
    assert len(set(int(i) for i in ids)) == m
    got = all_scores[np.asarray(ids, dtype=np.int64)]
    assert np.array_equal(got, sims), "a returned id does not carry the score reported for it"
    # where the oracle's ranking has no ties, ids must match position by position
    for i in range(m):
        s = want[i]
        tied = (i > 0 and want[i - 1] == s) or (i + 1 < n and all_scores[order[i + 1]] == s)
        if not tied:
            assert int(ids[i]) == int(order[i])
