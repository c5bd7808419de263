"""tests/golden/reference_test_vectors.json against the oracle (CPU) and the CUDA store (GPU)."""
import json
import os

import numpy as np
import pytest

from conftest import normalize_rows

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = json.load(open(os.path.join(HERE, "golden", "reference_test_vectors.json")))["cases"]
f32 = np.float32


def run_case(store, case):
    keys = np.array(case["keys"], f32)
    if case["normalize_keys"]:
        keys = normalize_rows(keys)
    vals = [v.encode() for v in case["values"]]
    store.StoresSet(keys, vals)
    q = keys[case["query_is_key"]] if "query_is_key" in case else np.array(case["query"], f32)
    ks, vs, sims = store.StoresFind(q, case["topk"])
    assert store.last_path == case["path"]
    e = case["expect"]
    if "keys" in e:
        for got, want in zip(ks, e["keys"]):
            assert np.array_equal(got, np.array(want, f32))
    if "first_value" in e:
        assert vs[0] == e["first_value"].encode()
    if "key_order" in e:
        for got, idx in zip(ks, e["key_order"]):
            assert np.array_equal(got, keys[idx])
    if "first_sim_approx" in e:
        assert abs(float(sims[0]) - e["first_sim_approx"][0]) <= e["first_sim_approx"][1]
    if "sims_exact" in e:
        assert np.array_equal(sims, np.array(e["sims_exact"], f32))
    if "sims_approx" in e:
        for s, (want, tol) in zip(sims, e["sims_approx"]):
            assert abs(float(s) - want) <= tol
    if "n_results" in e:
        assert len(ks) == e["n_results"]
    if "sims_in" in e:
        assert all(e["sims_in"][0] <= float(s) <= e["sims_in"][1] for s in sims)
    assert np.array_equal(sims, np.array(case["derived_sims"], f32)[:len(sims)])


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_oracle_on_reference_vectors(oracle, case):
    run_case(oracle.OracleStore(), case)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_cuda_store_on_reference_vectors(tfp, case):
    s = tfp.Store()
    run_case(s, case)
    s.close()
