"""tfp_scan_topk_noalloc: the find for hosts that must not allocate device memory while serving --
a runtime that has given all of HBM to its own arena, or a latency-critical loop that cannot take
cudaMalloc's device-wide synchronisation.  Its scratch is a fixed array inside the library.

Checked through the C ABI against the ordinary entry point (tfp_store_find) on the same keys AND
against the CPU oracle's scores of those keys, bit for bit like the other GPU tests, on two calls in a
row: the second call starts from whatever the first left in the shared scratch array.  One case uses
every CTA the entry point may launch with the largest top-k it accepts, i.e. the whole fixed array;
one runs on a second GPU when the box has one (the array exists once per device).  This file sorts
last so a `-x` run has already run every other test before it; it also passes alone in a fresh process."""
import ctypes as C

import numpy as np
import pytest

from conftest import check_find_against_scores, make_keys

pytestmark = pytest.mark.gpu


def noalloc_find(L, K_d, ids_d, q_d, topk):
    import torch
    n, d = K_d.shape
    dev = K_d.device
    sims = torch.full((topk,), float("nan"), dtype=torch.float32, device=dev)
    ids = torch.full((topk,), -7, dtype=torch.int64, device=dev)
    cnt = torch.full((1,), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    rc = L.tfp_scan_topk_noalloc(K_d.data_ptr(), n, d, ids_d.data_ptr(), q_d.data_ptr(), topk,
                                 sims.data_ptr(), ids.data_ptr(), cnt.data_ptr(), None)
    assert rc == 0, L.tfp_last_error().decode()
    torch.cuda.synchronize()
    m = int(cnt.item())
    return ids.cpu().numpy()[:m], sims.cpu().numpy()[:m]


def run_twice(tfp, oracle, n, d, topk, device=0):
    import torch
    L = tfp.load()
    K = make_keys(n, d, seed=n + d, normalized=True)
    Q = make_keys(2, d, seed=n + d + 1, normalized=True)
    store = tfp.Store(device=device)
    store.StoresSet(K, [b""] * n)                       # ids == row numbers
    with torch.cuda.device(device):                      # the entry point works on the CURRENT device
        K_d = torch.from_numpy(K).cuda()
        ids_d = torch.arange(n, dtype=torch.int64, device="cuda")
        for call in range(2):                            # two calls in a row share the scratch array
            q = Q[call]
            want_ids, want_sims = store.find_ids(q, topk)    # the existing entry point, tfp_store_find
            q_d = torch.from_numpy(q).cuda()
            free_before = torch.cuda.mem_get_info()[0]
            got_ids, got_sims = noalloc_find(L, K_d, ids_d, q_d, topk)
            if call == 1:                                # by now every lazy one-time set-up is done
                assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the amount of free device memory"
            assert np.array_equal(got_sims, want_sims), (got_sims, want_sims)
            assert np.array_equal(got_ids, want_ids)
            check_find_against_scores(got_ids, got_sims, oracle.scores_block(K, q, True), topk)   # and the oracle itself
    store.close()


# (20000, 64, 64): 157 tiles, so every CTA a launch may use publishes a list, with the largest top-k: the whole array
@pytest.mark.parametrize("n,d,topk", [(20000, 768, 10), (3000, 100, 64), (90, 48, 7), (500, 33, 5), (20000, 64, 64)])
def test_noalloc_matches_store_find_twice_in_a_row(tfp, oracle, n, d, topk):
    run_twice(tfp, oracle, n, d, topk)


def test_noalloc_on_a_second_gpu(tfp, oracle):
    """Every GPU that loads the library has its own copy of the fixed array, found on that device's first call."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    run_twice(tfp, oracle, 6000, 128, 20, device=1)
    run_twice(tfp, oracle, 6000, 128, 20, device=0)      # and device 0's copy is still its own


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch(tfp):
    import torch
    L = tfp.load()
    K_d = torch.zeros((256, 64), dtype=torch.float32, device="cuda")
    ids_d = torch.arange(256, dtype=torch.int64, device="cuda")
    q_d = torch.zeros(64, dtype=torch.float32, device="cuda")
    out_s = torch.zeros(128, dtype=torch.float32, device="cuda")
    out_i = torch.zeros(128, dtype=torch.int64, device="cuda")
    out_n = torch.zeros(1, dtype=torch.int32, device="cuda")

    def call(n, d, k):
        return L.tfp_scan_topk_noalloc(K_d.data_ptr(), n, d, ids_d.data_ptr(), q_d.data_ptr(), k,
                                       out_s.data_ptr(), out_i.data_ptr(), out_n.data_ptr(), None)

    assert call(256, 64, 65) > 0 and b"TFP_NOALLOC_TOPK_MAX" in L.tfp_last_error()
    assert call(256, 8193, 10) > 0 and b"TFP_NOALLOC_DIM_MAX" in L.tfp_last_error()
    assert call(256, 64, 0) > 0 and b"must be >= 1" in L.tfp_last_error()
    assert call(0, 64, 10) == 0                           # an empty key matrix is a valid, empty answer
    torch.cuda.synchronize()
    assert int(out_n.item()) == 0
