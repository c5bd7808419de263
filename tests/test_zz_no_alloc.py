"""tfp_scan_topk_noalloc: the find for hosts that must not allocate device memory while serving --
a runtime that has given all of HBM to its own arena, or a latency-critical loop that cannot take
cudaMalloc's device-wide synchronisation.  Its scratch is a fixed array inside the library.

Checked through the C ABI against the ordinary entry point (tfp_store_find) on the same keys,
bit for bit like the other GPU tests, on two calls in a row: the second call starts from whatever
the first left in the shared scratch array.  This file sorts last so a `-x` run has already run
every other test before it; it also passes alone in a fresh process."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_keys

pytestmark = pytest.mark.gpu


def noalloc_find(L, K_d, ids_d, q_d, topk):
    import torch
    n, d = K_d.shape
    sims = torch.full((topk,), float("nan"), dtype=torch.float32, device="cuda")
    ids = torch.full((topk,), -7, dtype=torch.int64, device="cuda")
    cnt = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = L.tfp_scan_topk_noalloc(K_d.data_ptr(), n, d, ids_d.data_ptr(), q_d.data_ptr(), topk,
                                 sims.data_ptr(), ids.data_ptr(), cnt.data_ptr(), None)
    assert rc == 0, L.tfp_last_error().decode()
    torch.cuda.synchronize()
    m = int(cnt.item())
    return ids.cpu().numpy()[:m], sims.cpu().numpy()[:m]


@pytest.mark.parametrize("n,d,topk", [(20000, 768, 10), (3000, 100, 64), (90, 48, 7), (500, 33, 5)])
def test_noalloc_matches_store_find_twice_in_a_row(tfp, n, d, topk):
    import torch
    L = tfp.load()
    K = make_keys(n, d, seed=n + d, normalized=True)
    Q = make_keys(2, d, seed=n + d + 1, normalized=True)
    store = tfp.Store()
    store.StoresSet(K, [b""] * n)                       # ids == row numbers
    K_d = torch.from_numpy(K).cuda()
    ids_d = torch.arange(n, dtype=torch.int64, device="cuda")
    for call in range(2):                                # two calls in a row share the scratch array
        q = Q[call]
        want_ids, want_sims = store.find_ids(q, topk)    # the existing entry point, tfp_store_find
        q_d = torch.from_numpy(q).cuda()
        free_before = torch.cuda.mem_get_info()[0]
        got_ids, got_sims = noalloc_find(L, K_d, ids_d, q_d, topk)
        if call == 1:                                    # by now every lazy one-time set-up is done
            assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the amount of free device memory"
        assert np.array_equal(got_sims, want_sims), (got_sims, want_sims)
        assert np.array_equal(got_ids, want_ids)
    store.close()


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
