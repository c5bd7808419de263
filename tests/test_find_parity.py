"""GPU parity of StoresFind against the oracle on seeded random data: every shape class the
kernels distinguish (TMA path, generic path, ragged last chunk), both arithmetic paths, batched
queries, fused and spilled top-k, ties and edge sizes.  Scores must be bit-identical."""
import numpy as np
import pytest

from conftest import check_find_against_scores, make_keys, normalize_rows

pytestmark = pytest.mark.gpu
f32 = np.float32


def load(tfp, keys):
    s = tfp.Store()
    s.StoresSet(keys, [b""] * len(keys))
    return s


@pytest.mark.parametrize("n,d", [(1, 32), (127, 32), (128, 64), (129, 36), (1000, 100), (5000, 384),
                                 (20000, 768), (3000, 1024), (700, 1536), (300, 4096)])
@pytest.mark.parametrize("normalized", [True, False])
def test_tma_path_bit_exact(tfp, oracle, n, d, normalized):
    K = make_keys(n, d, seed=n + d, normalized=normalized)
    q = make_keys(1, d, seed=n + d + 1, normalized=normalized)[0]
    s = load(tfp, K)
    scores = oracle.scores_block(K, q, normalized)
    for k in (1, 10, 100):
        ids, sims = s.find_ids(q, k)
        check_find_against_scores(ids, sims, scores, k)
    s.close()


@pytest.mark.parametrize("n,d", [(10, 3), (1000, 5), (513, 31), (2000, 33), (400, 130)])
@pytest.mark.parametrize("normalized", [True, False])
def test_generic_path_bit_exact(tfp, oracle, n, d, normalized):
    K = make_keys(n, d, seed=n + d, normalized=normalized)
    q = make_keys(1, d, seed=n + d + 1, normalized=normalized)[0]
    s = load(tfp, K)
    scores = oracle.scores_block(K, q, normalized)
    for k in (1, 7, 64):
        ids, sims = s.find_ids(q, k)
        check_find_against_scores(ids, sims, scores, k)
    s.close()


def test_path_dispatch_matches_reference(tfp, oracle):  # store.go:496-510
    K = make_keys(2000, 64, seed=1, normalized=True)
    s = load(tfp, K)
    o = oracle.OracleStore(); o.StoresSet(K, [b""] * len(K))
    q_unit = K[5]
    q_long = (K[5] * f32(3)).astype(f32)
    for q in (q_unit, q_long):
        a = s.StoresFind(q, 5); b = o.StoresFind(q, 5)
        assert s.last_path == o.last_path
        assert np.array_equal(a[2], b[2])
    s.StoresSet([(K[0] * f32(2)).astype(f32)], [b"x"]); o.StoresSet([(K[0] * f32(2)).astype(f32)], [b"x"])
    a = s.StoresFind(q_unit, 5); b = o.StoresFind(q_unit, 5)
    assert s.last_path == o.last_path == "fallback"
    assert np.array_equal(a[2], b[2])
    s.close()


@pytest.mark.parametrize("k", [1, 10, 100, 1000, 1024])
def test_fused_topk_sizes(tfp, oracle, k):
    K = make_keys(40000, 64, seed=k, normalized=True)
    q = make_keys(1, 64, seed=k + 1, normalized=True)[0]
    s = load(tfp, K)
    ids, sims = s.find_ids(q, k)
    check_find_against_scores(ids, sims, oracle.scores_block(K, q, True), k)
    s.close()


@pytest.mark.parametrize("k", [1025, 5000, 30000])
def test_spilled_topk_sizes(tfp, oracle, k):
    """k beyond TFP_FUSED_TOPK_MAX: score spill + radix sort, still on the GPU, same answer."""
    K = make_keys(30000, 64, seed=k, normalized=False)
    q = make_keys(1, 64, seed=k + 1, normalized=False)[0]
    s = load(tfp, K)
    ids, sims = s.find_ids(q, k)
    check_find_against_scores(ids, sims, oracle.scores_block(K, q, False), k)
    s.close()


def test_ascending_scores_worst_case_for_the_filter(tfp, oracle):
    """Rows arrive in increasing similarity: every row beats the running threshold."""
    d = 64
    q = normalize_rows(np.ones((1, d), f32))[0]
    t = np.linspace(-0.9, 0.9, 30000).astype(f32)
    e = np.zeros(d, f32); e[0] = 1; e[1] = -1; e = normalize_rows(e[None])[0]
    K = normalize_rows(t[:, None] * q[None, :] + np.sqrt(1 - t * t)[:, None].astype(f32) * e[None, :])
    s = load(tfp, K)
    scores = oracle.scores_block(K, q, True)
    for k in (10, 500):
        ids, sims = s.find_ids(q, k)
        check_find_against_scores(ids, sims, scores, k)
    s.close()


def test_heavy_ties(tfp, oracle):
    rng = np.random.default_rng(3)
    K = normalize_rows(rng.integers(-1, 2, size=(20000, 32)).astype(f32) + f32(0.001))
    K = np.unique(K, axis=0)
    rng.shuffle(K)
    q = K[0]
    s = load(tfp, K)
    scores = oracle.scores_block(K, q, True)
    for k in (5, 50, 300):
        ids, sims = s.find_ids(q, k)
        check_find_against_scores(ids, sims, scores, k)
        assert list(ids) == sorted(ids, key=lambda i: (-float(scores[i]), i))  # ties: older row first
    s.close()


def test_duplicate_rows_all_returned(tfp):
    s = tfp.Store()
    k = normalize_rows(np.ones((1, 32), f32))
    s.StoresSet(np.repeat(k, 6, axis=0), [b"%d" % i for i in range(6)])   # one batch: all six are stored
    ks, vs, sims = s.StoresFind(k[0], 10)
    assert len(ks) == 6 and sorted(vs) == [b"%d" % i for i in range(6)]
    s.close()


@pytest.mark.parametrize("nq", [1, 2, 3, 4, 7, 8, 9, 21])
@pytest.mark.parametrize("normalized", [True, False])
def test_batched_queries_equal_single_queries(tfp, oracle, nq, normalized):
    K = make_keys(9000, 128, seed=nq, normalized=normalized)
    Q = make_keys(nq, 128, seed=nq + 100, normalized=normalized)
    s = load(tfp, K)
    res = s.find_batch(Q, 10)
    for i in range(nq):
        check_find_against_scores(res[i][0], res[i][1], oracle.scores_block(K, Q[i], normalized), 10)
    s.close()


def test_mixed_paths_in_one_batch(tfp, oracle):
    K = make_keys(5000, 64, seed=11, normalized=True)
    Q = make_keys(6, 64, seed=12, normalized=True)
    Q[1] *= f32(2); Q[4] *= f32(0.5)   # these two take the fallback path
    s = load(tfp, K)
    res = s.find_batch(Q, 8)
    for i in range(6):
        norm = oracle.is_normalized(Q[i])
        check_find_against_scores(res[i][0], res[i][1], oracle.scores_block(K, Q[i], norm), 8)
    s.close()


def test_find_returns_key_rows(tfp):
    K = make_keys(3000, 96, seed=5, normalized=True)
    s = tfp.Store()
    s.StoresSet(K, [b"%d" % i for i in range(len(K))])
    ks, vs, sims = s.StoresFind(K[17], 5)
    assert vs[0] == b"17" and np.array_equal(ks[0], K[17])
    for k, v in zip(ks, vs):
        assert np.array_equal(k, K[int(v)])
    s.close()


def test_find_after_growth_and_delete(tfp, oracle):
    """Capacity doubling and delete compaction keep ids and scores right."""
    K = make_keys(6000, 64, seed=21, normalized=True)
    s = tfp.Store(capacity_rows=1024)
    o = oracle.OracleStore()
    for lo in range(0, 6000, 1500):
        vals = [b"%d" % i for i in range(lo, lo + 1500)]
        s.StoresSet(K[lo:lo + 1500], vals); o.StoresSet(K[lo:lo + 1500], vals)
    rng = np.random.default_rng(0)
    gone = rng.choice(6000, size=2500, replace=False)
    s.StoresDelete(K[gone]); o.StoresDelete(K[gone])
    assert len(s) == len(o) == 3500
    for qi in (1, 2, 3):
        q = K[int(gone[qi])]
        a = s.StoresFind(q, 20); b = o.StoresFind(q, 20)
        assert np.array_equal(a[2], b[2]) and a[1] == b[1]
    s.close()


def test_context_async_pipeline(tfp, oracle):
    K = make_keys(20000, 256, seed=31, normalized=True)
    Q = make_keys(16, 256, seed=32, normalized=True)
    s = load(tfp, K)
    c = s.find_context()
    c.upload(Q)
    for i in range(16):
        c.launch(i, 1, 10)
        ids, sims = c.download()[0]
        check_find_against_scores(ids, sims, oracle.scores_block(K, Q[i], True), 10)
    assert c.launch_count >= 16
    c.close(); s.close()


def test_c1_shape_10k_by_384(tfp, oracle):
    """BASELINE.json configs[0]: 10k keys x 384-d, top-10, full oracle comparison."""
    K = make_keys(10000, 384, seed=0x5EED0001 & 0xFFFF, normalized=True)
    Q = make_keys(8, 384, seed=77, normalized=True)
    s = load(tfp, K)
    o_rows = [oracle.scan_block(K, q, 10, True) for q in Q]
    for q, (rows, osims) in zip(Q, o_rows):
        ids, sims = s.find_ids(q, 10)
        assert np.array_equal(sims, osims) and np.array_equal(ids, rows)
    s.close()


def test_c2_shape_full_store_oracle_scan(tfp, oracle):
    """BASELINE.json configs[1] at full size against the oracle's scan of the WHOLE store ("full-N oracle
    is run for C1/C2"): 1M x 768 normalised, top-10, three queries; similarities bit for bit, ids position by
    position (exactly equal scores exempt)."""
    import torch
    n, d = 1_000_000, 768
    g = torch.Generator(device="cuda").manual_seed(0x5EED0002)
    K = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float32)
    K /= K.norm(dim=1, keepdim=True)
    s = tfp.Store(capacity_rows=n)
    s.set_device(K.data_ptr(), n, d)
    Kh = K.cpu().numpy()
    Q = make_keys(2, d, seed=99, normalized=True)
    for q in (Q[0], Q[1], Kh[777_777]):
        ids, sims = s.find_ids(q, 10)
        check_find_against_scores(ids, sims, oracle.scores_block(Kh, q, True), 10)
    res = s.find_batch(np.stack([Q[0], Q[1], Kh[5], Kh[6]]), 10)          # the tensor-core path over the same store
    for (ids, sims), q in zip(res, (Q[0], Q[1])):
        check_find_against_scores(ids, sims, oracle.scores_block(Kh, q, True), 10)
    s.close()


def test_c2_shape_subsample_1m_by_768(tfp, oracle):
    """BASELINE.json configs[1] at full size: 1M x 768 normalised, top-10.  The oracle scores the rows
    the GPU returned plus a 20k-row slice; size-independent properties cover the rest."""
    import torch
    n, d = 1_000_000, 768
    g = torch.Generator(device="cuda").manual_seed(0x5EED0002)
    K = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float32)
    K /= K.norm(dim=1, keepdim=True)
    s = tfp.Store(capacity_rows=n)
    ids0, new = s.set_device(K.data_ptr(), n, d)
    assert new.all() and np.array_equal(ids0, np.arange(n))
    assert s.keys_are_normalized
    q = K[123456].cpu().numpy()
    ids, sims = s.find_ids(q, 10)
    assert ids[0] == 123456 and list(sims) == sorted(sims, reverse=True)        # self-query ranks first
    Kh = K[torch.as_tensor(ids, device="cuda")].cpu().numpy()
    assert np.array_equal(oracle.scores_block(Kh, q, True), sims)                # returned scores are exact
    lo = 400_000
    sl = K[lo:lo + 20000].cpu().numpy()
    sc = oracle.scores_block(sl, q, True)
    assert (sc <= sims[-1]).all() or set(np.nonzero(sc > sims[-1])[0] + lo) <= set(ids.tolist())
    # idempotence and k-monotonicity
    ids2, sims2 = s.find_ids(q, 10)
    assert np.array_equal(ids, ids2) and np.array_equal(sims, sims2)
    ids5, sims5 = s.find_ids(q, 5)
    assert np.array_equal(ids5, ids[:5]) and np.array_equal(sims5, sims[:5])
    s.close()
