"""ONE logical store over several GPUs inside one process (tfp_store_open_devices, multi_store.inl): routing on
the CPU; on the GPU the same answers as the oracle and as the one-GPU store, through the same entry points."""
import threading

import numpy as np
import pytest

from conftest import check_find_against_scores, make_keys, normalize_rows, two_devices

f32 = np.float32


# ----------------------------------------------------------------------------- CPU: routing
def route(tfp, keys, shards):
    import ctypes as C
    L = tfp.load()
    k = np.ascontiguousarray(keys, dtype=f32)
    out = np.empty(len(k), np.int32)
    rc = L.tfp_route_rows(k.ctypes.data_as(C.POINTER(C.c_float)), len(k), k.shape[1], shards,
                          out.ctypes.data_as(C.POINTER(C.c_int)))
    assert rc == 0
    return out


def test_routing_is_deterministic_and_balanced(tfp):
    K = make_keys(40000, 96, seed=1)
    for shards in (1, 2, 3, 8):
        a, b = route(tfp, K, shards), route(tfp, K, shards)
        assert np.array_equal(a, b) and a.min() >= 0 and a.max() < shards
        counts = np.bincount(a, minlength=shards)
        assert counts.min() > 0.9 * len(K) / shards and counts.max() < 1.1 * len(K) / shards


def test_equal_keys_meet_in_one_shard(tfp):
    """Key equality is Go's cmp.Compare (store.go:48-52): -0 == +0 and NaN == NaN must route alike."""
    rng = np.random.default_rng(2)
    K = rng.standard_normal((500, 40)).astype(f32)
    K[:, 0] = 0.0
    K[::3, 5] = np.nan
    A = K.copy()
    A[:, 0] = -0.0
    A[::3, 5] = np.float32(np.nan).view(np.uint32).__or__(np.uint32(0x1234)).view(f32)   # another NaN payload
    assert np.array_equal(route(tfp, K, 8), route(tfp, A, 8))
    for d in (1, 2, 15, 16, 17, 1536):                                                   # every sampling regime
        X = rng.standard_normal((64, d)).astype(f32)
        assert np.array_equal(route(tfp, X, 4), route(tfp, X.copy(), 4))


def test_bad_routing_arguments_are_errors(tfp):
    import ctypes as C
    L = tfp.load()
    out = np.empty(4, np.int32)
    assert L.tfp_route_rows(None, 4, 8, 2, out.ctypes.data_as(C.POINTER(C.c_int))) == 2
    k = np.zeros((4, 8), f32)
    assert L.tfp_route_rows(k.ctypes.data_as(C.POINTER(C.c_float)), 4, 8, 0, out.ctypes.data_as(C.POINTER(C.c_int))) == 2


# ----------------------------------------------------------------------------- GPU
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("normalized", [True, False])
def test_multi_find_bit_exact(tfp, oracle, normalized):
    K = make_keys(30000, 128, seed=5, normalized=normalized)
    q = make_keys(1, 128, seed=6, normalized=normalized)[0]
    s = tfp.Store(devices=two_devices())
    assert s.shard_count == 2
    ids0, new = s.StoresSet(K, [b"%d" % i for i in range(len(K))])
    assert new.all()
    rows_of = {int(i): j for j, i in enumerate(ids0)}                  # ids carry the shard in their high bits
    st = [s.shard_stats(g) for g in range(2)]
    assert st[0].rows + st[1].rows == len(K) == len(s) and min(st[0].rows, st[1].rows) > 0.4 * len(K)
    scores = oracle.scores_block(K, q, normalized)
    for k in (1, 10, 100, 1000):                                       # device merge (2 x k <= 16384)
        ids, sims = s.find_ids(q, k)
        check_find_against_scores([rows_of[int(i)] for i in ids], sims, scores, k)
    for k in (9000, 30000, 50000):                                     # beyond the gather block: merged on the host
        ids, sims = s.find_ids(q, k)
        check_find_against_scores([rows_of[int(i)] for i in ids], sims, scores, k)
    ks, vs, sims = s.StoresFind(q, 7)                                  # the winners' key rows and values
    assert s.last_path == ("normalized" if normalized else "fallback")
    for krow, v, sim in zip(ks, vs, sims):
        assert np.array_equal(krow, K[int(v)]) and scores[int(v)] == sim
    s.close()


@gpu
def test_multi_batch_matches_single_store(tfp, oracle):
    """A batch large enough for the tensor-core path on every shard (>= 65536 rows each), and small exact batches."""
    K = make_keys(150000, 64, seed=9, normalized=True)
    Q = make_keys(12, 64, seed=10, normalized=True)
    m = tfp.Store(devices=two_devices())
    ids0, _ = m.StoresSet(K, [b""] * len(K))
    rows_of = {int(i): j for j, i in enumerate(ids0)}
    for nq, k in ((12, 10), (2, 5), (5, 100)):
        res, rows = m.find_batch(Q[:nq], k, with_keys=True)
        for i in range(nq):
            ids, sims = res[i]
            check_find_against_scores([rows_of[int(x)] for x in ids], sims, oracle.scores_block(K, Q[i], True), k)
            assert np.array_equal(rows[i, :len(ids)], K[[rows_of[int(x)] for x in ids]])
    m.close()


@gpu
def test_multi_set_device_routes_rows_that_are_already_in_hbm(tfp, oracle):
    import torch
    n, d = 40000, 96
    K = make_keys(n, d, seed=15, normalized=True)
    Kd = torch.from_numpy(K).cuda()
    m = tfp.Store(devices=two_devices())
    ids, new = m.set_device(Kd.data_ptr(), n, d)
    assert new.all() and len(set(ids.tolist())) == n and len(m) == n and m.keys_are_normalized
    assert min(m.shard_stats(0).rows, m.shard_stats(1).rows) > 0.4 * n
    rows_of = {int(i): j for j, i in enumerate(ids)}
    q = make_keys(1, d, seed=16, normalized=True)[0]
    fids, sims = m.find_ids(q, 10)
    check_find_against_scores([rows_of[int(i)] for i in fids], sims, oracle.scores_block(K, q, True), 10)
    ids2, new2 = m.set_device(Kd.data_ptr(), 100, d)                     # the same keys again: overwrites, same ids
    assert not new2.any() and np.array_equal(ids2, ids[:100]) and len(m) == n
    m.close()


@gpu
def test_multi_concurrent_finds_use_the_lanes(tfp, oracle):
    K = make_keys(20000, 96, seed=11, normalized=True)
    Q = make_keys(16, 96, seed=12, normalized=True)
    m = tfp.Store(devices=two_devices())
    ids0, _ = m.StoresSet(K, [b""] * len(K))
    rows_of = {int(i): j for j, i in enumerate(ids0)}
    want = [oracle.scores_block(K, q, True) for q in Q]
    errors = []

    def worker(t):
        try:
            for rep in range(25):
                i = (t + rep) % len(Q)
                ids, sims = m.find_ids(Q[i], 10)
                check_find_against_scores([rows_of[int(x)] for x in ids], sims, want[i], 10)
        except Exception as e:       # noqa: BLE001 - reported below
            errors.append(repr(e))

    ts = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors[:3]
    m.close()


@gpu
def test_multi_errors_and_handle_error_text(tfp):
    m = tfp.Store(devices=two_devices())
    with pytest.raises(tfp.StoreError) as e:
        m.StoresFind([1, 0, 0], 1)                                      # never-written store: store.go:480-482
    assert str(e.value) == "Try to find key with length 3 when existing length is -1"
    assert m.last_error() == str(e.value)                               # kept per handle, whichever thread asks
    seen = []
    t = threading.Thread(target=lambda: seen.append(m.last_error()))
    t.start(); t.join()
    assert seen == [str(e.value)]
    m.StoresSet([[1, 0, 0]], [b"a"])
    for call, text in ((lambda: m.StoresSet([[1, 0]], [b"x"]), "Try to add key with length 2 when existing length is 3"),
                       (lambda: m.StoresFind([1, 0, 0], 0), "opts.TopK = 0, must be >= 1"),
                       (lambda: m.StoresDelete([]), "no keys to delete"),
                       (lambda: m.StoresSet([[1, 0, 0]], []), "len(keys) = 1, len(values) = 0")):
        with pytest.raises(tfp.StoreError) as e:
            call()
        assert str(e.value) == text
    with pytest.raises(tfp.StoreError):
        m.find_context()                                                # the multi-GPU store runs its own lanes
    m.close()


@gpu
def test_multi_snapshot_and_restore(tfp, tmp_path):
    K = make_keys(5000, 48, seed=13, normalized=True)
    vals = [b"v%d" % i for i in range(len(K))]
    a = tfp.Store(devices=two_devices())
    a.StoresSet(K, vals)
    a.StoresDelete(K[10:500])
    path = str(tmp_path / "multi.snap")
    a.save(path)
    a.save(path)                                                        # a second save replaces the first pair
    import glob
    assert len(glob.glob(path + ".values.*")) == 1
    b = tfp.Store(devices=two_devices())
    b.load(path)
    assert len(b) == len(a) == 4510 and b.key_len == 48 and b.keys_are_normalized
    for q in (K[0], K[4999], K[100]):
        ra, rb = a.StoresFind(q, 10), b.StoresFind(q, 10)
        assert np.array_equal(ra[2], rb[2]) and ra[1] == rb[1]
    one = tfp.Store()
    with pytest.raises(tfp.StoreError):
        one.load(path)                                                  # a two-shard snapshot is not a one-GPU snapshot
    one.close(); a.close(); b.close()


@gpu
def test_snapshot_is_crash_consistent(tfp, tmp_path):
    """A save that dies before its commit point (the rename of the keys file) leaves the previous pair intact; a
    values file that does not belong to the keys file is refused."""
    import glob
    import os
    K = make_keys(300, 32, seed=14, normalized=True)
    a = tfp.Store()
    a.StoresSet(K, [b"old%d" % i for i in range(len(K))])
    path = str(tmp_path / "s.snap")
    a.save(path)
    old_values = glob.glob(path + ".values.*")
    # simulate the kill: new values file written, keys file only as far as its .tmp
    with open(path + ".values.00000000deadbeef", "wb") as f:
        f.write(b"TFPVALS1" + (0xdeadbeef).to_bytes(8, "little") + (0).to_bytes(8, "little"))
    with open(path + ".tmp", "wb") as f:
        f.write(b"TFPSTOR2 half written")
    b = tfp.Store()
    b.load(path)
    assert len(b) == 300 and b.StoresGet(K[:1])[1] == [b"old0"]
    b.close()
    os.remove(old_values[0])                                            # now the pair is broken
    c = tfp.Store()
    with pytest.raises((tfp.StoreError, OSError)):
        c.load(path)
    assert len(c) == 0
    c.close(); a.close()
