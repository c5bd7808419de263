"""The reference's own store tests (tests/integration/stores_test.go, core/http/app_test.go:904-974)
replayed against the CUDA store through the C ABI, each also compared call by call with the oracle."""
import numpy as np
import pytest

from conftest import normalize_rows
from test_oracle_kat import triangle_holds

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(params=["one-gpu", "multi"])
def pair(request, tfp, oracle):
    """Every case runs twice: on the one-GPU store and on ONE logical store over two shards in this process
    (tfp_store_open_devices; the two shards share GPU 0 when the box has a single GPU)."""
    from conftest import two_devices
    s = tfp.Store() if request.param == "one-gpu" else tfp.Store(devices=two_devices())
    yield s, oracle.OracleStore()
    s.close()


def both(pair, method, *args):
    """Run the same call on the CUDA store and on the oracle; compare what they return."""
    gpu, orc = pair
    a = getattr(gpu, method)(*args)
    b = getattr(orc, method)(*args)
    return a, b


def same_find(a, b):
    (ka, va, sa), (kb, vb, sb) = a, b
    assert np.array_equal(sa, sb), f"{sa} vs {sb}"
    assert va == vb
    assert len(ka) == len(kb) and all(np.array_equal(x, y, equal_nan=True) for x, y in zip(ka, kb))


def same_get(a, b):
    (ka, va), (kb, vb) = a, b
    assert va == vb
    assert len(ka) == len(kb) and all(np.array_equal(x, y, equal_nan=True) for x, y in zip(ka, kb))


def test_set_a_key(pair):  # stores_test.go:86-89
    both(pair, "StoresSet", [[0.1, 0.2, 0.3]], [b"test"])
    assert len(pair[0]) == len(pair[1]) == 1


def test_set_keys(pair):  # stores_test.go:91-97
    both(pair, "StoresSet", [[0.1, 0.2, 0.3], [0.4, 0.5, 0.6]], [b"test1", b"test2"])
    both(pair, "StoresSet", [[0.7, 0.8, 0.9], [0.10, 0.11, 0.12]], [b"test3", b"test4"])
    assert len(pair[0]) == len(pair[1]) == 4


def test_get_a_key(pair):  # stores_test.go:99-106
    both(pair, "StoresSet", [[0.1, 0.2, 0.3]], [b"test"])
    a, b = both(pair, "StoresGet", [[0.1, 0.2, 0.3]])
    same_get(a, b)
    assert a[1] == [b"test"]


def test_get_keys(pair):  # stores_test.go:108-146
    both(pair, "StoresSet", [[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    a, b = both(pair, "StoresGet", [[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]])
    same_get(a, b)
    assert len(a[0]) == 3
    a, b = both(pair, "StoresGet", [[0.7, 0.8, 0.9], [0.1, 0.2, 0.3]])
    same_get(a, b)
    assert a[1] == [b"test1", b"test3"]  # sorted-key order


def test_delete_a_key(pair):  # stores_test.go:148-157
    both(pair, "StoresSet", [[0.1, 0.2, 0.3]], [b"test"])
    both(pair, "StoresDelete", [[0.1, 0.2, 0.3]])
    a, b = both(pair, "StoresGet", [[0.1, 0.2, 0.3]])
    same_get(a, b)
    assert a == ([], [])


def test_delete_keys(pair):  # stores_test.go:159-181
    both(pair, "StoresSet", [[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    both(pair, "StoresDelete", [[0.1, 0.2, 0.3], [0.7, 0.8, 0.9]])
    a, b = both(pair, "StoresGet", [[0.4, 0.5, 0.6]])
    same_get(a, b)
    assert a[1] == [b"test2"] and np.array_equal(a[0][0], np.array([0.4, 0.5, 0.6], f32))
    a, b = both(pair, "StoresGet", [[0.1, 0.2, 0.3], [0.7, 0.8, 0.9]])
    same_get(a, b)
    assert a == ([], [])


def test_find_similar_keys(pair):  # stores_test.go:183-203
    both(pair, "StoresSet", [[0.5, 0.5, 0.5], [0.6, 0.6, -0.6], [0.7, -0.7, -0.7]], [b"test1", b"test2", b"test3"])
    a, b = both(pair, "StoresFind", [0.1, 0.3, 0.5], 2)
    same_find(a, b)
    ks, vs, sims = a
    assert np.array_equal(ks[0], np.array([0.5, 0.5, 0.5], f32)) and vs[0] == b"test1"
    assert np.array_equal(ks[1], np.array([0.6, 0.6, -0.6], f32))
    assert pair[0].last_path == pair[1].last_path == "fallback"
    assert sims[0] == f32(0.8783101) and sims[1] == f32(-0.09759001)


def test_find_similar_normalized_keys(pair):  # stores_test.go:205-232
    keys = normalize_rows([[0.1, 0.3, 0.5], [0.5, 0.5, 0.5], [0.6, 0.6, -0.6], [0.7, -0.7, -0.7]])
    both(pair, "StoresSet", keys, [b"test0", b"test1", b"test2", b"test3"])
    a, b = both(pair, "StoresFind", keys[0], 3)
    same_find(a, b)
    ks, vs, sims = a
    assert pair[0].last_path == "normalized"
    assert np.array_equal(ks[0], keys[0]) and vs[0] == b"test0" and abs(float(sims[0]) - 1) < 1e-4
    assert np.array_equal(ks[1], keys[1]) and vs[1] == b"test1"


def test_orthogonal_and_opposite_unit_vectors_exact(pair):  # stores_test.go:234-244
    keys = [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [-1.0, 0.0, 0.0]]
    both(pair, "StoresSet", keys, [b"x", b"y", b"z", b"-z"])
    (_, _, sims), (_, _, osims) = both(pair, "StoresFind", keys[0], 4)
    assert np.array_equal(sims, np.array([1.0, 0.0, 0.0, -1.0], f32))
    assert np.array_equal(sims, osims)


def test_orthogonal_and_opposite_vectors(pair):  # stores_test.go:246-259
    keys = [[1.0, 0.0, 1.0], [0.0, 2.0, 0.0], [0.0, 0.0, -1.0], [-1.0, 0.0, -1.0]]
    both(pair, "StoresSet", keys, [b"x", b"y", b"z", b"-z"])
    a, b = both(pair, "StoresFind", keys[0], 4)
    same_find(a, b)
    sims = a[2]
    assert abs(sims[0] - 1) < 0.1 and abs(sims[1]) < 0.1 and abs(sims[2] + 0.7) < 0.1 and abs(sims[3] + 1) < 0.1


def test_triangle_inequality_normalized(pair):  # stores_test.go:303-321
    keys = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, -1],
                     [2, 3, 4], [9, 7, 1], [0, -1.2, 2.3]], f32)
    keys[6:] = normalize_rows(keys[6:])
    vals = [b"x", b"y", b"z", b"-x", b"-y", b"-z", b"u", b"v", b"w"]
    both(pair, "StoresSet", keys, vals)
    assert triangle_holds(pair[0], keys, vals)


def test_triangle_inequality_random(pair):  # stores_test.go:323-348
    rng = np.random.default_rng(151)
    keys = rng.random((20, 768), dtype=f32)
    vals = [bytes([ord("a") + i]) for i in range(20)]
    both(pair, "StoresSet", keys, vals)
    assert triangle_holds(pair[0], keys, vals)
    for k in keys[:5]:  # and bit-exact against the oracle
        a, b = both(pair, "StoresFind", k, 9)
        assert np.array_equal(a[2], b[2])


def test_http_app_sequence(pair):  # core/http/app_test.go:904-974
    both(pair, "StoresSet", [[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    same_get(*both(pair, "StoresGet", [[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]]))
    both(pair, "StoresDelete", [[0.1, 0.2, 0.3]])
    a, b = both(pair, "StoresFind", [0.1, 0.3, 0.7], 10)
    same_find(a, b)
    assert len(a[0]) == 2 and a[1] == [b"test2", b"test3"]
    assert a[2][0] == f32(0.90502065) and a[2][1] == f32(0.8786198)


def test_error_texts_match_the_reference(tfp, oracle):
    s, o = tfp.Store(), oracle.OracleStore()

    def same_error(method, *args):
        with pytest.raises(tfp.StoreError) as eg:
            getattr(s, method)(*args)
        with pytest.raises(oracle.OracleError) as eo:
            getattr(o, method)(*args)
        assert str(eg.value) == str(eo.value)
        return str(eg.value)

    assert "existing length is -1" in same_error("StoresFind", [1, 0, 0], 1)   # store.go:480-482
    same_error("StoresSet", [], [])                                             # :107-109
    same_error("StoresSet", [[1, 0, 0]], [b"a", b"b"])                          # :111-113
    s.StoresSet([[1, 0, 0]], [b"a"]); o.StoresSet([[1, 0, 0]], [b"a"])
    same_error("StoresSet", [[1, 0]], [b"a"])                                   # :118-120
    same_error("StoresDelete", [])                                              # :201-203
    same_error("StoresDelete", [[1, 0]])                                        # :212-214
    same_error("StoresGet", [[1, 0]])                                           # :278-280
    same_error("StoresFind", [1, 0], 1)                                         # :480-482
    same_error("StoresFind", [1, 0, 0], 0)                                      # :484-486
    same_error("StoresFind", [1, 0, 0], -5)
    with pytest.raises(tfp.StoreError):   # documented deviation: the reference panics (store.go:276)
        s.StoresGet([])
    s.close()


def test_sticky_flag_overwrite_duplicates(pair):
    gpu, orc = pair
    both(pair, "StoresSet", [[1, 0, 0]], [b"a"])
    assert gpu.keys_are_normalized and orc.keys_are_normalized
    both(pair, "StoresSet", [[2, 0, 0]], [b"b"])
    assert not gpu.keys_are_normalized and not orc.keys_are_normalized
    both(pair, "StoresDelete", [[2, 0, 0]])
    assert not gpu.keys_are_normalized                     # never flips back (store.go:126-127)
    both(pair, "StoresSet", [[1, 0, 0]], [b"c"])            # overwrite (store.go:183-187)
    assert len(gpu) == len(orc) == 1
    same_get(*both(pair, "StoresGet", [[1, 0, 0]]))
    both(pair, "StoresSet", [[0, 1, 0], [0, 1, 0]], [b"d", b"d"])   # duplicates inside a batch are both kept
    assert len(gpu) == len(orc) == 3
    both(pair, "StoresSet", [[-0.0, 1, 0]], [b"d"])         # -0 == +0: replaces one of the two
    assert len(gpu) == len(orc) == 3
    a, b = both(pair, "StoresFind", [0, 1, 0], 10)          # TopK > N returns N
    assert len(a[0]) == len(b[0]) == 3 and np.array_equal(a[2], b[2])
    both(pair, "StoresDelete", [[0, 1, 0], [0, 1, 0], [0, 1, 0]])   # each stored row is consumed once
    assert len(gpu) == len(orc) == 1


def test_nan_keys_are_one_key(pair):  # cmp.Compare: NaN == NaN
    nan = float("nan")
    both(pair, "StoresSet", [[nan, 1, 0]], [b"a"])
    both(pair, "StoresSet", [[nan, 1, 0]], [b"b"])
    assert len(pair[0]) == len(pair[1]) == 1
    same_get(*both(pair, "StoresGet", [[nan, 1, 0]]))


def test_delete_latches_key_len_and_emptied_store_finds_nothing(pair):
    gpu, orc = pair
    both(pair, "StoresDelete", [[1, 2, 3]])                 # store.go:209-210
    assert gpu.key_len == orc.key_len == 3
    both(pair, "StoresSet", [[1, 0, 0]], [b"a"])
    both(pair, "StoresDelete", [[1, 0, 0]])
    a, b = both(pair, "StoresFind", [1, 0, 0], 3)
    assert a[0] == [] and b[0] == [] and len(a[2]) == 0


def test_random_mutation_stream_matches_oracle(pair):
    """Interleaved set / overwrite / delete / get / find on a small key universe."""
    gpu, orc = pair
    rng = np.random.default_rng(7)
    universe = normalize_rows(rng.standard_normal((60, 36)).astype(f32))
    for step in range(40):
        op = rng.integers(0, 4)
        pick = rng.choice(60, size=int(rng.integers(1, 9)), replace=False)
        ks = universe[pick]
        if op == 0 or len(gpu) == 0:
            vals = [b"v%d_%d" % (step, i) for i in pick]
            both(pair, "StoresSet", ks, vals)
        elif op == 1:
            both(pair, "StoresDelete", ks)
        elif op == 2:
            same_get(*both(pair, "StoresGet", ks))
        else:
            a, b = both(pair, "StoresFind", ks[0], int(rng.integers(1, 12)))
            same_find(a, b)
        assert len(gpu) == len(orc)


def test_snapshot_and_restore(tfp, oracle, tmp_path):
    """a saved store answers like the original after a restart (new process state = new Store)."""
    from conftest import make_keys
    K = make_keys(3000, 96, seed=3, normalized=True)
    vals = [b"v%d" % i for i in range(len(K))]
    a = tfp.Store()
    a.StoresSet(K, vals)
    a.StoresDelete(K[100:300])
    a.StoresSet([(K[0] * f32(2)).astype(f32)], [b"big"])       # flips the sticky flag
    a.StoresDelete([(K[0] * f32(2)).astype(f32)])
    path = str(tmp_path / "store.snap")
    a.save(path)
    b = tfp.Store()
    b.load(path)
    assert len(b) == len(a) == 2800 and b.key_len == 96 and not b.keys_are_normalized
    for q in (K[5], K[2999], K[150]):
        ra, rb = a.StoresFind(q, 10), b.StoresFind(q, 10)
        assert np.array_equal(ra[2], rb[2]) and ra[1] == rb[1]
    assert b.StoresGet(K[[7, 150, 2000]])[1] == [b"v7", b"v2000"] or b.StoresGet(K[[7, 150, 2000]])[1] == [b"v2000", b"v7"]
    b.StoresSet(K[100:110], [b"again"] * 10)                    # ids continue after the restored counter
    assert len(b) == 2810 and len(set(b.values)) == 2810
    with pytest.raises(tfp.StoreError):
        b.load(path)                                            # only an empty store can be restored into
    a.close(); b.close()


@pytest.mark.parametrize("kind", ["one-gpu", "multi"])
def test_finds_run_while_keys_are_appended(tfp, oracle, kind):
    """Appends do not fence the finds ("Finds are fenced ... never on an append"): reader threads keep
    finding while a writer appends and a deleter compacts; every answer must be the exact top-k of SOME state the store
    went through -- here checked as: similarities bit-identical to the oracle's score of each returned row, descending,
    no row twice -- and the final state must equal the oracle's."""
    import threading
    from conftest import make_keys, two_devices
    d, per, batches = 96, 1500, 24
    K = make_keys(per * batches, d, seed=41, normalized=True)
    Q = make_keys(8, d, seed=42, normalized=True)
    s = tfp.Store() if kind == "one-gpu" else tfp.Store(devices=two_devices())
    o = oracle.OracleStore()
    vals = [b"%d" % i for i in range(len(K))]
    s.StoresSet(K[:per], vals[:per]); o.StoresSet(K[:per], vals[:per])
    stop = threading.Event()
    errors, finds = [], [0]

    def reader(t):
        try:
            while not stop.is_set():
                q = Q[(t + finds[0]) % len(Q)]
                ks, vs, sims = s.StoresFind(q, 10)
                finds[0] += 1
                assert len(sims) == 10 and (np.diff(sims.astype(np.float64)) <= 0).all()
                assert len(set(vs)) == 10
                rows = np.stack(ks)
                assert np.array_equal(rows, K[[int(v) for v in vs]])             # the key rows are the stored rows of those values
                assert np.array_equal(oracle.scores_block(rows, q, True), sims)  # and the similarities are exact
        except Exception as e:      # noqa: BLE001
            errors.append(repr(e))
            stop.set()

    ts = [threading.Thread(target=reader, args=(t,)) for t in range(3)]
    [t.start() for t in ts]
    try:
        for b in range(1, batches):
            lo, hi = b * per, (b + 1) * per
            s.StoresSet(K[lo:hi], vals[lo:hi]); o.StoresSet(K[lo:hi], vals[lo:hi])
            if b % 6 == 0:                                                       # a compaction now and then (this one does fence)
                s.StoresDelete(K[lo - per:lo - per + 200]); o.StoresDelete(K[lo - per:lo - per + 200])
            if stop.is_set():
                break
    finally:
        stop.set()
        [t.join() for t in ts]
    assert not errors, errors[:2]
    assert finds[0] > 0 and len(s) == len(o)
    for q in Q[:3]:
        a, b2 = s.StoresFind(q, 10), o.StoresFind(q, 10)
        assert np.array_equal(a[2], b2[2]) and a[1] == b2[1]
    s.close()
