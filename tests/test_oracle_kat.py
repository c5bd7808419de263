"""Pins the CPU oracle (oracle/store_oracle.c) against every expectation the reference's own
tests hold for the store path, and cross-checks its arithmetic against an independent numpy
restatement.  All CPU; the reference itself cannot run here (no Go toolchain)."""
import numpy as np
import pytest

from conftest import make_keys, normalize_rows

f32 = np.float32


# ---- tests/integration/stores_test.go, restated case by case ---------------------------

def test_set_single_and_cols(oracle):  # stores_test.go:86-97
    s = oracle.OracleStore()
    s.StoresSet([[0.1, 0.2, 0.3]], [b"test"])
    s.StoresSet([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6]], [b"test1", b"test2"])
    s.StoresSet([[0.7, 0.8, 0.9], [0.10, 0.11, 0.12]], [b"test3", b"test4"])
    assert len(s) == 4


def test_get_keys_sorted_and_only_found(oracle):  # stores_test.go:99-146
    s = oracle.OracleStore()
    s.StoresSet([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    ks, vs = s.StoresGet([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]])
    assert vs == [b"test1", b"test2", b"test3"]
    ks, vs = s.StoresGet([[0.7, 0.8, 0.9], [0.1, 0.2, 0.3]])
    assert vs == [b"test1", b"test3"]  # sorted-key order, not request order
    assert np.array_equal(ks[0], np.array([0.1, 0.2, 0.3], f32))


def test_delete(oracle):  # stores_test.go:148-181
    s = oracle.OracleStore()
    s.StoresSet([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    s.StoresDelete([[0.1, 0.2, 0.3], [0.7, 0.8, 0.9]])
    ks, vs = s.StoresGet([[0.4, 0.5, 0.6]])
    assert vs == [b"test2"]
    ks, vs = s.StoresGet([[0.1, 0.2, 0.3], [0.7, 0.8, 0.9]])
    assert ks == [] and vs == []


def test_find_similar_keys(oracle):  # stores_test.go:183-203 ; SURVEY App. C row 1
    s = oracle.OracleStore()
    s.StoresSet([[0.5, 0.5, 0.5], [0.6, 0.6, -0.6], [0.7, -0.7, -0.7]], [b"test1", b"test2", b"test3"])
    ks, vs, sims = s.StoresFind([0.1, 0.3, 0.5], 2)
    assert s.last_path == "fallback"
    assert np.array_equal(ks[0], np.array([0.5, 0.5, 0.5], f32)) and vs[0] == b"test1"
    assert np.array_equal(ks[1], np.array([0.6, 0.6, -0.6], f32))
    assert sims[0] == f32(0.8783101) and sims[1] == f32(-0.09759001)


def test_find_similar_normalized_keys(oracle):  # stores_test.go:205-232 ; App. C row 2
    keys = normalize_rows([[0.1, 0.3, 0.5], [0.5, 0.5, 0.5], [0.6, 0.6, -0.6], [0.7, -0.7, -0.7]])
    s = oracle.OracleStore()
    s.StoresSet(keys, [b"test0", b"test1", b"test2", b"test3"])
    ks, vs, sims = s.StoresFind(keys[0], 3)
    assert s.last_path == "normalized"
    assert np.array_equal(ks[0], keys[0]) and vs[0] == b"test0"
    assert abs(float(sims[0]) - 1.0) < 1e-4
    assert np.array_equal(ks[1], keys[1]) and vs[1] == b"test1"
    assert sims[0] == f32(0.99999994) and sims[1] == f32(0.87830997) and sims[2] == f32(-0.09759)


def test_orthogonal_and_opposite_unit_vectors_exact(oracle):  # stores_test.go:234-244 (exact)
    keys = [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [-1.0, 0.0, 0.0]]
    s = oracle.OracleStore()
    s.StoresSet(keys, [b"x", b"y", b"z", b"-z"])
    _, _, sims = s.StoresFind(keys[0], 4)
    assert np.array_equal(sims, np.array([1.0, 0.0, 0.0, -1.0], f32))


def test_orthogonal_and_opposite_vectors(oracle):  # stores_test.go:246-259 ; App. C row 4
    keys = [[1.0, 0.0, 1.0], [0.0, 2.0, 0.0], [0.0, 0.0, -1.0], [-1.0, 0.0, -1.0]]
    s = oracle.OracleStore()
    s.StoresSet(keys, [b"x", b"y", b"z", b"-z"])
    _, _, sims = s.StoresFind(keys[0], 4)
    assert s.last_path == "fallback"
    assert abs(sims[0] - 1) < 0.1 and abs(sims[1]) < 0.1 and abs(sims[2] + 0.7) < 0.1 and abs(sims[3] + 1) < 0.1
    assert np.array_equal(sims, np.array([1.0, 0.0, -0.70710677, -1.0], f32))


def triangle_holds(store, keys, vals):  # stores_test.go:261-301 expectTriangleEq
    sims = {}
    for i, k in enumerate(keys):
        _, valsk, simsk = store.StoresFind(k, 9)
        for j, v in enumerate(valsk):
            sims.setdefault(vals[i], {})[v] = float(simsk[j])
    for u, simsu in sims.items():
        for w, simw in simsu.items():
            assert -1.0 <= simw <= 1.0 + 1e-6  # math.Acos must stay real
            uws = np.arccos(min(simw, 1.0))
            for v in simsu:
                if v not in sims or w not in sims[v]:
                    continue
                uvws = np.arccos(min(simsu[v], 1.0)) + np.arccos(min(sims[v][w], 1.0))
                assert uws <= uvws + 1e-6
    return True


def test_triangle_inequality_normalized(oracle):  # stores_test.go:303-321
    keys = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, -1],
                     [2, 3, 4], [9, 7, 1], [0, -1.2, 2.3]], f32)
    keys[6:] = normalize_rows(keys[6:])
    vals = [b"x", b"y", b"z", b"-x", b"-y", b"-z", b"u", b"v", b"w"]
    s = oracle.OracleStore()
    s.StoresSet(keys, vals)
    assert triangle_holds(s, keys, vals)


def test_triangle_inequality_random(oracle):  # stores_test.go:323-348 (Go's rand stream is not reproducible here)
    rng = np.random.default_rng(151)
    keys = rng.random((20, 768), dtype=f32)
    vals = [bytes([ord("a") + i]) for i in range(20)]
    s = oracle.OracleStore()
    s.StoresSet(keys, vals)
    assert triangle_holds(s, keys, vals)


def test_http_app_test_sequence(oracle):  # core/http/app_test.go:904-974 ; App. C row 5
    s = oracle.OracleStore()
    s.StoresSet([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    ks, vs = s.StoresGet([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]])
    assert len(ks) == 3
    s.StoresDelete([[0.1, 0.2, 0.3]])
    ks, vs, sims = s.StoresFind([0.1, 0.3, 0.7], 10)
    assert len(ks) == 2 and all(-1 <= x <= 1 for x in sims)
    assert vs == [b"test2", b"test3"]
    assert sims[0] == f32(0.90502065) and sims[1] == f32(0.8786198)


# ---- error texts and state quirks (store.go line cited per case) ------------------------

def test_error_texts(oracle):
    s = oracle.OracleStore()
    with pytest.raises(oracle.OracleError, match="Try to find key with length 3 when existing length is -1"):
        s.StoresFind([1, 0, 0], 1)  # store.go:480-482 on a never-written store
    with pytest.raises(oracle.OracleError, match="no keys to add"):
        s.StoresSet([], [])  # :107-109
    with pytest.raises(oracle.OracleError, match=r"len\(keys\) = 1, len\(values\) = 2"):
        s.StoresSet([[1, 0, 0]], [b"a", b"b"])  # :111-113
    s.StoresSet([[1, 0, 0]], [b"a"])
    with pytest.raises(oracle.OracleError, match="Try to add key with length 2 when existing length is 3"):
        s.StoresSet([[1, 0]], [b"a"])  # :118-120
    with pytest.raises(oracle.OracleError, match="no keys to delete"):
        s.StoresDelete([])  # :201-203
    with pytest.raises(oracle.OracleError, match="Trying to delete key with length 2 when existing length is 3"):
        s.StoresDelete([[1, 0]])  # :212-214
    with pytest.raises(oracle.OracleError, match="Try to get a key with length 2 when existing length is 3"):
        s.StoresGet([[1, 0]])  # :278-280
    with pytest.raises(oracle.OracleError, match=r"opts.TopK = 0, must be >= 1"):
        s.StoresFind([1, 0, 0], 0)  # :484-486
    with pytest.raises(oracle.OraclePanic):
        s.StoresGet([])  # :276 index out of range


def test_sticky_flag_overwrite_duplicates(oracle):
    s = oracle.OracleStore()
    s.StoresSet([[1, 0, 0]], [b"a"])
    assert s.keys_are_normalized
    s.StoresSet([[2, 0, 0]], [b"b"])
    assert not s.keys_are_normalized
    s.StoresDelete([[2, 0, 0]])
    assert not s.keys_are_normalized  # never flips back (store.go:126-127)
    s.StoresSet([[1, 0, 0]], [b"c"])  # overwrite (store.go:183-187)
    assert len(s) == 1 and s.StoresGet([[1, 0, 0]])[1] == [b"c"]
    s.StoresSet([[0, 1, 0], [0, 1, 0]], [b"d", b"e"])  # duplicates inside one batch are both kept
    assert len(s) == 3
    s.StoresSet([[-0.0, 1, 0]], [b"f"])  # -0 == +0 under cmp.Compare: replaces one of the two
    assert len(s) == 3
    ks, vs, sims = s.StoresFind([0, 1, 0], 10)
    assert len(ks) == 3  # TopK > N returns N (store.go:386)


def test_delete_latches_key_len_and_emptied_store_finds_nothing(oracle):
    s = oracle.OracleStore()
    s.StoresDelete([[1, 2, 3]])  # store.go:209-210 latches keyLen = 3
    assert s.key_len == 3
    s.StoresSet([[1, 0, 0]], [b"a"])
    s.StoresDelete([[1, 0, 0]])
    ks, vs, sims = s.StoresFind([1, 0, 0], 3)
    assert ks == [] and len(sims) == 0


# ---- arithmetic: C oracle vs an independent numpy restatement -------------------------------

def np_dot_seq_f32(K, q):
    acc = np.zeros(K.shape[0], f32)
    for i in range(K.shape[1]):
        acc = (acc + (K[:, i] * q[i]).astype(f32)).astype(f32)
    return acc


def np_cosine(K, q):
    mag1 = np.sqrt(np.cumsum((q * q).astype(f32).astype(np.float64))[-1])  # sequential fp64 sum
    dot = np.zeros(K.shape[0], np.float64)
    mag2 = np.zeros(K.shape[0], np.float64)
    for i in range(K.shape[1]):
        dot = dot + (K[:, i] * q[i]).astype(f32).astype(np.float64)
        mag2 = mag2 + (K[:, i] * K[:, i]).astype(f32).astype(np.float64)
    return (dot / (mag1 * np.sqrt(mag2))).astype(f32)


@pytest.mark.parametrize("d", [3, 33, 384, 768])
def test_scores_match_numpy_restatement(oracle, d):
    K = make_keys(500, d, seed=d, normalized=True)
    q = make_keys(1, d, seed=d + 1, normalized=True)[0]
    assert np.array_equal(oracle.scores_block(K, q, True), np_dot_seq_f32(K, q))
    K2 = make_keys(500, d, seed=d + 2, normalized=False)
    q2 = (q * f32(1.7)).astype(f32)
    assert np.array_equal(oracle.scores_block(K2, q2, False), np_cosine(K2, q2))


def go_heap_topk(sims, k):
    """container/heap restated in Python to cross-check the C heap, ties included."""
    h = []

    def less(i, j):
        return h[i][0] < h[j][0]

    def up(j):
        while True:
            i = (j - 1) // 2 if j > 0 else 0
            if i == j or not less(j, i):
                break
            h[i], h[j] = h[j], h[i]
            j = i

    def down(i0, n):
        i = i0
        while True:
            j1 = 2 * i + 1
            if j1 >= n:
                break
            j = j1
            if j1 + 1 < n and less(j1 + 1, j1):
                j = j1 + 1
            if not less(j, i):
                break
            h[i], h[j] = h[j], h[i]
            i = j

    def pop():
        n = len(h) - 1
        h[0], h[n] = h[n], h[0]
        down(0, n)
        return h.pop()

    for r, s in enumerate(sims):
        h.append((s, r))
        up(len(h) - 1)
        if len(h) > k:
            pop()
    out = [None] * len(h)
    for i in range(len(h) - 1, -1, -1):
        out[i] = pop()
    return out


@pytest.mark.parametrize("k", [1, 3, 10, 64])
def test_heap_mechanics_including_ties(oracle, k):
    rng = np.random.default_rng(k)
    K = normalize_rows(rng.integers(-1, 2, size=(300, 8)).astype(f32) + f32(0.001))  # many exact ties
    q = K[0]
    sims = oracle.scores_block(K, q, True)
    rows, top = oracle.scan_block(K, q, k, True)
    want = go_heap_topk([float(x) for x in sims], k)
    assert [r for _, r in want] == [int(r) for r in rows]
    assert np.array_equal(top, np.array([s for s, _ in want], f32))


def test_is_normalized_band(oracle):  # store.go:313-324
    assert oracle.is_normalized([1, 0, 0])
    assert oracle.is_normalized([0.99, 0, 0]) and oracle.is_normalized([1.01, 0, 0])
    assert not oracle.is_normalized([0.9899, 0, 0]) and not oracle.is_normalized([1.0101, 0, 0])
    assert not oracle.is_normalized([0, 0, 0])
