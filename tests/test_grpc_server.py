"""Wire-level test of the backend process: spawn `python -m localai_tfp_b200.server --addr ...` the way
LocalAI's loader spawns `local-store` (pkg/model/process.go:94-124: binary + `--addr`), poll Health
(pkg/model/initializers.go:375-387), LoadModel, then replay tests/integration/stores_test.go and
core/http/app_test.go:904-974 over loopback gRPC.  Same idiom as backend/python/rerankers/test.py:17-45."""
import socket
import subprocess
import sys
import time

import grpc
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(params=["one-gpu", "TFP_DEVICES"])
def client(request, tfp):
    """The backend process the way LocalAI's loader starts it: `--addr` and the environment, nothing else.  Every case
    runs twice: against the default one-GPU store, and with TFP_DEVICES naming two devices (the second GPU when the box
    has one, else GPU 0 twice), i.e. ONE logical store over two shards inside the one backend process."""
    import os
    from conftest import two_devices
    from localai_tfp_b200.server import Client
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    addr = f"127.0.0.1:{port}"
    env = dict(os.environ)
    env.pop("TFP_DEVICES", None)
    if request.param == "TFP_DEVICES":
        env["TFP_DEVICES"] = ",".join(str(d) for d in two_devices())
    proc = subprocess.Popen([sys.executable, "-m", "localai_tfp_b200.server", "--addr", addr], env=env)
    c = Client(addr)
    try:
        for attempt in range(60):                       # the loader polls Health up to 20 x 2 s
            try:
                if c.health() == b"OK":
                    break
            except grpc.RpcError:
                time.sleep(0.5)
        else:
            raise RuntimeError("backend did not become healthy")
        assert c.load_model()[1] is True
        yield c
    finally:
        c.close()
        proc.terminate()                                # SIGTERM must end the process
        assert proc.wait(timeout=20) is not None


def test_set_get_delete_over_grpc(client):              # stores_test.go:86-181
    assert client.set([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"]) == ("Set key", True)
    ks, vs = client.get([[0.7, 0.8, 0.9], [0.1, 0.2, 0.3]])
    assert vs == [b"test1", b"test3"] and np.array_equal(ks[0], np.array([0.1, 0.2, 0.3], f32))
    assert client.delete([[0.1, 0.2, 0.3], [0.7, 0.8, 0.9]]) == ("Deleted key", True)
    ks, vs = client.get([[0.4, 0.5, 0.6]])
    assert vs == [b"test2"]
    ks, vs = client.get([[0.1, 0.2, 0.3], [0.7, 0.8, 0.9]])
    assert ks == [] and vs == []


def test_find_over_grpc(client):                         # stores_test.go:183-203, 234-244
    client.set([[0.5, 0.5, 0.5], [0.6, 0.6, -0.6], [0.7, -0.7, -0.7]], [b"test1", b"test2", b"test3"])
    ks, vs, sims = client.find([0.1, 0.3, 0.5], 2)
    assert vs == [b"test1", b"test2"] and np.array_equal(ks[0], np.array([0.5, 0.5, 0.5], f32))
    assert sims[0] == f32(0.8783101) and sims[1] == f32(-0.09759001)


def test_exact_unit_vectors_over_grpc(client):
    keys = [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [-1.0, 0.0, 0.0]]
    client.set(keys, [b"x", b"y", b"z", b"-z"])
    _, _, sims = client.find(keys[0], 4)
    assert np.array_equal(sims, np.array([1.0, 0.0, 0.0, -1.0], f32))


def test_http_app_sequence_over_grpc(client):            # core/http/app_test.go:904-974
    client.set([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]], [b"test1", b"test2", b"test3"])
    assert len(client.get([[0.1, 0.2, 0.3], [0.4, 0.5, 0.6], [0.7, 0.8, 0.9]])[0]) == 3
    client.delete([[0.1, 0.2, 0.3]])
    ks, vs, sims = client.find([0.1, 0.3, 0.7], 10)
    assert vs == [b"test2", b"test3"] and all(-1 <= s <= 1 for s in sims)


def test_errors_are_grpc_unknown_with_the_reference_text(client):   # pkg/grpc/server.go:194-240
    with pytest.raises(grpc.RpcError) as e:
        client.find([1.0, 0.0, 0.0], 1)
    assert e.value.code() == grpc.StatusCode.UNKNOWN
    assert e.value.details() == "Try to find key with length 3 when existing length is -1"
    with pytest.raises(grpc.RpcError) as e:
        client.set([], [])
    assert e.value.details() == "no keys to add"
    client.set([[1.0, 0.0, 0.0]], [b"a"])
    with pytest.raises(grpc.RpcError) as e:
        client.find([1.0, 0.0, 0.0], 0)
    assert e.value.details() == "opts.TopK = 0, must be >= 1"


def test_status_reports_hbm(client):
    client.set([[1.0, 0.0, 0.0]], [b"a"])
    state, total, breakdown = client.status()
    assert state == 2 and breakdown["rows"] == 1 and total > 0


def test_c1_shape_over_grpc(client, oracle):             # BASELINE.json configs[0]: 10k x 384, top-10, via gRPC
    from conftest import make_keys
    K = make_keys(10000, 384, seed=1, normalized=True)
    for lo in range(0, 10000, 2500):
        client.set(K[lo:lo + 2500], [b"%d" % i for i in range(lo, lo + 2500)])
    q = make_keys(1, 384, seed=2, normalized=True)[0]
    rows, osims = oracle.scan_block(K, q, 10, True)
    ks, vs, sims = client.find(q, 10)
    assert np.array_equal(sims, osims) and [int(v) for v in vs] == [int(r) for r in rows]
    t0 = time.perf_counter()
    for _ in range(50):
        client.find(q, 10)
    print(f"C1 over loopback gRPC: {(time.perf_counter() - t0) / 50 * 1e6:.0f} us per StoresFind")


def test_snapshot_survives_a_respawn(tfp, tmp_path):
    """The reference loses the store when LocalAI's watchdog kills the backend; with
    --snapshot the respawned process answers as before."""
    from localai_tfp_b200.server import Client
    snap = str(tmp_path / "store.snap")

    def spawn():
        s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
        addr = f"127.0.0.1:{port}"
        proc = subprocess.Popen([sys.executable, "-m", "localai_tfp_b200.server", "--addr", addr, "--snapshot", snap])
        c = Client(addr)
        for attempt in range(60):
            try:
                if c.health() == b"OK":
                    return proc, c
            except grpc.RpcError:
                time.sleep(0.5)
        raise RuntimeError("backend did not become healthy")

    proc, c = spawn()
    c.set([[0.5, 0.5, 0.5], [0.6, 0.6, -0.6], [0.7, -0.7, -0.7]], [b"test1", b"test2", b"test3"])
    before = c.find([0.1, 0.3, 0.5], 2)
    c.close(); proc.terminate(); proc.wait(timeout=30)
    proc, c = spawn()
    after = c.find([0.1, 0.3, 0.5], 2)
    assert after[1] == before[1] == [b"test1", b"test2"] and np.array_equal(after[2], before[2])
    c.close(); proc.terminate(); proc.wait(timeout=30)


def test_concurrent_finds_are_batched_onto_the_tensor_cores(tfp, oracle):
    """64 clients call StoresFind at once against a 70k-row store: the server's micro-batcher must answer
    each exactly as a lone call would, and batches of >= 32 must have gone through the tensor-core path."""
    import threading
    from conftest import make_keys
    from localai_tfp_b200.server import Client, make_server
    server, backend, port = make_server("127.0.0.1:0")
    server.start()
    try:
        K = make_keys(70000, 64, seed=41, normalized=True)
        c0 = Client(f"127.0.0.1:{port}")
        for lo in range(0, len(K), 10000):
            c0.set(K[lo:lo + 10000], [b"%d" % i for i in range(lo, min(lo + 10000, len(K)))])
        Q = make_keys(64, 64, seed=42, normalized=True)
        want = [oracle.scan_block(K, q, 10, True) for q in Q]
        got = [None] * len(Q)
        clients = [Client(f"127.0.0.1:{port}") for _ in Q]
        for rounds in range(3):
            gate = threading.Barrier(len(Q))

            def work(i):
                gate.wait()
                got[i] = clients[i].find(Q[i], 10)
            ts = [threading.Thread(target=work, args=(i,)) for i in range(len(Q))]
            # hold the batcher briefly so that the requests pile up behind it, as they do behind a busy GPU:
            # it must then take the backlog as one batch
            backend.batcher.gate.acquire()
            [t.start() for t in ts]
            deadline = time.time() + 10
            while backend.batcher.q.qsize() < len(Q) - 8 and time.time() < deadline:
                time.sleep(0.01)
            backend.batcher.gate.release()
            [t.join() for t in ts]
            for i in range(len(Q)):
                ks, vs, sims = got[i]
                assert np.array_equal(sims, want[i][1])
                assert [int(v) for v in vs] == [int(r) for r in want[i][0]]
                assert all(np.array_equal(k, K[int(v)]) for k, v in zip(ks, vs))
        assert backend.batcher.largest >= 32, backend.batcher.largest
        assert backend.store.mma_counters()[0] >= 1
        [c.close() for c in clients]; c0.close()
    finally:
        server.stop(grace=0)


def test_find_batch_rpc_equals_single_finds(client):
    """StoresFindBatch: result i is exactly StoresFind(Keys[i], TopK), errors keep the reference's texts."""
    from conftest import make_keys
    K = make_keys(3000, 64, seed=21, normalized=True)
    client.set(list(K), [b"v%d" % i for i in range(len(K))])
    Q = [K[i] for i in (0, 5, 77, 1500, 2999)] + [(K[9] * f32(2)).astype(f32)]      # the last one takes the cosine path
    for topk in (1, 10):
        got = client.find_batch(Q, topk)
        assert len(got) == len(Q)
        for q, (ks, vs, sims) in zip(Q, got):
            k1, v1, s1 = client.find(q, topk)
            assert vs == v1 and np.array_equal(sims, s1) and all(np.array_equal(a, b) for a, b in zip(ks, k1))
    for keys, topk, text in (([], 3, "no keys to find"), ([K[0][:5]], 3, "Try to find key with length 5 when existing length is 64"),
                             ([K[0]], 0, "opts.TopK = 0, must be >= 1")):
        with pytest.raises(grpc.RpcError) as e:
            client.find_batch(keys, topk)
        assert e.value.code() == grpc.StatusCode.UNKNOWN and e.value.details() == text


def test_find_without_key_echo_and_packed_set(client):
    """SURVEY §8f.3 on the wire: StoresSetPacked stores what StoresSet would; StoresFind with NoKeys returns the same
    values and similarities with no keys in the reply."""
    rng = np.random.default_rng(3)
    K = rng.standard_normal((200, 48)).astype(f32)
    vals = [b"v%d" % i for i in range(len(K))]
    assert client.set_packed(K, vals) == ("Set key", True)
    assert client.set_packed(K[:5], [b"w%d" % i for i in range(5)]) == ("Set key", True)      # overwrite, like StoresSet
    ks, vs = client.get([K[3], K[100]])
    assert sorted(vs) == [b"v100", b"w3"]
    full = client.find(K[7], 10)
    lean = client.find(K[7], 10, no_keys=True)
    assert lean[0] == [] and lean[1] == full[1] and np.array_equal(lean[2], full[2]) and len(full[0]) == 10
    with pytest.raises(grpc.RpcError) as e:
        client.set_packed(K[:2, :7], [b"a", b"b"])
    assert e.value.details() == "Try to add key with length 7 when existing length is 48"
