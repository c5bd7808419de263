"""Worker body for the multi-process sharded-store tests (gloo on CPU, nccl on GPUs)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def scenario(store, single, rank):
    """Drive the sharded store and a single oracle store with the same calls; compare everything."""
    from conftest import make_keys
    f32 = np.float32
    K = make_keys(400, 48, seed=5, normalized=True)
    vals = [b"v%d" % i for i in range(len(K))]
    store.StoresSet(K[:300], vals[:300]); single.StoresSet(K[:300], vals[:300])
    assert len(store) == len(single) == 300
    assert store.keys_are_normalized

    def same_find(q, k):
        a = store.StoresFind(q, k); b = single.StoresFind(q, k)
        assert np.array_equal(a[2], b[2]), (rank, a[2], b[2])
        assert a[1] == b[1], (rank, a[1], b[1])
        assert len(a[0]) == len(b[0]) and all(np.array_equal(x, y) for x, y in zip(a[0], b[0]))

    for qi in (0, 7, 299, 350):
        same_find(K[qi], 10)
    same_find(K[3], 500)                      # TopK > N
    same_find((K[3] * f32(2)).astype(f32), 5)  # non-unit query: fallback path on every shard
    # get: sorted-key order across shards, missing keys dropped
    ask = K[[250, 10, 399, 77]]
    a = store.StoresGet(ask); b = single.StoresGet(ask)
    assert a[1] == b[1] and all(np.array_equal(x, y) for x, y in zip(a[0], b[0]))
    # overwrite + delete
    store.StoresSet(K[10:12], [b"new10", b"new11"]); single.StoresSet(K[10:12], [b"new10", b"new11"])
    store.StoresDelete(K[100:200]); single.StoresDelete(K[100:200])
    assert len(store) == len(single) == 200
    same_find(K[10], 3)
    same_find(K[150], 10)
    # one non-unit key on ONE shard flips the whole store to the fallback path (sticky)
    big = (K[399:400] * f32(3)).astype(f32)
    store.StoresSet(big, [b"big"]); single.StoresSet(big, [b"big"])
    assert not store.keys_are_normalized
    for qi in (0, 7, 399):
        same_find(K[qi], 10)
    store.StoresDelete(big); single.StoresDelete(big)
    assert not store.keys_are_normalized
    same_find(K[1], 4)
    # errors carry the reference's texts
    from localai_tfp_b200 import StoreError
    for fn, args, text in ((store.StoresFind, (K[0][:5], 3), "Try to find key with length 5 when existing length is 48"),
                           (store.StoresFind, (K[0], 0), "opts.TopK = 0, must be >= 1"),
                           (store.StoresSet, ([], []), "no keys to add"),
                           (store.StoresDelete, ([],), "no keys to delete")):
        try:
            fn(*args)
            raise AssertionError("no error")
        except StoreError as e:
            assert str(e) == text, str(e)


def native_path_matches(store, single, rank, world):
    """scan + ncclAllGather + merge inside libtfpstore (one call) must equal the torch.distributed path."""
    from conftest import make_keys
    K = make_keys(400, 48, seed=5, normalized=True)
    Q = np.ascontiguousarray(K[[1, 7, 150, 399]])
    store.ops.stage_queries(Q)
    ref = []
    for qi in range(len(Q)):
        sims, ids, shard, index, n, _ = store.find_device(qi, 10, with_keys=False)
        torch.cuda.synchronize()
        m = int(n.item())
        ref.append((ids[:m].cpu().numpy().copy(), sims[:m].cpu().numpy().copy()))
    store.ops.init_native_comm(rank, world)
    for qi in range(len(Q)):
        sims, ids, shard, index, n, _ = store.find_device(qi, 10, with_keys=False)
        torch.cuda.synchronize()
        m = int(n.item())
        assert np.array_equal(ids[:m].cpu().numpy(), ref[qi][0]) and np.array_equal(sims[:m].cpu().numpy(), ref[qi][1])
        got_ids, got_sims, got_shard, got_index = store.ops.lanes[0].ctx.download_merged()
        assert np.array_equal(got_ids, ref[qi][0]) and np.array_equal(got_sims, ref[qi][1])
        want = single.StoresFind(Q[qi], 10)
        assert np.array_equal(got_sims, want[2])


def p2p_path_matches(single, rank, world):
    """Sharded finds over peer memory (records pushed by the scan kernel, stream waits, merge) on a
    store that never had an NCCL communicator inside the library: a result at all proves the peer
    path ran; it must equal the torch.distributed path bit for bit, on both lanes, find after find."""
    from conftest import make_keys
    from localai_tfp_b200.sharded import CudaShardOps, ShardedStore
    ops = CudaShardOps(rank, rank, lanes=2)
    store = ShardedStore(ops, rank, world)
    K = make_keys(400, 48, seed=5, normalized=True)
    store.StoresSet(K[:300], [b"v%d" % i for i in range(300)])
    ref_store = __import__("oracle.oracle", fromlist=["x"]).OracleStore()
    ref_store.StoresSet(K[:300], [b"v%d" % i for i in range(300)])
    Q = np.ascontiguousarray(np.concatenate([K[[1, 7, 150, 299, 33, 34, 350]], (K[5:6] * np.float32(2)).astype(np.float32)]))
    ops.stage_queries(Q)

    def torch_path(qi, k, lane):
        sims, ids, shard, index, n, _ = store.find_device(qi, k, with_keys=False, lane=lane)
        torch.cuda.synchronize()
        m = int(n.item())
        return tuple(t[:m].cpu().numpy().copy() for t in (ids, sims, shard, index))

    ref = {(qi, k): torch_path(qi, k, 0) for qi in range(len(Q)) for k in (10, 1, 300)}
    assert not ops.native
    ops.init_p2p(rank, world, slot_bytes=64 << 10)
    for rep in range(3):                      # parity buffers are reused every second find
        for lane in (0, 1):
            ctx = ops.lanes[lane].ctx
            for qi in range(len(Q)):
                for k in (10, 1, 300):
                    ctx.launch_sharded(qi, k)
                    got = ctx.download_merged()
                    for a, b in zip(got, ref[(qi, k)]):
                        assert np.array_equal(a, b), (rank, rep, lane, qi, k, a, b)
    want = ref_store.StoresFind(Q[2], 10)
    assert np.array_equal(ref[(2, 10)][1], want[2])
    # the whole StoresFind (keys, values, similarities) through the peer exchange, against the oracle;
    # k = 500 does not fit the 64 KiB slot with its key rows and takes the general path
    for rep in range(2):
        for qi in (0, 2, 6, 7):
            for k in (10, 1, 300, 500):
                a = store.StoresFind(Q[qi], k); b = ref_store.StoresFind(Q[qi], k)
                assert np.array_equal(a[2], b[2]), (rank, qi, k, a[2], b[2])
                assert a[1] == b[1], (rank, qi, k)
                assert len(a[0]) == len(b[0]) and all(np.array_equal(x, y) for x, y in zip(a[0], b[0])), (rank, qi, k)
    ops.stage_queries(Q)
    # batches: 7 unit queries then one non-unit one (two paths -> several launches, the last one raises the words)
    ctx = ops.lanes[0].ctx
    for first, count in ((0, 8), (0, 3), (2, 5), (7, 1)):
        ctx.launch_sharded(first, 10, count=count)
        got = ctx.download_merged()
        got = [got] if count == 1 else got
        for j in range(count):
            for a, b in zip(got[j], ref[(first + j, 10)]):
                assert np.array_equal(a, b), (rank, first, count, j)
    # both lanes in flight at once, no host wait in between
    for i in range(20):
        ops.lanes[i % 2].ctx.launch_sharded(i % len(Q), 10)
    for lane, qi in ((0, 18 % len(Q)), (1, 19 % len(Q))):
        got = ops.lanes[lane].ctx.download_merged()
        for a, b in zip(got, ref[(qi, 10)]):
            assert np.array_equal(a, b), (rank, lane, qi)
    # a record too large for the slot and no NCCL communicator: a clean error on every rank, nothing enqueued
    from localai_tfp_b200 import StoreError
    try:
        ctx.launch_sharded(0, 8000)
        raise AssertionError("no error")
    except StoreError as e:
        assert "exceed the peer-exchange slot" in str(e), str(e)
    ctx.launch_sharded(1, 10)                  # and the context still works
    for a, b in zip(ctx.download_merged(), ref[(1, 10)]):
        assert np.array_equal(a, b)
    # a shard with no rows pushes its (empty) records through the stand-alone push kernel
    store.StoresDelete(K[:300])
    store.StoresSet(K[310:311], [b"one"])
    ops.stage_queries(Q)
    ctx.launch_sharded(0, 10)
    ids, sims, shard, index = ctx.download_merged()
    assert len(ids) == 1 and sims[0] == np.float32(ref_one(K[310], Q[0]))
    store.close()


def ref_one(k, q):
    """fp32 index-order dot (store.go:327-339)."""
    acc = np.float32(0)
    for a, b in zip(k, q):
        acc = np.float32(acc + np.float32(a * b))
    return acc


def run(rank, world, port, backend):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group(backend, rank=rank, world_size=world)
    try:
        from localai_tfp_b200.sharded import ShardedStore
        from oracle import oracle as orc
        if backend == "gloo":
            from sharded_double import CpuShardOps
            ops = CpuShardOps(rank)
        else:
            from localai_tfp_b200.sharded import CudaShardOps
            torch.cuda.set_device(rank)
            ops = CudaShardOps(rank, rank)
        store = ShardedStore(ops, rank, world)
        single = orc.OracleStore()
        scenario(store, single, rank)
        if backend == "nccl":
            native_path_matches(store, single, rank, world)
        store.close()
        if backend == "nccl":
            p2p_path_matches(single, rank, world)
        dist.barrier()
    finally:
        dist.destroy_process_group()
