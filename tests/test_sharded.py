"""Multi-shard store: routing, candidate all-gather, merge, result assembly.
CPU: world_size 2 over gloo with a CPU double of the device ops.  GPU: the merge kernel alone on
one GPU, and the full path over NCCL when the box has two GPUs."""
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import sharded_worker


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_owner_routing_is_deterministic_and_respects_key_equality():
    from localai_tfp_b200.sharded import owner_of
    rng = np.random.default_rng(0)
    K = rng.standard_normal((1000, 16)).astype(np.float32)
    a = owner_of(K, 8)
    assert np.array_equal(a, owner_of(K.copy(), 8))
    assert set(np.unique(a)) == set(range(8))
    assert np.bincount(a, minlength=8).min() > 60
    Z = K.copy(); Z[:, 3] = 0.0
    Zn = Z.copy(); Zn[:, 3] = -0.0
    assert np.array_equal(owner_of(Z, 8), owner_of(Zn, 8))   # -0 == +0 keys live on the same shard


def test_sharded_store_world2_gloo():
    mp.spawn(sharded_worker.run, args=(2, free_port(), "gloo"), nprocs=2, join=True)


@pytest.mark.gpu
def test_merge_kernel_matches_numpy(tfp):
    from localai_tfp_b200.sharded import pack_record, record_bytes
    from sharded_double import merge_records_numpy
    L = tfp.load()
    rng = np.random.default_rng(1)
    for world, k in ((2, 10), (8, 10), (8, 100), (3, 1), (8, 1024)):
        assert L.tfp_record_bytes(k) == record_bytes(k)
        sims = np.sort(rng.integers(-50, 50, size=(world, k)).astype(np.float32) / 50, axis=1)[:, ::-1].copy()  # ties
        ids = rng.integers(0, 1 << 40, size=(world, k)).astype(np.int64)
        n = rng.integers(0, k + 1, size=world).astype(np.int32)
        n[0] = k
        g = np.concatenate([pack_record(int(n[sh]), ids[sh], sims[sh], k) for sh in range(world)])
        want = merge_records_numpy(g, world, k)
        d = torch.from_numpy(g).cuda()
        o_sims = torch.empty(k, dtype=torch.float32, device="cuda"); o_ids = torch.empty(k, dtype=torch.int64, device="cuda")
        o_shard = torch.empty(k, dtype=torch.int32, device="cuda"); o_index = torch.empty(k, dtype=torch.int32, device="cuda")
        o_n = torch.empty(1, dtype=torch.int32, device="cuda")
        rc = L.tfp_merge_records(d.data_ptr(), world, 1, k, o_sims.data_ptr(), o_ids.data_ptr(), o_shard.data_ptr(),
                                 o_index.data_ptr(), o_n.data_ptr(), None)
        assert rc == 0, L.tfp_last_error()
        torch.cuda.synchronize()
        m = int(o_n.item())
        assert m == int(want[4].item())
        assert np.array_equal(o_sims.cpu().numpy()[:m], want[0].numpy()[:m])
        assert np.array_equal(o_ids.cpu().numpy()[:m], want[1].numpy()[:m])
        assert np.array_equal(o_shard.cpu().numpy()[:m], want[2].numpy()[:m])
        assert np.array_equal(o_index.cpu().numpy()[:m], want[3].numpy()[:m])


@pytest.mark.gpu
def test_sharded_store_world2_nccl(tfp):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(sharded_worker.run, args=(2, free_port(), "nccl"), nprocs=2, join=True)


@pytest.mark.gpu
def test_peer_exchange_argument_errors_leave_the_context_usable(tfp):
    """The peer-exchange set-up on ONE GPU: every misuse is a clean error (no hang, nothing half set up),
    and the find context keeps serving ordinary finds afterwards."""
    from conftest import make_keys
    K = make_keys(500, 32, seed=3, normalized=True)
    s = tfp.Store()
    s.StoresSet(K, [b""] * len(K))
    c = s.find_context()
    c.upload(K[:2])
    for world, slot, text in ((1, 1 << 16, "world = 1 outside [2, 8]"), (9, 1 << 16, "world = 9 outside [2, 8]"),
                              (2, 16, "slot_bytes outside [256, 64 MiB]")):
        with pytest.raises(tfp.StoreError) as e:
            c.p2p_export(world, slot)
        assert text in str(e.value)
    with pytest.raises(tfp.StoreError) as e:            # neither a peer exchange nor a communicator
        c.launch_sharded(0, 5)
    assert "tfp_find_ctx_p2p_attach or tfp_find_ctx_comm_init first" in str(e.value)
    with pytest.raises(tfp.StoreError) as e:            # attach before export
        c.p2p_attach(bytes(128), 0, 2)
    assert "tfp_find_ctx_p2p_export with the same world first" in str(e.value)
    handle = c.p2p_export(2, 1 << 16)
    assert len(handle) == 64
    with pytest.raises(RuntimeError) as e:              # rank 1's handle is garbage: the mapping must fail (a device-side
        c.p2p_attach(handle + bytes(64), 0, 2)          # failure, code < 0), not crash
    assert "cudaIpcOpenMemHandle" in str(e.value)
    with pytest.raises(tfp.StoreError):                 # still not attached
        c.launch_sharded(0, 5)
    with pytest.raises(tfp.StoreError) as e:
        c.launch_sharded_keys(0, 5)
    assert "tfp_find_ctx_p2p_attach first" in str(e.value)
    c.p2p_detach()
    c.launch(0, 2, 5)                                   # the context is intact
    res = c.download()
    assert res[0][0][0] == 0 and res[1][0][0] == 1
    c.close(); s.close()
