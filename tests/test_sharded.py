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
