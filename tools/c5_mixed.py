"""C5 (BASELINE.json configs[4]): mixed stream on the GPUs of one box, as ONE process over ONE logical
store (tfp_store_open_devices) -- the shape LocalAI runs a backend in.

    python tools/c5_mixed.py [--seconds 30] [--devices all|0,1,..] [--writers 4] [--readers 16]

Concurrently, for `--seconds`:
  * writer threads append 1536-d keys in 8192-key StoresSet calls (host memory in, as a Go host would hand them
    over), as fast as the store takes them (target of the config: 1M keys/s);
  * one deleter thread removes a uniformly random 10 % of every batch written, a few batches behind the writers;
  * reader threads issue StoresFind top-10 continuously, each timing its own calls.
Prints sustained set/s, delete/s, find/s, p50 / p99 find latency and a spot check of the final state: sampled live
keys must be found (rank 1, value intact, similarities bit-identical to the CPU oracle's score of the returned rows),
sampled deleted keys must be gone.  The oracle is used as the checker only, after the timed run."""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ap = argparse.ArgumentParser()
ap.add_argument("--seconds", type=float, default=30.0)
ap.add_argument("--devices", default="all")
ap.add_argument("--writers", type=int, default=4)
ap.add_argument("--readers", type=int, default=16)
ap.add_argument("--dim", type=int, default=1536)
ap.add_argument("--batch", type=int, default=8192)
ap.add_argument("--delete-group", type=int, default=16, help="batches whose victims go into one StoresDelete call")
args = ap.parse_args()

import torch
import localai_tfp_b200 as tfp
from oracle import oracle as orc

D, B = args.dim, args.batch
DEL_GROUP = max(1, args.delete_group)
devs = list(range(torch.cuda.device_count())) if args.devices == "all" else [int(x) for x in args.devices.split(",")]
if len(devs) == 1:
    devs = devs * 2            # one GPU: two shards on it, so that the multi-GPU code path is the one that runs
store = tfp.Store(devices=devs)
L, H = store._L, store._h
import ctypes as C
fp, i64p, u8p = C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_uint8)

rng = np.random.default_rng(5)
POOL = 3                                   # blocks per writer, reused with a per-use tag in columns 0 and 1 (unique keys)
pools = [[np.ascontiguousarray(rng.standard_normal((B, D), dtype=np.float32)) for _ in range(POOL)] for _ in range(args.writers)]
Q = np.ascontiguousarray(rng.standard_normal((64, D), dtype=np.float32))

stop = threading.Event()
lock = threading.Lock()
written = []                               # (writer, pool block, tag, ids) of every batch, for the deleter and the check
counts = {"set": 0, "del": 0, "find": 0}
lat = []
errors = []


def batch_rows(w, j, tag, rows=None):
    """Batch `tag` (a running number) of writer w: pool block j with the tag spread over columns 0 and 1, 0.25 apart per
    step, so that two uses of the same block are clearly different keys (cosine ~0.998), never near-ties."""
    k = pools[w][j] if rows is None else pools[w][j][rows].copy()
    k[:, 0] = np.float32((tag % 256) * 0.25 - 32.0)
    k[:, 1] = np.float32((tag // 256) * 0.25 - 32.0)
    return k


def writer(w):
    try:
        step = 0
        ids = np.empty(B, np.int64); new = np.empty(B, np.uint8)
        while not stop.is_set():
            j = step % POOL
            tag = (step + 1) * args.writers + w                    # unique per (writer, step); 0 is the warm-up batch
            k = batch_rows(w, j, tag)                               # in place: the call copies its input before it returns
            rc = L.tfp_store_set(H, k.ctypes.data_as(fp), B, D, B, ids.ctypes.data_as(i64p), new.ctypes.data_as(u8p))
            store._check(rc)
            with lock:
                counts["set"] += B
                written.append((w, j, tag, ids.copy()))
            step += 1
    except Exception as e:      # noqa: BLE001
        errors.append(f"writer {w}: {e!r}")
        stop.set()


deleted = []                               # (writer, block, tag, rows) of what was removed


def deleter():
    """Removes a uniformly random 10 % of every batch written, staying two batches behind the writers.  A delete
    compacts the shards it touches and so fences the finds for a moment: the victims of all batches that have
    piled up meanwhile (DEL_GROUP of them) go into ONE StoresDelete call."""
    r = np.random.default_rng(9)
    done = 0
    try:
        while not stop.is_set():
            with lock:
                todo = written[done:max(done, len(written) - 2)][:DEL_GROUP]
            if len(todo) < DEL_GROUP and not stop.is_set():
                time.sleep(0.002)            # let a full group pile up: one fence per DEL_GROUP batches
                continue
            if not todo:
                break
            picks = [np.sort(r.choice(B, size=B // 10, replace=False)) for _ in todo]
            k = np.concatenate([batch_rows(w, j, tag, rows) for (w, j, tag, ids), rows in zip(todo, picks)])
            want = np.concatenate([ids[rows] for (w, j, tag, ids), rows in zip(todo, picks)])
            out = np.empty(len(k), np.int64)
            store._check(L.tfp_store_delete(H, k.ctypes.data_as(fp), len(k), D, out.ctypes.data_as(i64p)))
            if not np.array_equal(out, want):
                errors.append("deleter: the ids removed are not the ids the set calls handed out")
                stop.set()
            with lock:
                counts["del"] += len(k)
                for (w, j, tag, ids), rows in zip(todo, picks):
                    deleted.append((w, j, tag, rows))
            done += len(todo)
    except Exception as e:      # noqa: BLE001
        errors.append(f"deleter: {e!r}")
        stop.set()


def reader(t):
    n = C.c_int()
    ids = np.empty(10, np.int64); sims = np.empty(10, np.float32)
    mine = []
    i = t
    try:
        while not stop.is_set():
            q = Q[i % len(Q)]
            t0 = time.perf_counter()
            rc = L.tfp_store_find(H, q.ctypes.data, D, 10, C.byref(n), ids.ctypes.data, sims.ctypes.data, None, None)
            dt = time.perf_counter() - t0
            store._check(rc)
            mine.append(dt)
            i += 1
    except Exception as e:      # noqa: BLE001
        errors.append(f"reader {t}: {e!r}")
        stop.set()
    with lock:
        counts["find"] += len(mine)
        lat.extend(mine)


# warm-up: one batch so that the key length is latched and every buffer exists before the clock starts
k0 = batch_rows(0, 0, 0)
ids0 = np.empty(B, np.int64); new0 = np.empty(B, np.uint8)
store._check(L.tfp_store_set(H, k0.ctypes.data_as(fp), B, D, B, ids0.ctypes.data_as(i64p), new0.ctypes.data_as(u8p)))
store.find_ids(Q[0], 10)

threads = [threading.Thread(target=writer, args=(w,)) for w in range(args.writers)] + [threading.Thread(target=deleter)] + \
          [threading.Thread(target=reader, args=(t,)) for t in range(args.readers)]
t0 = time.perf_counter()
[t.start() for t in threads]
time.sleep(args.seconds)
stop.set()
[t.join() for t in threads]
elapsed = time.perf_counter() - t0

# ---- spot check of the final state against the oracle (outside the timed run)
problems = list(errors)
r = np.random.default_rng(11)
gone = {(w, j, tag): set() for (w, j, tag, rows) in deleted}
for (w, j, tag, rows) in deleted:
    gone[(w, j, tag)].update(rows.tolist())
st = store.stats()
expect_rows = B + counts["set"] - counts["del"]
if int(st.rows) != expect_rows:
    problems.append(f"row count {int(st.rows)} != written - deleted = {expect_rows}")
checked_live = checked_gone = 0
for pick in r.choice(len(written), size=min(8, len(written)), replace=False):
    w, j, tag, ids = written[int(pick)]
    dead = gone.get((w, j, tag), set())
    live_rows = [x for x in r.choice(B, size=32, replace=False).tolist() if x not in dead][:2]
    for row in live_rows:
        key = batch_rows(w, j, tag, np.array([row]))[0]
        res, rows_out = store.find_batch(key[None, :], 10, with_keys=True)
        fids, fsims = res[0]
        if len(fids) == 0 or int(fids[0]) != int(ids[row]):
            problems.append(f"live key (writer {w}, tag {tag}, row {row}) is not its own nearest neighbour")
        sc = orc.scores_block(rows_out[0, :len(fids)], key, False)          # the store left the normalised path: cosine arithmetic
        if not np.array_equal(sc, fsims):
            problems.append("a returned similarity differs from the oracle's score of the returned row")
        checked_live += 1
    for row in list(dead)[:2]:
        key = batch_rows(w, j, tag, np.array([row]))
        out = np.empty(1, np.int64); order = np.empty(1, np.int64); nf = C.c_int64()
        store._check(L.tfp_store_get(H, key.ctypes.data_as(fp), 1, D, out.ctypes.data_as(i64p), order.ctypes.data_as(i64p), C.byref(nf)))
        if nf.value != 0:
            problems.append(f"deleted key (writer {w}, tag {tag}, row {row}) is still stored")
        checked_gone += 1

lat_ms = np.sort(np.array(lat)) * 1e3 if lat else np.array([0.0])
line = {
    "config": "C5: mixed stream, %d-d keys, one process, one logical store over GPUs %s" % (D, devs),
    "seconds": round(elapsed, 2), "writers": args.writers, "readers": args.readers, "set_batch": B,
    "set_per_s": counts["set"] / elapsed, "delete_per_s": counts["del"] / elapsed, "find_per_s": counts["find"] / elapsed,
    "find_p50_ms": float(lat_ms[len(lat_ms) // 2]), "find_p99_ms": float(lat_ms[min(len(lat_ms) - 1, int(len(lat_ms) * 0.99))]),
    "rows_at_end": int(st.rows), "hbm_gb_at_end": round(int(st.hbm_bytes) / 1e9, 1), "path": "cosine (keys are not unit length)",
    "spot_check": {"ok": not problems, "live_keys_checked": checked_live, "deleted_keys_checked": checked_gone,
                   "rule": "a live key is its own nearest neighbour with its id, similarities bit-identical to the oracle; a deleted key is gone"},
}
if problems:
    line["spot_check"]["problems"] = problems[:8]
print(json.dumps(line), flush=True)
store.close()
sys.exit(0 if not problems else 1)
