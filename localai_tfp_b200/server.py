"""The `local-store` backend process, speaking LocalAI's gRPC contract over the B200 store.

    python -m localai_tfp_b200.server --addr 127.0.0.1:50051

Stand-in for the Go binary of go/local-store/ (the build image has no Go toolchain): same flag
(`--addr`, backend/go/stores/main.go:14-16), same service and method paths
(`/backend.Backend/StoresSet` ..., backend/backend.proto:10-34), same replies as the reference's
gRPC shell (pkg/grpc/server.go:45-55, 185-240): Health -> Reply{message: "OK"}, LoadModel ->
Result{success: true}, StoresSet/StoresDelete -> Result{"Set key"/"Deleted key", true} or an error
whose text is the store's, StoresGet/StoresFind -> their result messages, errors as gRPC UNKNOWN.
The reference serialises every RPC with one mutex (base.SingleThread).  Here, like the Go host
(base.Base, no server mutex), only the mutations are serialised among themselves: finds are drained by the
batcher and run beside a StoresSet -- the library appends without stopping them.
"""
from __future__ import annotations

import argparse
import os
import queue
import signal
import sys
import threading
from concurrent import futures

import grpc

from . import wire
from .store import Store, StoreError

SERVICE = "backend.Backend"
STATE_READY, STATE_BUSY = 2, 1


def _ident(b):
    return b


class _FindBatcher:
    """Concurrent StoresFind requests become ONE batched find ("micro-batcher for concurrent
    StoresFind").  Whatever has queued up while the previous batch ran is taken together, grouped by
    top-k; 3 or more same-path queries on a large store ride the tensor-core path.  A lone request is
    served at once: nothing waits for company."""

    def __init__(self, backend: "Backend", max_batch: int = 256):
        self.backend = backend
        self.max_batch = max_batch
        self.q = queue.Queue()
        self.gate = threading.Lock()      # held while a backlog is collected and served (a test holds it to let one pile up)
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        self.batches = 0
        self.largest = 0

    def submit(self, key, topk, no_keys=False):
        fut = futures.Future()
        self.q.put((key, (topk, bool(no_keys)), fut))
        return fut

    def _run(self):
        while True:
            items = [self.q.get()]
            with self.gate:
                while len(items) < self.max_batch:
                    try:
                        items.append(self.q.get_nowait())
                    except queue.Empty:
                        break
                by_k = {}
                for it in items:
                    by_k.setdefault(it[1], []).append(it)
                for topk, group in by_k.items():
                    self._serve(topk, group)

    def _serve(self, topk_nokeys, group):
        """One group of finds (same top-k).  No server mutex: finds run beside the mutations."""
        b = self.backend
        topk, no_keys = topk_nokeys
        b.busy = True
        try:
            dims = {len(g[0]) for g in group}
            if len(group) == 1 or len(dims) != 1 or topk < 1 or b.store.key_len != next(iter(dims)) or len(b.store) == 0:
                for key, _, fut in group:        # singles and anything that will error: the plain call, per request
                    try:
                        fut.set_result(b.find_one(key, topk, no_keys))
                    except Exception as e:
                        fut.set_exception(e)
                return
            import numpy as np
            qs = np.stack([g[0] for g in group]).astype(np.float32)
            if no_keys:
                res, keys = b.store.find_batch(qs, topk), None
            else:
                res, keys = b.store.find_batch(qs, topk, with_keys=True)
            self.batches += 1
            self.largest = max(self.largest, len(group))
            for i, (key, k, fut) in enumerate(group):
                ids, sims = res[i]
                fut.set_result(([] if no_keys else [keys[i, j, :].copy() for j in range(len(ids))],
                                b.store._values_of([int(x) for x in ids]), sims))
        except Exception as e:
            for _, _, fut in group:
                if not fut.done():
                    fut.set_exception(e)
        finally:
            b.busy = False


def devices_from_env(default_device: int = 0):
    """The GPUs this backend spreads its store over.  LocalAI's loader gives a backend process only `--addr` and
    the inherited environment (pkg/model/process.go:112-113), so the device list travels in TFP_DEVICES:
    "0,1,2,3", or "all" for every visible GPU.  Unset: one GPU (`--device`)."""
    spec = os.environ.get("TFP_DEVICES", "").strip()
    if not spec:
        return None
    if spec.lower() == "all":
        import ctypes
        n = ctypes.c_int(0)
        try:
            ctypes.CDLL("libcudart.so").cudaGetDeviceCount(ctypes.byref(n))
        except OSError:
            import torch
            n = ctypes.c_int(torch.cuda.device_count())
        return list(range(max(n.value, 1)))
    return [int(x) for x in spec.split(",") if x.strip() != ""] or [default_device]


class Backend:
    def __init__(self, device: int = 0, snapshot: str | None = None, devices=None):
        self._open = (lambda: Store(devices=devices)) if devices and len(devices) > 1 else \
                     (lambda: Store(device=devices[0] if devices else device))
        self.store = self._open()
        self.snapshot = snapshot
        self.batcher = None
        if snapshot and os.path.exists(snapshot):
            try:
                self.store.load(snapshot)         # survive a watchdog kill / respawn
            except Exception as e:                # a damaged or mismatched snapshot must not crash-loop the backend
                print(f"local-store: snapshot {snapshot} not restored ({e}); starting empty", file=sys.stderr, flush=True)
                self.store.close()
                self.store = self._open()
        self.lock = threading.Lock()          # mutations (and un-batched finds) one at a time; batched finds do not take it
        self.busy = False

    # -- lifecycle ---------------------------------------------------------------------------
    def Health(self, request: bytes, context) -> bytes:
        return wire.enc_reply(b"OK")                                   # pkg/grpc/server.go:45-47

    def LoadModel(self, request: bytes, context) -> bytes:
        return wire.enc_result("Loading succeeded", True)              # store.go:101-103 Load is a no-op

    def Status(self, request: bytes, context) -> bytes:
        st = self.store.stats()
        state = STATE_BUSY if self.busy else STATE_READY
        breakdown = {"rows": int(st.rows)}
        for g in range(self.store.shard_count):                        # one entry per GPU (base.go:77-81 memory breakdown)
            sg = self.store.shard_stats(g)
            key = f"hbm-bytes-gpu{sg.device}"
            breakdown[key] = breakdown.get(key, 0) + int(sg.hbm_bytes)
        return wire.enc_status(state, int(st.hbm_bytes), breakdown)

    # -- the four store RPCs -------------------------------------------------------------------
    def _run(self, context, fn):
        with self.lock:
            self.busy = True
            try:
                return fn()
            except StoreError as e:                                    # err != nil -> gRPC status Unknown
                context.abort(grpc.StatusCode.UNKNOWN, str(e))
            finally:
                self.busy = False

    def StoresSet(self, request: bytes, context) -> bytes:
        keys, values = wire.dec_set(request)
        self._run(context, lambda: self.store.StoresSet(keys, values))
        return wire.enc_result("Set key", True)                        # pkg/grpc/server.go:203

    def StoresDelete(self, request: bytes, context) -> bytes:
        keys = wire.dec_keys(request)
        self._run(context, lambda: self.store.StoresDelete(keys))
        return wire.enc_result("Deleted key", True)                    # pkg/grpc/server.go:215

    def StoresGet(self, request: bytes, context) -> bytes:
        keys = wire.dec_keys(request)
        ks, vs = self._run(context, lambda: self.store.StoresGet(keys))
        return wire.enc_get_result(ks, vs)

    def StoresSetPacked(self, request: bytes, context) -> bytes:
        """Extension: the keys of one call as ONE block of n x Dim floats -- handed to the library as it
        arrived, no per-key message, no gather copy.  Same semantics and reply as StoresSet."""
        try:
            keys, values = wire.dec_set_packed(request)
        except ValueError as e:
            context.abort(grpc.StatusCode.UNKNOWN, str(e))
        self._run(context, lambda: self.store.StoresSet(keys, values))
        return wire.enc_result("Set key", True)

    def find_one(self, key, topk, no_keys):
        if not no_keys:
            return self.store.StoresFind(key, topk)
        ids, sims = self.store.find_ids(key, topk)        # no key rows gathered, copied or encoded
        return [], self.store._values_of([int(x) for x in ids]), sims

    def StoresFind(self, request: bytes, context) -> bytes:
        key, topk, no_keys = wire.dec_find_ext(request)
        if self.batcher is None:
            ks, vs, sims = self._run(context, lambda: self.find_one(key, topk, no_keys))
        else:
            try:
                ks, vs, sims = self.batcher.submit(key, topk, no_keys).result()
            except StoreError as e:
                context.abort(grpc.StatusCode.UNKNOWN, str(e))
        return wire.enc_find_result(ks, vs, sims)

    def StoresFindBatch(self, request: bytes, context) -> bytes:
        """Extension: many queries, one top-k, one reply.  Result i is exactly what
        StoresFind(Keys[i], TopK) returns, errors included (the first failing key aborts the call)."""
        keys, topk = wire.dec_find_batch(request)

        def serve():
            import numpy as np
            st = self.store
            if not keys:
                raise StoreError("no keys to find")
            for k in keys:                                             # store.go:480-486, per key
                if len(k) != st.key_len:
                    raise StoreError(f"Try to find key with length {len(k)} when existing length is {st.key_len}")
            if topk < 1:
                raise StoreError(f"opts.TopK = {topk}, must be >= 1")
            if len(st) == 0:
                return [([], [], np.zeros(0, np.float32)) for _ in keys]
            out = []
            for lo in range(0, len(keys), 256):                        # one library call per 256 queries
                qs = np.stack(keys[lo:lo + 256]).astype(np.float32)
                res, rows = st.find_batch(qs, topk, with_keys=True)
                for i, (ids, sims) in enumerate(res):
                    out.append(([rows[i, j, :].copy() for j in range(len(ids))],
                                st._values_of([int(x) for x in ids]), sims))
            return out
        return wire.enc_find_batch_result(self._run(context, serve))


def make_server(addr: str, device: int = 0, workers: int = 64, snapshot: str | None = None, batch_finds: bool = True,
                devices=None):
    backend = Backend(device, snapshot, devices if devices is not None else devices_from_env(device))
    if batch_finds:
        backend.batcher = _FindBatcher(backend)
    methods = {name: grpc.unary_unary_rpc_method_handler(getattr(backend, name), request_deserializer=_ident,
                                                         response_serializer=_ident)
               for name in ("Health", "LoadModel", "Status", "StoresSet", "StoresDelete", "StoresGet", "StoresFind",
                            "StoresFindBatch", "StoresSetPacked")}
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=workers),
                         options=[("grpc.max_receive_message_length", wire.MAX_MESSAGE),
                                  ("grpc.max_send_message_length", wire.MAX_MESSAGE)])
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, methods),))
    port = server.add_insecure_port(addr)
    if port == 0:
        raise RuntimeError(f"cannot listen on {addr}")
    return server, backend, port


class Client:
    """What pkg/grpc/client.go + pkg/store/client.go do, for tests and the C1 measurement."""

    def __init__(self, addr: str):
        self.ch = grpc.insecure_channel(addr, options=[("grpc.max_receive_message_length", wire.MAX_MESSAGE),
                                                       ("grpc.max_send_message_length", wire.MAX_MESSAGE)])
        self._m = {name: self.ch.unary_unary(f"/{SERVICE}/{name}", request_serializer=_ident,
                                             response_deserializer=_ident)
                   for name in ("Health", "LoadModel", "Status", "StoresSet", "StoresDelete", "StoresGet", "StoresFind",
                                "StoresFindBatch", "StoresSetPacked")}

    def health(self) -> bytes:
        return wire.dec_reply(self._m["Health"](b"", timeout=10))

    def load_model(self):
        return wire.dec_result(self._m["LoadModel"](b""))

    def status(self):
        return wire.dec_status(self._m["Status"](b""))

    def set(self, keys, values):
        return wire.dec_result(self._m["StoresSet"](wire.enc_set(keys, values)))

    def delete(self, keys):
        return wire.dec_result(self._m["StoresDelete"](wire.enc_keys(keys)))

    def get(self, keys):
        return wire.dec_get_result(self._m["StoresGet"](wire.enc_keys(keys)))

    def set_packed(self, keys, values):
        return wire.dec_result(self._m["StoresSetPacked"](wire.enc_set_packed(keys, values)))

    def find(self, key, topk: int, no_keys: bool = False):
        return wire.dec_find_result(self._m["StoresFind"](wire.enc_find(key, topk, no_keys)))

    def find_batch(self, keys, topk: int):
        return wire.dec_find_batch_result(self._m["StoresFindBatch"](wire.enc_find_batch(keys, topk)))

    def close(self):
        self.ch.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--addr", default="localhost:50051", help="the address to connect to")   # main.go:14-16
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--snapshot", default=os.environ.get("TFP_STORE_SNAPSHOT"),
                    help="file to restore the store from at start and to save it to on SIGTERM")
    args = ap.parse_args()
    server, backend, _ = make_server(args.addr, args.device, snapshot=args.snapshot)

    def on_term(signum, frame):
        if backend.snapshot:
            with backend.lock:
                backend.store.save(backend.snapshot)
        server.stop(grace=1)

    signal.signal(signal.SIGTERM, on_term)
    server.start()
    server.wait_for_termination()


if __name__ == "__main__":
    main()
