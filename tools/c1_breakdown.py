"""C1 (BASELINE.json configs[0]: 10k keys x 384-d, single query top-10, through gRPC): where the time of one StoresFind
goes, layer by layer -- the scan on the device, the C ABI call with host buffers, the Python host, the loopback gRPC
round trip with and without the key echo."""
import os, socket, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import grpc
import numpy as np
import localai_tfp_b200 as tfp
from localai_tfp_b200.server import Client

n, d, k, reps = 10_000, 384, 10, 2000
rng = np.random.default_rng(1)
K = rng.standard_normal((n, d)).astype(np.float32)
K /= np.linalg.norm(K, axis=1, keepdims=True)
Q = K[rng.choice(n, 64, replace=False)] + 0.01 * rng.standard_normal((64, d)).astype(np.float32)
Q = (Q / np.linalg.norm(Q, axis=1, keepdims=True)).astype(np.float32)
vals = [b"%d" % i for i in range(n)]


def per_call(fn):
    for i in range(50):
        fn(i)
    t0 = time.perf_counter()
    for i in range(reps):
        fn(i)
    return (time.perf_counter() - t0) / reps * 1e6


# ---- in process
s = tfp.Store()
s.StoresSet(K, vals)
c = s.find_context(); c.upload(Q); c.sync()


def dev_only(i):
    c.launch(i % 64, 1, k); c.sync()


us_dev = per_call(dev_only)
us_abi_ids = per_call(lambda i: s.find_ids(Q[i % 64], k))
us_abi_keys = per_call(lambda i: s.StoresFind(Q[i % 64], k))
c.close(); s.close()

# ---- over loopback gRPC, the way LocalAI's loader runs the backend
sock = socket.socket(); sock.bind(("127.0.0.1", 0)); port = sock.getsockname()[1]; sock.close()
addr = f"127.0.0.1:{port}"
proc = subprocess.Popen([sys.executable, "-m", "localai_tfp_b200.server", "--addr", addr])
cl = Client(addr)
for _ in range(120):
    try:
        if cl.health() == b"OK":
            break
    except grpc.RpcError:
        time.sleep(0.5)
cl.load_model()
cl.set_packed(K, vals)
us_grpc_keys = per_call(lambda i: cl.find(Q[i % 64], k))
us_grpc_lean = per_call(lambda i: cl.find(Q[i % 64], k, no_keys=True))
us_health = per_call(lambda i: cl.health())
reply_keys = len(cl._m["StoresFind"](__import__("localai_tfp_b200.wire", fromlist=["x"]).enc_find(Q[0], k)))
reply_lean = len(cl._m["StoresFind"](__import__("localai_tfp_b200.wire", fromlist=["x"]).enc_find(Q[0], k, True)))
cl.close(); proc.terminate(); proc.wait(timeout=20)

print(f"C1 {n} x {d}, top-{k}, one query at a time, microseconds per StoresFind ({reps} calls each):")
print(f"  scan + top-k on the device (query resident, launch + sync)     {us_dev:8.1f}")
print(f"  C ABI with host buffers, ids + similarities (tfp_store_find)     {us_abi_ids:8.1f}")
print(f"  C ABI with host buffers, + the k key rows (Store.StoresFind)     {us_abi_keys:8.1f}")
print(f"  loopback gRPC, reference wire format (keys echoed, {reply_keys} B reply) {us_grpc_keys:8.1f}")
print(f"  loopback gRPC, NoKeys extension ({reply_lean} B reply)               {us_grpc_lean:8.1f}")
print(f"  (an empty Health RPC on the same channel: the floor of the wire)  {us_health:8.1f}")
