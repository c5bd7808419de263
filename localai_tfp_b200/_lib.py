"""ctypes binding of libtfpstore.so -- one prototype per entry point of include/tfp_store.h."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libtfpstore.so")


class TfpCfg(C.Structure):
    _fields_ = [("device", C.c_int), ("capacity_rows", C.c_int64), ("flags", C.c_uint32), ("reserved", C.c_uint32),
                ("id_base", C.c_int64)]


class TfpStats(C.Structure):
    _fields_ = [("rows", C.c_int64), ("capacity_rows", C.c_int64), ("hbm_bytes", C.c_int64), ("dim", C.c_int),
                ("keys_are_normalized", C.c_int), ("device", C.c_int), ("reserved", C.c_int)]


fp = C.POINTER(C.c_float)
i64p = C.POINTER(C.c_int64)
i32p = C.POINTER(C.c_int)
u8p = C.POINTER(C.c_uint8)
vp = C.c_void_p

# name -> (restype, argtypes); kept in the order of include/tfp_store.h
PROTOTYPES = {
    "tfp_store_open": (C.c_int, [C.POINTER(TfpCfg), C.POINTER(vp)]),
    "tfp_store_close": (None, [vp]),
    "tfp_store_open_devices": (C.c_int, [C.POINTER(TfpCfg), i32p, C.c_int, C.POINTER(vp)]),
    "tfp_store_shard_count": (C.c_int, [vp]),
    "tfp_route_rows": (C.c_int, [fp, C.c_int64, C.c_int, C.c_int, i32p]),
    "tfp_last_error": (C.c_char_p, []),
    "tfp_store_last_error": (C.c_int, [vp, C.c_char_p, C.c_size_t]),
    "tfp_store_set": (C.c_int, [vp, fp, C.c_int64, C.c_int, C.c_int64, i64p, u8p]),
    "tfp_store_set_device": (C.c_int, [vp, vp, C.c_int64, C.c_int, i64p, u8p]),
    "tfp_store_delete": (C.c_int, [vp, fp, C.c_int64, C.c_int, i64p]),
    "tfp_store_get": (C.c_int, [vp, fp, C.c_int64, C.c_int, i64p, i64p, i64p]),
    "tfp_store_find": (C.c_int, [vp, vp, C.c_int, C.c_int, i32p, vp, vp, vp, i32p]),   # array arguments as plain addresses: the per-call path
    "tfp_store_find_batch": (C.c_int, [vp, fp, C.c_int, C.c_int, C.c_int, i32p, i64p, fp, fp]),
    "tfp_store_stats": (C.c_int, [vp, C.POINTER(TfpStats)]),
    "tfp_store_shard_stats": (C.c_int, [vp, C.c_int, C.POINTER(TfpStats)]),
    "tfp_store_save": (C.c_int, [vp, C.c_char_p]),
    "tfp_store_load": (C.c_int, [vp, C.c_char_p]),
    "tfp_store_save_tagged": (C.c_int, [vp, C.c_char_p, C.c_uint64]),
    "tfp_snapshot_tag": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint64)]),
    "tfp_store_mark_not_normalized": (C.c_int, [vp]),
    "tfp_store_default_ctx": (vp, [vp]),
    "tfp_find_ctx_create": (C.c_int, [vp, C.POINTER(vp)]),
    "tfp_find_ctx_destroy": (None, [vp]),
    "tfp_find_ctx_upload": (C.c_int, [vp, vp, C.c_int, C.c_int]),                       # per-query path: plain addresses
    "tfp_find_ctx_launch": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
    "tfp_find_ctx_download": (C.c_int, [vp, i32p, i64p, fp, fp]),
    "tfp_find_ctx_sync": (C.c_int, [vp]),
    "tfp_find_ctx_stream": (vp, [vp]),
    "tfp_find_ctx_device_results": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "tfp_find_ctx_gather_keys": (C.c_int, [vp, C.POINTER(vp)]),
    "tfp_find_ctx_mma_counters": (C.c_int, [vp, i64p, i64p]),
    "tfp_find_ctx_mma_escalated": (C.c_int64, [vp]),
    "tfp_find_ctx_launch_count": (C.c_int64, [vp]),
    "tfp_record_bytes": (C.c_size_t, [C.c_int]),
    "tfp_find_ctx_device_records": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]),
    "tfp_merge_records": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]),
    "tfp_nccl_unique_id": (C.c_int, [vp, C.c_size_t]),
    "tfp_find_ctx_comm_init": (C.c_int, [vp, vp, C.c_int, C.c_int]),
    "tfp_find_ctx_comm_share": (C.c_int, [vp, vp]),
    "tfp_find_ctx_launch_sharded": (C.c_int, [vp, C.c_int, C.c_int]),
    "tfp_find_ctx_launch_sharded_batch": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
    "tfp_find_ctx_merged_results": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "tfp_find_ctx_download_merged": (C.c_int, [vp, i32p, i64p, fp, i32p, i32p]),
    "tfp_find_ctx_p2p_export": (C.c_int, [vp, C.c_int, C.c_size_t, vp, C.c_size_t]),
    "tfp_find_ctx_p2p_attach": (C.c_int, [vp, vp, C.c_int, C.c_int]),
    "tfp_find_ctx_p2p_detach": (C.c_int, [vp]),
    "tfp_p2p_slot_bytes_with_keys": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "tfp_find_ctx_launch_sharded_keys": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
    "tfp_find_ctx_download_merged_keys": (C.c_int, [vp, vp, vp, vp, vp]),             # per-query path: plain addresses
    "tfp_host_alloc": (vp, [C.c_size_t]),
    "tfp_host_free": (None, [vp]),
    "tfp_scan_topk_noalloc": (C.c_int, [vp, C.c_int64, C.c_int, vp, vp, C.c_int, vp, vp, vp, vp]),
    "tfp_version": (C.c_char_p, []),
}

_lib = None


def lib_path() -> str:
    return _LIB


def load():
    """dlopen libtfpstore.so.  Raises if it has not been built: there is nothing to fall back to."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            raise RuntimeError(
                f"{_LIB} is missing: build it with `python -m localai_tfp_b200.build` "
                "(or __graft_entry__.build()); the store has no CPU fallback")
        L = C.CDLL(_LIB)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError here == the library does not match the header
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
