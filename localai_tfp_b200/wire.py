"""Protobuf wire codec for the store part of LocalAI's gRPC contract, without protoc.

Messages follow /root/reference/backend/backend.proto (proto3):
    StoresKey :69-71, StoresValue :73-75, StoresSetOptions :77-80, StoresDeleteOptions :82-84,
    StoresGetOptions :86-88, StoresGetResult :90-93, StoresFindOptions :95-98,
    StoresFindResult :100-104, Result :260-263, Reply :159-165, StatusResponse :360-369.
One extension the reference does not have ("batch Find on the wire"):
    message StoresFindBatchOptions { repeated StoresKey Keys = 1; int32 TopK = 2; }
    message StoresFindBatchResult  { repeated StoresFindResult Results = 1; }
    rpc StoresFindBatch(StoresFindBatchOptions) returns (StoresFindBatchResult) {}
Result i is what StoresFind(Keys[i], TopK) returns; one client can feed the tensor-core path directly.
Two more (no key echo, no per-key sub-messages):
    StoresFindOptions  { ...; bool NoKeys = 3; }      a decoder that predates the field skips it (proto3), so the
        request stays valid for the reference backend; this backend then leaves StoresFindResult.Keys empty --
        Values and Similarities identify the hits -- and the reply shrinks from k x (D x 4 + value) to k x (4 + value) bytes
        (core/http/endpoints/localai/stores.go:108-125 echoes the keys only because the backend returns them).
    message StoresSetPackedOptions { bytes Keys = 1; int32 Dim = 2; repeated StoresValue Values = 3; }
    rpc StoresSetPacked(StoresSetPackedOptions) returns (Result) {}
        Keys is n x Dim little-endian fp32 in one run: one memcpy instead of n sub-messages on both sides.
`repeated float` is packed (one length-delimited run of little-endian fp32), which numpy reads and
writes without a per-element loop; unpacked floats (legal on the wire) are accepted too.
Unknown fields are skipped, as a protobuf decoder must.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

MAX_MESSAGE = 50 * 1024 * 1024  # pkg/grpc/server.go:259-262


# ---- primitives -----------------------------------------------------------------------------
def _varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64  # int32/int64 negatives are 10-byte two's complement varints
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: memoryview, pos: int) -> Tuple[int, int]:
    shift = 0
    val = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _ld(field: int, payload: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def _fields(buf: memoryview):
    """Yield (field number, wire type, value) for every field; value is int or memoryview."""
    pos = 0
    n = len(buf)
    while pos < n:
        tag, pos = _read_varint(buf, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]; pos += 8
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            v = buf[pos:pos + ln]; pos += ln
            if len(v) != ln:
                raise ValueError("truncated length-delimited field")
        elif wt == 5:
            v = buf[pos:pos + 4]; pos += 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield field, wt, v


# ---- StoresKey / StoresValue ----------------------------------------------------------------
def enc_key(floats) -> bytes:
    a = np.ascontiguousarray(floats, dtype="<f4")
    return _ld(1, a.tobytes()) if a.size else b""


def dec_key(buf: memoryview) -> np.ndarray:
    parts = []
    for field, wt, v in _fields(buf):
        if field == 1 and wt == 2:
            parts.append(np.frombuffer(v, dtype="<f4"))
        elif field == 1 and wt == 5:
            parts.append(np.frombuffer(v, dtype="<f4"))
    if not parts:
        return np.zeros(0, np.float32)
    return parts[0].astype(np.float32) if len(parts) == 1 else np.concatenate(parts).astype(np.float32)


def enc_value(b: bytes) -> bytes:
    return _ld(1, bytes(b)) if b else b""


def dec_value(buf: memoryview) -> bytes:
    out = b""
    for field, wt, v in _fields(buf):
        if field == 1 and wt == 2:
            out = bytes(v)
    return out


# ---- requests -------------------------------------------------------------------------------
def enc_set(keys, values) -> bytes:
    return b"".join(_ld(1, enc_key(k)) for k in keys) + b"".join(_ld(2, enc_value(v)) for v in values)


def dec_set(data: bytes) -> Tuple[List[np.ndarray], List[bytes]]:
    keys, values = [], []
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            keys.append(dec_key(v))
        elif field == 2 and wt == 2:
            values.append(dec_value(v))
    return keys, values


def enc_keys(keys) -> bytes:           # StoresDeleteOptions / StoresGetOptions
    return b"".join(_ld(1, enc_key(k)) for k in keys)


def dec_keys(data: bytes) -> List[np.ndarray]:
    return [dec_key(v) for field, wt, v in _fields(memoryview(data)) if field == 1 and wt == 2]


def enc_find(key, topk: int, no_keys: bool = False) -> bytes:
    out = _ld(1, enc_key(key))
    if topk:
        out += _varint((2 << 3) | 0) + _varint(int(topk))
    if no_keys:
        out += _varint((3 << 3) | 0) + b"\x01"
    return out


def dec_find_ext(data: bytes) -> Tuple[np.ndarray, int, bool]:
    key, topk, no_keys = np.zeros(0, np.float32), 0, False
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            key = dec_key(v)
        elif field == 2 and wt == 0:
            topk = v & 0xFFFFFFFF
            if topk & 0x80000000:
                topk -= 1 << 32      # int32
        elif field == 3 and wt == 0:
            no_keys = bool(v)
    return key, topk, no_keys


def dec_find(data: bytes) -> Tuple[np.ndarray, int]:
    return dec_find_ext(data)[:2]


def enc_set_packed(keys, values) -> bytes:
    a = np.ascontiguousarray(keys, dtype="<f4")
    if a.ndim != 2:
        raise ValueError("packed keys are one [n, dim] block")
    out = _ld(1, a.tobytes()) if a.size else b""
    if a.shape[1]:
        out += _varint((2 << 3) | 0) + _varint(int(a.shape[1]))
    return out + b"".join(_ld(3, enc_value(v)) for v in values)


def dec_set_packed(data: bytes):
    """(keys [n, dim] fp32 -- a view of the message's bytes, no copy --, values)."""
    raw, dim, values = memoryview(b""), 0, []
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            raw = v
        elif field == 2 and wt == 0:
            dim = v & 0xFFFFFFFF
        elif field == 3 and wt == 2:
            values.append(dec_value(v))
    if dim <= 0 or len(raw) % (4 * dim):
        raise ValueError(f"packed keys: {len(raw)} bytes is not a whole number of {dim}-float keys")
    return np.frombuffer(raw, dtype="<f4").reshape(-1, dim), values


def enc_find_batch(keys, topk: int) -> bytes:
    out = b"".join(_ld(1, enc_key(k)) for k in keys)
    if topk:
        out += _varint((2 << 3) | 0) + _varint(int(topk))
    return out


def dec_find_batch(data: bytes) -> Tuple[List[np.ndarray], int]:
    keys, topk = [], 0
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            keys.append(dec_key(v))
        elif field == 2 and wt == 0:
            topk = v & 0xFFFFFFFF
            if topk & 0x80000000:
                topk -= 1 << 32      # int32
    return keys, topk


# ---- replies --------------------------------------------------------------------------------
def enc_get_result(keys, values) -> bytes:
    return enc_set(keys, values)        # same field layout (Keys = 1, Values = 2)


dec_get_result = dec_set


def enc_find_result(keys, values, sims) -> bytes:
    out = enc_set(keys, values)
    a = np.ascontiguousarray(sims, dtype="<f4")
    if a.size:
        out += _ld(3, a.tobytes())
    return out


def dec_find_result(data: bytes):
    keys, values, sims = [], [], []
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            keys.append(dec_key(v))
        elif field == 2 and wt == 2:
            values.append(dec_value(v))
        elif field == 3 and wt == 2:
            sims.append(np.frombuffer(v, dtype="<f4"))
        elif field == 3 and wt == 5:
            sims.append(np.frombuffer(v, dtype="<f4"))
    s = np.concatenate(sims).astype(np.float32) if sims else np.zeros(0, np.float32)
    return keys, values, s


def enc_find_batch_result(results) -> bytes:
    """results: one (keys, values, sims) triple per query, in request order."""
    return b"".join(_ld(1, enc_find_result(k, v, s)) for k, v, s in results)


def dec_find_batch_result(data: bytes):
    return [dec_find_result(bytes(v)) for field, wt, v in _fields(memoryview(data)) if field == 1 and wt == 2]


def enc_result(message: str, success: bool) -> bytes:
    out = _ld(1, message.encode()) if message else b""
    if success:
        out += _varint((2 << 3) | 0) + b"\x01"
    return out


def dec_result(data: bytes) -> Tuple[str, bool]:
    msg, ok = "", False
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            msg = bytes(v).decode(errors="replace")
        elif field == 2 and wt == 0:
            ok = bool(v)
    return msg, ok


def enc_reply(message: bytes) -> bytes:
    return _ld(1, message) if message else b""


def dec_reply(data: bytes) -> bytes:
    out = b""
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 2:
            out = bytes(v)
    return out


def enc_status(state: int, total: int, breakdown: dict) -> bytes:
    mem = b""
    if total:
        mem += _varint((1 << 3) | 0) + _varint(total)
    for k, v in breakdown.items():      # map<string, uint64> == repeated entry {key = 1, value = 2}
        entry = _ld(1, k.encode()) + _varint((2 << 3) | 0) + _varint(int(v))
        mem += _ld(2, entry)
    out = b""
    if state:
        out += _varint((1 << 3) | 0) + _varint(state)
    return out + _ld(2, mem)


def dec_status(data: bytes):
    state, total, breakdown = 0, 0, {}
    for field, wt, v in _fields(memoryview(data)):
        if field == 1 and wt == 0:
            state = v
        elif field == 2 and wt == 2:
            for f2, w2, v2 in _fields(v):
                if f2 == 1 and w2 == 0:
                    total = v2
                elif f2 == 2 and w2 == 2:
                    k, val = "", 0
                    for f3, w3, v3 in _fields(v2):
                        if f3 == 1 and w3 == 2:
                            k = bytes(v3).decode()
                        elif f3 == 2 and w3 == 0:
                            val = v3
                    breakdown[k] = val
    return state, total, breakdown
