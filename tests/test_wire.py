"""The hand-written protobuf codec (localai_tfp_b200/wire.py) against the real protobuf runtime.
The message classes are built at run time from a descriptor that restates the store part of
/root/reference/backend/backend.proto (no protoc in the image), so both sides must agree byte for
byte on encode and value for value on decode."""
import numpy as np
import pytest
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from localai_tfp_b200 import wire

F = descriptor_pb2.FieldDescriptorProto


def build_messages():
    fd = descriptor_pb2.FileDescriptorProto(name="tfp_backend_subset.proto", package="backend", syntax="proto3")

    def msg(name, *fields):
        m = fd.message_type.add(name=name)
        for fname, num, ftype, label, tname in fields:
            f = m.field.add(name=fname, number=num, type=ftype, label=label)
            if tname:
                f.type_name = ".backend." + tname
        return m

    R, O = F.LABEL_REPEATED, F.LABEL_OPTIONAL
    msg("StoresKey", ("Floats", 1, F.TYPE_FLOAT, R, None))                       # backend.proto:69-71
    msg("StoresValue", ("Bytes", 1, F.TYPE_BYTES, O, None))                      # :73-75
    msg("StoresSetOptions", ("Keys", 1, F.TYPE_MESSAGE, R, "StoresKey"), ("Values", 2, F.TYPE_MESSAGE, R, "StoresValue"))
    msg("StoresDeleteOptions", ("Keys", 1, F.TYPE_MESSAGE, R, "StoresKey"))
    msg("StoresGetOptions", ("Keys", 1, F.TYPE_MESSAGE, R, "StoresKey"))
    msg("StoresGetResult", ("Keys", 1, F.TYPE_MESSAGE, R, "StoresKey"), ("Values", 2, F.TYPE_MESSAGE, R, "StoresValue"))
    msg("StoresFindOptions", ("Key", 1, F.TYPE_MESSAGE, O, "StoresKey"), ("TopK", 2, F.TYPE_INT32, O, None))
    msg("StoresFindResult", ("Keys", 1, F.TYPE_MESSAGE, R, "StoresKey"), ("Values", 2, F.TYPE_MESSAGE, R, "StoresValue"),
        ("Similarities", 3, F.TYPE_FLOAT, R, None))
    # the batch-find extension (not in the reference proto)
    msg("StoresFindBatchOptions", ("Keys", 1, F.TYPE_MESSAGE, R, "StoresKey"), ("TopK", 2, F.TYPE_INT32, O, None))
    msg("StoresFindBatchResult", ("Results", 1, F.TYPE_MESSAGE, R, "StoresFindResult"))
    # the no-key-echo / packed-keys extensions (not in the reference proto)
    msg("StoresFindOptionsExt", ("Key", 1, F.TYPE_MESSAGE, O, "StoresKey"), ("TopK", 2, F.TYPE_INT32, O, None),
        ("NoKeys", 3, F.TYPE_BOOL, O, None))
    msg("StoresSetPackedOptions", ("Keys", 1, F.TYPE_BYTES, O, None), ("Dim", 2, F.TYPE_INT32, O, None),
        ("Values", 3, F.TYPE_MESSAGE, R, "StoresValue"))
    msg("Result", ("message", 1, F.TYPE_STRING, O, None), ("success", 2, F.TYPE_BOOL, O, None))   # :260-263
    msg("Reply", ("message", 1, F.TYPE_BYTES, O, None), ("tokens", 2, F.TYPE_INT32, O, None))     # :159-165
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return {m.name: message_factory.GetMessageClass(pool.FindMessageTypeByName("backend." + m.name))
            for m in fd.message_type}


PB = build_messages()
rng = np.random.default_rng(0)
KEYS = [rng.standard_normal(7).astype(np.float32) for _ in range(5)]
VALS = [b"a", b"", b"\x00\xff binary", b"test3", bytes(range(256))]


def pb_key(k):
    return PB["StoresKey"](Floats=[float(x) for x in k])


def test_set_options_bytes_identical():
    ref = PB["StoresSetOptions"](Keys=[pb_key(k) for k in KEYS], Values=[PB["StoresValue"](Bytes=v) for v in VALS])
    assert wire.enc_set(KEYS, VALS) == ref.SerializeToString()
    keys, vals = wire.dec_set(ref.SerializeToString())
    assert vals == VALS and all(np.array_equal(a, b) for a, b in zip(keys, KEYS))


def test_delete_get_options():
    ref = PB["StoresDeleteOptions"](Keys=[pb_key(k) for k in KEYS])
    assert wire.enc_keys(KEYS) == ref.SerializeToString()
    assert all(np.array_equal(a, b) for a, b in zip(wire.dec_keys(ref.SerializeToString()), KEYS))


@pytest.mark.parametrize("topk", [1, 10, 300, 2 ** 31 - 1, 0, -5])
def test_find_options(topk):
    ref = PB["StoresFindOptions"](Key=pb_key(KEYS[0]), TopK=topk)
    assert wire.enc_find(KEYS[0], topk) == ref.SerializeToString()
    k, t = wire.dec_find(ref.SerializeToString())
    assert t == topk and np.array_equal(k, KEYS[0])


def test_find_result_and_get_result():
    sims = np.array([1.0, 0.5, -0.25, 0.0, -1.0], np.float32)
    ref = PB["StoresFindResult"](Keys=[pb_key(k) for k in KEYS], Values=[PB["StoresValue"](Bytes=v) for v in VALS],
                                 Similarities=[float(s) for s in sims])
    assert wire.enc_find_result(KEYS, VALS, sims) == ref.SerializeToString()
    k, v, s = wire.dec_find_result(ref.SerializeToString())
    assert v == VALS and np.array_equal(s, sims) and all(np.array_equal(a, b) for a, b in zip(k, KEYS))
    parsed = PB["StoresFindResult"].FromString(wire.enc_find_result(KEYS, VALS, sims))
    assert list(parsed.Similarities) == [float(x) for x in sims] and [x.Bytes for x in parsed.Values] == VALS
    ref = PB["StoresGetResult"](Keys=[pb_key(k) for k in KEYS[:2]], Values=[PB["StoresValue"](Bytes=v) for v in VALS[:2]])
    assert wire.enc_get_result(KEYS[:2], VALS[:2]) == ref.SerializeToString()


@pytest.mark.parametrize("topk", [1, 100, 0, -1])
def test_find_batch_messages(topk):
    ref = PB["StoresFindBatchOptions"](Keys=[pb_key(k) for k in KEYS], TopK=topk)
    assert wire.enc_find_batch(KEYS, topk) == ref.SerializeToString()
    keys, t = wire.dec_find_batch(ref.SerializeToString())
    assert t == topk and all(np.array_equal(a, b) for a, b in zip(keys, KEYS))
    sims = [np.array([1.0, 0.5], np.float32), np.zeros(0, np.float32), np.array([-1.0], np.float32)]
    triples = [(KEYS[:2], VALS[:2], sims[0]), ([], [], sims[1]), (KEYS[4:5], VALS[4:5], sims[2])]
    ref = PB["StoresFindBatchResult"](Results=[
        PB["StoresFindResult"](Keys=[pb_key(k) for k in ks], Values=[PB["StoresValue"](Bytes=v) for v in vs],
                               Similarities=[float(x) for x in sm]) for ks, vs, sm in triples])
    data = wire.enc_find_batch_result(triples)
    assert data == ref.SerializeToString()
    back = wire.dec_find_batch_result(data)
    assert len(back) == 3 and back[1][0] == [] and back[1][1] == [] and len(back[1][2]) == 0
    for (ks, vs, sm), (k2, v2, s2) in zip(triples, back):
        assert v2 == list(vs) and np.array_equal(s2, sm) and all(np.array_equal(a, b) for a, b in zip(k2, ks))


def test_empty_messages():
    assert wire.enc_find_result([], [], np.zeros(0, np.float32)) == PB["StoresFindResult"]().SerializeToString() == b""
    assert wire.dec_find_result(b"")[0] == []
    assert wire.dec_keys(b"") == []


def test_result_and_reply():
    for m, ok in (("Set key", True), ("Error setting entry: no keys to add", False), ("", True)):
        ref = PB["Result"](message=m, success=ok)
        assert wire.enc_result(m, ok) == ref.SerializeToString()
        assert wire.dec_result(ref.SerializeToString()) == (m, ok)
    assert wire.enc_reply(b"OK") == PB["Reply"](message=b"OK").SerializeToString()
    assert wire.dec_reply(PB["Reply"](message=b"OK", tokens=3).SerializeToString()) == b"OK"   # unknown field skipped


def test_unpacked_floats_are_accepted():
    # a proto2-style writer may send repeated float unpacked: tag 0x0D + 4 bytes per element
    raw = b"".join(b"\x0d" + np.float32(x).tobytes() for x in (1.0, 2.0, 3.0))
    assert np.array_equal(wire.dec_key(memoryview(raw)), np.array([1, 2, 3], np.float32))


def test_status_roundtrip():
    data = wire.enc_status(2, 12345, {"hbm-bytes-gpu0": 12345, "rows": 7})
    assert wire.dec_status(data) == (2, 12345, {"hbm-bytes-gpu0": 12345, "rows": 7})


@pytest.mark.parametrize("no_keys", [False, True])
def test_find_options_no_keys_extension(no_keys):
    """Field 3 (NoKeys) encodes like protobuf's bool, and the REFERENCE's message -- which has no such field -- parses
    the same bytes to the same key and top-k: the extension does not break an old backend."""
    ref = PB["StoresFindOptionsExt"](Key=pb_key(KEYS[1]), TopK=10, NoKeys=no_keys)
    data = wire.enc_find(KEYS[1], 10, no_keys)
    assert data == ref.SerializeToString()
    k, t, nk = wire.dec_find_ext(data)
    assert t == 10 and nk == no_keys and np.array_equal(k, KEYS[1])
    old = PB["StoresFindOptions"]()
    old.ParseFromString(data)
    assert old.TopK == 10 and np.array_equal(np.array(old.Key.Floats, np.float32), KEYS[1])
    k2, t2 = wire.dec_find(data)
    assert t2 == 10 and np.array_equal(k2, KEYS[1])


def test_set_packed_options():
    block = np.stack(KEYS)
    ref = PB["StoresSetPackedOptions"](Keys=block.astype("<f4").tobytes(), Dim=block.shape[1],
                                       Values=[PB["StoresValue"](Bytes=v) for v in VALS])
    data = wire.enc_set_packed(block, VALS)
    assert data == ref.SerializeToString()
    keys, vals = wire.dec_set_packed(data)
    assert vals == VALS and keys.shape == block.shape and np.array_equal(keys, block)
    with pytest.raises(ValueError):
        wire.dec_set_packed(PB["StoresSetPackedOptions"](Keys=b"\x00" * 10, Dim=7).SerializeToString())
    with pytest.raises(ValueError):
        wire.dec_set_packed(PB["StoresSetPackedOptions"](Keys=b"\x00" * 28).SerializeToString())     # Dim missing
