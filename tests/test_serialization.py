"""Port of the reference's serialization tests (src/test/serialization_test.ts:8-49) + flat format."""
import numpy as np
import torch

from distriflow_b200 import (deserialize_var, flat_deserialize, flat_serialize, serialize_var, serialize_vars,
                             serialized_to_array, stack_serialized)


def test_round_trip_float_bool_int():
    f = torch.tensor([[[1.0, 2], [3, 4]], [[5, 6], [7, 8]]])
    b = torch.tensor([True, False])
    i = torch.tensor([[1, 2], [3, 4]], dtype=torch.int32)
    for t in (f, b, i):
        s = serialize_var(t)
        assert tuple(s.shape) == tuple(t.shape)
        back = deserialize_var(s)
        assert back.dtype == t.dtype and torch.equal(back, t)
    assert serialize_var(f).dtype == "float32" and serialize_var(b).dtype == "bool" and serialize_var(i).dtype == "int32"


def test_bfloat16_round_trip():
    t = torch.randn(3, 5).bfloat16()
    assert torch.equal(deserialize_var(serialize_var(t)), t)


def test_stack_serialized():
    f = lambda: torch.tensor([[[1.0, 2], [3, 4]], [[5, 6], [7, 8]]])
    i = lambda: torch.tensor([[1, 2], [3, 4]], dtype=torch.int32)
    updates = [serialize_vars([f(), i()]) for _ in range(3)]
    stacked = stack_serialized(updates)
    assert [s.dtype for s in stacked] == ["float32", "int32"]
    assert tuple(stacked[0].shape) == (3, 2, 2, 2) and tuple(stacked[1].shape) == (3, 2, 2)
    t = deserialize_var(stacked[0])
    assert torch.equal(t[1], f())
    assert torch.allclose(t.mean(0), f())


def test_serialized_to_array_accepts_bytearray():
    s = serialize_var(torch.arange(6, dtype=torch.float32).reshape(2, 3))
    s.data = bytearray(s.data)
    assert np.allclose(serialized_to_array(s), np.arange(6))


def test_flat_format():
    ts = [torch.randn(2, 3), torch.arange(4, dtype=torch.int32), torch.tensor([True, False, True])]
    data, meta = flat_serialize(ts)
    assert meta["byteOffsets"] == [0, 24, 40]
    assert [m["shape"] for m in meta["meta"]] == [[2, 3], [4], [3]]
    back = flat_deserialize(data, meta)
    for a, b in zip(ts, back):
        assert a.dtype == b.dtype and torch.equal(a, b)


# ---------------------------------------------------------------------------------------------- randomized round trips
from hypothesis import given, settings, strategies as st    # noqa: E402


@settings(max_examples=80, deadline=None)
@given(shape=st.lists(st.integers(min_value=0, max_value=5), min_size=0, max_size=4),
       dtype=st.sampled_from(["float32", "int32", "bool", "uint8", "float16", "int64"]), seed=st.integers(0, 10 ** 6))
def test_serialize_round_trip_any_shape_and_dtype(shape, dtype, seed):
    """`serializeVar` / `deserializeVar` (`/root/reference/src/common/utils.ts:32-51`) for every dtype of the table, scalars
    and empty tensors included; the flat format (`server/models.ts:236-267`) must give the same tensors back in order."""
    import numpy as np
    import torch
    from distriflow_b200 import deserialize_var, flat_deserialize, flat_serialize, serialize_var
    from distriflow_b200.serialization import DTYPES
    if dtype not in DTYPES:
        return
    rng = np.random.default_rng(seed)
    arr = (rng.random(shape) * 100).astype(np.dtype(DTYPES[dtype][0])) if dtype != "bool" else rng.random(shape) > 0.5
    arr = np.asarray(arr).reshape(shape)             # (ascontiguousarray would promote a scalar to 1-d)
    t = torch.from_numpy(arr.copy())
    s = serialize_var(t)
    assert s.dtype == dtype and list(s.shape) == list(shape) and len(bytes(s.data)) == arr.nbytes
    back = deserialize_var(s)
    assert back.dtype == t.dtype and tuple(back.shape) == tuple(t.shape) and torch.equal(back, t)
    data, meta = flat_serialize([t, t.reshape(-1)])
    a, b = flat_deserialize(data, meta)
    assert torch.equal(a, t) and torch.equal(b, t.reshape(-1))


def test_wire_codec_is_data_only_and_round_trips_messages():
    """The TCP frames carry a JSON header + raw blobs: every message type survives, nothing executable does."""
    import pickle
    import pytest as _pytest
    import torch as _torch
    from distriflow_b200.messages import (DataMsg, DownloadMsg, GradientMsg, ModelMsg, UploadMsg, WireFormatError, decode, encode)
    from distriflow_b200.serialization import deserialize_var, serialize_var
    w = _torch.arange(12, dtype=_torch.float32).reshape(3, 4)
    g = _torch.randn(5).to(_torch.bfloat16)
    down = DownloadMsg(ModelMsg("17", [serialize_var(w)]), {"learningRate": 0.5, "nested": {"a": [1, 2, (3, 4)]}},
                       DataMsg(3, 1, serialize_var(w), serialize_var(_torch.tensor([1, 0, 2], dtype=_torch.int32)), "bid", "eid"))
    up = UploadMsg("worker-1", GradientMsg("17", [serialize_var(g)]), batch=3, metrics=[0.25, float("inf")], epoch=1)
    for msg in (("downloadVars", down), ("uploadVars", up), True, None, {"k": b"\x00\x01raw"}):
        back = decode(encode(msg))
        assert type(back) is type(msg)
    ev, d2 = decode(encode(("downloadVars", down)))
    assert ev == "downloadVars" and d2.model.version == "17" and d2.hyperparams["nested"]["a"][2] == (3, 4)
    assert _torch.equal(deserialize_var(d2.model.vars[0]), w) and d2.data.batchId == "bid"
    _, u2 = decode(encode(("uploadVars", up)))
    assert _torch.equal(deserialize_var(u2.gradients.vars[0]), g) and u2.metrics[1] == float("inf")

    class Evil:
        def __reduce__(self):
            return (print, ("pwned",))

    with _pytest.raises(WireFormatError):
        encode(("uploadVars", Evil()))                       # arbitrary objects cannot be sent ...
    with _pytest.raises(WireFormatError):
        decode(pickle.dumps(Evil()))                         # ... and a pickle is not a frame
    with _pytest.raises(WireFormatError):
        decode(encode({"a": 1})[:9])                         # truncated header
