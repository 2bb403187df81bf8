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
