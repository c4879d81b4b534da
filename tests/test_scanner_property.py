"""Property test of the message scanner: for arbitrary messages built from the value kinds the wire
schema allows (btb/publisher.py:30-43), what bt_scan_message reports equals what pickle.loads builds,
for every pickle protocol a sender can choose."""
import pickle

import numpy as np
from hypothesis import given, settings, strategies as st
from hypothesis.extra import numpy as hnp

from blendtorch.btt import native as N

DTYPES = [np.uint8, np.int8, np.uint16, np.int16, np.uint32, np.int32, np.uint64, np.int64,
          np.float16, np.float32, np.float64, np.bool_]

arrays = st.sampled_from(DTYPES).flatmap(
    lambda dt: hnp.arrays(dt, hnp.array_shapes(min_dims=0, max_dims=4, min_side=0, max_side=6),
                          elements=(st.booleans() if dt is np.bool_ else
                                    st.integers(0, 100) if np.issubdtype(dt, np.integer) else
                                    st.floats(-8, 8, width=16))))
values = st.one_of(arrays, st.integers(-2 ** 63, 2 ** 63 - 1), st.floats(allow_nan=False), st.booleans(),
                   st.none(), st.text(max_size=12), st.lists(st.integers(0, 9), max_size=3))
messages = st.dictionaries(st.text(st.characters(min_codepoint=97, max_codepoint=122), min_size=1, max_size=10),
                           values, min_size=0, max_size=6)


@settings(max_examples=150, deadline=None)
@given(msg=messages, protocol=st.sampled_from([3, 4, 5]))
def test_scan_equals_pickle(msg, protocol):
    blob = pickle.dumps(msg, protocol=protocol)
    fields = {f.key.decode(): f for f in N.scan_message(blob)}
    assert list(fields) == list(msg)
    for k, v in msg.items():
        f = fields[k]
        if isinstance(v, np.ndarray):
            assert f.kind == N.BT_KIND_NDARRAY and N.NUMPY_DTYPES[f.dtype] == v.dtype.name
            assert [f.shape[d] for d in range(f.ndim)] == list(v.shape)
            assert bytes(blob[f.payload_off:f.payload_off + f.payload_len]) == v.tobytes()
        elif isinstance(v, bool):
            assert f.kind == N.BT_KIND_BOOL and bool(f.ival) == v
        elif isinstance(v, int):
            assert f.kind == N.BT_KIND_INT and f.ival == v
        elif isinstance(v, float):
            assert f.kind == N.BT_KIND_FLOAT and f.fval == v
        elif v is None:
            assert f.kind == N.BT_KIND_NONE
        elif isinstance(v, str):
            assert f.kind == N.BT_KIND_STR and bytes(blob[f.payload_off:f.payload_off + f.payload_len]).decode() == v
        else:
            assert f.kind == N.BT_KIND_OTHER
