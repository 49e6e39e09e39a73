"""W4: DPS1 frame codec, restricted pickle compat, fp16 compression (C5, C6)."""
import pickle

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from dps_b200.rpc import codec


def test_frame_roundtrip_mixed_dtypes():
    t = {"w": np.random.randn(3, 5).astype(np.float32), "g": np.random.randn(7).astype(np.float16),
         "n": np.array(12, dtype=np.int64), "e": np.zeros((0, 4), np.float32)}
    buf = codec.encode_frame(t, {"step": 4})
    out, meta = codec.decode_frame(buf)
    assert meta == {"step": 4}
    for k in t:
        assert out[k].dtype == t[k].dtype and out[k].shape == t[k].shape
        np.testing.assert_array_equal(out[k], t[k])


@settings(max_examples=25, deadline=None)
@given(st.lists(st.tuples(st.integers(0, 5), st.integers(0, 7)), min_size=1, max_size=5))
def test_frame_roundtrip_property(shapes):
    t = {f"t{i}": np.random.randn(a, b).astype(np.float32) for i, (a, b) in enumerate(shapes)}
    out, _ = codec.decode_frame(codec.encode_frame(t))
    assert list(out) == list(t)
    for k in t:
        np.testing.assert_array_equal(out[k], t[k])


def test_reference_pickle_is_accepted_only_when_enabled():
    ref_payload = pickle.dumps({"a": np.arange(6, dtype=np.float16).reshape(2, 3)})
    with pytest.raises(ValueError):
        codec.decode_any(ref_payload, allow_pickle=False)
    out, _ = codec.decode_any(ref_payload, allow_pickle=True)
    np.testing.assert_array_equal(out["a"], np.arange(6, dtype=np.float16).reshape(2, 3))


def test_restricted_unpickler_blocks_code_execution():
    class Evil:
        def __reduce__(self):
            import os
            return (os.system, ("echo pwned",))
    with pytest.raises(pickle.UnpicklingError):
        codec.restricted_pickle_loads(pickle.dumps({"x": Evil()}))
    with pytest.raises(pickle.UnpicklingError):
        codec.restricted_pickle_loads(pickle.dumps([1, 2, 3]))


def test_reference_pickle_dumps_is_loadable_by_plain_pickle():
    t = {"a": np.ones((2, 2), np.float32)}
    back = pickle.loads(codec.reference_pickle_dumps(t))
    np.testing.assert_array_equal(back["a"], t["a"])


def test_compression_halves_payload_and_matches_reference_rounding():
    g = {"w": (np.random.randn(1000) * 3).astype(np.float32)}
    c = codec.compress_gradients(g)
    assert c["w"].dtype == np.float16
    d = codec.decompress_gradients(c)
    np.testing.assert_array_equal(d["w"], g["w"].astype(np.float16).astype(np.float32))
    assert len(codec.encode_frame(c)) < 0.55 * len(codec.encode_frame(g))


def test_malformed_frames_rejected():
    good = codec.encode_frame({"a": np.ones(4, np.float32)})
    with pytest.raises(ValueError):
        codec.decode_frame(good[:20])
    with pytest.raises(ValueError):
        codec.decode_frame(b"NOPE" + good[4:])
