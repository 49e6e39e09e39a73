"""Payload codec for the ``bytes`` fields of PushRequest / FetchReply.

The reference serialises ``{name: np.ndarray}`` with ``pickle`` on both sides
(reference: src/parameter_server/server.py:222,245; src/workers/worker.py:241,289),
which lets any peer that can reach the port execute code.  The default
here is a flat, self-describing frame ("DPS1"): JSON header + 64-byte-aligned raw
tensor bytes, decoded zero-copy with ``np.frombuffer``.  Reference pickles are still
understood -- and can be emitted -- through a *restricted* unpickler that only
reconstructs numpy arrays, so a reference worker can talk to this server and
vice versa (``--compat-pickle``).

Gradient compression (reference: worker.py:264-268 compress, server.py:232-237
decompress) is fp32 -> fp16 on the wire, fp32 again at the server.
"""
from __future__ import annotations

import io
import json
import pickle
import struct
from typing import Dict, Mapping, Optional, Tuple

import numpy as np

MAGIC = b"DPS1"
_ALIGN = 64
_ALLOWED_DTYPES = {"float16", "float32", "float64", "int64", "int32", "uint8", "int8", "bool",
                   "uint16", "int16"}


def encode_frame(tensors: Mapping[str, np.ndarray], meta: Optional[dict] = None) -> bytes:
    entries = []
    offset = 0
    arrays = []
    for name, arr in tensors.items():
        a = np.asarray(arr)
        if not a.flags.c_contiguous:
            a = a.copy(order="C")
        if a.dtype.name not in _ALLOWED_DTYPES:
            raise TypeError(f"dtype {a.dtype} not allowed on the wire")
        entries.append({"name": name, "dtype": a.dtype.name, "shape": list(a.shape),
                        "offset": offset, "nbytes": int(a.nbytes)})
        arrays.append(a)
        offset += (a.nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
    header = json.dumps({"tensors": entries, "meta": meta or {}}, separators=(",", ":")).encode()
    prefix = len(MAGIC) + 4 + len(header)
    pad = (-prefix) % _ALIGN
    # one allocation, one copy: join the header and the tensors' own buffers (45 MB of parameters per pull; a zeroed
    # bytearray + slice assignment + bytes() touched every page three times)
    parts = [MAGIC, struct.pack("<I", len(header)), header, b"\0" * pad]
    for e, a in zip(entries, arrays):
        if e["nbytes"]:
            parts.append(a.reshape(-1).view(np.uint8).data)
            tail = (-e["nbytes"]) % _ALIGN
            if tail:
                parts.append(b"\0" * tail)
    return b"".join(parts)


def decode_frame(buf: bytes) -> Tuple[Dict[str, np.ndarray], dict]:
    if len(buf) < 8 or buf[:4] != MAGIC:
        raise ValueError("not a DPS1 frame")
    (hlen,) = struct.unpack_from("<I", buf, 4)
    if 8 + hlen > len(buf):
        raise ValueError("truncated DPS1 header")
    header = json.loads(bytes(buf[8:8 + hlen]).decode())
    prefix = 8 + hlen
    base = prefix + ((-prefix) % _ALIGN)
    out: Dict[str, np.ndarray] = {}
    for e in header["tensors"]:
        if e["dtype"] not in _ALLOWED_DTYPES:
            raise ValueError(f"dtype {e['dtype']} not allowed")
        dt = np.dtype(e["dtype"])
        shape = tuple(int(s) for s in e["shape"])
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if n * dt.itemsize != e["nbytes"] or e["offset"] < 0:
            raise ValueError("inconsistent DPS1 entry")
        start = base + e["offset"]
        if start + e["nbytes"] > len(buf):
            raise ValueError("truncated DPS1 payload")
        out[e["name"]] = np.frombuffer(buf, dtype=dt, count=n, offset=start).reshape(shape)
    return out, header.get("meta", {})


# --------------------------------------------------------------------------
# Reference-pickle compatibility (restricted)
# --------------------------------------------------------------------------
_SAFE_GLOBALS = {
    ("numpy.core.multiarray", "_reconstruct"),
    ("numpy._core.multiarray", "_reconstruct"),
    ("numpy.core.multiarray", "scalar"),
    ("numpy._core.multiarray", "scalar"),
    ("numpy", "ndarray"),
    ("numpy", "dtype"),
    ("collections", "OrderedDict"),
}


class _NumpyOnlyUnpickler(pickle.Unpickler):
    def find_class(self, module, name):  # noqa: D401
        if (module, name) in _SAFE_GLOBALS:
            return super().find_class(module, name)
        raise pickle.UnpicklingError(f"global {module}.{name} is forbidden on this wire")


def restricted_pickle_loads(buf: bytes) -> Dict[str, np.ndarray]:
    obj = _NumpyOnlyUnpickler(io.BytesIO(buf)).load()
    if not isinstance(obj, dict):
        raise pickle.UnpicklingError("expected a dict of numpy arrays")
    for k, v in obj.items():
        if not isinstance(k, str) or not isinstance(v, np.ndarray) or v.dtype.hasobject:
            raise pickle.UnpicklingError("expected {str: non-object ndarray}")
    return obj


def reference_pickle_dumps(tensors: Mapping[str, np.ndarray]) -> bytes:
    """Exactly what the reference emits: pickle of a plain dict of numpy arrays."""
    return pickle.dumps({k: np.asarray(v) for k, v in tensors.items()})


# --------------------------------------------------------------------------
# High-level helpers used by both roles
# --------------------------------------------------------------------------
def is_frame(buf: bytes) -> bool:
    return len(buf) >= 4 and buf[:4] == MAGIC


def decode_any(buf: bytes, allow_pickle: bool) -> Tuple[Dict[str, np.ndarray], dict]:
    """Decode a payload; pickles only when the operator opted in."""
    if is_frame(buf):
        return decode_frame(buf)
    if not allow_pickle:
        raise ValueError("payload is not a DPS1 frame and --compat-pickle is off")
    return restricted_pickle_loads(buf), {}


def encode_any(tensors: Mapping[str, np.ndarray], as_pickle: bool, meta: Optional[dict] = None) -> bytes:
    return reference_pickle_dumps(tensors) if as_pickle else encode_frame(tensors, meta)


def _torch_cast(a: np.ndarray, np_dtype, torch_dtype) -> np.ndarray:
    """dtype conversion through torch's vectorised kernels into a fresh numpy array (numpy's own fp32<->fp16
    casts are scalar code: 10 ms per 2 M elements against 1.3 ms)."""
    import warnings

    import torch
    a = np.ascontiguousarray(a)
    if a.size < 32768:                                # small tensors: the fixed cost of the torch path dominates
        return a.astype(np_dtype)
    out = np.empty(a.shape, dtype=np_dtype)
    if a.size:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")           # read-only source (a received frame): it is only read
            torch.from_numpy(out).copy_(torch.from_numpy(a) if a.flags.writeable
                                        else torch.frombuffer(a, dtype=_NP2TORCH[a.dtype.name]).view(a.shape))
    return out


_NP2TORCH: Dict[str, "object"] = {}


def _init_torch_map() -> None:
    if not _NP2TORCH:
        import torch
        _NP2TORCH.update({"float16": torch.float16, "float32": torch.float32, "float64": torch.float64,
                          "int64": torch.int64, "int32": torch.int32, "uint8": torch.uint8, "int8": torch.int8,
                          "bool": torch.bool, "int16": torch.int16})


def compress_gradients(grads: Mapping[str, np.ndarray], how: str = "fp16") -> Dict[str, np.ndarray]:
    """fp32 -> fp16 wire compression (reference: worker.py:264-268).  Round-to-nearest-even, same bits as
    ``astype(np.float16)``."""
    if how == "none":
        return {k: np.asarray(v, dtype=np.float32) for k, v in grads.items()}
    _init_torch_map()
    import torch
    return {k: (_torch_cast(np.asarray(v), np.float16, torch.float16) if np.asarray(v).dtype.name in _NP2TORCH
                else np.asarray(v).astype(np.float16)) for k, v in grads.items()}


def decompress_gradients(grads: Mapping[str, np.ndarray]) -> Dict[str, np.ndarray]:
    """wire -> fp32 (reference: server.py:232-237)."""
    _init_torch_map()
    import torch
    return {k: (_torch_cast(np.asarray(v), np.float32, torch.float32) if np.asarray(v).dtype.name in _NP2TORCH
                else np.asarray(v).astype(np.float32)) for k, v in grads.items()}
