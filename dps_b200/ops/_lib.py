"""ctypes binding of the in-tree native library (dps_b200/_native/libdps_b200.so).

The library is loaded lazily, on first use, so importing the package works on a CPU-only
box.  On a box WITH a GPU a missing/unloadable library is a hard error: device ops never
fall back silently to another implementation.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "_native", "libdps_b200.so")

MAX_RANKS = 32
MAX_SLOTS = 64
MAX_PLAN = 16
HIST_BINS = 16
FLAG_WORDS = 1024

# flag-block word offsets (mirror include/dps_b200.h)
F_ARRIVE, F_PULLCOUNT, F_ERROR, F_MISSING, F_TORN_PULLS, F_DROPPED, F_ALIVE, F_PULLEXPECT = 0, 32, 33, 34, 35, 36, 37, 38
F_UPDATE_NS, F_ROUND = 40, 39
F_HEARTBEAT, F_RING_TAIL, F_RING_HEAD, F_VERSION, F_PUBSEQ, F_APPLY_DONE = 64, 96, 97, 98, 99, 100
F_SLOTS = 128
F_STAT_ACCEPTED, F_STAT_REJECTED, F_STAT_SUM, F_STAT_MAX, F_STAT_HIST, F_STAT_LAST = 512, 513, 514, 515, 520, 544
F_PLAN, F_TICKET, F_SEEN, F_SEGCOUNT, F_SEGEXPECT = 640, 704, 800, 864, 880
F_CONSUMED, F_CONSUMED_ROUND, F_INCLUDE = 896, 928, 960
ERR_BARRIER_TIMEOUT, ERR_PULL_TIMEOUT, ERR_RING_FULL = 1, 2, 4

_P = C.c_void_p
_PtrArray = _P * MAX_RANKS


class DpsTables(C.Structure):
    _fields_ = [
        ("world", C.c_int32), ("rank_lo", C.c_int32), ("rank_hi", C.c_int32),
        ("block_elems", C.c_int32), ("ring_slots", C.c_int32), ("n_extra", C.c_int32),
        ("shard_elems", C.c_int64), ("n_small", C.c_int64),
        ("grad", _PtrArray), ("master", _PtrArray), ("mom", _PtrArray), ("stage", _PtrArray),
        ("ring", _PtrArray), ("wbf16", _PtrArray), ("small", _PtrArray), ("pub", _PtrArray),
        ("pubsmall", _PtrArray), ("flags", _PtrArray),
        ("step_ctr", _P), ("lr_ptr", _P),
        ("mc_wbf16", _P), ("mc_small", _P), ("mc_stage", _P),
    ]


class NativeError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None
_lock = threading.Lock()


def _declare(lib: C.CDLL) -> None:
    u64, i32, u32, f32 = C.c_uint64, C.c_int, C.c_uint, C.c_float
    TP = C.POINTER(DpsTables)
    sigs = {
        "dps_last_error": (C.c_char_p, []),
        "dps_tables_size": (i32, []),
        "dps_device_count": (i32, [C.POINTER(i32)]),
        "dps_arena_alloc": (i32, [i32, u64, C.POINTER(_P)]),
        "dps_arena_free": (i32, [i32, _P]),
        "dps_ipc_export": (i32, [_P, C.c_char_p]),
        "dps_ipc_open": (i32, [i32, C.c_char_p, C.POINTER(_P)]),
        "dps_ipc_close": (i32, [_P]),
        "dps_enable_peer": (i32, [i32, i32]),
        "dps_push_sync": (i32, [TP, u32, f32, f32, f32, i32, i32, u64, _P]),
        "dps_push_segment": (i32, [TP, u32, f32, f32, f32, i32, i32, u64, C.c_longlong, C.c_longlong, i32, u32, _P]),
        "dps_pull_bcast": (i32, [TP, u32, i32, _P]),
        "dps_wait_pull": (i32, [TP, i32, u32, i32, u64, i32, _P]),
        "dps_pull_segments": (i32, [TP, C.POINTER(C.c_longlong), i32, i32, _P]),
        "dps_wait_segment": (i32, [TP, i32, i32, u32, i32, u64, _P]),
        "dps_bump_step": (i32, [TP, _P]),
        "dps_push_async": (i32, [TP, i32, u32, i32, u64, _P]),
        "dps_async_apply": (i32, [TP, i32, i32, i32, f32, f32, i32, f32, f32, f32, i32, i32, i32, _P]),
        "dps_pull_async": (i32, [TP, i32, i32, i32, _P]),
        "dps_publish": (i32, [TP, i32, _P]),
        "dps_flag_update": (i32, [_P, u32, u32, _P]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    # optional groups (compute kernels) declare themselves in their own modules via declare()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise NativeError(
                        f"native library missing: {LIB_PATH}; run `python -m dps_b200.build` "
                        "(device ops have no fallback)")
                l = C.CDLL(LIB_PATH)
                _declare(l)
                if l.dps_tables_size() != C.sizeof(DpsTables):
                    raise NativeError(f"DpsTables: library has {l.dps_tables_size()} bytes, the ctypes mirror "
                                      f"{C.sizeof(DpsTables)}; rebuild with `python -m dps_b200.build`")
                _lib = l
    return _lib


def declare(name: str, restype, argtypes) -> None:
    fn = getattr(lib(), name)
    fn.restype, fn.argtypes = restype, argtypes


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = lib().dps_last_error()
        raise NativeError(f"{what} failed (code {code}): {msg.decode(errors='replace') if msg else ''}")


def available() -> bool:
    return os.path.exists(LIB_PATH)


def stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)
