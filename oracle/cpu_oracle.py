"""ctypes binding of oracle/nxs_oracle.cpp (the C++ oracle).  Test infrastructure.

Parity status: pinned against the reference's 8 golden scenarios
(tests/golden/supervisor_test_scenarios.json, extracted from
services/supervisor_test.go:46-540); nexus-core's IsFinished() beyond those
scenarios is unpinned and therefore a parameter (finished_mask).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libnxs_oracle.so")

JOB_ROW = np.dtype([("job_key", "<u8"), ("verdict", "<u4"), ("winner_idx", "<u4")])
FINISHED_MASK_DEFAULT = (1 << 3) | (1 << 4) | (1 << 5) | (1 << 6) | (1 << 7)


def build(force=False):
    src = os.path.join(_HERE, "nxs_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u8p, u16p, u32p, u64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint16, C.c_uint32, C.c_uint64))
        L.nxo_classify.restype = C.c_uint64
        L.nxo_classify.argtypes = [C.c_int, C.c_int, u8p, u16p, u8p, u8p, u64p, C.c_uint64, C.c_uint16,
                                   u32p, C.c_void_p]
        L.nxo_generate.restype = C.c_uint64
        L.nxo_generate.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                   u8p, u16p, u8p, u8p, u64p, C.c_uint64]
        L.nxo_splitmix64.restype = C.c_uint64
        L.nxo_splitmix64.argtypes = [C.c_uint64]
        L.nxo_event_strings.restype = C.c_int
        L.nxo_event_strings.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                        C.c_char_p, C.c_uint16] + [C.c_char_p] * 5 + [C.c_size_t]
        L.nxo_apply_action_string.restype = C.c_int
        L.nxo_apply_action_string.argtypes = [C.c_char_p, C.c_char_p, C.c_uint16]
        for f in (L.nxo_kind_name, L.nxo_reason_name, L.nxo_stage_name):
            f.restype = C.c_char_p
            f.argtypes = [C.c_uint32]
        L.nxo_bench_event_loop.restype = C.c_double
        L.nxo_bench_event_loop.argtypes = [C.c_int, u8p, u16p, u8p, u8p, u64p, C.c_uint64, C.c_uint16, C.c_int,
                                           u64p, u64p]
        L.nxo_encode_strings.restype = C.c_uint64
        L.nxo_encode_strings.argtypes = [u8p, u16p, u8p, u8p, u64p, C.c_uint64, C.c_uint16, C.c_void_p, C.c_uint64,
                                         C.c_void_p, C.c_void_p, u64p]
        L.nxo_hardware_threads.restype = C.c_uint
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _cols(cols):
    kind = np.ascontiguousarray(cols["kind"], dtype=np.uint8)
    reason = np.ascontiguousarray(cols["reason_id"], dtype=np.uint16)
    flags = np.ascontiguousarray(cols["flags"], dtype=np.uint8)
    stage = np.ascontiguousarray(cols["cur_stage"], dtype=np.uint8)
    key = np.ascontiguousarray(cols["job_key"], dtype=np.uint64)
    return kind, reason, flags, stage, key


def classify(cols, finished_mask=FINISHED_MASK_DEFAULT, mode="interned", threads=1, want_rows=True,
             want_verdict=True):
    """Returns (verdict[n] u32, rows[n_jobs] JOB_ROW).  mode: 'interned' | 'strings'."""
    kind, reason, flags, stage, key = _cols(cols)
    n = len(kind)
    verdict = np.zeros(n, dtype=np.uint32) if want_verdict else None
    rows = np.zeros(n, dtype=JOB_ROW) if want_rows else None
    nj = lib().nxo_classify(1 if mode == "strings" else 0, threads, _p(kind, C.c_uint8), _p(reason, C.c_uint16),
                            _p(flags, C.c_uint8), _p(stage, C.c_uint8), _p(key, C.c_uint64), n, finished_mask,
                            _p(verdict, C.c_uint32) if want_verdict else None,
                            rows.ctypes.data if want_rows else None)
    return verdict, (rows[:nj] if want_rows else nj)


def bench_event_loop(cols, threads=1, finished_mask=FINISHED_MASK_DEFAULT, repeats=3):
    """Timed stand-in for the Go loop: onEvent's string switches + one 7-string result per handled event +
    superviseAction's decision against the run's in-memory row; the events' strings are built before the
    clock starts (in Go they already exist in the *corev1.Event).  Returns (records/s, handled events)."""
    kind, reason, flags, stage, key = _cols(cols)
    handled, checksum = C.c_uint64(0), C.c_uint64(0)
    dt = lib().nxo_bench_event_loop(threads, _p(kind, C.c_uint8), _p(reason, C.c_uint16), _p(flags, C.c_uint8),
                                    _p(stage, C.c_uint8), _p(key, C.c_uint64), len(kind), finished_mask, repeats,
                                    C.byref(handled), C.byref(checksum))
    return len(kind) / dt, int(handled.value)


def encode_strings(cols, finished_mask=FINISHED_MASK_DEFAULT, bytes_per_record=48):
    """What the reference would have written, from the string-mode replay: (arena, offsets, update_job) with
    lifecycle_stage / algorithm_failure_cause / algorithm_failure_details of every upserted job back to back;
    event.Message of record i is "m<i>"."""
    kind, reason, flags, stage, key = _cols(cols)
    n = len(kind)
    arena = np.empty(n * bytes_per_record + 4096, np.uint8)
    offsets = np.empty(3 * n + 1, np.uint64)
    update_job = np.empty(n, np.uint32)
    nbytes = C.c_uint64(0)
    n_up = lib().nxo_encode_strings(_p(kind, C.c_uint8), _p(reason, C.c_uint16), _p(flags, C.c_uint8),
                                    _p(stage, C.c_uint8), _p(key, C.c_uint64), n, finished_mask, arena.ctypes.data,
                                    arena.nbytes, offsets.ctypes.data, update_job.ctypes.data, C.byref(nbytes))
    assert n_up != 2**64 - 1, "arena too small"
    return arena[:nbytes.value], offsets[:3 * n_up + 1], update_job[:n_up]


def generate(workload, n, seed=0x6E65787573, first=0, shard=0, n_shards=1):
    """Synthetic stream records [first, first+n) (optionally one job-hash shard of them)."""
    cap = n
    cols = {
        "kind": np.zeros(cap, np.uint8), "reason_id": np.zeros(cap, np.uint16), "flags": np.zeros(cap, np.uint8),
        "cur_stage": np.zeros(cap, np.uint8), "job_key": np.zeros(cap, np.uint64),
    }
    w = lib().nxo_generate(workload, seed, first, n, shard, n_shards, _p(cols["kind"], C.c_uint8),
                           _p(cols["reason_id"], C.c_uint16), _p(cols["flags"], C.c_uint8),
                           _p(cols["cur_stage"], C.c_uint8), _p(cols["job_key"], C.c_uint64), cap)
    assert w != 2**64 - 1
    return {k: v[:w] for k, v in cols.items()}


def splitmix64(x):
    return lib().nxo_splitmix64(x & (2**64 - 1))


def event_strings(kind, reason, stage, message="", supervised=True, found=True, checkpoint=True,
                  finished_mask=FINISHED_MASK_DEFAULT):
    """One event through the string-faithful onEvent + superviseAction."""
    bufs = [C.create_string_buffer(512) for _ in range(5)]
    rc = lib().nxo_event_strings(kind.encode(), reason.encode(), message.encode(), int(supervised), int(found),
                                 int(checkpoint), stage.encode(), finished_mask, *bufs, 512)
    stage_out, cause, details, action, msg = (b.value.decode() for b in bufs)
    return {"produced": bool(rc & 256), "delete_job": bool(rc & 1), "upsert": bool(rc & 2), "skipped": bool(rc & 4),
            "err": (rc >> 4) & 3, "stage": stage_out, "cause": cause, "details": details, "action": action,
            "message": msg}


def apply_action_string(action, stage="RUNNING", finished_mask=FINISHED_MASK_DEFAULT):
    return lib().nxo_apply_action_string(action.encode(), stage.encode(), finished_mask)


def hardware_threads():
    return int(lib().nxo_hardware_threads()) or os.cpu_count() or 1


def names():
    L = lib()
    return {"kind": [L.nxo_kind_name(i).decode() for i in range(3)],
            "reason": [L.nxo_reason_name(i).decode() for i in range(16)],
            "stage": [L.nxo_stage_name(i).decode() for i in range(8)]}
