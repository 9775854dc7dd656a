"""Classifier — thin host wrapper over the C ABI (include/nxs_classify.h).

`Classifier.classify` takes host (numpy) columns, `Classifier.classify_device`
takes device pointers (e.g. torch CUDA tensors) and is asynchronous on the
stream given to the constructor.  Both go through libnxs_classify.so.
"""
import ctypes as C

import numpy as np

from . import _abi

JOB_ROW = np.dtype([("job_key", "<u8"), ("verdict", "<u4"), ("winner_idx", "<u4")])

KIND = {"Other": 0, "Job": 1, "Pod": 2}
REASON = {"Unknown": 0, "FailedCreate": 1, "DeadlineExceeded": 2, "BackoffLimitExceeded": 3, "PodFailurePolicy": 4,
          "Started": 5, "Failed": 6, "BackOff": 7, "Scheduled": 8, "Pulling": 9, "Pulled": 10, "Created": 11,
          "Killing": 12, "SuccessfulCreate": 13, "Completed": 14, "FailedScheduling": 15}
STAGE = {"NEW": 0, "BUFFERED": 1, "RUNNING": 2, "COMPLETED": 3, "FAILED": 4, "SCHEDULING_FAILED": 5,
         "DEADLINE_EXCEEDED": 6, "CANCELLED": 7}
ACTION = {"": 0, "ToFailStuckInPending": 1, "ToFailFatalError": 2, "ToFailDeadlineExceeded": 3, "ToRunning": 4}
FLAG_SUPERVISED, FLAG_FOUND, FLAG_CHECKPOINT = 1, 2, 4
FINISHED_MASK_DEFAULT = sum(1 << STAGE[s] for s in
                            ("COMPLETED", "FAILED", "SCHEDULING_FAILED", "DEADLINE_EXCEEDED", "CANCELLED"))
NO_WINNER = 0xFFFFFFFF
WORKLOAD_UNIFORM, WORKLOAD_SKEWED = 1, 2
DEFAULT_SEED = 0x6E65787573


class NxsError(RuntimeError):
    def __init__(self, status, detail=""):
        self.status = status
        msg = _abi.load().nxs_strerror(status).decode()
        super().__init__(f"nxs error {status} ({msg})" + (f": {detail}" if detail else ""))


def unpack_verdict(v):
    """Packed u32 verdict -> dict of fields (include/nxs_classify.h bit layout)."""
    v = int(v)
    return {"action": v & 7, "new_stage": (v >> 3) & 15, "cause_tpl": (v >> 7) & 3, "msg_tpl": (v >> 9) & 7,
            "delete_job": (v >> 12) & 1, "upsert": (v >> 13) & 1, "skip_finished": (v >> 14) & 1,
            "err": (v >> 15) & 3}


def pack_columns(cols, out=None):
    """attr u16 per record = NXS_ATTR_PACK(kind, reason_id, flags, cur_stage) (nxs_pack_columns)."""
    kind = np.ascontiguousarray(cols["kind"], dtype=np.uint8)
    reason = np.ascontiguousarray(cols["reason_id"], dtype=np.uint16)
    flags = np.ascontiguousarray(cols["flags"], dtype=np.uint8)
    stage = np.ascontiguousarray(cols["cur_stage"], dtype=np.uint8)
    attr = out if out is not None else np.empty(len(kind), dtype=np.uint16)
    _abi.load().nxs_pack_columns(kind.ctypes.data, reason.ctypes.data, flags.ctypes.data, stage.ctypes.data, len(kind),
                                 attr.ctypes.data)
    return attr


def encode_checkpoint_rows(rows, msg_heap=None, msg_offsets=None, out=None):
    """nxs_encode_checkpoint_rows: (arena u8[], offsets u64[3u+1], update_row u32[u], update_flags u8[u]) for the
    job rows that upsert.  msg_heap / msg_offsets: the events' Message bytes per batch record.  `out`: four
    arrays to write into — arena u8[enough bytes, else NXS_ERR_CAPACITY], offsets u64[3 * len(rows) + 1],
    update_row u32[len(rows)], update_flags u8[len(rows)] (a caller that encodes batch after batch keeps its buffers) — without it the sizes are queried first and fresh arrays made."""
    lib = _abi.load()
    rows = np.ascontiguousarray(rows, dtype=JOB_ROW)
    n_rec = 0 if msg_offsets is None else len(msg_offsets) - 1
    heap_p = None if msg_heap is None else np.ascontiguousarray(msg_heap, dtype=np.uint8).ctypes.data
    offs_p = None if msg_offsets is None else np.ascontiguousarray(msg_offsets, dtype=np.uint64).ctypes.data
    n_up, n_bytes = C.c_uint64(0), C.c_uint64(0)
    if out is None:
        rc = lib.nxs_encode_checkpoint_rows(rows.ctypes.data, len(rows), heap_p, offs_p, n_rec, None, None, 0, None, None,
                                            None, C.byref(n_up), C.byref(n_bytes))
        if rc != _abi.NXS_OK:
            raise NxsError(rc, "nxs_encode_checkpoint_rows (size query)")
        arena = np.empty(n_bytes.value, dtype=np.uint8)
        offsets = np.empty(3 * n_up.value + 1, dtype=np.uint64)
        update_row = np.empty(n_up.value, dtype=np.uint32)
        update_flags = np.empty(n_up.value, dtype=np.uint8)
    else:
        arena, offsets, update_row, update_flags = out
        if len(offsets) < 3 * len(update_row) + 1 or len(update_flags) < len(update_row) or len(update_row) < len(rows):
            # (the C call checks the arena against the bytes it needs; the per-update arrays have no capacity
            # argument, so they must hold the worst case: one update per job row)
            raise ValueError("out: update_row / update_flags need len(rows) entries, offsets 3 * len(rows) + 1")
    rc = lib.nxs_encode_checkpoint_rows(rows.ctypes.data, len(rows), heap_p, offs_p, n_rec, None, arena.ctypes.data,
                                        arena.nbytes, offsets.ctypes.data, update_row.ctypes.data,
                                        update_flags.ctypes.data, C.byref(n_up), C.byref(n_bytes))
    if rc != _abi.NXS_OK:
        raise NxsError(rc, "nxs_encode_checkpoint_rows")
    u = n_up.value
    return arena[:n_bytes.value], offsets[:3 * u + 1], update_row[:u], update_flags[:u]


def shard_of(job_key, n_shards):
    return _abi.load().nxs_shard_of(int(job_key), int(n_shards))


def intern_kind(s):
    b = s.encode()
    return _abi.load().nxs_intern_kind(b, len(b))


def intern_reason(s):
    b = s.encode()
    return _abi.load().nxs_intern_reason(b, len(b))


def container_reason_to_event(s):
    """EXTENSION (no reference parity): (kind, reason_id) of the event through which a container-status reason
    reaches the supervisor, or None for reasons that have no such event."""
    b = s.encode()
    kind, reason = C.c_uint8(0), C.c_uint16(0)
    ok = _abi.load().nxs_ext_container_reason_to_event(b, len(b), C.byref(kind), C.byref(reason))
    return (kind.value, reason.value) if ok else None


def msg_template(i):
    return _abi.load().nxs_msg_template(i).decode()


def cause_template(i):
    t = _abi.load().nxs_cause_template(i)
    return None if t is None else t.decode()


def stage_name(i):
    return _abi.load().nxs_stage_name(i).decode()


def action_name(i):
    return _abi.load().nxs_action_name(i).decode()


def reason_name(i):
    return _abi.load().nxs_reason_name(i).decode()


class Classifier:
    """One nxs_ctx: a GPU, a stream and the scratch the kernel needs."""

    def __init__(self, device=0, max_records=1 << 20, finished_mask=FINISHED_MASK_DEFAULT, stream=None):
        self._lib = _abi.load()
        self._ctx = C.c_void_p()
        rc = self._lib.nxs_create(device, max_records, finished_mask, C.byref(self._ctx))
        if rc != _abi.NXS_OK:
            self._ctx = None
            raise NxsError(rc, f"nxs_create(device={device}, max_records={max_records})")
        self.device = device
        self.max_records = max_records
        self.finished_mask = finished_mask
        if stream is not None:
            self.set_stream(stream)

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.nxs_destroy(self._ctx)
            self._ctx = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != _abi.NXS_OK:
            raise NxsError(rc, self._lib.nxs_last_error(self._ctx).decode())

    def set_stream(self, cuda_stream):
        """cuda_stream: integer cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream)."""
        self._check(self._lib.nxs_set_stream(self._ctx, C.c_void_p(int(cuda_stream))))

    def sync(self):
        self._check(self._lib.nxs_sync(self._ctx))

    @property
    def launches(self):
        return int(self._lib.nxs_launch_count(self._ctx))

    @property
    def last_tile(self):
        """Records per tile of the kernel variant the last launch used."""
        return int(self._lib.nxs_last_tile(self._ctx))

    @property
    def last_wire_bytes(self):
        """(host->device, device->host) bytes the last host-pointer call moved over the link."""
        up, down = C.c_uint64(0), C.c_uint64(0)
        self._lib.nxs_last_wire_bytes(self._ctx, C.byref(up), C.byref(down))
        return int(up.value), int(down.value)

    # ------------------------------------------------------------- host path
    def classify(self, cols, out_verdict=None, out_rows=None, want_verdict=True):
        """cols: dict of numpy arrays kind u8, reason_id u16, flags u8, cur_stage u8, job_key u64.
        Returns (verdict u32[n], rows JOB_ROW[n_jobs]); verdict is None with want_verdict=False (the
        per-record verdicts are then neither produced nor copied back)."""
        if not want_verdict:
            return self._classify_rows_only(cols, out_rows)
        kind = np.ascontiguousarray(cols["kind"], dtype=np.uint8)
        reason = np.ascontiguousarray(cols["reason_id"], dtype=np.uint16)
        flags = np.ascontiguousarray(cols["flags"], dtype=np.uint8)
        stage = np.ascontiguousarray(cols["cur_stage"], dtype=np.uint8)
        key = np.ascontiguousarray(cols["job_key"], dtype=np.uint64)
        n = len(kind)
        if not (len(reason) == len(flags) == len(stage) == len(key) == n):
            raise ValueError("columns differ in length")
        verdict = out_verdict if out_verdict is not None else np.empty(n, dtype=np.uint32)
        rows = out_rows if out_rows is not None else np.empty(n, dtype=JOB_ROW)
        n_jobs = C.c_uint64(0)
        batch = _abi.Batch(kind.ctypes.data, reason.ctypes.data, flags.ctypes.data, stage.ctypes.data,
                           key.ctypes.data, n)
        out = _abi.Out(verdict.ctypes.data, rows.ctypes.data, C.addressof(n_jobs))
        self._check(self._lib.nxs_classify(self._ctx, C.byref(batch), C.byref(out)))
        return verdict[:n], rows[:n_jobs.value]

    def _classify_rows_only(self, cols, out_rows=None):
        kind = np.ascontiguousarray(cols["kind"], dtype=np.uint8)
        reason = np.ascontiguousarray(cols["reason_id"], dtype=np.uint16)
        flags = np.ascontiguousarray(cols["flags"], dtype=np.uint8)
        stage = np.ascontiguousarray(cols["cur_stage"], dtype=np.uint8)
        key = np.ascontiguousarray(cols["job_key"], dtype=np.uint64)
        n = len(kind)
        rows = out_rows if out_rows is not None else np.empty(n, dtype=JOB_ROW)
        n_jobs = C.c_uint64(0)
        batch = _abi.Batch(kind.ctypes.data, reason.ctypes.data, flags.ctypes.data, stage.ctypes.data,
                           key.ctypes.data, n)
        out = _abi.Out(None, rows.ctypes.data, C.addressof(n_jobs))
        self._check(self._lib.nxs_classify(self._ctx, C.byref(batch), C.byref(out)))
        return None, rows[:n_jobs.value]

    def classify_packed(self, job_key, attr, out_verdict=None, out_rows=None, want_verdict=False):
        """Packed batch (job_key u64 + attr u16, see pack_columns): 10 bytes per record go to the GPU; by
        default only the job rows come back.  Returns (verdict or None, rows)."""
        key = np.ascontiguousarray(job_key, dtype=np.uint64)
        attr = np.ascontiguousarray(attr, dtype=np.uint16)
        n = len(key)
        if len(attr) != n:
            raise ValueError("columns differ in length")
        verdict = None
        if want_verdict:
            verdict = out_verdict if out_verdict is not None else np.empty(n, dtype=np.uint32)
        rows = out_rows if out_rows is not None else np.empty(n, dtype=JOB_ROW)
        n_jobs = C.c_uint64(0)
        batch = _abi.BatchPacked(key.ctypes.data, attr.ctypes.data, n)
        out = _abi.Out(verdict.ctypes.data if want_verdict else None, rows.ctypes.data, C.addressof(n_jobs))
        self._check(self._lib.nxs_classify_packed(self._ctx, C.byref(batch), C.byref(out)))
        return (verdict[:n] if want_verdict else None), rows[:n_jobs.value]

    # ----------------------------------------------------------- device path
    def classify_device(self, n, kind, reason_id, flags, cur_stage, job_key, verdict, job_rows, n_jobs):
        """All arguments after n are device pointers (ints).  Asynchronous."""
        batch = _abi.Batch(kind, reason_id, flags, cur_stage, job_key, n)
        out = _abi.Out(verdict, job_rows, n_jobs)
        self._check(self._lib.nxs_classify_device(self._ctx, C.byref(batch), C.byref(out)))

    def generate_device(self, workload, n, kind, reason_id, flags, cur_stage, job_key, capacity,
                        seed=DEFAULT_SEED, first=0, shard=0, n_shards=1):
        """Fill device columns with a synthetic stream; returns the number of records written."""
        cols = _abi.Columns(kind, reason_id, flags, cur_stage, job_key, capacity)
        kept = C.c_uint64(0)
        self._check(self._lib.nxs_generate_device(self._ctx, workload, seed, first, n, shard, n_shards,
                                                  C.byref(cols), C.byref(kept)))
        return int(kept.value)


class ShardedClassifier:
    """One host batch on several GPUs (nxs_classify_sharded): records are partitioned by
    splitmix64(job_key) % n_shards without splitting a job, each shard is classified by its own
    ctx (shard s on devices[s % len(devices)]) and the results come back in stream order."""

    def __init__(self, n_shards, devices=None, max_records=1 << 20, finished_mask=FINISHED_MASK_DEFAULT):
        self._lib = _abi.load()
        self._h = C.c_void_p()
        if devices is None:
            devices = [0]
        arr = (C.c_int * len(devices))(*devices)
        rc = self._lib.nxs_sharded_create(n_shards, arr, len(devices), max_records, finished_mask, C.byref(self._h))
        if rc != _abi.NXS_OK:
            self._h = None
            raise NxsError(rc, f"nxs_sharded_create(n_shards={n_shards}, devices={list(devices)})")
        self.n_shards = n_shards
        self.devices = list(devices)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.nxs_sharded_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def classify(self, cols, out_verdict=None, out_rows=None, want_verdict=True):
        """Same contract as Classifier.classify: (verdict u32[n], rows JOB_ROW[n_jobs]) in stream order; verdict is
        None with want_verdict=False (job rows only: nothing per record is produced, downloaded or merged)."""
        kind = np.ascontiguousarray(cols["kind"], dtype=np.uint8)
        reason = np.ascontiguousarray(cols["reason_id"], dtype=np.uint16)
        flags = np.ascontiguousarray(cols["flags"], dtype=np.uint8)
        stage = np.ascontiguousarray(cols["cur_stage"], dtype=np.uint8)
        key = np.ascontiguousarray(cols["job_key"], dtype=np.uint64)
        n = len(kind)
        if not (len(reason) == len(flags) == len(stage) == len(key) == n):
            raise ValueError("columns differ in length")
        verdict = None
        if want_verdict:
            verdict = out_verdict if out_verdict is not None else np.empty(n, dtype=np.uint32)
        rows = out_rows if out_rows is not None else np.empty(n, dtype=JOB_ROW)
        n_jobs = C.c_uint64(0)
        batch = _abi.Batch(kind.ctypes.data, reason.ctypes.data, flags.ctypes.data, stage.ctypes.data,
                           key.ctypes.data, n)
        out = _abi.Out(verdict.ctypes.data if want_verdict else None, rows.ctypes.data, C.addressof(n_jobs))
        rc = self._lib.nxs_classify_sharded(self._h, C.byref(batch), C.byref(out))
        if rc != _abi.NXS_OK:
            raise NxsError(rc, self._lib.nxs_sharded_last_error(self._h).decode())
        return (verdict[:n] if want_verdict else None), rows[:n_jobs.value]
