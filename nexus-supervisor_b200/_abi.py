"""ctypes declarations for include/nxs_classify.h (the C ABI of libnxs_classify.so).

The library is the product: there is no Python or CPU fallback.  If the shared
object has not been built, loading fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NXS_LIB") or os.path.join(_HERE, "lib", "libnxs_classify.so")  # NXS_LIB: tuning builds only

NXS_OK = 0
NXS_ERR_ARG, NXS_ERR_CAPACITY, NXS_ERR_CUDA, NXS_ERR_NO_DEVICE, NXS_ERR_NOMEM, NXS_ERR_UNKNOWN_ACTION = (
    -1, -2, -3, -4, -5, -6)
NXS_ERR_NOALLOC_LIMIT = -7
NXS_NOALLOC_MAX_RECORDS = 1 << 24

u8p, u16p, u32p, u64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint16, C.c_uint32, C.c_uint64))


class JobRow(C.Structure):
    _fields_ = [("job_key", C.c_uint64), ("verdict", C.c_uint32), ("winner_idx", C.c_uint32)]


class Batch(C.Structure):
    _fields_ = [("kind", C.c_void_p), ("reason_id", C.c_void_p), ("flags", C.c_void_p), ("cur_stage", C.c_void_p),
                ("job_key", C.c_void_p), ("n", C.c_uint64)]


class BatchPacked(C.Structure):
    _fields_ = [("job_key", C.c_void_p), ("attr", C.c_void_p), ("n", C.c_uint64)]


class Out(C.Structure):
    _fields_ = [("verdict", C.c_void_p), ("job_rows", C.c_void_p), ("n_jobs", C.c_void_p)]


class Columns(C.Structure):
    _fields_ = [("kind", C.c_void_p), ("reason_id", C.c_void_p), ("flags", C.c_void_p), ("cur_stage", C.c_void_p),
                ("job_key", C.c_void_p), ("capacity", C.c_uint64)]


# name -> (restype, argtypes); every symbol include/nxs_classify.h declares
SIGNATURES = {
    "nxs_create": (C.c_int, [C.c_int, C.c_uint64, C.c_uint16, C.POINTER(C.c_void_p)]),
    "nxs_destroy": (None, [C.c_void_p]),
    "nxs_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nxs_sync": (C.c_int, [C.c_void_p]),
    "nxs_strerror": (C.c_char_p, [C.c_int]),
    "nxs_last_error": (C.c_char_p, [C.c_void_p]),
    "nxs_abi_version": (C.c_int, []),
    "nxs_classify": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(Out)]),
    "nxs_classify_packed": (C.c_int, [C.c_void_p, C.POINTER(BatchPacked), C.POINTER(Out)]),
    "nxs_pack_columns": (None, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "nxs_classify_device": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(Out)]),
    "nxs_launch_count": (C.c_uint64, [C.c_void_p]),
    "nxs_last_tile": (C.c_int, [C.c_void_p]),
    "nxs_last_wire_bytes": (None, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "nxs_classify_device_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_uint16, C.POINTER(Batch), C.POINTER(Out)]),
    "nxs_alloc_host": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "nxs_free_host": (None, [C.c_void_p]),
    "nxs_intern_kind": (C.c_uint8, [C.c_char_p, C.c_size_t]),
    "nxs_intern_reason": (C.c_uint16, [C.c_char_p, C.c_size_t]),
    "nxs_ext_container_reason_to_event": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_uint8), C.POINTER(C.c_uint16)]),
    "nxs_reason_name": (C.c_char_p, [C.c_uint32]),
    "nxs_action_name": (C.c_char_p, [C.c_uint32]),
    "nxs_stage_name": (C.c_char_p, [C.c_uint32]),
    "nxs_msg_template": (C.c_char_p, [C.c_uint32]),
    "nxs_cause_template": (C.c_char_p, [C.c_uint32]),
    "nxs_encode_checkpoint_rows": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                             C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                                             u64p, u64p]),
    "nxs_shard_of": (C.c_uint32, [C.c_uint64, C.c_uint32]),
    "nxs_sharded_create": (C.c_int, [C.c_uint32, C.POINTER(C.c_int), C.c_uint32, C.c_uint64, C.c_uint16,
                                     C.POINTER(C.c_void_p)]),
    "nxs_sharded_destroy": (None, [C.c_void_p]),
    "nxs_classify_sharded": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(Out)]),
    "nxs_sharded_last_error": (C.c_char_p, [C.c_void_p]),
    "nxs_generate_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32,
                                      C.c_uint32, C.POINTER(Columns), u64p]),
}

_lib = None


def load():
    """dlopen libnxs_classify.so and attach prototypes.  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C nexus-supervisor_b200/csrc).  There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
    return _lib
