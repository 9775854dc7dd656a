"""GPU test of nxs_classify_device_noalloc, the entry point for hosts that cannot let the library
allocate device memory (their own pool owns the HBM budget, or a cudaMalloc/cudaFree stall is not
acceptable on their loop).  It keeps the kernel's scratch in a static __device__ array, so every call
shares that array and must clear what it uses.  The test drives it through the C ABI several times in a row —
a large batch, then a smaller, different one, so stale descriptors of the first call would show, then the
largest batch it accepts — and
requires bit-identical output to the ordinary entry point (nxs_classify_device) AND to the CPU oracle run
on the same columns, an unchanged amount of free device memory across the second call, and a clean
refusal above NXS_NOALLOC_MAX_RECORDS.
The file name makes pytest collect it last; it also passes when run alone in a fresh process."""
import ctypes as C

import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle
from tests import util

pytestmark = pytest.mark.gpu


def _upload(torch, cols):
    """Device copies of the five columns (allocated by the TEST, not by the entry point)."""
    t = {k: torch.from_numpy(v.view(np.uint8).copy()).cuda() for k, v in cols.items()}
    return t, {k: x.data_ptr() for k, x in t.items()}


def _outputs(torch, n):
    return (torch.zeros(n, dtype=torch.int32, device="cuda"), torch.zeros((n, 2), dtype=torch.int64, device="cuda"),
            torch.zeros(1, dtype=torch.int64, device="cuda"))


def _call_noalloc(lib, stream, fm, ptr, n, verdict, rows, n_jobs):
    batch = nx._abi.Batch(ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"], n)
    out = nx._abi.Out(verdict.data_ptr(), rows.data_ptr(), n_jobs.data_ptr())
    return lib.nxs_classify_device_noalloc(0, C.c_void_p(stream), fm, C.byref(batch), C.byref(out))


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    import torch
    lib = nx._abi.load()
    stream = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(2024)
    # (the third batch is the largest the entry point takes: its last tile uses the last descriptor of the static scratch)
    cases = [(3_000_001, 7, util.FM_DEFAULT, True), (70_003, 900, 0x5A5A, False),
             (nx._abi.NXS_NOALLOC_MAX_RECORDS, 4, util.FM_DEFAULT, False)]
    with nx.Classifier(device=0, max_records=1 << 24) as plain, \
            nx.Classifier(device=0, max_records=1 << 22, finished_mask=0x5A5A) as masked:
        for call, (n, max_len, fm, wide) in enumerate(cases):
            cols = util.random_cols(rng, n, max_len, wide=wide)
            keep, ptr = _upload(torch, cols)
            want_v, want_r, want_n = _outputs(torch, n)
            ref = plain if fm == util.FM_DEFAULT else masked
            ref.set_stream(stream)
            ref.classify_device(n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"],
                                want_v.data_ptr(), want_r.data_ptr(), want_n.data_ptr())
            got_v, got_r, got_n = _outputs(torch, n)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            rc = _call_noalloc(lib, stream, fm, ptr, n, got_v, got_r, got_n)
            assert rc == nx._abi.NXS_OK, lib.nxs_strerror(rc)
            torch.cuda.synchronize()
            if call > 0:  # the first call may load the kernel image; later calls must not change free memory
                assert torch.cuda.mem_get_info()[0] == free_before
            jobs = int(want_n.item())
            assert jobs > 0 and int(got_n.item()) == jobs
            assert torch.equal(got_v, want_v), f"call {call}: verdicts differ"
            assert torch.equal(got_r[:jobs], want_r[:jobs]), f"call {call}: job rows differ"
            # ... and the oracle, directly (the columns are on the host already)
            ora_v, ora_r = cpu_oracle.classify(cols, fm, threads=4)
            host_r = got_r[:jobs].cpu().numpy().view(np.uint64).reshape(-1, 2).copy().view(cpu_oracle.JOB_ROW).reshape(-1)
            util.assert_same(got_v.cpu().numpy().view(np.uint32), host_r, ora_v, ora_r, f"no-alloc call {call} vs oracle")
            # job rows only (out->verdict == NULL), as on the other entry points
            rows_only, n_only = torch.zeros((n, 2), dtype=torch.int64, device="cuda"), torch.zeros(1, dtype=torch.int64, device="cuda")
            batch = nx._abi.Batch(ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"], n)
            out = nx._abi.Out(None, rows_only.data_ptr(), n_only.data_ptr())
            assert lib.nxs_classify_device_noalloc(0, C.c_void_p(stream), fm, C.byref(batch), C.byref(out)) == nx._abi.NXS_OK
            torch.cuda.synchronize()
            assert int(n_only.item()) == jobs and torch.equal(rows_only[:jobs], want_r[:jobs])
            del keep


def test_noalloc_refuses_batches_beyond_its_static_scratch():
    import torch
    lib = nx._abi.load()
    n_jobs = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    p = buf.data_ptr()
    batch = nx._abi.Batch(p, p, p, p, p, nx._abi.NXS_NOALLOC_MAX_RECORDS + 1)
    out = nx._abi.Out(p, p, n_jobs.data_ptr())
    rc = lib.nxs_classify_device_noalloc(0, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                                         util.FM_DEFAULT, C.byref(batch), C.byref(out))
    torch.cuda.synchronize()
    assert rc == nx._abi.NXS_ERR_NOALLOC_LIMIT
    assert b"NXS_NOALLOC_MAX_RECORDS" in lib.nxs_strerror(rc)
    assert int(n_jobs.item()) == -1 and int(buf.sum().item()) == 0  # nothing was launched or written
