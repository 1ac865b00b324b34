"""hovl_filter_records_noalloc (include/hovl.h, "no-allocation entry point"): the record compaction of
hovl_run with its scan scratch in a file-scope __device__ array, for callers that may not cudaMalloc.
Checked through the C ABI against the existing entry point (hovl_run with filter_valid = 1 / 0), bit for
bit, on two calls in a row that share the static scratch.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from conftest import synth_reads

pytestmark = pytest.mark.gpu


def _noalloc(hovl, torch, recs, filter_valid, device=0):
    """records (host) -> device, one call of the entry point on torch's current stream -> (kept records, counts[5])"""
    L = hovl.lib()
    n = len(recs)
    raw = np.ascontiguousarray(recs).view(np.uint8).reshape(-1)
    d_in = torch.from_numpy(raw.copy()).cuda(device) if n else torch.zeros(64, dtype=torch.uint8, device="cuda:%d" % device)
    d_out = torch.full((max(n, 1) * hovl.RECORD_DTYPE.itemsize,), 0xEE, dtype=torch.uint8, device=d_in.device)
    d_cnt = torch.full((5,), -1, dtype=torch.int64, device=d_in.device)
    torch.cuda.synchronize(device)
    free0, _ = torch.cuda.mem_get_info(device)
    stream = torch.cuda.current_stream(device).cuda_stream
    rc = L.hovl_filter_records_noalloc(device, C.c_void_p(stream), C.c_void_p(d_in.data_ptr()), n, filter_valid,
                                       C.c_void_p(d_out.data_ptr()), C.c_void_p(d_cnt.data_ptr()))
    assert rc == hovl.HOVL_OK, (rc, L.hovl_last_error(None))
    torch.cuda.synchronize(device)
    free1, _ = torch.cuda.mem_get_info(device)
    assert free1 >= free0, "the call allocated device memory"
    cnt = d_cnt.cpu().numpy()
    kept = int(cnt[0])
    out = d_out.cpu().numpy()[:kept * hovl.RECORD_DTYPE.itemsize].view(hovl.RECORD_DTYPE)
    return out, cnt


def _check(hovl, torch, ctx, settings_kw, bases, offs):
    every = ctx.run(hovl.default_settings(filter_valid=0, **settings_kw), bases, offs)
    valid = ctx.run(hovl.default_settings(filter_valid=1, **settings_kw), bases, offs)
    assert len(every.records) > len(valid.records) > 0
    got, cnt = _noalloc(hovl, torch, every.records, 1)
    assert got.tobytes() == np.ascontiguousarray(valid.records).tobytes()          # bit-exact, same order
    st = valid.stats
    assert [int(x) for x in cnt] == [st["n_records"], st["n_duds"], st["n_align_valid"], st["n_ovl_valid"], 0]
    got0, cnt0 = _noalloc(hovl, torch, every.records, 0)                          # filter off: identity
    assert got0.tobytes() == np.ascontiguousarray(every.records).tobytes()
    assert int(cnt0[0]) == len(every.records) and [int(x) for x in cnt0[1:]] == [int(x) for x in cnt[1:]]
    return len(every.records), len(valid.records)


def test_noalloc_filter_matches_hovl_run_twice_in_a_row(crp177):
    import torch
    import hovl
    with hovl.Context(0) as ctx:
        # call 1 and 2: a synthetic set that spans several scan tiles (wide ids, canonical order)
        bases, offs = synth_reads(77, 1200, 150, 9000, sub=0.02, indel=0.004)
        n1 = _check(hovl, torch, ctx, dict(k=12, id_mode=1, order=1, max_ignore=150), bases, offs)
        assert n1[0] > 4096
        # call 3 and 4 straight after, smaller and in reference order: the shared scratch still holds the
        # previous call's tile sums and status word
        n2 = _check(hovl, torch, ctx, dict(k=12), *crp177)
        assert n2[0] < n1[0]
        # and the first input once more
        assert _check(hovl, torch, ctx, dict(k=12, id_mode=1, order=1, max_ignore=150), bases, offs) == n1


def test_noalloc_empty_and_limits():
    import torch
    import hovl
    L = hovl.lib()
    got, cnt = _noalloc(hovl, torch, np.zeros(0, hovl.RECORD_DTYPE), 1)
    assert len(got) == 0 and [int(x) for x in cnt] == [0, 0, 0, 0, 0]
    d = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
    too_many = (1 << 24) + 1
    rc = L.hovl_filter_records_noalloc(0, None, C.c_void_p(d.data_ptr()), too_many, 1, C.c_void_p(d.data_ptr()),
                                       C.c_void_p(d.data_ptr()))
    assert rc == hovl.E_UNSUPPORTED
    assert b"HOVL_NOALLOC_MAX_RECORDS" in L.hovl_last_error(None)
    assert L.hovl_filter_records_noalloc(0, None, None, 5, 1, None, C.c_void_p(d.data_ptr())) == hovl.E_ARG
    assert L.hovl_filter_records_noalloc(99, None, None, 0, 1, None, C.c_void_p(d.data_ptr())) == hovl.E_ARG
    # a record with status flags is reported, not dropped silently
    recs = np.zeros(3, hovl.RECORD_DTYPE)
    recs["flags"] = [hovl.F_ALIGN_VALID | hovl.F_OVL_VALID, hovl.F_SHORT_B, hovl.F_DUD]
    got, cnt = _noalloc(hovl, torch, recs, 1)
    assert len(got) == 1 and [int(x) for x in cnt] == [1, 1, 1, 1, 1]
