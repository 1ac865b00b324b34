"""ctypes loader for the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product package (sequence-aligner_b200/hovl) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Settings(C.Structure):
    """hovl_settings (include/hovl.h) = AlignSettings, src/ObjectStore.scala:17-36."""
    _fields_ = [("k", C.c_int32), ("cost", (C.c_int32 * 4) * 4),
                ("gap_open", C.c_int32), ("gap_extend", C.c_int32),
                ("min_overlap", C.c_int32), ("min_identity_f32", C.c_uint32),
                ("max_ignore_f32", C.c_uint32), ("min_coll", C.c_int32), ("max_coll", C.c_int32),
                ("head_f32", C.c_uint32), ("tail_f32", C.c_uint32),
                ("mid_lead_f32", C.c_uint32), ("mid_tail_f32", C.c_uint32),
                ("aligner", C.c_int32), ("order", C.c_int32), ("id_mode", C.c_int32),
                ("filter_valid", C.c_int32)]


RECORD_DTYPE = np.dtype([(n, "<i4") for n in (
    "a", "b", "count", "flags", "start_i", "start_j", "end_i", "end_j",
    "correct", "error", "score", "ahg", "bhg")])

STATS_FIELDS = ["n_reads", "n_bases", "n_kmers", "n_buckets", "max_bucket", "n_pair_instances",
                "n_distinct_pairs", "n_dispatched", "n_leads", "n_duds", "n_align_valid",
                "n_ovl_valid", "dp_cells", "n_records", "ns_h2d", "ns_pack", "ns_encode",
                "ns_kmer_sort", "ns_group", "ns_expand", "ns_pair_sort", "ns_pair_reduce",
                "ns_order", "ns_dp", "ns_d2h", "ns_total", "gpu_launches", "n_list_entries", "kmer_tuple_bytes", "dp_cells_skipped",
                "dp_cells_phase1"]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in STATS_FIELDS]


class _Result(C.Structure):
    _fields_ = [("recs", C.c_void_p), ("n_recs", C.c_uint64), ("st", Stats),
                ("kmer_code", C.POINTER(C.c_uint32)), ("kmer_rid", C.POINTER(C.c_uint32)),
                ("kmer_loc", C.POINTER(C.c_uint32)), ("n_kmers", C.c_uint64),
                ("pair_key", C.POINTER(C.c_uint64)), ("pair_count", C.POINTER(C.c_uint32)),
                ("n_pairs", C.c_uint64),
                ("disp_a", C.POINTER(C.c_uint32)), ("disp_b", C.POINTER(C.c_uint32)),
                ("disp_count", C.POINTER(C.c_uint32)), ("n_disp", C.c_uint64),
                ("ns_kmer", C.c_uint64), ("ns_table", C.c_uint64), ("ns_pairs", C.c_uint64),
                ("ns_dispatch", C.c_uint64), ("ns_align", C.c_uint64), ("ns_total", C.c_uint64)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        src = os.path.join(_HERE, "hovl_oracle.c")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            build()
        L = C.CDLL(path)
        L.orc_run.restype = C.c_int
        L.orc_run.argtypes = [C.POINTER(Settings), C.c_void_p, C.c_void_p, C.c_uint32, C.c_int,
                              C.c_int, C.POINTER(C.POINTER(_Result))]
        L.orc_result_free.argtypes = [C.POINTER(_Result)]
        L.orc_read_fasta.restype = C.c_int
        L.orc_read_fasta.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_uint32)]
        L.orc_read_hoxd.restype = C.c_int
        L.orc_read_hoxd.argtypes = [C.c_char_p, C.c_void_p]
        L.orc_settings_default.argtypes = [C.POINTER(Settings)]
        L.orc_settings_set_sections.argtypes = [C.POINTER(Settings), C.c_float, C.c_float]
        L.orc_align_pairs.restype = C.c_int
        L.orc_align_pairs.argtypes = [C.POINTER(Settings), C.c_void_p, C.c_void_p, C.c_uint32,
                                      C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p,
                                      C.POINTER(C.c_uint64)]
        L.orc_format_ovl.restype = C.c_uint64
        L.orc_format_ovl.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
        L.trove_new.restype = C.c_void_p
        L.trove_free.argtypes = [C.c_void_p]
        L.trove_put_if_absent.restype = C.POINTER(C.c_int32)
        L.trove_put_if_absent.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int)]
        L.trove_size.restype = C.c_int32
        L.trove_size.argtypes = [C.c_void_p]
        L.trove_capacity.restype = C.c_int32
        L.trove_capacity.argtypes = [C.c_void_p]
        L.trove_iterate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.free = C.CDLL(None).free
        L.free.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def f32_bits(x):
    return int(np.float32(x).view(np.uint32))


def default_settings(**kw):
    """Project4 defaults (src/Project4.scala:103-114); keyword overrides mirror the CLI."""
    s = Settings()
    lib().orc_settings_default(C.byref(s))
    edge = kw.pop("kmer_edge", None)
    center = kw.pop("kmer_center", None)
    if edge is not None or center is not None:
        lib().orc_settings_set_sections(C.byref(s), C.c_float(0.4 if edge is None else edge),
                                        C.c_float(0.4 if center is None else center))
    if "min_identity" in kw:
        v = np.float32(kw.pop("min_identity"))
        if v >= 1:
            v = np.float32(v * np.float32(0.01))          # src/Project4.scala:144-146
        s.min_identity_f32 = int(v.view(np.uint32))
    if "max_ignore" in kw:
        s.max_ignore_f32 = f32_bits(float(abs(int(kw.pop("max_ignore")))))
    if "cost" in kw:
        m = kw.pop("cost")
        for i in range(4):
            for j in range(4):
                s.cost[i][j] = int(m[i][j])
    for key, val in kw.items():
        if not hasattr(s, key):
            raise KeyError(key)
        setattr(s, key, int(val))
    return s


def read_fasta(path):
    """readSeq, src/BioLibs.scala:26-50 -> (bases uint8[], offsets uint64[n+1])."""
    L = lib()
    pb, po, n = C.c_void_p(), C.c_void_p(), C.c_uint32()
    rc = L.orc_read_fasta(os.fsencode(path), C.byref(pb), C.byref(po), C.byref(n))
    if rc != 0:
        raise ValueError("Invalid Sequence File: %s" % path)
    offs = np.ctypeslib.as_array(C.cast(po, C.POINTER(C.c_uint64)), (n.value + 1,)).copy()
    total = int(offs[-1])
    bases = (np.ctypeslib.as_array(C.cast(pb, C.POINTER(C.c_uint8)), (max(total, 1),))[:total]).copy()
    L.free(pb)
    L.free(po)
    return bases, offs


def read_hoxd(path):
    m = ((C.c_int32 * 4) * 4)()
    rc = lib().orc_read_hoxd(os.fsencode(path), m)
    if rc != 0:
        raise ValueError("bad HOXD file: %s" % path)
    return [[m[i][j] for j in range(4)] for i in range(4)]


def _arr(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, (n,)).astype(dtype, copy=True)


class OracleResult:
    pass


def run(settings, bases, offsets, threads=1, keep_stages=False):
    """Whole path on the CPU.  Returns an object with .records (structured array), .stats
    (dict), .rc and, with keep_stages, the stage products."""
    L = lib()
    bases = np.ascontiguousarray(bases, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n = len(offsets) - 1
    pr = C.POINTER(_Result)()
    rc = L.orc_run(C.byref(settings), bases.ctypes.data, offsets.ctypes.data, n, threads,
                   1 if keep_stages else 0, C.byref(pr))
    if not pr:
        raise RuntimeError("oracle failed rc=%d" % rc)
    r = pr.contents
    out = OracleResult()
    out.rc = rc
    nrec = int(r.n_recs)
    if nrec:
        buf = C.string_at(r.recs, nrec * RECORD_DTYPE.itemsize)
        out.records = np.frombuffer(buf, RECORD_DTYPE).copy()
    else:
        out.records = np.zeros(0, RECORD_DTYPE)
    out.stats = {n_: int(getattr(r.st, n_)) for n_ in STATS_FIELDS}
    out.ns = {k: int(getattr(r, k)) for k in ("ns_kmer", "ns_table", "ns_pairs", "ns_dispatch",
                                               "ns_align", "ns_total")}
    if keep_stages:
        nk, np_, nd = int(r.n_kmers), int(r.n_pairs), int(r.n_disp)
        out.kmer_code = _arr(r.kmer_code, nk, np.uint32)
        out.kmer_rid = _arr(r.kmer_rid, nk, np.uint32)
        out.kmer_loc = _arr(r.kmer_loc, nk, np.uint32)
        out.pair_key = _arr(r.pair_key, np_, np.uint64)
        out.pair_count = _arr(r.pair_count, np_, np.uint32)
        out.disp_a = _arr(r.disp_a, nd, np.uint32)
        out.disp_b = _arr(r.disp_b, nd, np.uint32)
        out.disp_count = _arr(r.disp_count, nd, np.uint32)
    L.orc_result_free(pr)
    return out


def align_pairs(settings, bases, offsets, pa, pb, threads=1):
    L = lib()
    bases = np.ascontiguousarray(bases, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    pa = np.ascontiguousarray(pa, np.uint32)
    pb = np.ascontiguousarray(pb, np.uint32)
    out = np.zeros(len(pa), RECORD_DTYPE)
    cells = C.c_uint64()
    rc = L.orc_align_pairs(C.byref(settings), bases.ctypes.data, offsets.ctypes.data,
                           len(offsets) - 1, pa.ctypes.data, pb.ctypes.data, len(pa), threads,
                           out.ctypes.data, C.byref(cells))
    if rc != 0:
        raise ValueError("orc_align_pairs rc=%d" % rc)
    return out, int(cells.value)


def format_ovl(records):
    L = lib()
    records = np.ascontiguousarray(records, RECORD_DTYPE)
    need = L.orc_format_ovl(records.ctypes.data, len(records), None, 0)
    buf = C.create_string_buffer(int(need) + 1)
    L.orc_format_ovl(records.ctypes.data, len(records), buf, need)
    return buf.raw[:need]
