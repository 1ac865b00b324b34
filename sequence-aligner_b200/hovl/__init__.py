"""hovl — host-side mirror of the reference's hash-overlap interface over libhovl.so.

The reference (rohit507/Sequence-Aligner, Scala) exposes this path as
    generateKmerTable(file)                 src/Project4.scala:508
    genAlignment(table, settings, filter)   src/Project4.scala:567
consumed by calcOverlaps (src/Project4.scala:795).  Here the same thing is one call into the
C ABI of include/hovl.h (CUDA, sm_100a).  There is NO CPU fallback: importing works anywhere,
but every compute call needs libhovl.so and a CUDA device and raises otherwise.
"""
import ctypes as C
import weakref
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libhovl.so")

HOVL_OK, E_ARG, E_UNSUPPORTED, E_SHORT_READ, E_BAD_BASE, E_CUDA, E_NOMEM, E_NCCL, E_NO_DEVICE = (
    0, -1, -2, -3, -4, -5, -6, -7, -8)
F_DUD, F_ALIGN_VALID, F_OVL_VALID, F_SHORT_B, F_BAD_BASE = 1, 2, 4, 8, 16


class Settings(C.Structure):
    """hovl_settings = AlignSettings, src/ObjectStore.scala:17-36."""
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


class HovlError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("hovl error %d: %s" % (code, msg))
        self.code = code


_LIB = None
EXPORTS = ["hovl_settings_default", "hovl_settings_set_sections", "hovl_create", "hovl_destroy",
           "hovl_last_error", "hovl_run", "hovl_run_device", "hovl_records", "hovl_stats",
           "hovl_result_free", "hovl_format_ovl", "hovl_format_test_overlaps", "hovl_format_float", "hovl_set_option", "hovl_stage_table",
           "hovl_stage_pairs", "hovl_stage_align", "hovl_free", "hovl_trove_order",
           "hovl_kernel_times", "hovl_shard_of_code", "hovl_shard_of_read", "hovl_shard_encode",
           "hovl_shard_pairs", "hovl_shard_fetch_pairs", "hovl_shard_align",
           "hovl_shard_scatter", "hovl_peer_alloc", "hovl_peer_open", "hovl_peer_close", "hovl_peer_free",
           "hovl_reads_open", "hovl_reads_bases", "hovl_reads_offsets", "hovl_reads_free",
           "hovl_read_hoxd", "hovl_parse_int", "hovl_write_ovl", "hovl_io_last_error",
           "hovl_filter_records_noalloc", "hovl_stage_cover", "hovl_shard_of_code_range", "hovl_shard_sort",
           "hovl_shard_gather_peer", "hovl_shard_pairs_local", "hovl_shard_fetch_records", "hovl_trove_order_device",
           "hovl_shard_select", "hovl_shard_split", "hovl_shard_fetch_selected", "hovl_shard_align_pairs"]
CLI_PATH = os.path.join(os.path.dirname(_HERE), "hovl_cli")


def lib():
    """Load libhovl.so (built in-tree by __graft_entry__.build()).  Fails loudly if absent."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise HovlError(E_NO_DEVICE, "libhovl.so not built (%s); run __graft_entry__.build(). "
                        "There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    PS = C.POINTER(Settings)
    L.hovl_settings_default.argtypes = [PS]
    L.hovl_settings_default.restype = None
    L.hovl_settings_set_sections.argtypes = [PS, u32, u32]
    L.hovl_settings_set_sections.restype = None
    L.hovl_create.argtypes = [C.POINTER(vp), C.POINTER(C.c_int), i32]
    L.hovl_create.restype = i32
    L.hovl_destroy.argtypes = [vp]
    L.hovl_destroy.restype = None
    L.hovl_last_error.argtypes = [vp]
    L.hovl_last_error.restype = C.c_char_p
    L.hovl_run.argtypes = [vp, PS, vp, vp, u32, C.POINTER(vp)]
    L.hovl_run.restype = i32
    L.hovl_run_device.argtypes = [vp, PS, vp, vp, u32, u64, C.POINTER(vp)]
    L.hovl_run_device.restype = i32
    L.hovl_records.argtypes = [vp, C.POINTER(u64)]
    L.hovl_records.restype = vp
    L.hovl_stats.argtypes = [vp, C.POINTER(Stats)]
    L.hovl_stats.restype = i32
    L.hovl_result_free.argtypes = [vp]
    L.hovl_result_free.restype = None
    L.hovl_format_ovl.argtypes = [vp, u64, vp, u64]
    L.hovl_format_ovl.restype = u64
    L.hovl_format_test_overlaps.argtypes = [vp, u64, vp, vp, C.c_uint32, vp, u64]
    L.hovl_format_test_overlaps.restype = u64
    L.hovl_format_float.argtypes = [C.c_float, vp, u64]
    L.hovl_format_float.restype = u64
    L.hovl_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.hovl_set_option.restype = i32
    L.hovl_stage_table.argtypes = [vp, PS, vp, vp, u32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(u64)]
    L.hovl_stage_table.restype = i32
    L.hovl_stage_pairs.argtypes = [vp, PS, vp, vp, u32, C.POINTER(vp), C.POINTER(vp), C.POINTER(u64), C.POINTER(u64)]
    L.hovl_stage_pairs.restype = i32
    L.hovl_stage_align.argtypes = [vp, PS, vp, vp, u32, vp, vp, u64, vp]
    L.hovl_stage_align.restype = i32
    L.hovl_free.argtypes = [vp]
    L.hovl_free.restype = None
    L.hovl_kernel_times.argtypes = [vp, u32, C.c_char_p, u32, C.POINTER(u64), C.POINTER(u64)]
    L.hovl_kernel_times.restype = i32
    L.hovl_trove_order.argtypes = [vp, u64, vp]
    L.hovl_trove_order.restype = i32
    L.hovl_shard_select.argtypes = [vp, PS, vp, vp, u64, vp, u32, C.POINTER(u64), C.POINTER(u64)]
    L.hovl_shard_select.restype = i32
    L.hovl_shard_split.argtypes = [vp, u64, u64, u32, vp]
    L.hovl_shard_split.restype = i32
    L.hovl_shard_fetch_selected.argtypes = [vp, vp, u64]
    L.hovl_shard_fetch_selected.restype = i32
    L.hovl_shard_align_pairs.argtypes = [vp, PS, vp, u64, vp, u64, vp, vp, u32, vp, u64, C.POINTER(vp)]
    L.hovl_shard_align_pairs.restype = i32
    L.hovl_trove_order_device.argtypes = [vp, vp, u64, vp]
    L.hovl_trove_order_device.restype = i32
    L.hovl_shard_of_code.argtypes = [u32, u32]
    L.hovl_shard_of_code.restype = u32
    L.hovl_shard_of_read.argtypes = [u32, u32, u32]
    L.hovl_shard_of_read.restype = u32
    L.hovl_shard_encode.argtypes = [vp, PS, vp, vp, u32, u64, u32, u32, u32, vp, vp, u64, vp, u64, vp, vp, vp, C.POINTER(u64)]
    L.hovl_shard_encode.restype = i32
    L.hovl_shard_pairs.argtypes = [vp, PS, vp, vp, u64, u32, u32, vp, C.POINTER(u64)]
    L.hovl_shard_pairs.restype = i32
    L.hovl_shard_fetch_pairs.argtypes = [vp, vp, vp, u64]
    L.hovl_shard_fetch_pairs.restype = i32
    L.hovl_shard_align.argtypes = [vp, PS, vp, vp, u64, vp, u64, vp, vp, u32, vp, u64, C.POINTER(vp)]
    L.hovl_shard_align.restype = i32
    L.hovl_shard_scatter.argtypes = [vp, u32, vp, vp, vp]
    L.hovl_shard_scatter.restype = i32
    L.hovl_peer_alloc.argtypes = [vp, u64, C.POINTER(vp), vp]
    L.hovl_peer_alloc.restype = i32
    L.hovl_peer_open.argtypes = [vp, vp, C.POINTER(vp)]
    L.hovl_peer_open.restype = i32
    L.hovl_peer_close.argtypes = [vp, vp]
    L.hovl_peer_close.restype = i32
    L.hovl_peer_free.argtypes = [vp, vp]
    L.hovl_peer_free.restype = i32
    L.hovl_reads_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.hovl_reads_open.restype = i32
    L.hovl_reads_bases.argtypes = [vp, C.POINTER(u64)]
    L.hovl_reads_bases.restype = vp
    L.hovl_reads_offsets.argtypes = [vp, C.POINTER(u32)]
    L.hovl_reads_offsets.restype = vp
    L.hovl_reads_free.argtypes = [vp]
    L.hovl_reads_free.restype = None
    L.hovl_read_hoxd.argtypes = [C.c_char_p, vp]
    L.hovl_read_hoxd.restype = i32
    L.hovl_parse_int.argtypes = [C.c_char_p, C.POINTER(C.c_int32)]
    L.hovl_parse_int.restype = i32
    L.hovl_write_ovl.argtypes = [C.c_char_p, vp, u64]
    L.hovl_write_ovl.restype = i32
    L.hovl_io_last_error.argtypes = []
    L.hovl_io_last_error.restype = C.c_char_p
    L.hovl_shard_of_code_range.argtypes = [u32, C.c_int32, u32]
    L.hovl_shard_of_code_range.restype = u32
    L.hovl_shard_sort.argtypes = [vp, PS, vp, vp, u64]
    L.hovl_shard_sort.restype = i32
    L.hovl_shard_gather_peer.argtypes = [vp, u32, vp, vp, vp, vp, vp]
    L.hovl_shard_gather_peer.restype = i32
    L.hovl_shard_pairs_local.argtypes = [vp, PS, vp, vp, u64, vp, u64, vp, vp, u32, u32, u32, C.POINTER(vp)]
    L.hovl_shard_pairs_local.restype = i32
    L.hovl_shard_fetch_records.argtypes = [vp, vp, u64]
    L.hovl_shard_fetch_records.restype = i32
    L.hovl_stage_cover.argtypes = [vp, PS, vp, vp, u32, C.POINTER(u64), C.POINTER(vp), C.POINTER(vp), C.POINTER(u64)]
    L.hovl_stage_cover.restype = i32
    L.hovl_filter_records_noalloc.argtypes = [i32, vp, vp, u64, i32, vp, vp]
    L.hovl_filter_records_noalloc.restype = i32
    _LIB = L
    return L


class Reads:
    """hovl_reads: a FASTA file read natively with readSeq's semantics (src/BioLibs.scala:26-50) into
    page-locked host memory.  `bases` / `offsets` are zero-copy views that live as long as this object."""

    def __init__(self, path):
        L = lib()
        self._h = C.c_void_p()
        rc = L.hovl_reads_open(os.fsencode(path), C.byref(self._h))
        if rc != HOVL_OK:
            raise HovlError(rc, (L.hovl_io_last_error() or b"").decode())
        nb, nr = C.c_uint64(), C.c_uint32()
        pb = L.hovl_reads_bases(self._h, C.byref(nb))
        po = L.hovl_reads_offsets(self._h, C.byref(nr))
        self.n_reads = int(nr.value)
        self.bases = np.ctypeslib.as_array(C.cast(pb, C.POINTER(C.c_uint8)), shape=(max(int(nb.value), 1),))[:int(nb.value)]
        self.offsets = np.ctypeslib.as_array(C.cast(po, C.POINTER(C.c_uint64)), shape=(self.n_reads + 1,))

    def close(self):
        if getattr(self, "_h", None) and _LIB is not None:
            self.bases = self.offsets = None
            _LIB.hovl_reads_free(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def read_hoxd_native(path):
    """hovl_read_hoxd = readHOXD (src/BioLibs.scala:66-114); 4x4 list, A,C,G,T order."""
    m = ((C.c_int32 * 4) * 4)()
    rc = lib().hovl_read_hoxd(os.fsencode(path), C.cast(m, C.c_void_p))
    if rc != HOVL_OK:
        raise HovlError(rc, (lib().hovl_io_last_error() or b"").decode())
    return [[int(m[i][j]) for j in range(4)] for i in range(4)]


def trove_order(keys):
    """Iteration order (indices into `keys`) of a Trove 3.0.3 TIntObjectHashMap filled in this order."""
    keys = np.ascontiguousarray(keys, np.int32)
    out = np.zeros(len(keys), np.uint32)
    rc = lib().hovl_trove_order(keys.ctypes.data, len(keys), out.ctypes.data)
    if rc != HOVL_OK:
        raise HovlError(rc, "hovl_trove_order")
    return out


def f32_bits(x):
    return int(np.float32(x).view(np.uint32))


def default_settings(**kw):
    """Project4 defaults (src/Project4.scala:103-114, :41); keyword overrides mirror the CLI flags:
    k, cost (4x4 A,C,G,T), gap_open, gap_extend, min_overlap, min_identity, max_ignore, min_coll,
    max_coll, kmer_edge, kmer_center, order, id_mode, filter_valid."""
    s = Settings()
    lib().hovl_settings_default(C.byref(s))
    edge, center = kw.pop("kmer_edge", None), kw.pop("kmer_center", None)
    if edge is not None or center is not None:
        lib().hovl_settings_set_sections(C.byref(s), f32_bits(abs(0.4 if edge is None else edge)),
                                         f32_bits(abs(0.4 if center is None else center)))
    if "min_identity" in kw:
        v = np.float32(kw.pop("min_identity"))
        if v >= 1:
            v = np.float32(v * np.float32(0.01))           # src/Project4.scala:144-146
        s.min_identity_f32 = int(v.view(np.uint32))
    if "max_ignore" in kw:
        s.max_ignore_f32 = f32_bits(float(abs(int(kw.pop("max_ignore")))))
    if "cost" in kw:
        m = kw.pop("cost")
        for i in range(4):
            for j in range(4):
                s.cost[i][j] = int(m[i][j])
    for key in ("gap_open", "gap_extend"):
        if key in kw:
            setattr(s, key, -abs(int(kw.pop(key))))         # src/Project4.scala:160-165
    for key, val in kw.items():
        if not hasattr(s, key):
            raise KeyError(key)
        setattr(s, key, int(val))
    return s


class Result:
    """Seq[Alignment] of genAlignment as a structured array (+ run statistics)."""

    def __init__(self, records, stats, rc, message=""):
        self.records, self.stats, self.rc, self.message = records, stats, rc, message

    def ovl(self):
        return format_ovl(self.records)


def format_ovl(records):
    """calcOverlaps + Overlap.print (src/Project4.scala:806-820, src/ObjectStore.scala:127-135)."""
    L = lib()
    records = np.ascontiguousarray(records, RECORD_DTYPE)
    need = L.hovl_format_ovl(records.ctypes.data, len(records), None, 0)
    buf = C.create_string_buffer(int(need) + 1)
    L.hovl_format_ovl(records.ctypes.data, len(records), buf, need)
    return buf.raw[:need]


def format_test_overlaps(records, bases, offsets):
    """testOverlaps (src/Project4.scala:484-504, --test-overlaps) for records made with filter_valid = 0."""
    L = lib()
    records = np.ascontiguousarray(records, RECORD_DTYPE)
    bases = np.ascontiguousarray(bases, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    args = (records.ctypes.data, len(records), bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1)
    need = L.hovl_format_test_overlaps(*args, None, 0)
    buf = C.create_string_buffer(int(need) + 1)
    L.hovl_format_test_overlaps(*args, buf, need)
    return buf.raw[:need]


def format_float(x):
    """java.lang.Float.toString."""
    buf = C.create_string_buffer(64)
    lib().hovl_format_float(float(np.float32(x)), buf, 64)
    return buf.value.decode()


def _take(ptr, n, dtype):
    n = int(n)
    out = np.zeros(n, dtype)
    if n:
        C.memmove(out.ctypes.data, ptr.value, n * out.itemsize)
    lib().hovl_free(ptr)
    return out


class Context:
    """hovl_ctx: one CUDA device, or (device = a list of ordinals) several devices driven from this process — `run` then
    shards the work over them through peer memory (csrc/multi.cu; needs id_mode = 1, order = 1).
    Raises HovlError if no device / library."""

    def __init__(self, device=0):
        L = lib()
        self._h = C.c_void_p()
        devs = list(device) if isinstance(device, (list, tuple)) else [int(device)]
        arr = (C.c_int * len(devs))(*devs)
        rc = L.hovl_create(C.byref(self._h), arr, len(devs))
        if rc != HOVL_OK:
            raise HovlError(rc, (L.hovl_last_error(None) or b"").decode())

    def close(self):
        if getattr(self, "_h", None) and _LIB is not None:      # _LIB is gone at interpreter shutdown
            _LIB.hovl_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _err(self):
        return (lib().hovl_last_error(self._h) or b"").decode()

    def trove_order(self, keys):
        """trove_order computed on the device (hovl_trove_order_device)."""
        keys = np.ascontiguousarray(keys, np.int32)
        out = np.zeros(len(keys), np.uint32)
        rc = lib().hovl_trove_order_device(self._h, keys.ctypes.data, len(keys), out.ctypes.data)
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return out

    def set_option(self, name, value):
        rc = lib().hovl_set_option(self._h, name.encode(), int(value))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    def kernel_times(self):
        """Per-kernel device time of the last run (after set_option("kernel_timing", 1)):
        {"stage:kernel": (launches, ns)}."""
        L, out = lib(), {}
        n = L.hovl_kernel_times(self._h, 0xffffffff, None, 0, None, None)
        for i in range(max(n, 0)):
            buf, cnt, ns = C.create_string_buffer(160), C.c_uint64(), C.c_uint64()
            L.hovl_kernel_times(self._h, i, buf, 160, C.byref(cnt), C.byref(ns))
            out[buf.value.decode()] = (int(cnt.value), int(ns.value))
        return out

    def _collect(self, rc, pres, strict):
        L = lib()
        if not pres:
            raise HovlError(rc, self._err())
        n = C.c_uint64()
        p = L.hovl_records(pres, C.byref(n))
        st = Stats()
        L.hovl_stats(pres, C.byref(st))
        msg = self._err()
        if rc != HOVL_OK and strict:
            L.hovl_result_free(pres)
            raise HovlError(rc, msg)
        # the records stay in the library's page-locked buffer (as a JVM host would read them
        # through a direct ByteBuffer); hovl_result_free runs when the array is dropped
        nbytes = int(n.value) * RECORD_DTYPE.itemsize
        if nbytes:
            raw = (C.c_uint8 * nbytes).from_address(p)
            recs = np.frombuffer(raw, dtype=RECORD_DTYPE)
            recs.flags.writeable = False
            weakref.finalize(raw, L.hovl_result_free, C.c_void_p(pres.value))
        else:
            recs = np.zeros(0, RECORD_DTYPE)
            L.hovl_result_free(pres)
        return Result(recs, {k: int(getattr(st, k)) for k in STATS_FIELDS}, rc, msg)

    def run(self, settings, bases, offsets, strict=True):
        """generateKmerTable + genAlignment on host arrays (bases uint8 ASCII, offsets uint64[n+1])."""
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        pres = C.c_void_p()
        rc = lib().hovl_run(self._h, C.byref(settings), bases.ctypes.data, offsets.ctypes.data,
                            len(offsets) - 1, C.byref(pres))
        return self._collect(rc, pres, strict)

    def run_device(self, settings, d_bases, d_offsets, n_reads, n_bases, strict=True):
        """Same with the inputs already in device memory (raw device pointers as ints)."""
        pres = C.c_void_p()
        rc = lib().hovl_run_device(self._h, C.byref(settings), C.c_void_p(d_bases), C.c_void_p(d_offsets),
                                   int(n_reads), int(n_bases), C.byref(pres))
        return self._collect(rc, pres, strict)

    # ---- sharded path (device pointers as ints; see hovl/dist.py) ----
    def shard_encode(self, settings, d_bases, d_offsets, n_local, n_bases, id_base, n_total, n_parts,
                     d_codes, d_vals, tuple_cap, d_packed, packed_cap, d_len, d_bad):
        counts = (C.c_uint64 * n_parts)()
        nw = C.c_uint64()
        rc = lib().hovl_shard_encode(self._h, C.byref(settings), d_bases, d_offsets, n_local, n_bases, id_base, n_total,
                                     n_parts, d_codes, d_vals, tuple_cap, d_packed, packed_cap, d_len, d_bad,
                                     C.cast(counts, C.c_void_p), C.byref(nw))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return [int(x) for x in counts], int(nw.value)

    def shard_scatter(self, n_parts, peer_codes, peer_vals, peer_base):
        a = (C.c_uint64 * n_parts)(*[int(x) for x in peer_codes])
        b = (C.c_uint64 * n_parts)(*[int(x) for x in peer_vals])
        o = (C.c_uint64 * n_parts)(*[int(x) for x in peer_base])
        rc = lib().hovl_shard_scatter(self._h, n_parts, C.cast(a, C.c_void_p), C.cast(b, C.c_void_p), C.cast(o, C.c_void_p))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    def peer_alloc(self, nbytes):
        """(device address, 64-byte IPC handle) of a buffer other ranks of the node can map."""
        p, hd = C.c_void_p(), (C.c_uint8 * 64)()
        rc = lib().hovl_peer_alloc(self._h, int(nbytes), C.byref(p), C.cast(hd, C.c_void_p))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return int(p.value), bytes(hd)

    def peer_open(self, handle):
        p, hd = C.c_void_p(), (C.c_uint8 * 64)(*handle)
        rc = lib().hovl_peer_open(self._h, C.cast(hd, C.c_void_p), C.byref(p))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return int(p.value)

    def peer_close(self, ptr):
        lib().hovl_peer_close(self._h, C.c_void_p(ptr))

    def peer_free(self, ptr):
        lib().hovl_peer_free(self._h, C.c_void_p(ptr))

    def shard_pairs(self, settings, d_codes, d_vals, n_tuples, n_total, n_parts):
        counts = (C.c_uint64 * n_parts)()
        npart = C.c_uint64()
        rc = lib().hovl_shard_pairs(self._h, C.byref(settings), d_codes, d_vals, n_tuples, n_total, n_parts,
                                    C.cast(counts, C.c_void_p), C.byref(npart))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return [int(x) for x in counts], int(npart.value)

    def shard_fetch_pairs(self, d_keys, d_counts, cap):
        rc = lib().hovl_shard_fetch_pairs(self._h, d_keys, d_counts, cap)
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    # alignment re-balance by DP cells (include/hovl.h)
    def shard_select(self, settings, d_keys, d_counts, n_partial, d_len, n_total):
        n, cost = C.c_uint64(), C.c_uint64()
        rc = lib().hovl_shard_select(self._h, C.byref(settings), d_keys, d_counts, n_partial, d_len, n_total, C.byref(n), C.byref(cost))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return int(n.value), int(cost.value)

    def shard_split(self, cost_before, cost_total, n_parts):
        counts = (C.c_uint64 * n_parts)()
        rc = lib().hovl_shard_split(self._h, int(cost_before), int(cost_total), n_parts, C.cast(counts, C.c_void_p))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return [int(x) for x in counts]

    def shard_fetch_selected(self, d_abc, cap_pairs):
        rc = lib().hovl_shard_fetch_selected(self._h, d_abc, cap_pairs)
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    def shard_align_pairs(self, settings, d_abc, n_pairs, d_packed, n_words, d_len, d_bad, n_total, d_records=None,
                          records_cap=0, strict=True):
        pres = C.c_void_p()
        rc = lib().hovl_shard_align_pairs(self._h, C.byref(settings), d_abc, n_pairs, d_packed, n_words, d_len, d_bad, n_total,
                                          d_records, records_cap, C.byref(pres))
        return self._collect(rc, pres, strict)

    # replicated-table variant (include/hovl.h)
    def shard_sort(self, settings, d_codes, d_vals, n_tuples):
        rc = lib().hovl_shard_sort(self._h, C.byref(settings), d_codes, d_vals, n_tuples)
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    def shard_gather_peer(self, n_parts, peer_codes, peer_vals, counts, d_codes_all, d_vals_all):
        a = (C.c_uint64 * n_parts)(*[int(x) for x in peer_codes])
        b = (C.c_uint64 * n_parts)(*[int(x) for x in peer_vals])
        n = (C.c_uint64 * n_parts)(*[int(x) for x in counts])
        rc = lib().hovl_shard_gather_peer(self._h, n_parts, C.cast(a, C.c_void_p), C.cast(b, C.c_void_p), C.cast(n, C.c_void_p),
                                          d_codes_all, d_vals_all)
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    def shard_pairs_local(self, settings, d_codes_all, d_vals_all, n_all, d_packed, n_words, d_len, d_bad, n_total,
                          read_lo, n_local, strict=True):
        """Statistics of the rank's share; its records wait on the device for shard_fetch_records."""
        pres = C.c_void_p()
        rc = lib().hovl_shard_pairs_local(self._h, C.byref(settings), d_codes_all, d_vals_all, n_all, d_packed, n_words,
                                          d_len, d_bad, n_total, read_lo, n_local, C.byref(pres))
        return self._collect(rc, pres, strict)

    def shard_fetch_records(self, dst, cap):
        rc = lib().hovl_shard_fetch_records(self._h, dst, cap)
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())

    def shard_align(self, settings, d_keys, d_counts, n_partial, d_packed, n_words, d_len, d_bad, n_total,
                    d_records=None, records_cap=0, strict=True):
        pres = C.c_void_p()
        rc = lib().hovl_shard_align(self._h, C.byref(settings), d_keys, d_counts, n_partial, d_packed, n_words,
                                    d_len, d_bad, n_total, d_records, records_cap, C.byref(pres))
        return self._collect(rc, pres, strict)

    def stage_table(self, settings, bases, offsets):
        """KmerData after addKmerSet: (codes, read_ids, loc_bits) sorted by (code, id, position)."""
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        pc, pr, pl, n = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_uint64()
        rc = lib().hovl_stage_table(self._h, C.byref(settings), bases.ctypes.data, offsets.ctypes.data,
                                    len(offsets) - 1, C.byref(pc), C.byref(pr), C.byref(pl), C.byref(n))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return _take(pc, n.value, np.uint32), _take(pr, n.value, np.uint32), _take(pl, n.value, np.uint32)

    def stage_cover(self, settings, bases, offsets):
        """uniqueKmers + kmerCollisionHistogram (src/KmerTable.scala:189-221): (n_buckets, {bucket size: k-mers})."""
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        nb, ps, pc, n = C.c_uint64(), C.c_void_p(), C.c_void_p(), C.c_uint64()
        rc = lib().hovl_stage_cover(self._h, C.byref(settings), bases.ctypes.data, offsets.ctypes.data,
                                    len(offsets) - 1, C.byref(nb), C.byref(ps), C.byref(pc), C.byref(n))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        sizes, counts = _take(ps, n.value, np.uint64), _take(pc, n.value, np.uint64)
        return int(nb.value), {int(a): int(b) for a, b in zip(sizes, counts)}

    def stage_pairs(self, settings, bases, offsets):
        """PairData: keys (fst<<32|snd, ascending), counts, and the number of counted instances."""
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        pk, pc, n, ni = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint64()
        rc = lib().hovl_stage_pairs(self._h, C.byref(settings), bases.ctypes.data, offsets.ctypes.data,
                                    len(offsets) - 1, C.byref(pk), C.byref(pc), C.byref(n), C.byref(ni))
        if rc != HOVL_OK:
            raise HovlError(rc, self._err())
        return _take(pk, n.value, np.uint64), _take(pc, n.value, np.uint32), int(ni.value)

    def stage_align(self, settings, bases, offsets, pair_a, pair_b, strict=False):
        """generateFastDovetailAlignment for the given (a,b) id pairs; one record per pair, in order."""
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        pa = np.ascontiguousarray(pair_a, np.uint32)
        pb = np.ascontiguousarray(pair_b, np.uint32)
        out = np.zeros(len(pa), RECORD_DTYPE)
        rc = lib().hovl_stage_align(self._h, C.byref(settings), bases.ctypes.data, offsets.ctypes.data,
                                    len(offsets) - 1, pa.ctypes.data, pb.ctypes.data, len(pa), out.ctypes.data)
        if rc not in (HOVL_OK, E_SHORT_READ, E_BAD_BASE) or (rc != HOVL_OK and strict):
            raise HovlError(rc, self._err())
        return out
