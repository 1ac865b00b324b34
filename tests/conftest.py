"""Shared test plumbing.  `-m "not gpu"` runs on the CPU box (oracle, host logic, ABI symbols);
`-m gpu` runs the parity tests proper through the C ABI on a B200."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "sequence-aligner_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def emul(oracle):
    """Host lane emulator of the DP kernel (tests/cpu/dp_emul.cpp)."""
    out = os.path.join(ROOT, "build", "libdp_emul.so")
    src = os.path.join(ROOT, "tests", "cpu", "dp_emul.cpp")
    hdr = os.path.join(PKG, "csrc", "dp_core.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-Wno-unknown-pragmas", "-o", out, src])
    E = C.CDLL(out)
    E.emul_align_pairs.argtypes = [C.POINTER(oracle.Settings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]

    def run(s, bases, offsets, pa, pb, G, D):
        out_ = np.zeros(len(pa), oracle.RECORD_DTYPE)
        cells = C.c_uint64()
        pa_ = np.ascontiguousarray(pa, np.uint32)
        pb_ = np.ascontiguousarray(pb, np.uint32)
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        rc = E.emul_align_pairs(C.byref(s), bases.ctypes.data, offsets.ctypes.data, pa_.ctypes.data, pb_.ctypes.data,
                                len(pa_), G, D, out_.ctypes.data, C.byref(cells))
        assert rc == 0
        return out_, int(cells.value)
    return run


@pytest.fixture(scope="session")
def tp_emul(oracle):
    """Host harness of the thread-per-pair DP (tests/cpu/tp_emul.cpp over csrc/dp_tp.cuh)."""
    out = os.path.join(ROOT, "build", "libtp_emul.so")
    src = os.path.join(ROOT, "tests", "cpu", "tp_emul.cpp")
    hdrs = [os.path.join(PKG, "csrc", h) for h in ("dp_core.cuh", "dp_tp.cuh", "dp_tp2.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(f) for f in [src] + hdrs):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-Wno-unknown-pragmas", "-o", out, src])
    E = C.CDLL(out)
    E.tp_emul_align_pairs.argtypes = [C.POINTER(oracle.Settings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_uint64, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]

    E.tp_emul_probe.argtypes = [C.POINTER(oracle.Settings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_uint64, C.c_void_p, C.c_void_p]

    def probe(s, bases, offsets, pa, pb):
        """(probe_valid, doveStart of the phase-1 DP, doveStart of the exact-prefix probe) per pair"""
        pa_ = np.ascontiguousarray(pa, np.uint32)
        pb_ = np.ascontiguousarray(pb, np.uint32)
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        d0, d1 = np.zeros(len(pa_), np.int32), np.zeros(len(pa_), np.int32)
        rc = E.tp_emul_probe(C.byref(s), bases.ctypes.data, offsets.ctypes.data, pa_.ctypes.data, pb_.ctypes.data,
                             len(pa_), d0.ctypes.data, d1.ctypes.data)
        return bool(rc), d0, d1

    E.tp2_emul_align_pairs.argtypes = [C.POINTER(oracle.Settings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_uint64, C.c_int, C.c_void_p]

    def run(s, bases, offsets, pa, pb, fast=0):
        """fast: 0 = 32-bit cells, 1 = 32-bit cells with packed arg-max keys, 2 = phase 1 with two pairs per
        thread (16x2 cells), 3 = both phases with two pairs per thread (phase 2 in moving frames)"""
        out_ = np.zeros(len(pa), oracle.RECORD_DTYPE)
        cells = C.c_uint64()
        pa_ = np.ascontiguousarray(pa, np.uint32)
        pb_ = np.ascontiguousarray(pb, np.uint32)
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        if fast >= 2:
            rc = E.tp2_emul_align_pairs(C.byref(s), bases.ctypes.data, offsets.ctypes.data, pa_.ctypes.data,
                                        pb_.ctypes.data, len(pa_), fast - 2, out_.ctypes.data)
            assert rc == 0, rc
            assert E.tp2_emul_wraps() == 0, "a 16-bit half word wrapped"
            return out_, None
        rc = E.tp_emul_align_pairs(C.byref(s), bases.ctypes.data, offsets.ctypes.data, pa_.ctypes.data,
                                   pb_.ctypes.data, len(pa_), fast, out_.ctypes.data, C.byref(cells))
        assert rc == 0, rc
        return out_, int(cells.value)
    run.probe = probe
    return run


@pytest.fixture(scope="session")
def wb_emul(oracle):
    """Host lane emulator of the wide-band aligner (tests/cpu/wb_emul.cpp over csrc/dp_wb.cuh)."""
    out = os.path.join(ROOT, "build", "libwb_emul.so")
    src = os.path.join(ROOT, "tests", "cpu", "wb_emul.cpp")
    hdrs = [os.path.join(PKG, "csrc", h) for h in ("dp_core.cuh", "dp_tp.cuh", "dp_wb.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(f) for f in [src] + hdrs):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-Wno-unknown-pragmas", "-o", out, src])
    E = C.CDLL(out)
    E.wb_emul_align_pairs.argtypes = [C.POINTER(oracle.Settings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_uint64, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]

    def run(s, bases, offsets, pa, pb, D):
        out_ = np.zeros(len(pa), oracle.RECORD_DTYPE)
        cells = C.c_uint64()
        pa_ = np.ascontiguousarray(pa, np.uint32)
        pb_ = np.ascontiguousarray(pb, np.uint32)
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        rc = E.wb_emul_align_pairs(C.byref(s), bases.ctypes.data, offsets.ctypes.data, pa_.ctypes.data,
                                   pb_.ctypes.data, len(pa_), D, out_.ctypes.data, C.byref(cells))
        assert rc == 0, rc
        return out_, int(cells.value)
    return run


@pytest.fixture(scope="session")
def quad_emul(oracle):
    """Host lane emulator of the quadratic aligner (tests/cpu/quad_emul.cpp over csrc/quad_core.cuh)."""
    out = os.path.join(ROOT, "build", "libquad_emul.so")
    src = os.path.join(ROOT, "tests", "cpu", "quad_emul.cpp")
    hdrs = [os.path.join(PKG, "csrc", h) for h in ("dp_core.cuh", "quad_core.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(f) for f in [src] + hdrs):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-Wno-unknown-pragmas", "-o", out, src])
    E = C.CDLL(out)
    E.quad_emul_align_pairs.argtypes = [C.POINTER(oracle.Settings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_uint64, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]

    def run(s, bases, offsets, pa, pb, D=4):
        out_ = np.zeros(len(pa), oracle.RECORD_DTYPE)
        cells = C.c_uint64()
        pa_ = np.ascontiguousarray(pa, np.uint32)
        pb_ = np.ascontiguousarray(pb, np.uint32)
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        rc = E.quad_emul_align_pairs(C.byref(s), bases.ctypes.data, offsets.ctypes.data, pa_.ctypes.data,
                                     pb_.ctypes.data, len(pa_), D, out_.ctypes.data, C.byref(cells))
        assert rc == 0, rc
        return out_, int(cells.value)
    return run


@pytest.fixture(scope="session")
def crp177():
    """amos/small/crp177.seq of the reference, committed as a golden fixture (223 x 100 bp)."""
    from hovl import io as hio
    return hio.read_seq(os.path.join(GOLDEN, "crp177.seq"))


@pytest.fixture(scope="session")
def hovl_ctx():
    import hovl
    ctx = hovl.Context(0)
    yield ctx
    ctx.close()


def records_equal(a, b, fields=None):
    fields = fields or [f for f in a.dtype.names]
    if len(a) != len(b):
        return False
    return all(np.array_equal(a[f], b[f]) for f in fields)


def synth_reads(seed, n_reads, read_len, genome_len, sub=0.01, indel=0.0, ragged=0):
    """Reads sampled from a random genome, forward strand only (the reference has no
    reverse complement), substitutions / indels at the given rates."""
    rng = np.random.default_rng(seed)
    g = rng.integers(0, 4, genome_len)
    reads = []
    for _ in range(n_reads):
        L = read_len - (int(rng.integers(0, ragged + 1)) if ragged else 0)
        st = int(rng.integers(0, max(1, genome_len - L - 64)))
        seg = g[st:st + L + 64]
        if indel > 0:
            out = []
            for x in seg:
                u = rng.random()
                if u < indel / 2:
                    continue
                if u < indel:
                    out.append(int(rng.integers(0, 4)))
                out.append(int(x))
            seg = np.array(out[:L])
        else:
            seg = seg[:L].copy()
        flip = rng.random(len(seg)) < sub
        seg = np.where(flip, (seg + rng.integers(1, 4, len(seg))) % 4, seg)
        reads.append(seg)
    offs = np.zeros(n_reads + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in reads])
    bases = np.frombuffer(b"ACGT", np.uint8)[np.concatenate(reads)] if reads else np.zeros(0, np.uint8)
    return np.ascontiguousarray(bases), offs


def java_float(x):
    """Float.toString of a positive float32: shortest digits that read back, plain for 1e-3 <= x < 1e7."""
    x = np.float32(x)
    if x == 0:
        return "0.0"
    if 1e-3 <= x < 1e7:
        s = np.format_float_positional(x, unique=True, trim="0")
        return s if "." in s and not s.endswith(".") else s.rstrip(".") + ".0"
    s = np.format_float_scientific(x, unique=True, trim="0", exp_digits=1)      # d.ddde-N
    mant, exp = s.split("e")
    if "." not in mant:
        mant += ".0"
    if mant.endswith("."):
        mant += "0"
    return "%sE%d" % (mant, int(exp))


def want_test_overlaps(recs, bases, offs):
    """The text testOverlaps prints (src/Project4.scala:484-504) for `recs` (filter_valid = 0), restated line by line."""
    out = []
    text = bytes(bases).decode()
    for i, r in enumerate(recs):
        if r["flags"] & 1:                                            # BioLibs.dud (src/BioLibs.scala:22)
            ida = idb = ahg = bhg = 0
            sa = sb = ""
            st = en = (0, 0)
            ratio = np.float32(0) / (np.float32(0) + np.float32(1))
            valid = False
        else:
            ida, idb, ahg, bhg = int(r["a"]), int(r["b"]), int(r["ahg"]), int(r["bhg"])
            sa = text[int(offs[ida - 1]):int(offs[ida])]
            sb = text[int(offs[idb - 1]):int(offs[idb])]
            st, en = (int(r["start_i"]), int(r["start_j"])), (int(r["end_i"]), int(r["end_j"]))
            ratio = np.float32(r["correct"]) / (np.float32(r["correct"]) + np.float32(r["error"]))
            valid = bool(r["flags"] & 4)
        out.append(" Overlap %d : %d <-> %d" % (i + 1, ida, idb))
        if ahg >= 0:
            out += ["   Seq A   : " + sa + "-" * max(bhg, 0), "   Seq B   : " + "-" * ahg + sb]
        else:
            out += ["   Seq A   : " + "-" * (-ahg) + sa, "   Seq B   : " + sb + "-" * max(-bhg, 0)]
        out += ["   Ahg     : %d" % ahg, "   Bhg     : %d" % bhg, "   Start   : (%d,%d)" % st, "   End     : (%d,%d)" % en,
                "   Error   : " + java_float(ratio), "   Valid?  : " + ("true" if valid else "false")]
    return "".join(l + "\n" for l in out)
