"""Test-side loader for the CPU oracle (oracle/liboracle.so) and the compiled reference
(oracle/_ref/poregen_ref). Test infrastructure only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from poregen_b200 import _capi as capi
from poregen_b200.batch import GmoveOptions, GmoveResult, ReadBatch, _np_from

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_BIN = os.path.join(ORACLE_DIR, "_ref", "poregen_ref")


class OracleResult(C.Structure):
    _fields_ = [
        ("n_slots", C.c_uint32), ("n_reads", C.c_uint32),
        ("n_events_examined", C.c_uint64), ("n_events_pass", C.c_uint64),
        ("n_sel_events", C.c_uint64), ("n_sel_samples", C.c_uint64),
        ("slot_sel", C.c_void_p), ("slot_total", C.c_void_p),
        ("ev_len", C.c_void_p), ("ev_read", C.c_void_p), ("ev_start", C.c_void_p),
        ("samples", C.c_void_p), ("read_flags", C.c_void_p),
        ("text", C.c_void_p), ("slot_text_off", C.c_void_p),
    ]


def build_oracle() -> str:
    src = os.path.join(ORACLE_DIR, "gmove_oracle.c")
    if (not os.path.exists(ORACLE_SO)) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-g", "-std=c99", "-Wall", "-ffp-contract=off", "-fPIC", "-shared",
                               src, "-o", ORACLE_SO, "-lm"])
    return ORACLE_SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_oracle())
        _lib.oracle_gmove_batch.argtypes = [C.POINTER(capi.PgmOpts), C.POINTER(capi.PgmBatch), C.c_void_p,
                                            C.POINTER(OracleResult)]
        _lib.oracle_gmove_batch.restype = C.c_int
        _lib.oracle_free_result.argtypes = [C.POINTER(OracleResult)]
        _lib.oracle_median.argtypes = [C.c_void_p, C.c_size_t]
        _lib.oracle_median.restype = C.c_double
        _lib.oracle_mad.argtypes = [C.c_void_p, C.c_size_t, C.c_double]
        _lib.oracle_mad.restype = C.c_double
        _lib.oracle_model_row.argtypes = [C.c_void_p, C.c_size_t, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _lib.oracle_model_row.restype = C.c_int
        _lib.oracle_model_row_text.argtypes = [C.c_char_p, C.c_size_t, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_char_p, C.c_char_p]
        _lib.oracle_model_row_text.restype = C.c_size_t
        _lib.oracle_dwell_median_text.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_uint64)]
        _lib.oracle_dwell_median_text.restype = C.c_size_t
        _lib.oracle_svb_zd_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p]
        _lib.oracle_svb_zd_decode.restype = C.c_int
    return _lib


def dwell_median_text(text: bytes):
    """scripts/poregen.sh:33-52 on one dump file's text: (values, twice the median); (0, None) for an empty file."""
    tm = C.c_uint64(0)
    n = lib().oracle_dwell_median_text(text, len(text), C.byref(tm))
    return int(n), (int(tm.value) if n else None)


def svb_zd_decode(stream: bytes, expect: int):
    """oracle_svb_zd_decode: int16 samples of one svb-zd stream, or None where slow5lib reports an error."""
    buf = np.frombuffer(stream, dtype=np.uint8)
    out = np.empty(max(expect, 1), dtype=np.int16)
    rc = lib().oracle_svb_zd_decode(buf.ctypes.data if buf.size else None, buf.size, expect, out.ctypes.data)
    return out[:expect] if rc == 0 else None


SVBREF_SO = os.path.join(ORACLE_DIR, "_ref", "libsvbref.so")
_svbref = None


def svbref():
    """slow5lib's own svb-zd compress/depress (oracle/_ref/libsvbref.so, built by oracle/build_ref.sh), or None."""
    global _svbref
    if _svbref is None and os.path.exists(SVBREF_SO):
        L = C.CDLL(SVBREF_SO)
        L.svbref_compress.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.svbref_compress.restype = C.c_void_p
        L.svbref_depress.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.svbref_depress.restype = C.c_void_p
        L.svbref_free.argtypes = [C.c_void_p]
        _svbref = L
    return _svbref


def svbref_compress(x: np.ndarray) -> bytes:
    L = svbref()
    x = np.ascontiguousarray(x, dtype=np.int16)
    n = C.c_size_t(0)
    p = L.svbref_compress(x.ctypes.data, x.size, C.byref(n))
    out = C.string_at(p, n.value)
    L.svbref_free(p)
    return out


def svbref_depress(stream: bytes):
    L = svbref()
    n = C.c_size_t(0)
    p = L.svbref_depress(stream, len(stream), C.byref(n))
    if not p:
        return None
    out = np.frombuffer(C.string_at(p, n.value), dtype=np.int16).copy()
    L.svbref_free(p)
    return out


def model_row_text(text: bytes, limit: float):
    """scripts/poregen.sh:54-85 on the text of one dump file: (values used, median, stddev, '%.14Lg' strings)."""
    med, sd = C.c_double(), C.c_double()
    ms, ss = C.create_string_buffer(64), C.create_string_buffer(64)
    n = lib().oracle_model_row_text(text, len(text), limit, C.byref(med), C.byref(sd), ms, ss)
    return int(n), med.value, sd.value, ms.value.decode(), ss.value.decode()


def oracle_run(opts: GmoveOptions, batch: ReadBatch, totals: np.ndarray | None = None, want_text: bool = True) -> GmoveResult:
    """One pass of the CPU restatement over `batch`; `totals` (uint32[n_slots]) is updated in place.
    want_text=False skips the "%.8f" formatting (full-size comparisons look at the doubles)."""
    lib().oracle_set_text(1 if want_text else 0)
    co, keep = opts.to_c()
    cb = batch.to_c()
    if totals is None:
        totals = np.zeros(co.n_slots, np.uint32)
    assert totals.dtype == np.uint32 and totals.size == co.n_slots
    r = OracleResult()
    rc = lib().oracle_gmove_batch(C.byref(co), C.byref(cb), totals.ctypes.data, C.byref(r))
    if rc != 0:
        raise RuntimeError(f"oracle failed rc={rc}")
    res = GmoveResult.from_c(r, want_text=want_text)
    lib().oracle_free_result(C.byref(r))
    lib().oracle_set_text(1)
    del keep
    return res


def cli_args(opts: GmoveOptions) -> list:
    """`poregen gmove` option list equivalent to `opts` (order matters: gmove.cpp:285-308)."""
    a = ["-k", str(opts.kmer_size), "-m", str(opts.sig_move_offset)]
    if opts.kmer_start_offset:
        a += ["-s", str(opts.kmer_start_offset)]
    a += ["--scaling", str(opts.scaling), "--margin", str(opts.signal_print_margin),
          "--sample_limit", str(opts.sample_limit), "--index_start", str(opts.index_start),
          "--index_end", str(opts.index_end),
          "--max_dur", str(opts.max_dur), "--min_dur", str(opts.min_dur), "--pa_min", repr(opts.pa_min),
          "--pa_max", repr(opts.pa_max), "--kmer_pick_margin", str(opts.kmer_pick_margin)]
    if opts.rna:
        a += ["--rna"]
    if opts.delimit_files:
        a += ["-d"]
    return a


def run_reference(opts: GmoveOptions, slow5: str, align: str, outdir: str, fastq: str | None = None,
                  kmer_file: str | None = None, binary: str = REF_BIN) -> dict:
    """Runs the UNMODIFIED compiled reference; returns {kmer: bytes} for dump/ plus 'freq.txt'."""
    cmd = [binary, "gmove"] + cli_args(opts)
    if kmer_file:
        cmd += ["--kmer_file", kmer_file]
    if fastq:
        cmd += ["--fastq", fastq]
    cmd += [slow5, align, outdir]
    p = subprocess.run(cmd, capture_output=True)
    if p.returncode != 0:
        raise RuntimeError(f"reference failed rc={p.returncode}: {p.stderr.decode()[-2000:]}")
    out = {}
    d = os.path.join(outdir, "dump")
    for name in os.listdir(d):
        with open(os.path.join(d, name), "rb") as f:
            out[name] = f.read()
    with open(os.path.join(outdir, "freq.txt"), "rb") as f:
        out["freq.txt"] = f.read()
    return out


def result_as_files(opts: GmoveOptions, res: GmoveResult) -> dict:
    """dump/<kmer> contents and freq.txt as the host writer would produce them (no -d)."""
    out = {}
    names = opts.slot_kmers()
    for s, name in enumerate(names):
        out[name] = res.slot_text(s)
    out["freq.txt"] = "".join(f"{n}\t{int(res.slot_total[s])}\n" for s, n in enumerate(names)).encode()
    return out
