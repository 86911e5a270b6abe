"""ctypes mirror of include/pgm_gmove.h (the C ABI of the gmove hot path).

The structures here are byte-for-byte the C ones; tests/test_abi.py checks the sizes and that
the shared library exports every declared symbol.  There is no Python fallback: if the CUDA
library is missing, `load()` raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PGM_LIB") or os.path.join(HERE, "libpgm_gmove.so")   # PGM_LIB: a kernel-variant build (profiles/micro)

PGM_OK, PGM_ERR_ARG, PGM_ERR_CUDA, PGM_ERR_NOMEM, PGM_ERR_INPUT, PGM_ERR_STATE = range(6)
PGM_MODE_MOVES_CHAR, PGM_MODE_SS, PGM_MODE_MOVES_I8 = 0, 1, 2
PGM_READ_RNA, PGM_READ_SKIP = 1, 2
PGM_RF_SKIPPED, PGM_RF_BAD_ALIGN, PGM_RF_BAD_CALIB = 1, 2, 4
PGM_OUT_F64, PGM_OUT_TEXT, PGM_OUT_TEXT_ONLY = 0, 1, 2
PGM_ABI_VERSION = 4


class PgmOpts(C.Structure):
    _fields_ = [
        ("kmer_size", C.c_uint32),
        ("sig_move_offset", C.c_uint32),
        ("kmer_start_offset", C.c_uint32),
        ("signal_print_margin", C.c_uint32),
        ("sample_limit", C.c_uint32),
        ("max_dur", C.c_uint32),
        ("min_dur", C.c_uint32),
        ("kmer_pick_margin", C.c_uint32),
        ("flag_rna", C.c_uint32),
        ("scaling", C.c_uint32),
        ("output", C.c_uint32),
        ("n_slots", C.c_uint32),
        ("index_lo", C.c_uint32),
        ("reserved0", C.c_uint32),
        ("pa_min", C.c_double),
        ("pa_max", C.c_double),
        ("kmer_list", C.c_char_p),
    ]


class PgmRead(C.Structure):
    _fields_ = [
        ("raw_off", C.c_uint64),
        ("align_off", C.c_uint64),
        ("seq_off", C.c_uint64),
        ("raw_len", C.c_uint32),
        ("align_len", C.c_uint32),
        ("seq_len", C.c_uint32),
        ("trim", C.c_uint32),
        ("stride", C.c_uint32),
        ("fastq_len", C.c_int32),
        ("query_start", C.c_uint32),
        ("st_k", C.c_int32),
        ("end_k", C.c_int32),
        ("flags", C.c_uint32),
        ("digitisation", C.c_double),
        ("offset", C.c_double),
        ("range", C.c_double),
    ]


class PgmBatch(C.Structure):
    _fields_ = [
        ("n_reads", C.c_uint32),
        ("mode", C.c_uint32),
        ("n_raw", C.c_uint64),
        ("n_align", C.c_uint64),
        ("n_bases", C.c_uint64),
        ("raw", C.c_void_p),
        ("align", C.c_void_p),
        ("bases", C.c_void_p),
        ("reads", C.c_void_p),
    ]


class PgmSvb(C.Structure):
    _fields_ = [("n_bytes", C.c_uint64), ("bytes", C.c_void_p), ("off", C.c_void_p)]


class PgmResult(C.Structure):
    _fields_ = [
        ("n_slots", C.c_uint32),
        ("n_reads", C.c_uint32),
        ("n_events_examined", C.c_uint64),
        ("n_events_pass", C.c_uint64),
        ("n_sel_events", C.c_uint64),
        ("n_sel_samples", C.c_uint64),
        ("slot_sel", C.c_void_p),
        ("slot_total", C.c_void_p),
        ("ev_len", C.c_void_p),
        ("ev_read", C.c_void_p),
        ("ev_start", C.c_void_p),
        ("samples", C.c_void_p),
        ("read_flags", C.c_void_p),
        ("text", C.c_void_p),
        ("slot_text_off", C.c_void_p),
        ("ev_text_off", C.c_void_p),
    ]


# numpy dtype with the same layout as PgmRead (88 bytes, no padding surprises: checked in tests)
READ_DTYPE = [
    ("raw_off", "<u8"), ("align_off", "<u8"), ("seq_off", "<u8"),
    ("raw_len", "<u4"), ("align_len", "<u4"), ("seq_len", "<u4"), ("trim", "<u4"),
    ("stride", "<u4"), ("fastq_len", "<i4"), ("query_start", "<u4"), ("st_k", "<i4"),
    ("end_k", "<i4"), ("flags", "<u4"),
    ("digitisation", "<f8"), ("offset", "<f8"), ("range", "<f8"),
]

# every symbol include/pgm_gmove.h declares
SYMBOLS = [
    "pgm_create", "pgm_destroy", "pgm_last_error", "pgm_abi_version", "pgm_process", "pgm_upload",
    "pgm_attach_device", "pgm_count", "pgm_emit", "pgm_download", "pgm_sync", "pgm_counts_device",
    "pgm_totals_device", "pgm_set_prior", "pgm_reset", "pgm_stream", "pgm_launch_count",
    "pgm_profile_enable", "pgm_profile_read", "pgm_stage_name", "pgm_debug_format",
    "pgm_counts_host", "pgm_totals_host", "pgm_set_prior_host", "pgm_alloc_pinned", "pgm_free_pinned", "pgm_device_count",
    "pgm_set_gathered", "pgm_model_begin", "pgm_model_finalize", "pgm_warmup", "pgm_upload_svb", "pgm_debug_raw", "pgm_model_dwell", "pgm_summary", "pgm_submit", "pgm_wait", "pgm_create_sibling",
    "pgm_model_log", "pgm_model_log_size", "pgm_model_log_entry", "pgm_model_log_append",
    "pgm_format_noalloc", "pgm_model_log_reserve",
]
PGM_NOALLOC_MAX_VALUES = 1 << 24
PGM_N_STAGES = 11
# pgm_model_row (include/pgm_gmove.h)
MODEL_ROW_DTYPE = [("median", "<f8"), ("stddev", "<f8"), ("median_num", "<i8"), ("n", "<u8"), ("flags", "<u4"), ("pad", "<u4")]
PGM_MODEL_EMPTY, PGM_MODEL_NO_SD, PGM_MODEL_RANGE, PGM_MODEL_CLAMPED = 1, 2, 4, 8

_lib = None


def load() -> C.CDLL:
    """Load the CUDA library. Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the gmove hot path)")
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    lib.pgm_create.argtypes = [P(PgmOpts), C.c_int, C.c_void_p, P(C.c_void_p)]
    lib.pgm_create.restype = C.c_int
    lib.pgm_destroy.argtypes = [C.c_void_p]
    lib.pgm_destroy.restype = None
    lib.pgm_last_error.argtypes = [C.c_void_p]
    lib.pgm_last_error.restype = C.c_char_p
    lib.pgm_abi_version.argtypes = []
    lib.pgm_abi_version.restype = C.c_int
    for name in ("pgm_process",):
        getattr(lib, name).argtypes = [C.c_void_p, P(PgmBatch), P(PgmResult)]
        getattr(lib, name).restype = C.c_int
    for name in ("pgm_upload", "pgm_attach_device"):
        getattr(lib, name).argtypes = [C.c_void_p, P(PgmBatch)]
        getattr(lib, name).restype = C.c_int
    lib.pgm_upload_svb.argtypes = [C.c_void_p, P(PgmBatch), P(PgmSvb)]
    lib.pgm_upload_svb.restype = C.c_int
    for name in ("pgm_count", "pgm_emit", "pgm_sync", "pgm_reset"):
        getattr(lib, name).argtypes = [C.c_void_p]
        getattr(lib, name).restype = C.c_int
    lib.pgm_download.argtypes = [C.c_void_p, P(PgmResult)]
    lib.pgm_download.restype = C.c_int
    lib.pgm_submit.argtypes = [C.c_void_p, P(PgmBatch), C.c_void_p, P(C.c_uint64)]
    lib.pgm_submit.restype = C.c_int
    lib.pgm_wait.argtypes = [C.c_void_p, C.c_uint64, P(PgmResult)]
    lib.pgm_wait.restype = C.c_int
    lib.pgm_create_sibling.argtypes = [C.c_void_p, C.c_void_p, P(C.c_void_p)]
    lib.pgm_create_sibling.restype = C.c_int
    lib.pgm_model_log.argtypes = [C.c_void_p]
    lib.pgm_model_log.restype = C.c_int
    lib.pgm_model_log_size.argtypes = [C.c_void_p, P(C.c_uint32)]
    lib.pgm_model_log_size.restype = C.c_int
    lib.pgm_model_log_entry.argtypes = [C.c_void_p, C.c_uint32, P(C.c_void_p), P(C.c_void_p)]
    lib.pgm_model_log_entry.restype = C.c_int
    lib.pgm_model_log_append.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.pgm_model_log_append.restype = C.c_int
    lib.pgm_summary.argtypes = [C.c_void_p, P(PgmResult)]
    lib.pgm_summary.restype = C.c_int
    lib.pgm_counts_device.argtypes = [C.c_void_p, P(C.c_void_p)]
    lib.pgm_counts_device.restype = C.c_int
    lib.pgm_totals_device.argtypes = [C.c_void_p, P(C.c_void_p)]
    lib.pgm_totals_device.restype = C.c_int
    lib.pgm_set_prior.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.pgm_set_prior.restype = C.c_int
    lib.pgm_set_gathered.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    lib.pgm_set_gathered.restype = C.c_int
    lib.pgm_model_begin.argtypes = [C.c_void_p]
    lib.pgm_model_begin.restype = C.c_int
    lib.pgm_model_finalize.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_void_p)]
    lib.pgm_model_finalize.restype = C.c_int
    lib.pgm_stream.argtypes = [C.c_void_p]
    lib.pgm_stream.restype = C.c_void_p
    lib.pgm_launch_count.argtypes = [C.c_void_p]
    lib.pgm_launch_count.restype = C.c_uint64
    lib.pgm_profile_enable.argtypes = [C.c_void_p, C.c_int]
    lib.pgm_profile_enable.restype = C.c_int
    lib.pgm_profile_read.argtypes = [C.c_void_p, P(C.c_float), P(C.c_uint32)]
    lib.pgm_profile_read.restype = C.c_int
    lib.pgm_stage_name.argtypes = [C.c_int]
    lib.pgm_stage_name.restype = C.c_char_p
    lib.pgm_model_dwell.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.pgm_model_dwell.restype = C.c_int
    lib.pgm_debug_raw.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    lib.pgm_debug_raw.restype = C.c_int
    lib.pgm_debug_format.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, P(C.c_uint64)]
    lib.pgm_debug_format.restype = C.c_int
    lib.pgm_format_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.pgm_format_noalloc.restype = C.c_int
    lib.pgm_model_log_reserve.argtypes = [C.c_void_p, C.c_uint64]
    lib.pgm_model_log_reserve.restype = C.c_int
    _lib = lib
    return lib
