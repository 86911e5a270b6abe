"""CPU-side checks of the C ABI: struct layouts, exported symbols, loud failure without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from poregen_b200 import _capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_struct_layouts():
    assert C.sizeof(capi.PgmRead) == 88
    assert np.dtype(capi.READ_DTYPE).itemsize == 88
    for (name, _), f in zip(capi.READ_DTYPE, capi.PgmRead._fields_):
        assert name == f[0]
        assert np.dtype(capi.READ_DTYPE).fields[name][1] == getattr(capi.PgmRead, name).offset
    assert C.sizeof(capi.PgmOpts) == 14 * 4 + 2 * 8 + 8
    assert C.sizeof(capi.PgmBatch) == 8 + 3 * 8 + 4 * 8


def test_header_symbols_exported(built):
    hdr = open(os.path.join(ROOT, "include", "pgm_gmove.h")).read()
    declared = set(re.findall(r"\b(pgm_[a-z_]+)\s*\(", hdr))
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    lib = C.CDLL(capi.LIB_PATH)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.pgm_abi_version() == capi.PGM_ABI_VERSION


def test_no_cpu_fallback(built):
    """Without a CUDA device pgm_create must fail with PGM_ERR_CUDA and say so."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from poregen_b200.batch import GmoveOptions
    from poregen_b200.gmove import GmoveEngine, PgmError
    with pytest.raises(PgmError) as e:
        GmoveEngine(GmoveOptions(kmer_size=5).resolve_range())
    assert e.value.code == capi.PGM_ERR_CUDA
    assert "no CPU fallback" in str(e.value)
