"""PGM_OUT_TEXT: the device formatter must print exactly what the reference's fprintf("%.8f") prints
(reference gmove.cpp:674-678) - for every finite double, and the grouped dump text of a batch must equal
the oracle's (which formats with the C library, as the reference does)."""
import copy
import struct

import numpy as np
import pytest

import oracle_lib as ol
from cases import _o, parity_cases
from poregen_b200 import _capi as capi, synth
from poregen_b200.gmove import GmoveEngine

pytestmark = pytest.mark.gpu


def _expected(vals):
    return (",".join("%.8f" % v for v in vals) + ";").encode()


def test_formatter_matches_printf(built):
    rng = np.random.default_rng(7)
    vals = []
    # pA-like and scaled values
    vals += list(rng.normal(90, 30, 4000)) + list(rng.normal(0, 3, 4000))
    # exact ties of the 8th decimal: odd multiples of 2^-9 (x * 1e8 ends in .5) -> round half to even
    vals += [m / 512.0 for m in range(-401, 402, 2)] + [1000.0 + m / 512.0 for m in range(1, 200, 2)]
    # every binade, random mantissas, both signs
    bits = (rng.integers(0, 2047, 6000, dtype=np.uint64) << np.uint64(52)) | rng.integers(0, 1 << 52, 6000, dtype=np.uint64)
    bits |= rng.integers(0, 2, 6000, dtype=np.uint64) << np.uint64(63)
    vals += list(bits.view(np.float64))
    # edges
    vals += [0.0, -0.0, 5e-324, -5e-324, 4.9999999e-9, 5e-9, 5.0000001e-9, 1.5e-8, 2.5e-8, 0.999999995, 9.999999995, 99999999.999999995,
             1e8, 184467440737.09551, 184467440737.09552, 184467440738.0, 2.0 ** 52, 2.0 ** 53, 2.0 ** 63, 2.0 ** 64, 1e22, 1e23, 1.7976931348623157e308,
             -1e300, 123456789012345.678, 4503599627370495.5, 0.1, 0.2, 0.3, 1.0 / 3.0]
    vals = np.array(vals, dtype=np.float64)
    with GmoveEngine(_o(kmer_size=3)) as eng:
        for lo in range(0, vals.size, 997):
            chunk = vals[lo:lo + 997]
            assert eng.format_doubles(chunk) == _expected(chunk)
        assert eng.format_doubles(np.array([np.inf, -np.inf])) == b"inf,-inf;"
        nan_pos = struct.unpack("<d", struct.pack("<Q", 0x7ff8000000000000))[0]
        nan_neg = struct.unpack("<d", struct.pack("<Q", 0xfff8000000000000))[0]
        assert eng.format_doubles(np.array([nan_pos, nan_neg])) == b"nan,-nan;"   # glibc prints the sign of a NaN


_CASES = parity_cases()[:8] + parity_cases()[12:16]


@pytest.mark.parametrize("label,cfg,opts", _CASES, ids=[c[0] for c in _CASES])
def test_dump_text_equals_oracle(built, label, cfg, opts):
    batch = synth.make_batch(cfg)
    ref = ol.oracle_run(opts, batch)
    topts = copy.copy(opts)
    topts.output = capi.PGM_OUT_TEXT
    with GmoveEngine(topts) as eng:
        res = eng.process(batch)
    np.testing.assert_array_equal(res.slot_total, ref.slot_total)
    np.testing.assert_array_equal(res.ev_len, ref.ev_len)
    np.testing.assert_array_equal(res.slot_text_off, ref.slot_text_off)
    assert res.text == ref.text
    assert res.samples.size == 0


def test_text_only_output_leaves_event_arrays_on_the_device(built):
    """PGM_OUT_TEXT_ONLY: the same text and per-slot offsets, no per-event arrays on the host."""
    from cases import _o
    from poregen_b200.synth import SynthConfig
    opts = _o(kmer_size=4, sample_limit=30, scaling=1)
    batch = synth.make_batch(SynthConfig(n_reads=40, median_samples=3000, seed=55))
    ref = ol.oracle_run(opts, batch)
    topts = copy.copy(opts)
    topts.output = capi.PGM_OUT_TEXT_ONLY
    with GmoveEngine(topts) as eng:
        res = eng.process(batch)
        res2 = eng.process(batch)          # second batch on the carried totals: the buffer of the first one is reused
    assert res.text == ref.text
    np.testing.assert_array_equal(res.slot_text_off, ref.slot_text_off)
    np.testing.assert_array_equal(res.slot_sel, ref.slot_sel)
    assert res.ev_len.size == 0 and res.ev_read.size == 0 and res.ev_start.size == 0 and res.samples.size == 0
    ref2 = ol.oracle_run(opts, batch, totals=ref.slot_total.copy())
    assert res2.text == ref2.text
