"""GPU parity: the CUDA path through the C ABI vs the CPU oracle, bit-exact (integers, indices,
order) and bit-exact on the fp64 sample values (the reference computes in double; the
north star's 1e-4 relative tolerance is therefore met with 0 ulp)."""
import numpy as np
import pytest

import oracle_lib as ol
from cases import parity_cases
from poregen_b200 import synth
from poregen_b200.gmove import GmoveEngine

pytestmark = pytest.mark.gpu


def assert_same(res, ref, label=""):
    assert res.n_slots == ref.n_slots and res.n_reads == ref.n_reads
    np.testing.assert_array_equal(res.read_flags, ref.read_flags, err_msg=label + " read_flags")
    np.testing.assert_array_equal(res.slot_sel, ref.slot_sel, err_msg=label + " slot_sel")
    np.testing.assert_array_equal(res.slot_total, ref.slot_total, err_msg=label + " slot_total")
    np.testing.assert_array_equal(res.ev_len, ref.ev_len, err_msg=label + " ev_len")
    np.testing.assert_array_equal(res.ev_read, ref.ev_read, err_msg=label + " ev_read")
    np.testing.assert_array_equal(res.ev_start, ref.ev_start, err_msg=label + " ev_start")
    assert res.samples.size == ref.samples.size
    # bit patterns, not values: -0.0 vs 0.0 and NaNs would slip through ==
    np.testing.assert_array_equal(res.samples.view(np.uint64), ref.samples.view(np.uint64), err_msg=label + " samples")
    assert res.n_events_examined == ref.n_events_examined, label + " examined"


@pytest.mark.parametrize("label,cfg,opts", parity_cases(), ids=[c[0] for c in parity_cases()])
def test_single_batch(built, label, cfg, opts):
    batch = synth.make_batch(cfg)
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(opts) as eng:
        res = eng.process(batch)
    assert_same(res, ref, label)


@pytest.mark.parametrize("label,cfg,opts", parity_cases()[:6], ids=[c[0] for c in parity_cases()[:6]])
def test_streamed_batches(built, label, cfg, opts):
    """Totals carried across batches: 4 consecutive slices must reproduce the oracle's single run
    slice by slice (sample_limit is enforced in record order across batches)."""
    batch = synth.make_batch(cfg)
    totals = np.zeros(opts.index_end - opts.index_start + 1, np.uint32)
    cuts = np.linspace(0, batch.n_reads, 5).astype(int)
    with GmoveEngine(opts) as eng:
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            sl = batch.slice(lo, hi)
            ref = ol.oracle_run(opts, sl, totals)
            res = eng.process(sl)
            assert_same(res, ref, f"{label}[{lo}:{hi}]")
