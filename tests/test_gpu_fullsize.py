"""Full-size batches of BASELINE.json's configurations (one GPU's share), checked through properties that do
not need the CPU oracle to finish 1e8..6e8 samples:

  * streaming invariance: the batch in one piece == the batch in two pieces with the per-k-mer totals carried
    (the dump files of the reference grow the same way whatever the batching);
  * the quota and the filters: per-k-mer counts, order inside a k-mer file, durations, event bounds;
  * the oracle, bit for bit, on a prefix of the same records run as their own batch.

C3's --sample_limit is lowered (1000 -> 8) so that the selected samples of 12,500 reads x 50,000 samples fit a test's
host memory; the event stream that is decoded, counted and sorted (two digit passes over 18-bit k-mer ids) is full size."""
import dataclasses

import numpy as np
import pytest

import oracle_lib as ol
from poregen_b200 import synth, workloads
from poregen_b200.gmove import GmoveEngine

pytestmark = pytest.mark.gpu


def _slot_of_event(res):
    return np.repeat(np.arange(res.slot_sel.size, dtype=np.int64), res.slot_sel.astype(np.int64))


def _check_invariants(opts, batch, res):
    limit = opts.sample_limit
    assert int(res.slot_sel.sum()) == res.ev_len.size
    assert int(res.ev_len.astype(np.int64).sum()) == res.samples.size
    assert (res.slot_sel <= limit).all() and (res.slot_total <= limit).all()
    assert (res.slot_total >= res.slot_sel).all()
    if opts.signal_print_margin == 0:
        assert (res.ev_len >= opts.min_dur).all() and (res.ev_len <= opts.max_dur).all()
    rd = batch.reads
    n = rd["raw_len"].astype(np.int64) - rd["trim"].astype(np.int64)
    assert (res.ev_start.astype(np.int64) + res.ev_len <= n[res.ev_read]).all()
    # inside a k-mer file: record order, then position order
    slot = _slot_of_event(res)
    key = (slot << 32) | res.ev_read.astype(np.int64)
    assert (np.diff(key) >= 0).all()
    same = np.diff(key) == 0
    assert (np.diff(res.ev_start.astype(np.int64))[same] > 0).all()
    assert np.isfinite(res.samples).all()


def _merge(parts, n_first):
    """Results of consecutive pieces -> what one pass over all records must give."""
    slot = np.concatenate([_slot_of_event(p) for p in parts])
    piece = np.concatenate([np.full(p.ev_len.size, i, np.int64) for i, p in enumerate(parts)])
    order = np.lexsort((np.arange(slot.size), piece, slot))
    ev_len = np.concatenate([p.ev_len for p in parts])[order]
    ev_start = np.concatenate([p.ev_start for p in parts])[order]
    ev_read = np.concatenate([p.ev_read.astype(np.int64) + (n_first if i else 0) for i, p in enumerate(parts)])[order]
    # samples: event e of the concatenation starts at off[e]
    lens = np.concatenate([p.ev_len for p in parts]).astype(np.int64)
    base = np.concatenate([[0], np.cumsum([p.samples.size for p in parts])[:-1]])
    off = np.concatenate([np.concatenate(([0], np.cumsum(p.ev_len.astype(np.int64))[:-1])) + base[i] if p.ev_len.size else np.zeros(0, np.int64)
                          for i, p in enumerate(parts)])
    samples = np.concatenate([p.samples for p in parts])
    o_off, o_len = off[order], lens[order]
    idx = np.repeat(o_off - np.concatenate(([0], np.cumsum(o_len)[:-1])), o_len) + np.arange(int(o_len.sum()))
    return ev_len, ev_start, ev_read, samples[idx]


def _run_fullsize(opts, batch, prefix_reads):
    n = batch.n_reads
    with GmoveEngine(opts) as eng:
        whole = eng.process(batch)
    _check_invariants(opts, batch, whole)
    with GmoveEngine(opts) as eng:
        parts = [eng.process(batch.slice(0, n // 2)), eng.process(batch.slice(n // 2, n))]
    ev_len, ev_start, ev_read, samples = _merge(parts, n // 2)
    np.testing.assert_array_equal(parts[1].slot_total, whole.slot_total)
    np.testing.assert_array_equal(ev_len, whole.ev_len)
    np.testing.assert_array_equal(ev_start, whole.ev_start)
    np.testing.assert_array_equal(ev_read, whole.ev_read)
    np.testing.assert_array_equal(samples.view(np.uint64), whole.samples.view(np.uint64))
    # the oracle on a prefix, as its own batch
    head = batch.slice(0, prefix_reads)
    ref = ol.oracle_run(opts, head)
    with GmoveEngine(opts) as eng:
        got = eng.process(head)
    np.testing.assert_array_equal(got.slot_total, ref.slot_total)
    np.testing.assert_array_equal(got.ev_len, ref.ev_len)
    np.testing.assert_array_equal(got.ev_read, ref.ev_read)
    np.testing.assert_array_equal(got.samples.view(np.uint64), ref.samples.view(np.uint64))
    return whole


def test_c2_full_size(built):
    cfg, opts = workloads.c2()
    whole = _run_fullsize(opts, synth.make_batch(cfg), 300)
    assert whole.ev_len.size > 500000


def test_c3_one_gpu_share(built):
    cfg, opts = workloads.c3(n_reads=12500)
    opts = dataclasses.replace(opts, sample_limit=8)
    whole = _run_fullsize(opts, synth.make_batch(cfg), 40)
    assert whole.n_events_pass > 30_000_000            # the sort really saw the full event stream
    assert (whole.slot_total == 8).mean() > 0.9        # and the quota really cut it


@pytest.mark.parametrize("rna", [False, True], ids=["dna", "rna"])
def test_c4_full_size(built, rna):
    cfg, opts = workloads.c4(rna=rna)
    whole = _run_fullsize(opts, synth.make_batch(cfg), 300)
    assert whole.ev_len.size > 100000
