"""Full-size batches of BASELINE.json's configurations.

Against the oracle, bit for bit, with the quota binding: ALL of C2 (10,000 records), all of C4 (DNA and RNA), and a share of
C3 (6,250 of the 100,000 records of bench.py's generator, --sample_limit scaled with the share so that the same two thirds
of the 262,144 k-mers fill up) - as one batch, streamed in blocks with the totals carried, and cut into 4 and 8 rank
ranges with the multi-GPU quota exchange.

And through properties that do not need the CPU oracle to finish 6e8 samples:

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


def _assert_same(got, ref, label):
    """GPU result == oracle result, every array bit for bit."""
    np.testing.assert_array_equal(got.slot_total, ref.slot_total, err_msg=f"{label}: per-k-mer totals")
    np.testing.assert_array_equal(got.slot_sel, ref.slot_sel, err_msg=f"{label}: per-k-mer selected counts")
    np.testing.assert_array_equal(got.read_flags, ref.read_flags, err_msg=f"{label}: record flags")
    assert got.n_events_examined == ref.n_events_examined, label
    # n_events_pass is a batch-level diagnostic (events that passed the filters and whose k-mer was not full when the BATCH
    # started): it equals the oracle's running count only where the quota does not bind inside the batch
    assert got.n_events_pass >= ref.n_events_pass, label
    np.testing.assert_array_equal(got.ev_read, ref.ev_read, err_msg=f"{label}: emission order (records)")
    np.testing.assert_array_equal(got.ev_start, ref.ev_start, err_msg=f"{label}: event starts")
    np.testing.assert_array_equal(got.ev_len, ref.ev_len, err_msg=f"{label}: event lengths")
    assert got.samples.size == ref.samples.size
    assert np.array_equal(got.samples.view(np.uint64), ref.samples.view(np.uint64)), f"{label}: sample bits differ from the oracle"


def _run_ranks(opts, batch, n_ranks):
    """The batch cut into n_ranks contiguous record ranges, one context per range on this device, the quota exchanged on the
    host exactly like the ranks of a multi-GPU job do (count -> all counts -> prior per rank -> emit): returns the ranks'
    results and what their concatenation per k-mer must look like (= one context over the whole batch)."""
    from poregen_b200.gmove import quota_exchange
    import ctypes as C
    import torch
    n = batch.n_reads
    cuts = [n * g // n_ranks for g in range(n_ranks + 1)]
    engines = [GmoveEngine(opts) for _ in range(n_ranks)]
    try:
        counts = []
        for g, e in enumerate(engines):
            e.upload(synth.rebase(batch, cuts[g], cuts[g + 1]))
            e.count()
            e.sync()
            t = torch.as_tensor(type("A", (), {"__cuda_array_interface__": {"shape": (e.n_slots,), "typestr": "<u4", "data": (e.counts_ptr(), False), "version": 3}})(), device="cuda:0")
            counts.append(t.cpu().numpy().astype(np.uint32))
        prior, consumed = quota_exchange(np.stack(counts), np.zeros(engines[0].n_slots, np.uint32), opts.sample_limit)
        keep = []
        outs = []
        for g, e in enumerate(engines):
            p = torch.from_numpy(prior[g].astype(np.int32)).cuda()
            cns = torch.from_numpy(consumed.astype(np.int32)).cuda()
            keep += [p, cns]
            e.set_prior(p.data_ptr(), cns.data_ptr())
            e.emit()
            outs.append(e.download())
    finally:
        for e in engines:
            e.close()
    return outs, cuts


def _concat_ranks(parts, cuts):
    slot = np.concatenate([_slot_of_event(p) for p in parts])
    rank = np.concatenate([np.full(p.ev_len.size, i, np.int64) for i, p in enumerate(parts)])
    order = np.lexsort((np.arange(slot.size), rank, slot))
    ev_len = np.concatenate([p.ev_len for p in parts])[order]
    ev_start = np.concatenate([p.ev_start for p in parts])[order]
    ev_read = np.concatenate([p.ev_read.astype(np.int64) + cuts[i] for i, p in enumerate(parts)])[order]
    lens = np.concatenate([p.ev_len for p in parts]).astype(np.int64)
    base = np.concatenate([[0], np.cumsum([p.samples.size for p in parts])[:-1]])
    off = np.concatenate([np.concatenate(([0], np.cumsum(p.ev_len.astype(np.int64))[:-1])) + base[i] if p.ev_len.size else np.zeros(0, np.int64)
                          for i, p in enumerate(parts)])
    samples = np.concatenate([p.samples for p in parts])
    o_off, o_len = off[order], lens[order]
    idx = np.repeat(o_off - np.concatenate(([0], np.cumsum(o_len)[:-1])), o_len) + np.arange(int(o_len.sum()))
    return ev_len, ev_start, ev_read, samples[idx]


def test_c2_whole_batch_against_the_oracle(built):
    """All 10,000 C2 records (not a prefix). The --min_dur 20 --max_dur 40 window leaves ~840 events per 5-mer, so C2's
    quota of 5,000 never binds whatever the batch; C4 and the C3 share below are the quota-bound cases."""
    cfg, opts = workloads.c2()
    batch = synth.make_batch(cfg)
    ref = ol.oracle_run(opts, batch, want_text=False)
    with GmoveEngine(opts) as eng:
        got = eng.process(batch)
    _assert_same(got, ref, "C2")


@pytest.mark.parametrize("rna", [False, True], ids=["dna", "rna"])
def test_c4_whole_batch_against_the_oracle(built, rna):
    cfg, opts = workloads.c4(rna=rna)
    batch = synth.make_batch(cfg)
    ref = ol.oracle_run(opts, batch, want_text=False)
    with GmoveEngine(opts) as eng:
        got = eng.process(batch)
    assert (ref.slot_total == opts.sample_limit).mean() > 0.5
    _assert_same(got, ref, "C4")


C3_SHARE = 6250          # a sixteenth of the 100,000 C3 reads: what the oracle decodes in about a minute


@pytest.fixture(scope="module")
def c3_share(built):
    """6,250 C3 reads (3.3e8 samples, 2.6e7 events, 262,144 k-mers) with --sample_limit scaled like the prefix (1000 / 16 = 63),
    so that, as in the whole job, practically every k-mer fills up and about two thirds of the events are written; the oracle's result, and the bench's generator."""
    cfg, opts = workloads.c3(n_reads=100000)
    opts = dataclasses.replace(opts, sample_limit=63)
    batch = synth.make_moves_fast(cfg, 0, C3_SHARE)
    ref = ol.oracle_run(opts, batch, want_text=False)
    return opts, batch, ref


def test_c3_share_against_the_oracle(c3_share):
    opts, batch, ref = c3_share
    assert (ref.slot_total == opts.sample_limit).mean() > 0.9 and 0.5 < ref.ev_len.size / ref.n_events_examined < 0.8   # like the whole job: every k-mer fills up, two thirds of the events are written
    with GmoveEngine(opts) as eng:
        got = eng.process(batch)
    _assert_same(got, ref, "C3 share, one batch")
    # the same records streamed in 5 blocks with the totals carried (what bench.py and the command line do)
    with GmoveEngine(opts) as eng:
        parts = [eng.process(synth.rebase(batch, lo, min(lo + 1250, C3_SHARE))) for lo in range(0, C3_SHARE, 1250)]
    ev_len, ev_start, ev_read, samples = _concat_ranks(parts, list(range(0, C3_SHARE, 1250)))
    np.testing.assert_array_equal(parts[-1].slot_total, ref.slot_total)
    np.testing.assert_array_equal(ev_len, ref.ev_len)
    np.testing.assert_array_equal(ev_start, ref.ev_start)
    np.testing.assert_array_equal(ev_read, ref.ev_read)
    assert np.array_equal(samples.view(np.uint64), ref.samples.view(np.uint64))


@pytest.mark.parametrize("n_ranks", [4, 8])
def test_c3_share_rank_split(c3_share, n_ranks):
    """4 and 8 contiguous read ranges (one context each, quota exchanged like the ranks of a multi-GPU job): the ranks'
    slices, concatenated per k-mer in rank order, are the oracle's single loop."""
    opts, batch, ref = c3_share
    parts, cuts = _run_ranks(opts, batch, n_ranks)
    ev_len, ev_start, ev_read, samples = _concat_ranks(parts, cuts)
    np.testing.assert_array_equal(parts[0].slot_total, ref.slot_total)          # every rank reports the totals after the round
    np.testing.assert_array_equal(sum(p.slot_sel.astype(np.int64) for p in parts), ref.slot_sel)
    np.testing.assert_array_equal(ev_len, ref.ev_len)
    np.testing.assert_array_equal(ev_start, ref.ev_start)
    np.testing.assert_array_equal(ev_read, ref.ev_read)
    assert np.array_equal(samples.view(np.uint64), ref.samples.view(np.uint64))


def test_c3_block_at_the_real_limit(built):
    """One bench block (3,125 reads) at the real --sample_limit 1000 against the oracle (the quota cannot bind on a
    thirty-second of the job: this is the all-selected path the first twenty blocks of a bench step take)."""
    cfg, opts = workloads.c3(n_reads=100000)
    batch = synth.make_moves_fast(cfg, 50000, 1200)
    ref = ol.oracle_run(opts, batch, want_text=False)
    with GmoveEngine(opts) as eng:
        got = eng.process(batch)
    _assert_same(got, ref, "C3 block")


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
