"""pgm_submit / pgm_wait and sibling contexts (include/pgm_gmove.h): two batches in flight, the per-k-mer totals carried on
the device, results identical to the one-batch-at-a-time loop (and so to the oracle's single loop)."""
import copy

import numpy as np
import pytest

import oracle_lib as ol
from cases import parity_cases
from poregen_b200 import _capi as capi, synth
from poregen_b200.gmove import GmoveEngine, PgmError

pytestmark = pytest.mark.gpu

CASES = [c for c in parity_cases() if c[0] in ("tsv_k5_sc1", "ss_dna_k6_sc1", "sam_k5_sc0", "tsv_k9_window_sc0")]


def _cuts(n, parts):
    return [n * i // parts for i in range(parts + 1)]


@pytest.mark.parametrize("label,cfg,opts", CASES, ids=[c[0] for c in CASES])
def test_submit_wait_equals_the_sequential_loop(built, label, cfg, opts):
    batch = synth.make_batch(cfg)
    cuts = _cuts(batch.n_reads, 5)
    pieces = [batch.slice(cuts[i], cuts[i + 1]) for i in range(5)]
    with GmoveEngine(opts) as e:
        want = [e.process(p) for p in pieces]
    with GmoveEngine(opts) as e:
        got = []
        ids = [e.submit(pieces[0]), e.submit(pieces[1])]
        for i in range(5):
            got.append(e.wait(ids[i]))
            if i + 2 < 5:
                ids.append(e.submit(pieces[i + 2]))
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g.slot_total, w.slot_total)
        np.testing.assert_array_equal(g.slot_sel, w.slot_sel)
        np.testing.assert_array_equal(g.ev_len, w.ev_len)
        np.testing.assert_array_equal(g.ev_read, w.ev_read)
        np.testing.assert_array_equal(g.ev_start, w.ev_start)
        assert np.array_equal(g.samples.view(np.uint64), w.samples.view(np.uint64))
    # and the last totals are the oracle's after the whole file
    ref = ol.oracle_run(opts, batch)
    np.testing.assert_array_equal(got[-1].slot_total, ref.slot_total)


def test_submit_wait_text_and_svb(built):
    label, cfg, opts = [c for c in parity_cases() if c[0] == "tsv_k5_sc1"][0]
    topts = copy.copy(opts)
    topts.output = capi.PGM_OUT_TEXT
    batch = synth.align_raw(synth.make_batch(cfg), 8)
    cuts = _cuts(batch.n_reads, 3)
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(topts) as e:
        texts = [b""] * e.n_slots
        ids = []
        for i in range(3):
            piece = synth.rebase(batch, cuts[i], cuts[i + 1])
            blob, off = synth.svb_streams(piece)
            ids.append(e.submit(piece, svb=(blob, off)) if i % 2 else e.submit(piece))
            if len(ids) == 2:
                r = e.wait(ids.pop(0))
                texts = [t + r.slot_text(s) for s, t in enumerate(texts)]
        while ids:
            r = e.wait(ids.pop(0))
            texts = [t + r.slot_text(s) for s, t in enumerate(texts)]
    for s in range(len(texts)):
        assert texts[s] == ref.slot_text(s)


def test_submit_wait_order_is_enforced(built):
    label, cfg, opts = CASES[0]
    batch = synth.make_batch(cfg)
    with GmoveEngine(opts) as e:
        a = e.submit(batch.slice(0, 5))
        b = e.submit(batch.slice(5, 10))
        with pytest.raises(PgmError) as ei:
            e.submit(batch.slice(10, 15))              # two in flight already
        assert ei.value.code == capi.PGM_ERR_STATE
        with pytest.raises(PgmError):
            e.wait(b)                                  # out of order
        e.wait(a)
        e.wait(b)
        with pytest.raises(PgmError):
            e.wait(b + 1)                              # never submitted


def test_siblings_share_the_totals(built):
    """Split-phase calls on two sibling contexts, interleaved: count(j+1) may be queued before emit(j) was even called on the
    host only if the batches are counted in order; here the natural order upload/count/emit per batch, alternating contexts."""
    label, cfg, opts = [c for c in parity_cases() if c[0] == "ss_dna_k6_sc1"][0]
    batch = synth.make_batch(cfg)
    cuts = _cuts(batch.n_reads, 4)
    with GmoveEngine(opts) as e:
        want = [e.process(batch.slice(cuts[i], cuts[i + 1])) for i in range(4)]
    with GmoveEngine(opts) as a:
        b = a.sibling()
        try:
            got = []
            ctx = [a, b]
            pending = []
            for i in range(4):
                c = ctx[i % 2]
                if len(pending) == 2:
                    got.append(pending.pop(0).download())
                c.upload(batch.slice(cuts[i], cuts[i + 1])); c.count(); c.emit()
                pending.append(c)
            while pending:
                got.append(pending.pop(0).download())
        finally:
            b.close()
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g.slot_total, w.slot_total)
        np.testing.assert_array_equal(g.ev_read, w.ev_read)
        assert np.array_equal(g.samples.view(np.uint64), w.samples.view(np.uint64))
