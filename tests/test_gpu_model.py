"""K7 — per-k-mer model rows (reference scripts/poregen.sh:54-85) on the device against the script's
arithmetic restated on the dump TEXT (oracle_model_row_text: strtold + long double, like GNU datamash).

Parity of this row is UNPINNED (datamash is not installed, no reference test holds a model row), so the
bar is: the median, an exact quantity, must agree digit for digit; the stddev within 1e-12 relative of the
long-double evaluation."""
import math

import numpy as np
import pytest

import oracle_lib as ol
from cases import _o
from poregen_b200 import _capi as capi, synth
from poregen_b200.batch import GmoveOptions
from poregen_b200.gmove import GmoveEngine
from poregen_b200.synth import SynthConfig

pytestmark = pytest.mark.gpu


def _slot_texts(res):
    off = np.concatenate(([0], np.cumsum(res.slot_sel.astype(np.int64))))
    soff = np.concatenate(([0], np.cumsum(res.ev_len.astype(np.int64))))
    out = []
    for s in range(res.slot_sel.size):
        parts = []
        for e in range(off[s], off[s + 1]):
            parts.append(",".join("%.8f" % v for v in res.samples[soff[e]:soff[e + 1]]) + ";")
        out.append("".join(parts).encode())
    return out


def _check(rows, texts, limit):
    assert rows.size == len(texts)
    for s, text in enumerate(texts):
        n, med, sd, med_s, sd_s = ol.model_row_text(text, limit)
        r = rows[s]
        if n == 0:
            assert r["flags"] & capi.PGM_MODEL_EMPTY and math.isnan(r["median"])
            continue
        assert int(r["n"]) == n
        # the median is exact: median_num / 2e8, printed like datamash prints it
        assert "%.14g" % (int(r["median_num"]) / 2e8) == med_s, (s, r, med_s)
        assert r["median"] == med
        if n < 2:
            assert r["flags"] & capi.PGM_MODEL_NO_SD and math.isnan(r["stddev"])
        else:
            assert abs(r["stddev"] - sd) <= 1e-12 * abs(sd), (s, r["stddev"], sd)
            assert bool(r["flags"] & capi.PGM_MODEL_CLAMPED) == (sd == limit and r["stddev"] == limit)


@pytest.mark.parametrize("scaling,limit", [(0, 1e9), (1, 1e9), (1, 0.9)], ids=["pA", "medmad", "clamped"])
def test_model_rows_one_batch(built, scaling, limit):
    opts = _o(kmer_size=4, sample_limit=60, scaling=scaling)
    batch = synth.make_batch(SynthConfig(n_reads=80, median_samples=4000, seed=301))
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(opts) as eng:
        eng.model_begin()
        eng.process(batch)
        rows = eng.model_finalize(limit)
    _check(rows, _slot_texts(ref), limit)


def test_model_rows_accumulate_over_batches(built):
    """The dump files grow over the record loop; so does the model log. Four batches, totals carried:
    the rows must be those of the whole run."""
    opts = _o(kmer_size=3, sample_limit=500, scaling=1)
    batch = synth.make_batch(SynthConfig(n_reads=64, median_samples=3000, seed=302))
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(opts) as eng:
        eng.model_begin()
        for lo in range(0, 64, 16):
            eng.process(batch.slice(lo, lo + 16))
        rows = eng.model_finalize(1e9)
    _check(rows, _slot_texts(ref), 1e9)


def test_model_rows_few_values(built):
    """k-mers with zero, one or two values: datamash prints nothing / a median and nan."""
    opts = _o(kmer_size=6, sample_limit=1, scaling=0, min_dur=1, max_dur=2)
    batch = synth.make_batch(SynthConfig(n_reads=6, median_samples=1500, seed=303))
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(opts) as eng:
        eng.model_begin()
        eng.process(batch)
        rows = eng.model_finalize(1e9)
    _check(rows, _slot_texts(ref), 1e9)
    assert (rows["flags"] & capi.PGM_MODEL_EMPTY).any()


def _check_dwell(tm, texts):
    assert tm.size == len(texts)
    for s, text in enumerate(texts):
        n, want = ol.dwell_median_text(text)
        if n == 0:
            assert tm[s] == -1, (s, tm[s])
        else:
            assert int(tm[s]) == want, (s, int(tm[s]), want, n)


@pytest.mark.parametrize("nbatches", [1, 4])
def test_dwell_medians(built, nbatches):
    """calculate_dwell_times_medians (reference scripts/poregen.sh:33-52): median of the per-event comma counts (+ the empty last
    field) of every dump file, from the logged event lengths; one batch and a run cut into four (totals carried)."""
    opts = _o(kmer_size=4, sample_limit=37, scaling=1)
    batch = synth.make_batch(SynthConfig(n_reads=64, median_samples=3000, seed=305))
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(opts) as eng:
        eng.model_begin()
        step = 64 // nbatches
        for lo in range(0, 64, step):
            eng.process(batch.slice(lo, lo + step))
        tm = eng.model_dwell()
        rows = eng.model_finalize(1e9)          # the level rows are still available afterwards
        tm2 = eng.model_dwell()
    _check_dwell(tm, _slot_texts(ref))
    assert np.array_equal(tm, tm2)
    _check(rows, _slot_texts(ref), 1e9)


def test_dwell_medians_sparse(built):
    opts = _o(kmer_size=6, sample_limit=2, scaling=0, min_dur=1, max_dur=12)
    batch = synth.make_batch(SynthConfig(n_reads=6, median_samples=1500, seed=306))
    ref = ol.oracle_run(opts, batch)
    with GmoveEngine(opts) as eng:
        eng.model_begin()
        eng.process(batch)
        tm = eng.model_dwell()
    _check_dwell(tm, _slot_texts(ref))
    assert (tm == -1).any()          # k-mers without events: the file is empty, datamash prints nothing


@pytest.mark.parametrize("n_ranks", [2, 3, 8])
def test_model_rows_across_ranks_equal_one_context(built, n_ranks):
    """K7 across GPUs (BASELINE.json configs[4]), the ranks' contexts on one device: every round the records are dealt to the
    ranks in contiguous ranges with the quota exchange of a multi-GPU job, every rank logs its own values, the owners of the
    k-mer ranges collect them round by round and rank by rank. The rows must be the ones a single context that saw every
    record in file order produces - including which value is the file's first (the one `tail -n +2` drops)."""
    import torch
    from poregen_b200.gmove import model_merge_local, quota_exchange
    cfg = synth.SynthConfig(n_reads=96, median_samples=5000, seed=77)
    opts = GmoveOptions(kmer_size=5, sample_limit=40, file_limit=1024, index_end=1024).resolve_range()
    batch = synth.make_batch(cfg)
    n_rounds = 3
    rcut = [batch.n_reads * i // n_rounds for i in range(n_rounds + 1)]
    # one context over the whole file, streamed in the same rounds
    with GmoveEngine(opts) as one:
        one.model_begin()
        for r in range(n_rounds):
            one.process(batch.slice(rcut[r], rcut[r + 1]))
        want = one.model_finalize(3.1)
    engines = [GmoveEngine(opts) for _ in range(n_ranks)]
    try:
        for e in engines:
            e.model_begin()
        totals = np.zeros(engines[0].n_slots, np.uint32)
        for r in range(n_rounds):
            n = rcut[r + 1] - rcut[r]
            cuts = [rcut[r] + n * g // n_ranks for g in range(n_ranks + 1)]
            counts = []
            for g, e in enumerate(engines):
                e.upload(batch.slice(cuts[g], cuts[g + 1])); e.count(); e.sync()
                t = torch.as_tensor(type("A", (), {"__cuda_array_interface__": {"shape": (e.n_slots,), "typestr": "<u4", "data": (e.counts_ptr(), False), "version": 3}})(), device="cuda:0")
                counts.append(t.cpu().numpy().astype(np.uint32))
            prior, consumed = quota_exchange(np.stack(counts), totals, opts.sample_limit)
            keep = []
            for g, e in enumerate(engines):
                p = torch.from_numpy(prior[g].astype(np.int32)).cuda(); cns = torch.from_numpy(consumed.astype(np.int32)).cuda()
                keep += [p, cns]
                e.set_prior(p.data_ptr(), cns.data_ptr())
                e.emit()
                e.model_log()
            torch.cuda.synchronize()
            totals = consumed
        got = model_merge_local(engines, opts, 3.1)
    finally:
        for e in engines:
            e.close()
    for f in ("n", "flags", "median_num"):
        np.testing.assert_array_equal(got[f], want[f], err_msg=f)
    assert np.array_equal(got["median"].view(np.uint64), want["median"].view(np.uint64))
    assert np.array_equal(got["stddev"].view(np.uint64), want["stddev"].view(np.uint64))
