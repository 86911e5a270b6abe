"""The counter-based generator behind bench.py's C3 workload (poregen_b200/synth_fast.c): a read is a function of
(seed, global read index) only, the batch is well-formed, and the oracle accepts it."""
import numpy as np

import oracle_lib as ol
from poregen_b200 import synth, workloads


def test_reads_do_not_depend_on_the_cut(built):
    cfg, _ = workloads.c3(n_reads=100000)
    a = synth.make_moves_fast(cfg, 0, 40)
    b = synth.make_moves_fast(cfg, 25, 10)
    for i in range(10):
        ra, rb = a.reads[25 + i], b.reads[i]
        for f in ("raw_len", "align_len", "seq_len", "offset", "stride"):
            assert ra[f] == rb[f]
        assert np.array_equal(a.raw[int(ra["raw_off"]):int(ra["raw_off"]) + int(ra["raw_len"])],
                              b.raw[int(rb["raw_off"]):int(rb["raw_off"]) + int(rb["raw_len"])])
        assert np.array_equal(a.align[int(ra["align_off"]):int(ra["align_off"]) + int(ra["align_len"])],
                              b.align[int(rb["align_off"]):int(rb["align_off"]) + int(rb["align_len"])])
        assert np.array_equal(a.bases[int(ra["seq_off"]):int(ra["seq_off"]) + int(ra["seq_len"])],
                              b.bases[int(rb["seq_off"]):int(rb["seq_off"]) + int(rb["seq_len"])])
    assert a.read_ids[25:35] == b.read_ids


def test_shape_and_oracle(built):
    cfg, opts = workloads.c3(n_reads=100000)
    b = synth.make_moves_fast(cfg, 1000, 24)
    rd = b.reads
    assert (rd["raw_off"] % 8 == 0).all()
    ones = np.array([int((b.align[int(r["align_off"]):int(r["align_off"]) + int(r["align_len"])] == ord("1")).sum()) for r in rd])
    assert np.array_equal(ones, rd["seq_len"])                      # one set move per base
    assert (rd["raw_len"] >= rd["align_len"] * cfg.stride).all() and (rd["raw_len"] < (rd["align_len"] + 1) * cfg.stride).all()
    spb = rd["raw_len"].sum() / rd["seq_len"].sum()
    assert 11.5 < spb < 13.5
    res = ol.oracle_run(opts, b)
    assert res.n_events_examined == int(rd["seq_len"].sum()) - rd.size      # the last event of a record is never emitted
    assert res.ev_len.size > 0.95 * res.n_events_examined                 # durations sit inside [5, 70], every 9-mer is dumped
    zero = float((res.samples == 0.0).mean())
    assert 0.0002 < zero < 0.005                                         # the spikes outside [pa_min, pa_max] print as 0.0
    sub = synth.rebase(b, 5, 9)
    r2 = ol.oracle_run(opts, sub)
    r1 = ol.oracle_run(opts, b.slice(5, 9))
    assert np.array_equal(r1.samples, r2.samples) and np.array_equal(r1.ev_start, r2.ev_start)
