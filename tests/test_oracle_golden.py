"""Pins the CPU oracle to the UNMODIFIED reference: its dump/ + freq.txt must hash to the
outputs the compiled reference produced for (a) the reference's own gmove fixture in TSV, PAF and
SAM form (the runnable cases of test/test_gmove.sh) and (b) every seeded synthetic case
(tests/golden/reference_outputs.json, made by tests/golden/make_golden.py). Runs on CPU."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from cases import parity_cases
from fixture import FIXTURE_CASES, fixture_batch
from golden_util import assert_matches_golden, digest, format_result_files, load_golden
from poregen_b200 import synth

GOLD = load_golden()


@pytest.mark.parametrize("name,form,opts,kfile", FIXTURE_CASES(), ids=[c[0] for c in FIXTURE_CASES()])
def test_fixture_matches_reference(name, form, opts, kfile):
    batch = fixture_batch(form, opts.kmer_size)
    res = ol.oracle_run(opts, batch)
    assert_matches_golden(digest(ol.result_as_files(opts, res)), GOLD["fixture"][name], name)
    # the doubles -> text path of the host writer gives the same bytes as the oracle's own text
    assert format_result_files(opts, res) == ol.result_as_files(opts, res)


def test_known_answer_case_1_1():
    """-k 6 -m 0 --kmer_file single_kmer_file.txt, no scaling."""
    name = "tsv_single_kmer_file_pm2_sc0"
    case = [c for c in FIXTURE_CASES() if c[0] == name][0]
    res = ol.oracle_run(case[2], fixture_batch("tsv", 6))
    assert res.slot_text(0) == (b"143.00816337,118.41790281,115.53301191,119.24215736,114.43400585,117.18152100,"
                                b"113.33499979,115.53301191,128.44633310,162.51552091;")
    assert GOLD["fixture"][name]["freq"] == "ATGTTG\t1\n"


def test_tsv_paf_sam_equivalence():
    """test_gmove.sh 1.1 == 1.2 == 1.3 and 2.1 == 2.2 == 2.3 (PAF at --kmer_pick_margin 0)."""
    for lst in ("single_kmer_file", "kmer_file"):
        for sc in (0, 1):
            a = GOLD["fixture"][f"tsv_{lst}_pm2_sc{sc}"]
            assert a == GOLD["fixture"][f"sam_{lst}_pm2_sc{sc}"]
            assert a == GOLD["fixture"][f"paf_{lst}_pm0_sc{sc}"]


@pytest.mark.parametrize("label,cfg,opts", parity_cases(), ids=[c[0] for c in parity_cases()])
def test_synthetic_matches_reference(label, cfg, opts):
    batch = synth.make_batch(cfg)
    res = ol.oracle_run(opts, batch)
    assert_matches_golden(digest(ol.result_as_files(opts, res)), GOLD["synthetic"][label], label)


@pytest.mark.skipif(not os.path.exists(ol.REF_BIN), reason="oracle/_ref not built here")
def test_oracle_vs_live_reference(tmp_path):
    """Where the compiled reference is present, run it live on a fresh seed (not in the goldens)."""
    from poregen_b200 import _capi as capi
    from poregen_b200.batch import GmoveOptions
    cfg = synth.SynthConfig(n_reads=25, median_samples=3000, mode=capi.PGM_MODE_SS, p_del=0.05, p_ins=0.05, rna_orientation=True,
                            seed=int.from_bytes(os.urandom(2), "little"))
    opts = GmoveOptions(kmer_size=5, rna=True, scaling=1, sample_limit=7, file_limit=1024, index_end=1024).resolve_range()
    batch = synth.make_batch(cfg)
    synth.write_blow5(batch, str(tmp_path / "reads.blow5"))
    fl = synth.write_alignment(batch, str(tmp_path))
    ref = ol.run_reference(opts, str(tmp_path / "reads.blow5"), fl["align"], str(tmp_path / "out"), fastq=fl["fastq"])
    mine = ol.result_as_files(opts, ol.oracle_run(opts, batch))
    assert ref == mine, f"seed {cfg.seed}"


def test_median_mad_units():
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 10, 101, 1000):
        x = rng.normal(0, 3, n)
        s = np.sort(x)
        med = ol.lib().oracle_median(x.ctypes.data, n)
        assert med == s[n // 2]                       # upper median
        mad = ol.lib().oracle_mad(x.ctypes.data, n, med)
        want = 0.0 if n == 1 else np.sort(np.abs(x - med))[n // 2] * 1.4826
        assert mad == want
