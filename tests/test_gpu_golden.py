"""GPU path vs the reference's own outputs (committed goldens), without going through the oracle:
the doubles returned through the C ABI, printed with "%.8f", must hash to the dump/ files the
unmodified reference wrote."""
import pytest

from cases import parity_cases
from fixture import FIXTURE_CASES, fixture_batch
from golden_util import assert_matches_golden, digest, format_result_files, load_golden
from poregen_b200 import synth
from poregen_b200.gmove import GmoveEngine

pytestmark = pytest.mark.gpu
GOLD = load_golden()


@pytest.mark.parametrize("name,form,opts,kfile", FIXTURE_CASES(), ids=[c[0] for c in FIXTURE_CASES()])
def test_fixture(built, name, form, opts, kfile):
    with GmoveEngine(opts) as eng:
        res = eng.process(fixture_batch(form, opts.kmer_size))
    assert_matches_golden(digest(format_result_files(opts, res)), GOLD["fixture"][name], name)


@pytest.mark.parametrize("label,cfg,opts", parity_cases(), ids=[c[0] for c in parity_cases()])
def test_synthetic(built, label, cfg, opts):
    with GmoveEngine(opts) as eng:
        res = eng.process(synth.make_batch(cfg))
    assert_matches_golden(digest(format_result_files(opts, res)), GOLD["synthetic"][label], label)
