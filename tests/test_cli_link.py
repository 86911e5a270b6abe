"""The link-level boundary: libporegen_gmove_cli.so exports `int gmove(int, char**)` and
`int reform(int, char**)` with the C++ linkage the reference's dispatcher wants (its main.c is compiled as C++,
Makefile:5,39), and the reference's UNMODIFIED src/main.c, linked against that library (oracle/build_refmain.sh ->
oracle/_ref/poregen_refmain, built where /root/reference exists and carried to the GPU box), runs the cases of the
reference's own test/test_gmove.sh."""
import os
import re
import shutil
import subprocess
import tempfile

import pytest

import oracle_lib as ol
from fixture import FIXTURE_CASES, write_fixture_files
from golden_util import assert_matches_golden, digest, load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLILIB = os.path.join(ROOT, "poregen_b200", "libporegen_gmove_cli.so")
REFMAIN = os.path.join(ROOT, "oracle", "_ref", "poregen_refmain")
HAVE_REF = os.path.isdir("/root/reference/src")


def _need_refmain():
    if not os.path.exists(REFMAIN):
        if HAVE_REF:
            pytest.fail(f"{REFMAIN} missing although /root/reference is here: run oracle/build_refmain.sh")
        pytest.skip("oracle/_ref/poregen_refmain was not built (needs /root/reference)")


def test_library_exports_the_reference_symbols(built):
    assert os.path.exists(CLILIB), f"{CLILIB} missing: poregen_b200/csrc/host/build_cli.sh builds it"
    syms = subprocess.run(["nm", "-D", "--defined-only", CLILIB], capture_output=True, text=True, check=True).stdout
    for s in ("_Z5gmoveiPPc", "_Z6reformiPPc", "pgm_gmove_main", "pgm_reform_main", "pgm_cli_build_info"):
        assert re.search(rf"\sT\s{s}$", syms, re.M), f"{s} not exported"
    hdr = open(os.path.join(ROOT, "include", "pgm_gmove_cli.h")).read()
    for s in ("int gmove(int argc, char *argv[])", "int reform(int argc, char *argv[])", "pgm_gmove_main", "pgm_reform_main", "pgm_cli_build_info"):
        assert s in hdr


def test_cli_binary_matches_its_sources(built):
    """a stale poregen binary next to fresh sources must not pass silently."""
    cli = os.path.join(ROOT, "poregen_b200", "bin", "poregen")
    assert os.path.exists(cli)
    import hashlib
    h = hashlib.sha256()
    for f in ("poregen_b200/csrc/host/gmove_cli.cpp", "poregen_b200/csrc/host/bam_reader.hpp", "poregen_b200/csrc/host/reform.inc", "include/pgm_gmove.h"):
        h.update(open(os.path.join(ROOT, f), "rb").read())
    got = subprocess.run([cli, "--build-info"], capture_output=True, text=True).stdout.strip()
    assert got == h.hexdigest(), "poregen_b200/bin/poregen was built from other sources than the tree holds: rebuild it"


def test_reference_main_dispatches_to_this_build(built):
    _need_refmain()
    run = lambda *a: subprocess.run([REFMAIN, *a], capture_output=True, text=True)
    p = run("--version")
    assert p.returncode == 0 and p.stdout.strip() == "poregen 0.1.0"          # reference src/main.c:84-86
    p = run()
    assert p.returncode == 1 and "Usage: poregen <command>" in p.stderr        # :69-70
    p = run("nonsense")
    assert p.returncode == 1 and "Unrecognised command nonsense" in p.stderr   # :90-91
    p = run("gmove", "--help")
    assert p.returncode == 0 and p.stdout.startswith("Usage: poregen gmove reads.blow5")   # gmove.cpp:80, help on stdout
    p = run("gmove")                                                            # test_gmove.sh case 0.1: no arguments must fail
    assert p.returncode == 1 and "Usage: poregen gmove" in p.stderr
    p = run("gmove", "-h")                                                      # short -h is not in the optstring (gmove.cpp:217)
    assert p.returncode == 1
    p = run("kmer_freq")
    assert p.returncode == 1 and "not part of the gmove hot-path build" in p.stderr
    p = run("reform", "--help")
    assert p.returncode == 0


@pytest.mark.gpu
def test_reference_main_runs_test_gmove_sh(built):
    """reference test/test_gmove.sh, cases 0.2-2.3, through the reference's own main(): expected exit statuses, and the
    outputs against the digests of the unmodified reference (tests/golden/reference_outputs.json)."""
    _need_refmain()
    gold = load_golden()["fixture"]
    d = tempfile.mkdtemp(prefix="pgm_link_")
    try:
        fx = write_fixture_files(os.path.join(d, "single_read"))

        def ex(*args, out):
            return subprocess.run([REFMAIN, "gmove", *args, os.path.join(d, out)], capture_output=True, text=True, timeout=300)

        def files(out):
            o = os.path.join(d, out)
            r = {n: open(os.path.join(o, "dump", n), "rb").read() for n in os.listdir(os.path.join(o, "dump"))}
            r["freq.txt"] = open(os.path.join(o, "freq.txt"), "rb").read()
            return r
        s5, tsv, paf, bam, fq = fx["reads.slow5"], fx["guppy_move"], fx["guppy_move.paf"], fx["guppy_move.bam"], fx["read_0.fastq"]
        p = ex(s5, tsv, "--file_limit", "50", out="0.2")
        assert p.returncode == 0, p.stderr[-2000:]
        assert_matches_golden(digest(files("0.2")), gold["t0.2_tsv_k9_sc0"], "0.2")
        assert ex(s5, tsv, "--kmer_file", fx["kmer_file.txt"], out="0.3").returncode != 0       # k = 9 against 6-mer lines
        assert ex("-k", "6", "-m", "0", s5, tsv, "--kmer_file", fx["kmer_file.txt"], out="0.4").returncode == 0
        assert ex(s5, paf, "--file_limit", "50", out="0.5").returncode != 0                      # PAF needs --fastq
        assert ex(s5, paf, "--file_limit", "50", "--fastq", fq, out="0.6").returncode == 0
        for case, lst in (("1", "single_kmer_file"), ("2", "kmer_file")):
            kf = fx[lst + ".txt"]
            assert ex("-k", "6", "-m", "0", s5, tsv, "--kmer_file", kf, out=case + ".1").returncode == 0
            assert_matches_golden(digest(files(case + ".1")), gold[f"tsv_{lst}_pm2_sc0"], case + ".1")
            assert ex("-k", "6", s5, paf, "--fastq", fq, "--kmer_file", kf, out=case + ".2").returncode == 0
            assert_matches_golden(digest(files(case + ".2")), gold[f"paf_{lst}_pm2_sc0"], case + ".2")
            # the TSV == PAF identity the script asserts holds at --kmer_pick_margin 0
            assert ex("-k", "6", s5, paf, "--fastq", fq, "--kmer_file", kf, "--kmer_pick_margin", "0", out=case + ".2m0").returncode == 0
            assert files(case + ".2m0") == files(case + ".1")
            assert ex("-k", "6", s5, bam, "--kmer_file", kf, out=case + ".3").returncode == 0
            assert files(case + ".3") == files(case + ".1")
    finally:
        shutil.rmtree(d, ignore_errors=True)
