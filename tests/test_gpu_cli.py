"""The C++ command line (poregen_b200/bin/poregen gmove ...) against the reference: same files in, same
options, dump/ and freq.txt byte for byte. The expected outputs are the committed digests of the UNMODIFIED
reference (tests/golden/reference_outputs.json); where the compiled reference travelled with the tree
(oracle/_ref) extra cases (-d delimiters, several device batches, no .fai) are diffed against it live."""
import os
import shutil
import subprocess
import tempfile

import pytest

import oracle_lib as ol
from cases import _o, parity_cases
from golden_util import assert_matches_golden, digest, load_golden
from poregen_b200 import _capi as capi, synth
from poregen_b200.synth import SynthConfig

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "poregen_b200", "bin", "poregen")


def _make_files(cfg, opts, d):
    batch = synth.make_batch(cfg)
    synth.write_blow5(batch, os.path.join(d, "reads.blow5"))
    fl = synth.write_alignment(batch, d)
    kf = None
    if opts.kmer_list is not None:
        kf = os.path.join(d, "kmers.txt")
        open(kf, "w").write("".join(k + "\n" for k in opts.kmer_list))
    return fl, kf


def _run_cli(opts, d, fl, kf, outname="o", extra=()):
    cmd = [CLI, "gmove"] + ol.cli_args(opts) + list(extra)
    if kf:
        cmd += ["--kmer_file", kf]
    if fl["fastq"]:
        cmd += ["--fastq", fl["fastq"]]
    out = os.path.join(d, outname)
    cmd += [os.path.join(d, "reads.blow5"), fl["align"], out]
    p = subprocess.run(cmd, capture_output=True, timeout=300, env=dict(os.environ, PGM_CLI_SHARE_DEVICES="1"))
    assert p.returncode == 0, p.stderr.decode()[-3000:]
    files = {}
    for name in os.listdir(os.path.join(out, "dump")):
        files[name] = open(os.path.join(out, "dump", name), "rb").read()
    files["freq.txt"] = open(os.path.join(out, "freq.txt"), "rb").read()
    return files


@pytest.fixture(scope="module")
def cli(built):
    if not os.path.exists(CLI):
        pytest.fail(f"{CLI} is missing: build it with poregen_b200/csrc/host/build_cli.sh")
    return CLI


@pytest.mark.parametrize("label,cfg,opts", parity_cases(), ids=[c[0] for c in parity_cases()])
def test_cli_matches_reference_digests(cli, label, cfg, opts):
    want = load_golden()["synthetic"][label]
    d = tempfile.mkdtemp(prefix="pgm_cli_")
    try:
        fl, kf = _make_files(cfg, opts, d)
        got = _run_cli(opts, d, fl, kf)
        assert_matches_golden(digest(got), want, label)
        # the same run cut into several device batches must not change a byte
        got2 = _run_cli(opts, d, fl, kf, outname="o2", extra=["--batch-samples", "30000"])
        assert got2 == got
    finally:
        shutil.rmtree(d, ignore_errors=True)


_LIVE = [
    ("tsv_delimit", SynthConfig(n_reads=40, median_samples=3000, short_read_frac=0.1, seed=31), dict(kmer_size=4, sample_limit=15, scaling=1, delimit_files=True)),
    ("tsv_delimit_limit0", SynthConfig(n_reads=10, median_samples=3000, seed=32), dict(kmer_size=3, sample_limit=0, delimit_files=True)),
    ("ss_rna_delimit", SynthConfig(n_reads=40, median_samples=3000, mode=capi.PGM_MODE_SS, p_del=0.05, p_ins=0.05, rna_orientation=True, seed=33),
     dict(kmer_size=5, sample_limit=10, scaling=1, rna=True, delimit_files=True)),
    ("sam_delimit", SynthConfig(n_reads=40, median_samples=3000, mode=capi.PGM_MODE_MOVES_I8, seed=34), dict(kmer_size=5, sample_limit=25, delimit_files=True)),
    # a --kmer_file that repeats lines: the reference keys files and counts by the k-mer string (std::map, gmove.cpp:460-473), so the
    # repeats share one file and one count, and freq.txt prints that count on every one of the lines
    ("tsv_kmer_file_repeated_lines", SynthConfig(n_reads=30, median_samples=3000, seed=35),
     dict(kmer_size=3, sample_limit=12, kmer_list=["AAC", "ACG", "AAC", "GGT", "ACG", "TTT", "AAC"], file_limit=7, index_end=7)),
]   # (with files still open at the end - e.g. --sample_limit 0 - the reference fclose()s the repeated string's FILE* once per line and
    #  aborts in free(): undefined there; every k-mer of the case above reaches its limit, which closes each file once)


@pytest.mark.parametrize("label,cfg,kw", _LIVE, ids=[c[0] for c in _LIVE])
def test_cli_matches_live_reference(cli, label, cfg, kw):
    if not os.path.exists(ol.REF_BIN):
        pytest.skip("compiled reference (oracle/_ref) not present")
    opts = _o(**kw)
    d = tempfile.mkdtemp(prefix="pgm_cli_")
    try:
        fl, kf = _make_files(cfg, opts, d)
        want = ol.run_reference(opts, os.path.join(d, "reads.blow5"), fl["align"], os.path.join(d, "ref"), fastq=fl["fastq"], kmer_file=kf)
        for extra in ((), ("--batch-samples", "20000")):
            got = _run_cli(opts, d, fl, kf, outname="o" + str(len(extra)), extra=extra)
            assert got.keys() == want.keys()
            bad = [k for k in want if got[k] != want[k]]
            assert not bad, f"{label}: {bad[:5]} differ from the reference"
        if fl["fastq"]:   # no .fai next to the FASTQ: the index is built on the fly
            os.remove(fl["fastq"] + ".fai")
            got = _run_cli(opts, d, fl, kf, outname="nofai")
            assert got == want
    finally:
        shutil.rmtree(d, ignore_errors=True)


def test_cli_usage_errors(cli):
    assert subprocess.run([CLI, "gmove"], capture_output=True).returncode == 1              # help on stderr
    p = subprocess.run([CLI, "gmove", "--help"], capture_output=True)
    assert p.returncode == 0 and b"Usage: poregen gmove" in p.stdout
    assert subprocess.run([CLI, "gmove", "-h"], capture_output=True).returncode == 1        # -h is not in the optstring (gmove.cpp:217)
    p = subprocess.run([CLI, "gmove", "--version"], capture_output=True)
    assert p.returncode == 0 and p.stdout.startswith(b"gmove ")


def test_cli_model_rows(cli):
    """--model FILE: the rows calculate_mean_stddev_all (reference scripts/poregen.sh:54-85) would compute from
    the dump files the same run wrote, several device batches; checked against the script's arithmetic
    restated on those files' text (oracle_model_row_text; parity of this row is unpinned)."""
    opts = _o(kmer_size=3, sample_limit=300, scaling=1)
    cfg = SynthConfig(n_reads=50, median_samples=3000, seed=77)
    with tempfile.TemporaryDirectory(prefix="pgm_cli_model_") as d:
        fl, kf = _make_files(cfg, opts, d)
        model = os.path.join(d, "model.tsv")
        dwell = os.path.join(d, "dwell.tsv")
        files = _run_cli(opts, d, fl, kf, extra=["--model", model, "--stddev_limit", "1.05", "--batch-samples", "40000", "--dwell", dwell])
        lines = open(model).read().splitlines()
        dlines = open(dwell).read().splitlines()
    names = sorted(n for n in files if n != "freq.txt")
    assert [ln.split("\t")[0] for ln in lines] == names
    # --dwell: calculate_dwell_times_medians (scripts/poregen.sh:33-52) on the files the same run wrote
    assert [ln.split("\t")[0] for ln in dlines] == names
    for ln in dlines:
        name, med_s = ln.split("\t")
        n, tm = ol.dwell_median_text(files[name])
        assert med_s == ("" if n == 0 else "%.14g" % (tm / 2.0))
    for ln in lines:
        name, med_s, sd_s = ln.split("\t")
        n, med, sd, ms, ss = ol.model_row_text(files[name], 1.05)
        if n == 0:
            assert med_s == "" and sd_s == ""
            continue
        assert med_s == ms
        if n == 1:
            assert sd_s == "nan"
        else:
            assert abs(float(sd_s) - sd) <= 1e-12 * sd


def test_cli_reads_bam(cli):
    """guppy/dorado hand over BAM: the same records as BAM (BGZF, records straddling members) must give the
    files the text SAM gives, which are pinned to the reference by the digests above (the reference's own
    BAM case, test_gmove.sh 1.4/2.4, needs htslib)."""
    label, cfg, opts = next(c for c in parity_cases() if c[0] == "sam_k5_sc1")
    with tempfile.TemporaryDirectory(prefix="pgm_cli_bam_") as d:
        fl, kf = _make_files(cfg, opts, d)
        bam = os.path.join(d, "moves.bam")
        synth.write_bam(synth.make_batch(cfg), bam, block=2000)
        sam_files = _run_cli(opts, d, fl, kf, outname="o_sam")
        bam_files = _run_cli(opts, d, dict(fl, align=bam), kf, outname="o_bam")
    assert bam_files == sam_files
    assert_matches_golden(digest(bam_files), load_golden()["synthetic"][label], label)


@pytest.mark.parametrize("record_press", ["none", "zlib"])
def test_cli_decodes_svb_zd_signal_on_the_device(cli, record_press):
    """BLOW5 files as slow5tools writes them (signal_press svb-zd, records zlib-compressed or not): the command line
    ships the streams and k_svb_decode unpacks them; the files must equal the reference digests, the host-decoded run
    (--host-signal-decode: slow5lib decompresses) and a run cut into several batches and two device contexts."""
    for label in ("tsv_k5_sc1", "ss_rna_k5_sc1"):
        _, cfg, opts = next(c for c in parity_cases() if c[0] == label)
        with tempfile.TemporaryDirectory(prefix="pgm_cli_svb_") as d:
            fl, kf = _make_files(cfg, opts, d)
            os.remove(os.path.join(d, "reads.blow5"))
            for f in os.listdir(d):
                if f.endswith(".idx"):
                    os.remove(os.path.join(d, f))
            synth.write_blow5(synth.make_batch(cfg), os.path.join(d, "reads.blow5"), signal_press="svb-zd", record_press=record_press)
            dev = _run_cli(opts, d, fl, kf, outname="o_dev")
            host = _run_cli(opts, d, fl, kf, outname="o_host", extra=["--host-signal-decode"])
            cut = _run_cli(opts, d, fl, kf, outname="o_cut", extra=["--batch-samples", "30000", "--gpus", "2"])
        assert dev == host and cut == dev
        assert_matches_golden(digest(dev), load_golden()["synthetic"][label], label)
