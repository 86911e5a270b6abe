"""`poregen reform` (poregen_b200/csrc/host/reform.inc + bam_reader.hpp; host-side, no GPU) against the
reference's src/reform.cpp: committed digests of the unmodified reference's output on synthetic SAM files,
the same records as BAM (BGZF members that split records), and — where the reference tree is present — the
reference's own 14 runnable cases of test/test_reform.sh with their expected files."""
import hashlib
import json
import os
import subprocess
import tempfile

import pytest

from poregen_b200 import synth
from reform_cases import CASES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "poregen_b200", "bin", "poregen")
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "reform_outputs.json")))
REF_TREE = "/root/reference/test/data"

pytestmark = pytest.mark.skipif(not os.path.exists(CLI), reason="poregen_b200/bin/poregen has not been built (build_cli.sh)")


def _run(args, path):
    return subprocess.run([CLI, "reform"] + args + [path], capture_output=True, timeout=120)


@pytest.mark.parametrize("label,cfg,args", CASES, ids=[c[0] for c in CASES])
def test_reform_matches_reference_digest_from_sam_and_bam(label, cfg, args):
    batch = synth.make_batch(cfg)
    with tempfile.TemporaryDirectory(prefix="pgm_reform_") as d:
        fl = synth.write_alignment(batch, d)
        bam = os.path.join(d, "moves.bam")
        synth.write_bam(batch, bam, block=1500)
        for path in (fl["align"], bam):
            p = _run(args, path)
            assert p.returncode == 0, p.stderr.decode()[-2000:]
            assert len(p.stdout) == GOLD[label]["bytes"]
            assert hashlib.sha256(p.stdout).hexdigest() == GOLD[label]["sha256"], (label, path)


def test_reform_rejects_what_the_reference_rejects():
    cfg = CASES[0][1]
    with tempfile.TemporaryDirectory(prefix="pgm_reform_") as d:
        fl = synth.write_alignment(synth.make_batch(cfg), d)
        assert _run(["-k", "0", "-m", "1", "-c"], fl["align"]).returncode != 0      # test_reform.sh case 2
        assert _run(["-k", "9", "-m", "10", "-c"], fl["align"]).returncode != 0     # case 4
        assert subprocess.run([CLI, "reform"], capture_output=True).returncode != 0  # case 1
        six = synth.make_batch(type(cfg)(**{**cfg.__dict__, "stride": 6}))
        fl6 = synth.write_alignment(six, os.path.join(d, "six"))
        p = _run(["-k", "9", "-m", "0"], fl6["align"])
        assert p.returncode != 0 and b"expected stride of 5 is missing" in p.stderr   # reform.cpp:241
        assert _run(["-k", "9", "-m", "0", "--stride", "0"], fl6["align"]).returncode == 0


def test_bam_reader_rejects_garbage():
    with tempfile.TemporaryDirectory(prefix="pgm_reform_") as d:
        bad = os.path.join(d, "x.bam")
        good = os.path.join(d, "g.bam")
        synth.write_bam(synth.make_batch(CASES[0][1]), good)
        data = open(good, "rb").read()
        open(bad, "wb").write(data[:len(data) // 2])              # truncated in the middle of a member
        assert _run(["-k", "9", "-m", "0"], bad).returncode != 0
        flipped = bytearray(data)
        flipped[200] ^= 0x55                                       # corrupts the deflate stream or its checksum
        open(bad, "wb").write(bytes(flipped))
        assert _run(["-k", "9", "-m", "0"], bad).returncode != 0


_REF_CASES = [("guppy_one_read.bam", n, ["-k", k, "-m", m]) for n, k, m in
              [("r1k1m0", "1", "0"), ("r1k9m0", "9", "0"), ("r1k9m1", "9", "1"), ("r1k9m6", "9", "6"), ("r1k9m8", "9", "8")]] + \
             [("slow5-dorado.sam", n, ["-k", k, "-m", m]) for n, k, m in [("dr2k9m8", "9", "8"), ("dr2k1m0", "1", "0")]]


@pytest.mark.skipif(not os.path.isdir(REF_TREE), reason="the reference tree is not on this machine")
@pytest.mark.parametrize("inp,name,args", _REF_CASES, ids=[c[1] for c in _REF_CASES])
def test_reference_own_reform_cases(inp, name, args):
    """test/test_reform.sh cases 5-18: real guppy BAM and dorado SAM, expected files shipped by the reference."""
    for extra, ext in ((["-c"], ".paf"), ([], ".tsv")):
        p = _run(args + extra, os.path.join(REF_TREE, "raw", "reform", inp))
        assert p.returncode == 0, p.stderr.decode()[-2000:]
        assert p.stdout == open(os.path.join(REF_TREE, "exp", "reform", name + ext), "rb").read()
