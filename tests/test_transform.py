"""apply_transformation (reference scripts/poregen.sh:87-129): raw model -> ONT .model text.

The product is csrc/host/transform.inc behind `poregen transform` / pgm_transform_main / `gmove --ont_model`.
The checker here is an independent restatement on Python's decimal module (ROUND_DOWN) of bc's published rules.
Pinned: level_mean, on the reference's own model file (tests/golden/model_transform_pin.tsv, made by
tests/golden/make_model_pin.py). Unpinned: level_stdv (no bc/datamash here, no reference input/output pair).
"""
import ctypes as C
import os
import random
import re
import subprocess
from decimal import Decimal, getcontext, ROUND_DOWN

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "poregen_b200", "bin", "poregen")
CLI_LIB = os.path.join(ROOT, "poregen_b200", "libporegen_gmove_cli.so")
GOLD = os.path.join(ROOT, "tests", "golden", "model_transform_pin.tsv")
HEADER = ("#ont_model_name\tnone\n#kit\tnone\n#strand\ttemplate\n#k\t5\n#alphabet\tnucleotide\n#original_file\tnone\n"
          "kmer\tlevel_mean\tlevel_stdv\tsd_mean\tsd_stdv\tweight\n")
getcontext().prec = 200
NUM = re.compile(r"-?(\d+\.?\d*|\.\d+)$")
WORD = re.compile(r"-?[a-z][a-z0-9_]*$")

pytestmark = pytest.mark.skipif(not os.path.exists(BIN), reason="CLI not built (python -c 'import __graft_entry__ as g; g.build()')")


def scale(x: Decimal) -> int:
    return max(0, -x.as_tuple().exponent)


def trunc(x: Decimal, s: int) -> Decimal:
    return x.quantize(Decimal(1).scaleb(-s), rounding=ROUND_DOWN)


def bc_mul(a, b):
    return trunc(a * b, min(scale(a) + scale(b), max(20, scale(a), scale(b))))


def bc_str(x: Decimal) -> str:
    if x == 0:
        return "0"
    t = format(abs(x), "f")
    if t.startswith("0."):
        t = t[1:]
    return ("-" if x < 0 else "") + t


def operand(t: str):
    if NUM.match(t):
        return Decimal(t if not t.endswith(".") else t + "0").quantize(Decimal(1).scaleb(-len(t.split(".")[1]) if "." in t else 0))
    if WORD.match(t):
        return Decimal(0)
    return None


def expected(text: str, A, B, Cc, D) -> str:
    lines = text.split("\n")[:-1]                      # a last line without '\n' is not processed by `read`
    cut3, rows = [], []
    for ln in lines:
        parts = ln.split("\t")
        cut3.append(ln if len(parts) == 1 else (parts[2] if len(parts) > 2 else ""))
        f = [x for x in parts if x != ""]              # IFS white space: runs of tabs are one separator
        rows.append((f + ["", "", ""])[:3])
    rng = None
    try:
        vals = [float(c) for c in cut3]
        if cut3 and all(c != "" and v == v for c, v in zip(cut3, vals)):
            rng = (operand("%.14g" % min(vals)), operand("%.14g" % max(vals)))
            if rng[0] is None or rng[1] is None:
                rng = None
    except ValueError:
        rng = None
    a, b, c, d = (operand(x) for x in (A, B, Cc, D))
    out = HEADER
    for kmer, mean, sd in rows:
        om = osd = ""
        m, s = operand(mean), operand(sd)
        if m is not None and a is not None and b is not None:
            om = bc_str(bc_mul(m, a) + b)
        if s is not None and rng and c is not None and d is not None and rng[1] - rng[0] != 0:
            osd = bc_str(trunc(bc_mul(s - rng[0], d - c) / (rng[1] - rng[0]), 20) + c)
        out += f"{kmer}\t{om}\t{osd}\n"
    return out


def run_transform(tmp_path, text, A, B, Cc, D):
    src, dst = tmp_path / "raw_model", tmp_path / "out.model"
    src.write_text(text)
    p = subprocess.run([BIN, "transform", str(src), str(dst), A, B, Cc, D], capture_output=True)
    assert p.returncode == 0, p.stderr.decode()
    return dst.read_text()


WORKFLOW = ("17.569300789355", "84.112089074928", "2.5", "4")


def test_golden_level_mean(tmp_path):
    rows = [ln.rstrip("\n").split("\t") for ln in open(GOLD) if not ln.startswith("#")]
    assert len(rows) == 1024
    text = "".join(f"{k}\t{m}\t{1 + i % 7}.5\n" for i, (k, m, _) in enumerate(rows))
    got = run_transform(tmp_path, text, *WORKFLOW)
    assert got.startswith(HEADER)
    body = got[len(HEADER):].split("\n")[:-1]
    assert [ln.split("\t")[1] for ln in body] == [r[2] for r in rows]          # digit for digit the reference's column
    assert got == expected(text, *WORKFLOW)


def fmt14(v):
    return "%.14g" % v


def test_random_models(tmp_path):
    rnd = random.Random(7)
    for trial in range(30):
        n = rnd.choice([1, 2, 5, 64, 300])
        lines = []
        for i in range(n):
            mean = rnd.choice([rnd.uniform(-3, 3), rnd.uniform(40, 160), rnd.randint(-5, 200), rnd.uniform(0, 1e-3), 0.5])
            sd = rnd.choice([rnd.uniform(0.1, 9), 3.1, rnd.randint(1, 4), rnd.uniform(0, 0.01)])
            lines.append(f"K{i}\t{fmt14(mean)}\t{fmt14(sd)}\n")
        text = "".join(lines)
        abcd = rnd.choice([WORKFLOW, ("1", "0", "0", "1"), ("-2.25", "-.5", "1.000", "7"), ("0.1", "100", "4", "2.5"),
                           ("3.14159265358979323846264338", "2", "0.00000000000000000000001", "5")])
        assert run_transform(tmp_path, text, *abcd) == expected(text, *abcd), (trial, abcd)


def test_edge_rows(tmp_path):
    cases = [
        "A\t1.5\t2\nC\t\t\nG\t3\t4\n",              # an empty row (k-mer without data): datamash fails, stddev column empty
        "A\t1.5\t2\nC\t2.5\t2\n",                    # max == min: bc divide by zero, stddev column empty
        "A\t1e-05\t2\nC\t2\t3\n",                    # exponent notation is a bc syntax error
        "A\tnan\t2\nC\t2\t3\n",                      # a lower-case word is a bc variable: 0
        "A\t1\tnan\nC\t2\t3\n",                      # nan in the stddev column: no range
        "A\t1.25\t2\nC\t2\t3",                       # last line without newline is not read
        "",                                          # empty input: header only
        "A\t-0.75\t2\nC\t.5\t3.5\nG\t007.50\t2.75\n",
    ]
    for text in cases:
        assert run_transform(tmp_path, text, *WORKFLOW) == expected(text, *WORKFLOW), repr(text)
    out = run_transform(tmp_path, cases[1], *WORKFLOW)
    assert out.split("\n")[7] == "A\t110.4660402589605\t"           # 1.5 * A = 26.3539511840325 exactly (scale 13), + B


def test_usage_and_errors(tmp_path):
    p = subprocess.run([BIN, "transform", "x"], capture_output=True)
    assert p.returncode == 1 and b"Usage: apply_transformation <input_file> <output_file> <A> <B> <C> <D>" in p.stderr
    p = subprocess.run([BIN, "transform", str(tmp_path / "missing"), str(tmp_path / "o"), *WORKFLOW], capture_output=True)
    assert p.returncode == 1 and b"Error in opening" in p.stderr


def test_library_entry(tmp_path):
    lib = C.CDLL(CLI_LIB)
    lib.pgm_transform_main.restype = C.c_int
    src, dst = tmp_path / "raw", tmp_path / "out"
    text = "AAAAA\t79.8712345678\t3.1\nAAAAC\t0.5\t1.25\nAAAAG\t-3\t2\n"
    src.write_text(text)
    args = [b"transform", str(src).encode(), str(dst).encode()] + [x.encode() for x in WORKFLOW]
    argv = (C.c_char_p * len(args))(*args)
    assert lib.pgm_transform_main(len(args), argv) == 0
    got = dst.read_text()
    assert got == expected(text, *WORKFLOW)
    assert got.split("\n")[7:10] == ["AAAAA\t1487.39383361373490226576\t4.00000000000000000000",
                                     "AAAAC\t92.8967394696055\t2.50000000000000000000",
                                     "AAAAG\t31.404186706863\t3.10810810810810810810"]
