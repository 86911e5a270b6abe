#!/usr/bin/env python
"""Makes tests/golden/model_transform_pin.tsv from the reference's own model file (test/data/poregen_5mer.model).

The reference ships the OUTPUT of apply_transformation (scripts/poregen.sh:87-129) but not the raw model that went in.
Every level_mean there is trunc(mean * A) + B in `bc -l` arithmetic with the workflow's A and B (poregen.sh:175) and a
mean that datamash printed with at most 14 significant digits. This script inverts each row: it searches the few
candidate means next to (out - B) / A and keeps the one that reproduces the output digit for digit. A 22-digit output
has such a 14-digit preimage by chance about once in 10^8, so 1024 hits out of 1024 pin the multiplication, the
truncation and the scale rule of tests/test_transform.py's restatement and of csrc/host/transform.inc.
The level_stdv column cannot be inverted (the minimum and maximum of the input column are unknown): it stays unpinned.

Run in the build container only (needs /root/reference): python tests/golden/make_model_pin.py
"""
import os
from decimal import Decimal, getcontext, ROUND_DOWN

getcontext().prec = 80
A, B = Decimal("17.569300789355"), Decimal("84.112089074928")
SRC = "/root/reference/test/data/poregen_5mer.model"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "model_transform_pin.tsv")


def forward(mean: Decimal) -> Decimal:
    sa = max(0, -mean.as_tuple().exponent)
    rs = min(sa + 12, max(20, sa, 12))
    return (mean * A).quantize(Decimal(1).scaleb(-rs), rounding=ROUND_DOWN) + B


def main():
    rows = []
    for ln in open(SRC):
        if ln.startswith("#") or ln.startswith("kmer\t"):
            continue
        kmer, out_mean = ln.rstrip("\n").split("\t")[:2]
        out = Decimal(out_mean)
        sc = -out.as_tuple().exponent
        hit = None
        for sa in ([sc - 12] if sc < 20 else range(8, 16)):
            q = Decimal(1).scaleb(-sa)
            base = ((out - B) / A).quantize(q, rounding=ROUND_DOWN)
            for d in (0, 1, -1):
                c = base + d * q
                if len(c.as_tuple().digits) <= 14 and str(forward(c)) == out_mean and (sa == 0 or not str(c).endswith("0")):   # "%.14g" leaves no trailing zero after the point
                    hit = c
                    break
            if hit is not None:
                break
        assert hit is not None, (kmer, out_mean)
        rows.append((kmer, str(hit), out_mean))
    with open(DST, "w") as f:
        f.write("# kmer\tmean (recovered preimage)\tlevel_mean of the reference's poregen_5mer.model; A=%s B=%s\n" % (A, B))
        for r in rows:
            f.write("\t".join(r) + "\n")
    print(len(rows), "rows ->", DST)


if __name__ == "__main__":
    main()
