"""The script-literal model row (oracle_model_row_text; reference scripts/poregen.sh:54-85) on hand-checked texts."""
import math

import oracle_lib as ol


def test_drop_first_median_sstdev():
    n, med, sd, ms, ss = ol.model_row_text(b"1.00000000,2.00000000;3.00000000,5.00000000;", 1e9)
    assert n == 3 and med == 3.0 and ms == "3"
    assert abs(sd - math.sqrt(7.0 / 3.0)) < 1e-15 and ss == "1.5275252316519"


def test_even_count_takes_mean_of_middles_and_clamps():
    n, med, sd, ms, ss = ol.model_row_text(b"9.00000000;1.00000000,2.00000001,4.00000000,10.00000000;", 2.5)
    assert n == 4 and ms == "3.000000005" and med == 3.000000005
    assert sd == 2.5 and ss == "2.5"


def test_too_few_values():
    assert ol.model_row_text(b"", 1.0)[0] == 0
    assert ol.model_row_text(b"1.50000000;", 1.0)[0] == 0
    n, med, sd, ms, ss = ol.model_row_text(b"1.50000000,-2.25000000;", 1.0)
    assert n == 1 and med == -2.25 and math.isnan(sd)


def test_dwell_median_known_answers():
    """scripts/poregen.sh:33-52 by hand: fields of "1,2,3;4,5;6;" are "1,2,3" "4,5" "6" "" -> comma counts 2,1,0,0 -> median 0.5."""
    assert ol.dwell_median_text(b"1.0,2.0,3.0;4.0,5.0;6.0;") == (4, 1)
    assert ol.dwell_median_text(b"") == (0, None)
    assert ol.dwell_median_text(b"1.0,2.0,3.0;") == (2, 2)            # counts 2 and 0 -> median 1
    assert ol.dwell_median_text(b"1,2,3,4;1,2,3,4;1,2;1;") == (5, 2)   # counts 3,3,1,0,0 -> median 1


# ------------------------------------------------------------------------------------------------
# A second, independent restatement of the script's pipeline, in exact rational arithmetic with
# an explicit model of the x87 long double (64-bit significand, round to nearest even) that GNU datamash computes in:
#   tr ';,' '\n' < file | tail -n +2 | datamash median 1      (poregen.sh:73)
#   tr ';,' '\n' < file | tail -n +2 | datamash sstdev 1      (poregen.sh:74; a separate run: values in FILE order)
#   awk -F';' '{for(i=1;i<=NF;i++) print gsub(",", "", $i)}' file | datamash median 1      (poregen.sh:49)
# and datamash's "%.14Lg" output, formatted here digit by digit from the exact value.
# This pins the C oracle's arithmetic and printing against a different implementation; it does NOT pin either against
# GNU datamash itself (not installed, version not fixed by the reference): parity for K7 stays "unpinned".
from fractions import Fraction  # noqa: E402
import random  # noqa: E402


def ld(x: Fraction) -> Fraction:
    """x rounded to the nearest 64-bit-significand binary float (ties to even); exponent range ignored."""
    if x == 0:
        return x
    s, a = (-1 if x < 0 else 1), abs(x)
    e = a.numerator.bit_length() - a.denominator.bit_length() - 64
    while a / Fraction(2) ** e >= 2 ** 64:
        e += 1
    while a / Fraction(2) ** e < 2 ** 63:
        e -= 1
    q = a / Fraction(2) ** e
    n, r = divmod(q.numerator, q.denominator)
    twice = 2 * r
    if twice > q.denominator or (twice == q.denominator and n & 1):
        n += 1
    return s * n * Fraction(2) ** e


def ld_sqrt(x: Fraction) -> Fraction:
    import math
    if x == 0:
        return x
    k = 200                                             # floor(sqrt(x) * 2^k): far more bits than the 64 kept
    t = (x.numerator << (2 * k)) // x.denominator
    r = math.isqrt(t)
    exact = r * r == t and (x.numerator << (2 * k)) % x.denominator == 0
    return ld(Fraction(r, 1 << k) if exact else Fraction(2 * r + 1, 1 << (k + 1)))   # inexact: any point strictly inside (r, r+1) rounds alike


def fmt14g(x: Fraction) -> str:
    """printf("%.14Lg") of an exact value."""
    if x == 0:
        return "0"
    sign, a = ("-" if x < 0 else ""), abs(x)
    e10 = len(str(a.numerator)) - len(str(a.denominator))
    while Fraction(10) ** e10 > a:
        e10 -= 1
    while Fraction(10) ** (e10 + 1) <= a:
        e10 += 1
    q = a / Fraction(10) ** (e10 - 13)                  # 14 significant digits in front of the point
    n, r = divmod(q.numerator, q.denominator)
    if 2 * r > q.denominator or (2 * r == q.denominator and n & 1):
        n += 1
    if n == 10 ** 14:
        n, e10 = 10 ** 13, e10 + 1
    digits = str(n)
    if -4 <= e10 < 14:
        if e10 >= 0:
            ip, fp = digits[:e10 + 1], digits[e10 + 1:]
        else:
            ip, fp = "0", "0" * (-e10 - 1) + digits
        fp = fp.rstrip("0")
        return sign + ip + ("." + fp if fp else "")
    fp = digits[1:].rstrip("0")
    return f"{sign}{digits[0]}{'.' + fp if fp else ''}e{'-' if e10 < 0 else '+'}{abs(e10):02d}"


def script_row(text: str, limit: str):
    toks = [t for t in text.replace(";", "\n").replace(",", "\n").split("\n")]
    if toks and toks[-1] == "":
        toks = toks[:-1]                                # tr leaves a final newline after the last ';': no extra line
    vals = [ld(Fraction(t)) for t in toks[1:]]          # tail -n +2
    if not vals:
        return 0, "", ""
    s = sorted(vals)
    m = len(s)
    med = s[m // 2] if m & 1 else ld(ld(s[m // 2 - 1] + s[m // 2]) / 2)
    if m < 2:
        return m, fmt14g(med), "nan"
    total = Fraction(0)
    for v in vals:
        total = ld(total + v)
    mean = ld(total / m)
    ss = Fraction(0)
    for v in vals:
        d = ld(v - mean)
        ss = ld(ss + ld(d * d))
    sd = ld_sqrt(ld(ss / (m - 1)))
    sd_txt = fmt14g(sd)
    if Fraction(sd_txt) > Fraction(limit):              # `echo "$stddev > $limit" | bc -l` compares the PRINTED text
        sd_txt = limit
    return m, fmt14g(med), sd_txt


def _check(text: str, limit="1000000"):
    n, med, sd, ms, ss = ol.model_row_text(text.encode(), float(limit))
    want = script_row(text, limit)
    assert (n, ms, ss) == want, (text[:200], (n, ms, ss), want)


def _f8(v):
    return "%.8f" % v


def test_independent_restatement_hand_cases():
    _check("1.00000000,2.00000000;3.00000000,5.00000000;")
    _check("9.00000000;1.00000000,2.00000001,4.00000000,10.00000000;", "2.5")
    _check("1.50000000,-2.25000000;")
    _check("7.00000000;")
    _check("")
    _check("0.10000000,0.10000000,0.10000000,0.10000000;")                       # all equal: sstdev 0 (or a rounding residue)
    _check("5.00000000;0.00000001,0.00000002;0.00000004;")                       # tiny values: %g switches to exponent form
    _check("1.0;123456789012.12345678,123456789012.12345679;0.00000001;")        # a sum that a double cannot hold
    _check("3.0;" + ",".join(["100000000.00000001"] * 3 + ["-100000000.00000001"] * 3) + ";")


def test_independent_restatement_random_texts():
    rnd = random.Random(11)
    for trial in range(300):
        n = rnd.choice([2, 3, 4, 5, 8, 9, 33, 100, 257])
        kind = rnd.randrange(5)
        vals = []
        for _ in range(n):
            if kind == 0:
                vals.append(rnd.gauss(90, 15))
            elif kind == 1:
                vals.append(rnd.choice([80.5, 80.50000001, 80.49999999, 120.25]))           # ties at the middles
            elif kind == 2:
                vals.append(rnd.gauss(0, 1) * 10 ** rnd.randint(-6, 9))                     # wide dynamic range
            elif kind == 3:
                vals.append(rnd.randint(-3, 3) + rnd.choice([0.0, 0.5, 0.00000001]))
            else:
                vals.append(rnd.uniform(40, 180))
        text, i = "", 0
        while i < n:                                                                          # events of 1..12 values
            ln = min(n - i, rnd.randint(1, 12))
            text += ",".join(_f8(v) for v in vals[i:i + ln]) + ";"
            i += ln
        _check(text, rnd.choice(["1000000", "3.1", "0.5", "15"]))


def test_sum_order_is_file_order():
    """datamash sstdev runs on its own copy of the stream: the sum goes in file order, not in the sorted order the
    median pass leaves behind. A text where the two orders round differently in long double."""
    big, small = "4611686018427387904.00000000", "0.75000000"       # 2^62 and 0.75: 2^62 + 0.75 needs 65 bits
    _check("1.0;" + ",".join([small, big, small, "-" + big, small]) + ";")
    _check("1.0;" + ",".join([big, small, small, small, "-" + big]) + ";")


def test_dwell_independent_restatement():
    rnd = random.Random(5)
    for trial in range(200):
        n_ev = rnd.choice([0, 1, 2, 3, 4, 7, 50])
        text = "".join(",".join("1.0" for _ in range(rnd.randint(1, 9))) + ";" for _ in range(n_ev))
        if text == "":
            assert ol.dwell_median_text(b"") == (0, None)
            continue
        counts = sorted(f.count(",") for f in text.split(";"))      # awk -F';': the piece after the last ';' is a field too
        m = len(counts)
        twice = 2 * counts[m // 2] if m & 1 else counts[m // 2 - 1] + counts[m // 2]
        assert ol.dwell_median_text(text.encode()) == (m, twice), text
