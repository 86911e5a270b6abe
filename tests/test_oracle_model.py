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
