"""Shared parity cases: (label, SynthConfig, GmoveOptions). Sized so the oracle takes well under a second each."""
from poregen_b200 import _capi as capi
from poregen_b200.batch import GmoveOptions
from poregen_b200.synth import SynthConfig


def _o(**kw):
    o = GmoveOptions(**kw)
    if "index_end" not in kw and "file_limit" not in kw:
        o.file_limit = o.index_end = 4 ** o.kmer_size
    return o.resolve_range()


def parity_cases():
    S, T, B = capi.PGM_MODE_SS, capi.PGM_MODE_MOVES_CHAR, capi.PGM_MODE_MOVES_I8
    cases = []
    for sc in (0, 1):
        cases += [
            (f"tsv_k5_sc{sc}", SynthConfig(n_reads=60, median_samples=4000, trim_max=40, short_read_frac=0.05, n_frac=0.01, seed=11),
             _o(kmer_size=5, sample_limit=20, scaling=sc)),
            (f"tsv_k9_window_sc{sc}", SynthConfig(n_reads=40, median_samples=6000, stride=5, seed=12),
             GmoveOptions(kmer_size=9, sample_limit=3, scaling=sc, index_start=1000, index_end=19000, file_limit=18001).resolve_range()),
            (f"sam_k5_sc{sc}", SynthConfig(n_reads=60, median_samples=4000, mode=B, trim_max=40, spike_frac=0.00002, seed=13),
             _o(kmer_size=5, sample_limit=20, scaling=sc)),
            (f"ss_dna_k6_sc{sc}", SynthConfig(n_reads=60, median_samples=4000, mode=S, p_del=0.06, p_ins=0.06, slice_max=7, seed=14),
             _o(kmer_size=6, sample_limit=20, scaling=sc)),
            (f"ss_dna_k6_margin0_sc{sc}", SynthConfig(n_reads=30, median_samples=4000, mode=S, p_del=0.06, p_ins=0.06, seed=15),
             _o(kmer_size=6, sample_limit=50, scaling=sc, kmer_pick_margin=0)),
            (f"ss_rna_k5_sc{sc}", SynthConfig(n_reads=60, median_samples=4000, mode=S, p_del=0.06, p_ins=0.06, slice_max=7,
                                             rna_orientation=True, seed=16),
             _o(kmer_size=5, sample_limit=20, scaling=sc, rna=True, min_dur=20, max_dur=40)),
        ]
    cases += [
        ("tsv_offsets", SynthConfig(n_reads=30, median_samples=3000, seed=17),
         _o(kmer_size=6, sample_limit=1000, sig_move_offset=2, kmer_start_offset=1, signal_print_margin=3, scaling=1)),
        ("tsv_limit0", SynthConfig(n_reads=10, median_samples=3000, seed=18), _o(kmer_size=4, sample_limit=0)),
        ("tsv_k1", SynthConfig(n_reads=10, median_samples=3000, seed=19), _o(kmer_size=1, sample_limit=100000, scaling=1)),
        ("tsv_wide_pa", SynthConfig(n_reads=20, median_samples=3000, seed=20, spike_frac=0.01),
         _o(kmer_size=5, sample_limit=30, scaling=1, pa_min=-1e9, pa_max=1e9)),
        ("tsv_narrow_pa", SynthConfig(n_reads=20, median_samples=3000, seed=21),
         _o(kmer_size=5, sample_limit=30, scaling=1, pa_min=85.0, pa_max=95.0)),
        ("tsv_negative_range", SynthConfig(n_reads=20, median_samples=3000, seed=22, range_=-1467.61, offset_lo=-1200, offset_hi=-900),
         _o(kmer_size=5, sample_limit=30, scaling=1, pa_min=40.0, pa_max=180.0)),
        ("ss_m1_margin", SynthConfig(n_reads=30, median_samples=3000, mode=S, p_del=0.03, p_ins=0.03, seed=23),
         _o(kmer_size=6, sample_limit=100, sig_move_offset=1, signal_print_margin=2, kmer_pick_margin=3, scaling=1)),
        ("tsv_kmer_list", SynthConfig(n_reads=40, median_samples=4000, seed=24),
         GmoveOptions(kmer_size=6, sample_limit=5, kmer_list=["ATGTTG", "CAGTTA", "GGGGGG", "ACGTAC", "TTTTTT", "NNNNNN"],
                      file_limit=6, index_end=6).resolve_range()),
    ]
    return cases
