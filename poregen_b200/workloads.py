"""The BASELINE.json configurations as (SynthConfig, GmoveOptions) pairs.

C2 is the configuration the bench is quoted on at one GPU; the others are parity-test shapes and
multi-GPU scale points.
"""
from __future__ import annotations

from . import _capi as capi
from .batch import GmoveOptions
from .synth import SynthConfig


def c2(n_reads: int = 10000, seed: int = 0xC2) -> tuple:
    """configs[1]: synthetic RNA004 reads, 10k reads x ~10k samples, ss records in RNA orientation
    (target_start > target_end) with ',' tokens only; gmove -k 5 --rna --scaling 1
    --sample_limit 5000 --min_dur 20 --max_dur 40 --file_limit 5000."""
    cfg = SynthConfig(n_reads=n_reads, median_samples=10000, sigma_len=0.3, min_samples=2000, max_samples=40000,
                      stride=6, mean_dwell_strides=5.0, mode=capi.PGM_MODE_SS, rna_orientation=True,
                      digitisation=8192.0, range_=1467.61, offset_lo=-20, offset_hi=30, spike_frac=0.001, seed=seed)
    opts = GmoveOptions(kmer_size=5, rna=True, scaling=1, sample_limit=5000, min_dur=20, max_dur=40,
                        file_limit=5000, index_start=1, index_end=5000).resolve_range()
    return cfg, opts


def c3(n_reads: int = 100000, seed: int = 0xC3) -> tuple:
    """configs[2]: synthetic DNA R10.4.1 5 kHz move-table reads, ~50k samples each; gmove -k 9
    --sample_limit 1000 --file_limit 262144 (no scaling: the reference's real default)."""
    cfg = SynthConfig(n_reads=n_reads, median_samples=50000, sigma_len=0.3, min_samples=5000, max_samples=250000,
                      stride=6, mean_dwell_strides=12.5 / 6.0, mode=capi.PGM_MODE_MOVES_CHAR,
                      digitisation=2048.0, range_=281.345551, offset_lo=-110, offset_hi=-90, level_k=6, seed=seed)
    opts = GmoveOptions(kmer_size=9, sample_limit=1000, file_limit=262144, index_start=1, index_end=262144).resolve_range()
    return cfg, opts


def c4(n_reads: int = 10000, rna: bool = False, margin: int = 2, seed: int = 0xC4) -> tuple:
    """configs[3]: eventalign-style ss strings with dense indels; gmove -k 6 --kmer_pick_margin 2."""
    cfg = SynthConfig(n_reads=n_reads, median_samples=10000, sigma_len=0.0, min_samples=10000, max_samples=10000,
                      mode=capi.PGM_MODE_SS, rna_orientation=rna, p_del=0.06, p_ins=0.06, seed=seed)
    opts = GmoveOptions(kmer_size=6, rna=rna, kmer_pick_margin=margin, file_limit=4096, index_end=4096).resolve_range()
    return cfg, opts
