"""Host-side containers for the gmove hot path: options, record batches, results.

Mirrors the reference's option block (src/poregen.h:48-79, defaults src/poregen.cpp:209-237 and
gmove()'s own handling src/gmove.cpp:240-327,428-440) and the data the three record loops read
(src/gmove.cpp:539-706, 707-975, 1061-1266), laid out as include/pgm_gmove.h wants it.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import _capi as capi


@dataclass
class GmoveOptions:
    """`poregen gmove` options after gmove()'s parsing. Defaults = init_opt + gmove.cpp:229
    (scaling really defaults to 0)."""

    kmer_size: int = 9
    sig_move_offset: int = 0
    kmer_start_offset: int = 0
    signal_print_margin: int = 0
    sample_limit: int = 100
    file_limit: int = 500
    index_start: int = 1
    index_end: int = 500
    max_dur: int = 70
    min_dur: int = 5
    pa_min: float = 40.0
    pa_max: float = 180.0
    kmer_pick_margin: int = 2
    rna: bool = False
    scaling: int = 0
    delimit_files: bool = False
    kmer_list: Optional[List[str]] = None  # --kmer_file lines (whole file), else generated
    output: int = capi.PGM_OUT_F64

    def num_kmers(self) -> int:
        return len(self.kmer_list) if self.kmer_list is not None else 4 ** self.kmer_size

    def resolve_range(self) -> "GmoveOptions":
        """The file_limit / index_start / index_end reconciliation of gmove.cpp:428-440."""
        n = self.num_kmers()
        if self.file_limit < n:
            pass
        elif self.file_limit > n - self.index_start + 1:
            if self.index_end > n:
                self.file_limit = n - self.index_start + 1
                self.index_end = self.index_start + self.file_limit - 1
            else:
                self.file_limit = self.index_end - self.index_start + 1
        return self

    def slot_kmers(self) -> List[str]:
        """K-mer strings of the dumped range kmers[index_start-1 .. index_end-1] (gmove.cpp:462)."""
        lo, hi = self.index_start - 1, self.index_end
        if self.kmer_list is not None:
            return list(self.kmer_list[lo:hi])
        al = "ACGU" if self.rna else "ACGT"
        k = self.kmer_size
        out = []
        for i in range(lo, hi):
            out.append("".join(al[(i >> (2 * (k - 1 - j))) & 3] for j in range(k)))
        return out

    def to_c(self):
        """Returns (PgmOpts, keepalive). Duplicate --kmer_file lines are not representable
        (the reference's std::map collapses them); callers de-duplicate first."""
        o = capi.PgmOpts()
        o.kmer_size = self.kmer_size
        o.sig_move_offset = self.sig_move_offset
        o.kmer_start_offset = self.kmer_start_offset
        o.signal_print_margin = self.signal_print_margin
        o.sample_limit = self.sample_limit
        o.max_dur = self.max_dur & 0xFFFFFFFF
        o.min_dur = self.min_dur & 0xFFFFFFFF
        o.kmer_pick_margin = self.kmer_pick_margin & 0xFFFFFFFF
        o.flag_rna = 1 if self.rna else 0
        o.scaling = self.scaling
        o.output = self.output
        o.n_slots = self.index_end - self.index_start + 1
        o.index_lo = self.index_start - 1
        o.pa_min = self.pa_min
        o.pa_max = self.pa_max
        keep = None
        if self.kmer_list is not None:
            sl = self.slot_kmers()
            if len(set(sl)) != len(sl):
                raise ValueError("duplicate k-mers in the dumped range")
            keep = "".join(sl).encode()
            o.kmer_list = keep
        else:
            o.kmer_list = None
        return o, keep


@dataclass
class ReadBatch:
    """Consecutive alignment records in file order, CSR-style."""

    mode: int
    raw: np.ndarray      # int16 [n_raw]
    align: np.ndarray    # uint8 [n_align]
    bases: np.ndarray    # uint8 [n_bases]
    reads: np.ndarray    # capi.READ_DTYPE [n_reads]
    read_ids: List[str] = field(default_factory=list)

    def __post_init__(self):
        self.raw = np.ascontiguousarray(self.raw, dtype=np.int16)
        self.align = np.ascontiguousarray(self.align, dtype=np.uint8)
        self.bases = np.ascontiguousarray(self.bases, dtype=np.uint8)
        self.reads = np.ascontiguousarray(self.reads, dtype=np.dtype(capi.READ_DTYPE))
        assert self.reads.dtype.itemsize == C.sizeof(capi.PgmRead)

    @property
    def n_reads(self) -> int:
        return int(self.reads.shape[0])

    def to_c(self) -> capi.PgmBatch:
        b = capi.PgmBatch()
        b.n_reads = self.n_reads
        b.mode = self.mode
        b.n_raw = self.raw.size
        b.n_align = self.align.size
        b.n_bases = self.bases.size
        b.raw = self.raw.ctypes.data
        b.align = self.align.ctypes.data
        b.bases = self.bases.ctypes.data
        b.reads = self.reads.ctypes.data
        return b

    def slice(self, lo: int, hi: int) -> "ReadBatch":
        """Records [lo, hi) as a batch sharing the big arrays (offsets stay absolute)."""
        return ReadBatch(self.mode, self.raw, self.align, self.bases, self.reads[lo:hi].copy(),
                         self.read_ids[lo:hi])

    def algorithmic_bytes(self, n_slots: int, n_sel_events: int, n_sel_samples: int) -> int:
        """2·S + T + N_b + 24·R + 8·S_sel + 12·E_sel + 4·n_slots."""
        r = self.reads
        s = int(r["raw_len"].astype(np.int64).sum())
        t = int(r["align_len"].astype(np.int64).sum())
        nb = int(r["seq_len"].astype(np.int64).sum())
        return 2 * s + t + nb + 24 * self.n_reads + 8 * n_sel_samples + 12 * n_sel_events + 4 * n_slots


def _np_from(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    dt = np.dtype(dtype)
    buf = (C.c_char * (n * dt.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dt, count=n).copy()


@dataclass
class GmoveResult:
    n_slots: int
    n_reads: int
    n_events_examined: int
    n_events_pass: int
    slot_sel: np.ndarray
    slot_total: np.ndarray
    ev_len: np.ndarray
    ev_read: np.ndarray
    ev_start: np.ndarray
    samples: np.ndarray
    read_flags: np.ndarray
    text: Optional[bytes] = None
    slot_text_off: Optional[np.ndarray] = None

    @staticmethod
    def from_c(r, want_text: bool) -> "GmoveResult":
        ne, ns = int(r.n_sel_events), int(r.n_sel_samples)
        text = None
        off = None
        if want_text and r.slot_text_off:
            off = _np_from(r.slot_text_off, r.n_slots + 1, np.uint64)
            tl = int(off[-1])
            text = C.string_at(r.text, tl) if tl else b""
        return GmoveResult(
            int(r.n_slots), int(r.n_reads), int(r.n_events_examined), int(r.n_events_pass),
            _np_from(r.slot_sel, r.n_slots, np.uint32), _np_from(r.slot_total, r.n_slots, np.uint32),
            _np_from(r.ev_len, ne, np.uint32), _np_from(r.ev_read, ne, np.uint32),
            _np_from(r.ev_start, ne, np.uint32), _np_from(r.samples, ns, np.float64),
            _np_from(r.read_flags, r.n_reads, np.uint8), text, off)

    def slot_text(self, s: int) -> bytes:
        return self.text[int(self.slot_text_off[s]):int(self.slot_text_off[s + 1])]
