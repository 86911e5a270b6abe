"""Seeded synthetic gmove workloads (C2..C5 shapes) as in-memory record batches,
and optionally as the files the reference CLI reads (ASCII SLOW5 + TSV / PAF+FASTQ / SAM).

Nothing here is on the hot path: it only manufactures inputs. numpy-vectorised so that the
1e8-sample C2 workload is generated in seconds.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import _capi as capi
from .batch import ReadBatch


@dataclass
class SynthConfig:
    n_reads: int = 100
    median_samples: int = 10000      # per read, log-normal
    sigma_len: float = 0.3
    min_samples: int = 2000
    max_samples: int = 40000
    stride: int = 6
    mean_dwell_strides: float = 5.0  # mean samples/base = stride * this
    max_dwell_strides: int = 24
    mode: int = capi.PGM_MODE_MOVES_CHAR
    rna_orientation: bool = False    # ss records with target_start > target_end
    digitisation: float = 8192.0
    range_: float = 1467.61
    offset_lo: int = -20
    offset_hi: int = 30
    level_mean: float = 90.0
    level_sd: float = 15.0
    noise_sd: float = 3.0
    spike_frac: float = 0.001        # samples pushed outside [pa_min, pa_max]
    level_k: int = 5
    p_del: float = 0.0               # ss: probability that a target base is deleted
    p_ins: float = 0.0               # ss: probability of an insertion token before a match
    trim_max: int = 0                # moves: trim_offset ~ U[0, trim_max]
    slice_max: int = 0               # ss: st_k ~ U[0, slice_max]
    short_read_frac: float = 0.0     # moves: fraction of reads with < 10 bases
    n_frac: float = 0.0              # fraction of bases replaced by 'N'
    seed: int = 0x5EED


def _lengths(cfg: SynthConfig, rng) -> np.ndarray:
    ln = rng.lognormal(np.log(cfg.median_samples), cfg.sigma_len, cfg.n_reads)
    return np.clip(ln, cfg.min_samples, cfg.max_samples).astype(np.int64)


def make_batch(cfg: SynthConfig) -> ReadBatch:
    rng = np.random.default_rng(cfg.seed)
    R = cfg.n_reads
    target = _lengths(cfg, rng)
    spb = cfg.stride * cfg.mean_dwell_strides
    nb = np.maximum((target / spb).astype(np.int64), 12)                  # bases (=events) per read
    if cfg.short_read_frac > 0 and cfg.mode != capi.PGM_MODE_SS:
        short = rng.random(R) < cfg.short_read_frac
        nb = np.where(short, rng.integers(3, 10, R), nb)
    B = int(nb.sum())
    base_off = np.zeros(R + 1, np.int64)
    np.cumsum(nb, out=base_off[1:])
    read_of_base = np.repeat(np.arange(R), nb)

    codes = rng.integers(0, 4, B, dtype=np.int64)
    dwell = np.minimum(rng.geometric(1.0 / cfg.mean_dwell_strides, B), cfg.max_dwell_strides).astype(np.int64)
    ev_samples = dwell * cfg.stride

    # ---- alignment tokens ----
    is_ss = cfg.mode == capi.PGM_MODE_SS
    deleted = np.zeros(B, bool)
    ins_len = np.zeros(B, np.int64)
    if is_ss:
        if cfg.p_del > 0:
            deleted = rng.random(B) < cfg.p_del
            deleted[base_off[:-1]] = False
            deleted[base_off[1:] - 1] = False
        if cfg.p_ins > 0:
            ins = (rng.random(B) < cfg.p_ins) & ~deleted
            ins[base_off[:-1]] = False
            ins_len = np.where(ins, rng.integers(1, 31, B), 0)
    sig_per_base = np.where(deleted, 0, ev_samples + ins_len)             # raw samples consumed per target base

    trim = rng.integers(0, cfg.trim_max + 1, R).astype(np.int64) if (cfg.trim_max > 0 and not is_ss) else np.zeros(R, np.int64)
    qstart = rng.integers(0, 50, R).astype(np.int64) if is_ss else np.zeros(R, np.int64)
    lead = trim + qstart
    body = np.add.reduceat(sig_per_base, base_off[:-1])
    tail = rng.integers(0, cfg.stride, R)
    raw_len = lead + body + tail
    raw_off = np.zeros(R + 1, np.int64)
    # every read starts on a 16-byte boundary (8 samples) in the batch
    padded = (raw_len + 7) // 8 * 8
    np.cumsum(padded, out=raw_off[1:])
    S = int(raw_off[-1])

    # ---- signal ----
    table = rng.normal(cfg.level_mean, cfg.level_sd, 4 ** cfg.level_k)
    kid = np.zeros(B, np.int64)
    for j in range(cfg.level_k):
        kid = kid * 4 + np.roll(codes, -j)
    level = table[kid]
    offset = rng.integers(cfg.offset_lo, cfg.offset_hi + 1, R).astype(np.float64)
    offset[offset == 0] = 1.0  # slow5lib's ASCII parser rejects "0" when errno is a stale ERANGE (slow5_misc.c:369)
    scale = cfg.range_ / cfg.digitisation
    pa = rng.normal(cfg.level_mean, cfg.noise_sd * 2, S).astype(np.float32)   # lead / tail / padding filler
    # positions of each base's samples
    base_sig_off = np.zeros(B + 1, np.int64)
    np.cumsum(sig_per_base, out=base_sig_off[1:])
    within = base_sig_off[:-1] - base_sig_off[base_off[:-1]][read_of_base]
    base_start = raw_off[:-1][read_of_base] + lead[read_of_base] + within
    tot = int(sig_per_base.sum())
    idx = np.repeat(base_start - base_sig_off[:-1], sig_per_base) + np.arange(tot)
    pa[idx] = np.repeat(level, sig_per_base).astype(np.float32) + rng.normal(0, cfg.noise_sd, tot).astype(np.float32)
    if cfg.spike_frac > 0:
        ns = int(S * cfg.spike_frac)
        sp = rng.integers(0, S, ns)
        pa[sp] = np.where(rng.random(ns) < 0.5, 15.0, 230.0).astype(np.float32)
    off_s = np.repeat(offset, padded).astype(np.float32)
    raw = np.clip(np.rint(pa / np.float32(scale) - off_s), -32768, 32767).astype(np.int16)
    del pa, off_s, idx

    # ---- bases ----
    alpha = np.frombuffer(b"ACGT", np.uint8)
    chars = alpha[codes]
    if cfg.n_frac > 0:
        chars = np.where(rng.random(B) < cfg.n_frac, np.uint8(ord("N")), chars)

    reads = np.zeros(R, np.dtype(capi.READ_DTYPE))
    reads["raw_off"] = raw_off[:-1]
    reads["raw_len"] = raw_len
    reads["digitisation"] = cfg.digitisation
    reads["offset"] = offset
    reads["range"] = cfg.range_
    reads["stride"] = cfg.stride
    read_ids = [f"read_{cfg.seed & 0xffff:04x}_{i:07d}" for i in range(R)]

    if not is_ss:
        # move table: '1' then dwell-1 '0's per base
        M = int(dwell.sum())
        mv = np.zeros(M, np.uint8)
        mstart = np.zeros(B + 1, np.int64)
        np.cumsum(dwell, out=mstart[1:])
        mv[mstart[:-1]] = 1
        if cfg.mode == capi.PGM_MODE_MOVES_CHAR:
            mv = mv + np.uint8(ord("0"))
        reads["align_off"] = mstart[base_off[:-1]]
        reads["align_len"] = mstart[base_off[1:]] - mstart[base_off[:-1]]
        reads["seq_off"] = base_off[:-1]
        reads["seq_len"] = nb
        reads["fastq_len"] = nb
        reads["trim"] = trim
        return ReadBatch(cfg.mode, raw, mv, chars, reads, read_ids)

    # ---- ss strings ----
    # per target base, tokens: [<ins>I] then <n>, ; deleted runs become one <n>D token
    if cfg.rna_orientation:
        # events run 3'->5': event j <-> target base nb-1-j. Build tokens in event order by
        # walking each read's bases backwards.
        order = (base_off[1:][read_of_base] - 1) - (np.arange(B) - base_off[:-1][read_of_base])
    else:
        order = np.arange(B)
    del_o, ins_o, evs_o = deleted[order], ins_len[order], ev_samples[order]
    prev_del = np.concatenate(([False], del_o[:-1]))
    prev_del[base_off[:-1]] = False
    run_start = del_o & ~prev_del
    # length of each deletion run
    run_id = np.cumsum(run_start) - 1
    run_len = np.bincount(run_id[del_o], minlength=int(run_start.sum())) if del_o.any() else np.zeros(0, np.int64)
    pieces: List[np.ndarray] = []
    tok_read: List[np.ndarray] = []
    tok_key: List[np.ndarray] = []
    pos = np.arange(B)
    # insertion tokens (key = 3*pos), match tokens (3*pos+1), deletion tokens (3*pos+2 at run start)
    mi = ins_o > 0
    if mi.any():
        pieces.append(np.char.add(ins_o[mi].astype(str), "I")); tok_read.append(read_of_base[mi]); tok_key.append(3 * pos[mi])
    mm = ~del_o
    pieces.append(np.char.add(evs_o[mm].astype(str), ",")); tok_read.append(read_of_base[mm]); tok_key.append(3 * pos[mm] + 1)
    if run_start.any():
        pieces.append(np.char.add(run_len.astype(str), "D")); tok_read.append(read_of_base[run_start]); tok_key.append(3 * pos[run_start] + 2)
    toks = np.concatenate(pieces)
    tr = np.concatenate(tok_read)
    tk = np.concatenate(tok_key)
    o = np.argsort(tk, kind="stable")
    toks, tr = toks[o], tr[o]
    tcount = np.bincount(tr, minlength=R)
    toff = np.zeros(R + 1, np.int64)
    np.cumsum(tcount, out=toff[1:])
    ss_list = ["".join(toks[toff[i]:toff[i + 1]]) for i in range(R)]
    ss_len = np.array([len(s) for s in ss_list], np.int64)
    ss_off = np.zeros(R + 1, np.int64)
    np.cumsum(ss_len, out=ss_off[1:])
    align = np.frombuffer("".join(ss_list).encode(), np.uint8)

    st_k = rng.integers(0, cfg.slice_max + 1, R).astype(np.int64) if cfg.slice_max > 0 else np.zeros(R, np.int64)
    reads["align_off"] = ss_off[:-1]
    reads["align_len"] = ss_len
    reads["seq_off"] = base_off[:-1]
    reads["seq_len"] = nb                      # the fetched slice [st_k, end_k)
    reads["fastq_len"] = nb
    reads["query_start"] = qstart
    reads["st_k"] = st_k
    reads["end_k"] = st_k + nb
    reads["flags"] = capi.PGM_READ_RNA if cfg.rna_orientation else 0
    return ReadBatch(cfg.mode, raw, align, chars, reads, read_ids)


# ---------------------------------------------------------------------------------------------
# File forms of a batch, for the reference CLI (`poregen gmove reads.slow5 <alignment> out`).
# ---------------------------------------------------------------------------------------------

_SLOW5_HDR = (
    "#slow5_version\t0.2.0\n#num_read_groups\t1\n@asic_id\tsynthetic\n@sample_frequency\t4000\n"
    "#char*\tuint32_t\tdouble\tdouble\tdouble\tdouble\tuint64_t\tint16_t*\n"
    "#read_id\tread_group\tdigitisation\toffset\trange\tsampling_rate\tlen_raw_signal\traw_signal\n")


def _fmt_g(x: float) -> str:
    return repr(float(x)) if float(x) != int(x) else str(int(x))


def write_slow5(batch: ReadBatch, path: str) -> None:
    """ASCII SLOW5 (slow5lib reads it and builds the .idx on first open)."""
    with open(path, "w") as f:
        f.write(_SLOW5_HDR)
        for i, rid in enumerate(batch.read_ids):
            r = batch.reads[i]
            sig = batch.raw[int(r["raw_off"]):int(r["raw_off"]) + int(r["raw_len"])]
            f.write(f"{rid}\t0\t{_fmt_g(r['digitisation'])}\t{_fmt_g(r['offset'])}\t{_fmt_g(r['range'])}\t4000\t{int(r['raw_len'])}\t")
            f.write(",".join(map(str, sig.tolist())))
            f.write("\n")


def _seq(batch: ReadBatch, i: int) -> str:
    r = batch.reads[i]
    return batch.bases[int(r["seq_off"]):int(r["seq_off"]) + int(r["seq_len"])].tobytes().decode()


def _align(batch: ReadBatch, i: int) -> np.ndarray:
    r = batch.reads[i]
    return batch.align[int(r["align_off"]):int(r["align_off"]) + int(r["align_len"])]


def write_alignment(batch: ReadBatch, outdir: str) -> dict:
    """Writes the alignment in the form matching batch.mode; returns {'align': path, 'fastq': path|None}."""
    os.makedirs(outdir, exist_ok=True)
    res = {"fastq": None}
    if batch.mode == capi.PGM_MODE_MOVES_CHAR:
        p = os.path.join(outdir, "moves.tsv")
        with open(p, "w") as f:
            for i, rid in enumerate(batch.read_ids):
                r = batch.reads[i]
                f.write(f"{rid}\t{int(r['fastq_len'])}\t{_seq(batch, i)}\t{int(r['stride'])}\t"
                        f"{_align(batch, i).tobytes().decode()}\t{int(r['raw_len'])}\t{int(r['trim'])}\n")
        res["align"] = p
    elif batch.mode == capi.PGM_MODE_MOVES_I8:
        p = os.path.join(outdir, "moves.sam")
        with open(p, "w") as f:
            f.write("@HD\tVN:1.6\tSO:unknown\n@PG\tID:synth\tPN:synth\n")
            for i, rid in enumerate(batch.read_ids):
                r = batch.reads[i]
                mv = ",".join(map(str, _align(batch, i).astype(np.int8).tolist()))
                s = _seq(batch, i)
                f.write(f"{rid}\t4\t*\t0\t0\t*\t*\t0\t0\t{s}\t*\tmv:B:c,{int(r['stride'])},{mv}\t"
                        f"ns:i:{int(r['raw_len'])}\tts:i:{int(r['trim'])}\n")
        res["align"] = p
    else:
        p = os.path.join(outdir, "events.paf")
        fq = os.path.join(outdir, "reads.fastq")
        with open(p, "w") as f, open(fq, "w") as g, open(fq + ".fai", "w") as h:
            off = 0
            for i, rid in enumerate(batch.read_ids):
                r = batch.reads[i]
                st, en = int(r["st_k"]), int(r["end_k"])
                full = ("ACGT" * (st // 4 + 1))[:st] + _seq(batch, i) + "TTGCA"
                ts, te = (en, st) if (int(r["flags"]) & capi.PGM_READ_RNA) else (st, en)
                ss = _align(batch, i).tobytes().decode()
                f.write(f"{rid}\t{int(r['raw_len'])}\t{int(r['query_start'])}\t{int(r['raw_len'])}\t+\t{rid}\t{len(full)}\t"
                        f"{ts}\t{te}\t{en - st}\t{en - st}\t255\tss:Z:{ss}\n")
                hdr = f"@{rid}\n"
                g.write(hdr + full + "\n+\n" + "I" * len(full) + "\n")
                h.write(f"{rid}\t{len(full)}\t{off + len(hdr)}\t{len(full)}\t{len(full) + 1}\t{off + len(hdr) + len(full) + 3}\n")
                off += len(hdr) + 2 * (len(full) + 1) + 2
        res["align"] = p
        res["fastq"] = fq
    return res


def svb_zd_encode(x: np.ndarray) -> bytes:
    """One signal as slow5lib stores it under signal_press svb-zd (reference slow5lib/src/slow5_press.c:1034-1083:
    u32 count, then StreamVByte keys and data of the zig-zag coded differences, thirdparty/streamvbyte/src/
    streamvbyte_encode.c and streamvbyte_zigzag.c:19-32). numpy, for test and bench inputs."""
    x = np.ascontiguousarray(x, dtype=np.int16).astype(np.int32)
    n = x.size
    d = np.diff(x, prepend=np.int32(0))
    z = ((d << 1) ^ (d >> 31)).astype(np.uint32)
    code = (z > 0xFF).astype(np.uint8) + (z > 0xFFFF).astype(np.uint8) + (z > 0xFFFFFF).astype(np.uint8)
    kc = np.zeros((n + 3) // 4 * 4, dtype=np.uint8)
    kc[:n] = code
    kc = kc.reshape(-1, 4)
    keys = (kc[:, 0] | (kc[:, 1] << 2) | (kc[:, 2] << 4) | (kc[:, 3] << 6)).astype(np.uint8)
    b4 = z.astype("<u4").view(np.uint8).reshape(-1, 4)
    keep = np.arange(4, dtype=np.uint8)[None, :] <= code[:, None]
    return np.uint32(n).tobytes() + keys.tobytes() + b4[keep].tobytes()


def svb_streams(batch: ReadBatch):
    """(bytes uint8[], off uint64[R+1]): every record's signal as its svb-zd stream, concatenated."""
    parts, off = [], [0]
    for r in batch.reads:
        o, n = int(r["raw_off"]), int(r["raw_len"])
        parts.append(svb_zd_encode(batch.raw[o:o + n]))
        off.append(off[-1] + len(parts[-1]))
    blob = np.frombuffer(b"".join(parts), dtype=np.uint8).copy() if parts else np.zeros(0, np.uint8)
    return blob, np.asarray(off, dtype=np.uint64)


def align_raw(batch: ReadBatch, align: int = 8) -> ReadBatch:
    """The same records with every raw_off rounded up to a multiple of `align` samples (the gaps are zero): the layout
    a host that assembles batches for pgm_upload_svb picks, so that the decoder's stores are 16-byte aligned."""
    reads = batch.reads.copy()
    lens = reads["raw_len"].astype(np.int64)
    padded = (lens + align - 1) // align * align
    offs = np.concatenate([[0], np.cumsum(padded)[:-1]]) if lens.size else np.zeros(0, np.int64)
    raw = np.zeros(int(padded.sum()), np.int16)
    for r, o in zip(reads, offs):
        n = int(r["raw_len"])
        raw[o:o + n] = batch.raw[int(r["raw_off"]):int(r["raw_off"]) + n]
    reads["raw_off"] = offs.astype(np.uint64)
    return ReadBatch(batch.mode, raw, batch.align, batch.bases, reads, batch.read_ids)


def write_blow5(batch: ReadBatch, path: str, signal_press: str = "none", record_press: str = "none") -> None:
    """Binary BLOW5 as slow5lib serialises it (reference slow5lib/src/slow5.c:946-1147 header, :3907-4033 record, EOF
    marker slow5_defs.h:133). Default: nothing compressed (fast to write and read for the CPU-baseline sample).
    signal_press "svb-zd": raw_signal holds the svb-zd stream and len_raw_signal its byte length (slow5lib's default
    for new files; on-disk code 1, slow5_press.c:46-66); record_press "zlib": every record is one zlib stream (code 1)."""
    import struct
    import zlib
    sp = {"none": 0, "svb-zd": 1}[signal_press]
    rp = {"none": 0, "zlib": 1}[record_press]
    hdr_txt = ("@asic_id\tsynthetic\n@sample_frequency\t4000\n"
               "#char*\tuint32_t\tdouble\tdouble\tdouble\tdouble\tuint64_t\tint16_t*\n"
               "#read_id\tread_group\tdigitisation\toffset\trange\tsampling_rate\tlen_raw_signal\traw_signal\n").encode()
    with open(path, "wb") as f:
        head = b"BLOW5\x01" + bytes([0, 2, 0]) + bytes([rp]) + struct.pack("<I", 1) + bytes([sp])
        f.write(head + b"\0" * (64 - len(head)))
        f.write(struct.pack("<I", len(hdr_txt)))
        f.write(hdr_txt)
        for i, rid in enumerate(batch.read_ids):
            r = batch.reads[i]
            n = int(r["raw_len"])
            rb = rid.encode()
            sig = batch.raw[int(r["raw_off"]):int(r["raw_off"]) + n]
            payload = svb_zd_encode(sig) if sp else sig.tobytes()
            rec = struct.pack("<H", len(rb)) + rb + struct.pack("<IddddQ", 0, float(r["digitisation"]), float(r["offset"]),
                                                                 float(r["range"]), 4000.0, len(payload) if sp else n) + payload
            if rp:
                rec = zlib.compress(rec)
            f.write(struct.pack("<Q", len(rec)))
            f.write(rec)
        f.write(b"5WOLB")


def write_bam(batch: ReadBatch, path: str, block: int = 3000):
    """The SAM records of write_alignment (PGM_MODE_MOVES_I8) as a BAM file: BGZF members of at most `block`
    uncompressed bytes (small on purpose: records straddle members), unmapped reads with SEQ and the tags
    mv:B:c, ns:i, ts:i; ns/ts are stored in the smallest integer type like htslib does. Test helper."""
    import struct
    import zlib
    assert batch.mode == capi.PGM_MODE_MOVES_I8
    codes = {c: i for i, c in enumerate("=ACMGRSVTWYHKDBN")}

    def aux_int(tag, v):
        if 0 <= v < 256:
            return tag + b"C" + struct.pack("<B", v)
        if 0 <= v < 65536:
            return tag + b"S" + struct.pack("<H", v)
        return tag + b"I" + struct.pack("<I", v)

    text = b"@HD\tVN:1.6\tSO:unknown\n@PG\tID:synth\tPN:synth\n"
    raw = bytearray(b"BAM\x01" + struct.pack("<i", len(text)) + text + struct.pack("<i", 0))
    for i, rid in enumerate(batch.read_ids):
        r = batch.reads[i]
        seq = _seq(batch, i)
        name = rid.encode() + b"\x00"
        packed = bytearray((len(seq) + 1) // 2)
        for j, ch in enumerate(seq):
            packed[j >> 1] |= codes.get(ch, 15) << (4 if (j & 1) == 0 else 0)
        mv = np.concatenate(([int(r["stride"])], _align(batch, i).astype(np.int8))).astype(np.int8)
        aux = b"mvBc" + struct.pack("<I", mv.size) + mv.tobytes() + aux_int(b"ns", int(r["raw_len"])) + aux_int(b"ts", int(r["trim"]))
        body = struct.pack("<iiBBHHHIiii", -1, -1, len(name), 0, 4680, 0, 4, len(seq), -1, -1, 0) + name + bytes(packed) + b"\xff" * len(seq) + aux
        raw += struct.pack("<i", len(body)) + body

    def member(data: bytes) -> bytes:
        co = zlib.compressobj(6, zlib.DEFLATED, -15)
        comp = co.compress(data) + co.flush()
        bsize = 12 + 6 + len(comp) + 8
        return (b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff" + struct.pack("<H", 6) + b"BC" + struct.pack("<HH", 2, bsize - 1) + comp +
                struct.pack("<II", zlib.crc32(data) & 0xffffffff, len(data)))

    with open(path, "wb") as f:
        for o in range(0, len(raw), block):
            f.write(member(bytes(raw[o:o + block])))
        f.write(member(b""))        # the end-of-file marker block


# ---------------------------------------------------------------------------------------------
# Fast counter-based generator for the big move-table shapes (C3 / C5): synth_fast.c, built by
# __graft_entry__.build() as libpgm_synth.so. Every read is a function of (seed, global read index).
# ---------------------------------------------------------------------------------------------
_FAST_LIB = None


def _fast_lib():
    global _FAST_LIB
    if _FAST_LIB is None:
        import ctypes as C
        here = os.path.dirname(os.path.abspath(__file__))
        path = os.path.join(here, "libpgm_synth.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run __graft_entry__.build() first")
        _FAST_LIB = C.CDLL(path)
    return _FAST_LIB


def make_moves_fast(cfg: SynthConfig, read_lo: int = 0, n_reads: Optional[int] = None, alloc=None) -> ReadBatch:
    """Reads [read_lo, read_lo + n_reads) of the move-table workload `cfg` describes (cfg.n_reads = the whole
    job). `alloc(nbytes)` returns a uint8 numpy array to put the big arrays in (e.g. pinned memory)."""
    import ctypes as C
    assert cfg.mode in (capi.PGM_MODE_MOVES_CHAR, capi.PGM_MODE_MOVES_I8)
    lib = _fast_lib()
    n = cfg.n_reads - read_lo if n_reads is None else n_reads
    hi = read_lo + n
    alloc = alloc or (lambda nbytes: np.empty(nbytes, np.uint8))
    # per-read scalars: prefixes of fixed streams, so a read does not depend on how the job is cut
    ln = np.random.default_rng([cfg.seed, 1]).lognormal(np.log(cfg.median_samples), cfg.sigma_len, hi)[read_lo:]
    target = np.clip(ln, cfg.min_samples, cfg.max_samples)
    nb = np.maximum((target / (cfg.stride * cfg.mean_dwell_strides)).astype(np.int64), 12).astype(np.uint32)
    offset = np.random.default_rng([cfg.seed, 2]).integers(cfg.offset_lo, cfg.offset_hi + 1, hi)[read_lo:].astype(np.int32)
    offset[offset == 0] = 1
    tail = np.random.default_rng([cfg.seed, 3]).integers(0, cfg.stride, hi)[read_lo:].astype(np.int64)
    scale = cfg.range_ / cfg.digitisation
    q = 1.0 - 1.0 / cfg.mean_dwell_strides
    thr = np.array([min(2 ** 32 - 1, int(q ** j * 2 ** 32)) for j in range(1, cfg.max_dwell_strides)], np.uint32)
    level = np.rint(np.random.default_rng([cfg.seed, 4]).normal(cfg.level_mean, cfg.level_sd, 4096) / scale).astype(np.int32)

    class P(C.Structure):
        _fields_ = [("seed", C.c_uint64), ("stride", C.c_uint32), ("n_thr", C.c_uint32), ("dwell_thr", C.c_void_p),
                    ("level", C.c_void_p), ("filler", C.c_int32), ("lo_code", C.c_int32), ("hi_code", C.c_int32),
                    ("spike_thr", C.c_uint32), ("noise_num", C.c_int32), ("moves_char", C.c_uint32)]
    p = P(cfg.seed, cfg.stride, thr.size, thr.ctypes.data, level.ctypes.data, int(round(cfg.level_mean / scale)),
          int(round(15.0 / scale)), int(round(230.0 / scale)), int(cfg.spike_frac * 2 ** 32),
          int(round(cfg.noise_sd / scale / 37837.2 * 2 ** 20)), 1 if cfg.mode == capi.PGM_MODE_MOVES_CHAR else 0)
    moves = np.zeros(n, np.uint64)
    lib.pgm_synth_moves_len(C.byref(p), C.c_uint64(read_lo), C.c_uint32(n), C.c_void_p(nb.ctypes.data), C.c_void_p(moves.ctypes.data))
    moves = moves.astype(np.int64)
    raw_len = moves * cfg.stride + tail
    padded = (raw_len + 7) // 8 * 8
    raw_off = np.zeros(n + 1, np.int64); np.cumsum(padded, out=raw_off[1:])
    al_off = np.zeros(n + 1, np.int64); np.cumsum(moves, out=al_off[1:])
    sq_off = np.zeros(n + 1, np.int64); np.cumsum(nb.astype(np.int64), out=sq_off[1:])
    raw = alloc(int(raw_off[-1]) * 2).view(np.int16)
    align = alloc(int(al_off[-1]))
    bases = alloc(int(sq_off[-1]))
    u64 = lambda a: np.ascontiguousarray(a, dtype=np.uint64)
    ro, rc, ao, so = u64(raw_off[:-1]), u64(padded), u64(al_off[:-1]), u64(sq_off[:-1])
    lib.pgm_synth_moves_fill(C.byref(p), C.c_uint64(read_lo), C.c_uint32(n), C.c_void_p(nb.ctypes.data), C.c_void_p(offset.ctypes.data),
                             C.c_void_p(ro.ctypes.data), C.c_void_p(rc.ctypes.data), C.c_void_p(ao.ctypes.data), C.c_void_p(so.ctypes.data),
                             C.c_void_p(raw.ctypes.data), C.c_void_p(align.ctypes.data), C.c_void_p(bases.ctypes.data))
    reads = np.zeros(n, np.dtype(capi.READ_DTYPE))
    reads["raw_off"] = raw_off[:-1]; reads["raw_len"] = raw_len
    reads["align_off"] = al_off[:-1]; reads["align_len"] = moves
    reads["seq_off"] = sq_off[:-1]; reads["seq_len"] = nb; reads["fastq_len"] = nb
    reads["digitisation"] = cfg.digitisation; reads["offset"] = offset.astype(np.float64); reads["range"] = cfg.range_
    reads["stride"] = cfg.stride
    ids = [f"read_{cfg.seed & 0xffff:04x}_{i:07d}" for i in range(read_lo, hi)]
    return ReadBatch(cfg.mode, raw, align, bases, reads, ids)


def rebase(batch: ReadBatch, lo: int, hi: int) -> ReadBatch:
    """Records [lo, hi) as a self-contained batch: views of the big arrays from the first record's data on, offsets
    rebased (the records of a generated batch lie in order). What a host that streams a file in pieces hands over."""
    rd = batch.reads[lo:hi].copy()
    r0, a0, s0 = int(rd["raw_off"][0]), int(rd["align_off"][0]), int(rd["seq_off"][0])
    last = rd[-1]
    r1 = (int(last["raw_off"]) + int(last["raw_len"]) + 7) // 8 * 8
    a1, s1 = int(last["align_off"]) + int(last["align_len"]), int(last["seq_off"]) + int(last["seq_len"])
    rd["raw_off"] -= r0; rd["align_off"] -= a0; rd["seq_off"] -= s0
    return ReadBatch(batch.mode, batch.raw[r0:min(r1, batch.raw.size)], batch.align[a0:a1], batch.bases[s0:s1], rd, batch.read_ids[lo:hi])
