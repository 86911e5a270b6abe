"""The reference's single gmove fixture (test/data/raw/gmove/single_read) as batches.

`load_fixture_from_reference` parses the reference's files (only in the build container, used by
tests/golden/make_golden.py); the tests themselves read the committed tests/golden/single_read.npz.
"""
import os

import numpy as np

from poregen_b200 import _capi as capi
from poregen_b200.batch import GmoveOptions, ReadBatch

HERE = os.path.dirname(os.path.abspath(__file__))


def load_fixture_from_reference(d: str) -> dict:
    rec = [l for l in open(os.path.join(d, "reads.slow5")) if not l.startswith(("#", "@"))][0].rstrip("\n").split("\t")
    raw = np.array(rec[7].split(","), dtype=np.int16)
    tsv = open(os.path.join(d, "guppy_move")).readline().rstrip("\n").split("\t")
    paf = open(os.path.join(d, "guppy_move.paf")).readline().rstrip("\n").split("\t")
    ss = [f for f in paf[12:] if f.startswith("ss:Z:")][0][5:]
    sam = [l for l in open(os.path.join(d, "guppy_move.sam")) if not l.startswith("@")][0].rstrip("\n").split("\t")
    mv = [f for f in sam[11:] if f.startswith("mv:B:c,")][0][7:].split(",")
    fq = open(os.path.join(d, "read_0.fastq")).read().split("\n")
    return dict(read_id=np.array(rec[0]), digitisation=float(rec[2]), offset=float(rec[3]), range=float(rec[4]), raw=raw,
                tsv_fastq_len=int(tsv[1]), tsv_seq=np.array(tsv[2]), tsv_stride=int(tsv[3]), tsv_moves=np.array(tsv[4]),
                tsv_signal_len=int(tsv[5]), tsv_trim=int(tsv[6]),
                paf_qstart=int(paf[2]), paf_tstart=int(paf[7]), paf_tend=int(paf[8]), paf_ss=np.array(ss), fastq_seq=np.array(fq[1]),
                sam_seq=np.array(sam[9]), sam_stride=int(mv[0]), sam_moves=np.array(mv[1:], dtype=np.int8),
                sam_ns=int([f for f in sam[11:] if f.startswith("ns:i:")][0][5:]),
                sam_ts=int([f for f in sam[11:] if f.startswith("ts:i:")][0][5:]))


def fixture_batch(form: str, k: int) -> ReadBatch:
    fx = np.load(os.path.join(HERE, "golden", "single_read.npz"))
    reads = np.zeros(1, np.dtype(capi.READ_DTYPE))
    raw = fx["raw"]
    reads["raw_len"] = raw.size
    reads["digitisation"] = float(fx["digitisation"]); reads["offset"] = float(fx["offset"]); reads["range"] = float(fx["range"])
    if form == "tsv":
        mode, align, seq = capi.PGM_MODE_MOVES_CHAR, np.frombuffer(str(fx["tsv_moves"]).encode(), np.uint8), str(fx["tsv_seq"])
        reads["stride"] = int(fx["tsv_stride"]); reads["trim"] = int(fx["tsv_trim"]); reads["fastq_len"] = int(fx["tsv_fastq_len"])
    elif form == "sam":
        mode, align, seq = capi.PGM_MODE_MOVES_I8, fx["sam_moves"].view(np.uint8), str(fx["sam_seq"])
        reads["stride"] = int(fx["sam_stride"]); reads["trim"] = int(fx["sam_ts"]); reads["fastq_len"] = len(seq)
    else:
        mode, align = capi.PGM_MODE_SS, np.frombuffer(str(fx["paf_ss"]).encode(), np.uint8)
        ts, te = int(fx["paf_tstart"]), int(fx["paf_tend"])
        st, en = min(ts, te), max(ts, te)
        seq = str(fx["fastq_seq"])[st:en]          # faidx_fetch_seq(tid, st_k, end_k-1)
        reads["query_start"] = int(fx["paf_qstart"]); reads["st_k"] = st; reads["end_k"] = en
        reads["flags"] = (capi.PGM_READ_RNA if ts > te else 0) | (capi.PGM_READ_SKIP if len(seq) < k else 0)
    bases = np.frombuffer(seq.encode(), np.uint8)
    reads["align_len"] = align.size
    reads["seq_len"] = bases.size
    return ReadBatch(mode, raw, align, bases, reads, [str(fx["read_id"])])




def _kmer_lists():
    p = os.path.join(HERE, "golden", "kmer_lists.json")
    import json
    return json.load(open(p))


def FIXTURE_CASES():
    """(name, form, options, kmer file name in the reference fixture dir or None) — the runnable
    cases of the reference's test/test_gmove.sh (0.2, 0.4/2.1, 1.1, 1.2/2.2 at margin 0 and at the
    default margin 2, 1.3/2.3 via text SAM) plus Med-MAD variants."""
    lists = _kmer_lists()
    out = []
    for sc in (0, 1):
        out.append((f"t0.2_tsv_k9_sc{sc}", "tsv", GmoveOptions(kmer_size=9, file_limit=50, index_end=50, scaling=sc).resolve_range(), None))
        for form in ("tsv", "sam", "paf"):
            for lname in ("single_kmer_file.txt", "kmer_file.txt"):
                for margin in ((0, 2) if form == "paf" else (2,)):
                    kl = lists[lname]
                    o = GmoveOptions(kmer_size=6, kmer_list=kl, kmer_pick_margin=margin, scaling=sc)
                    o.resolve_range()
                    out.append((f"{form}_{lname.split('.')[0]}_pm{margin}_sc{sc}", form, o, lname))
    return out


def write_fixture_files(d: str) -> dict:
    """The reference fixture test/data/raw/gmove/single_read re-created from the committed arrays (same content for
    gmove, not the same bytes): reads.slow5, guppy_move (TSV), guppy_move.paf, guppy_move.sam, guppy_move.bam,
    read_0.fastq(+.fai), kmer_file.txt, single_kmer_file.txt. Returns {name: path}."""
    from poregen_b200 import synth
    os.makedirs(d, exist_ok=True)
    fx = np.load(os.path.join(HERE, "golden", "single_read.npz"))
    rid = str(fx["read_id"])
    p = {n: os.path.join(d, n) for n in ("reads.slow5", "guppy_move", "guppy_move.paf", "guppy_move.sam", "guppy_move.bam",
                                          "read_0.fastq", "kmer_file.txt", "single_kmer_file.txt")}
    tsv = fixture_batch("tsv", 6)
    synth.write_slow5(tsv, p["reads.slow5"])
    with open(p["guppy_move"], "w") as f:
        f.write(f"{rid}\t{int(fx['tsv_fastq_len'])}\t{str(fx['tsv_seq'])}\t{int(fx['tsv_stride'])}\t{str(fx['tsv_moves'])}\t"
                f"{int(fx['tsv_signal_len'])}\t{int(fx['tsv_trim'])}\n")
    seq = str(fx["fastq_seq"])
    n = int(fx["raw"].size)
    with open(p["guppy_move.paf"], "w") as f:
        ts, te = int(fx["paf_tstart"]), int(fx["paf_tend"])
        f.write(f"{rid}\t{n}\t{int(fx['paf_qstart'])}\t{n}\t+\t{rid}\t{len(seq)}\t{ts}\t{te}\t{abs(te - ts)}\t{abs(te - ts)}\t0\t"
                f"sc:f:1.000000\tsh:f:0.000000\tss:Z:{str(fx['paf_ss'])}\n")
    hdr = f"@{rid}\n"
    with open(p["read_0.fastq"], "w") as f:
        f.write(hdr + seq + "\n+\n" + "I" * len(seq) + "\n")
    with open(p["read_0.fastq"] + ".fai", "w") as f:
        f.write(f"{rid}\t{len(seq)}\t{len(hdr)}\t{len(seq)}\t{len(seq) + 1}\t{len(hdr) + len(seq) + 3}\n")
    sam = fixture_batch("sam", 6)
    fl = synth.write_alignment(sam, os.path.join(d, "_sam"))
    os.replace(fl["align"], p["guppy_move.sam"])
    synth.write_bam(sam, p["guppy_move.bam"])
    for name, lst in _kmer_lists().items():
        with open(os.path.join(d, name), "w") as f:
            f.write("".join(k + "\n" for k in lst))
    return p
