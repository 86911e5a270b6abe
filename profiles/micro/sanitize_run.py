"""Small batches through every kernel family of libpgm_gmove.so, for compute-sanitizer:
    compute-sanitizer --tool memcheck  python profiles/micro/sanitize_run.py > profiles/sanitizer_r02_memcheck.txt
    compute-sanitizer --tool racecheck python profiles/micro/sanitize_run.py > profiles/sanitizer_r02_racecheck.txt
One tool per gpurun call. Sizes are tiny on purpose (the tools run kernels 10-100x slower); no oracle here: the
parity tests check values, this run checks addresses, barriers and shared-memory hazards.
Covers: move table '0'/'1' and int8 (k=5: one-digit sort; k=9: MSD scatter + rank), ss:Z DNA / RNA with indels,
Med-MAD scaling, text output, svb-zd upload, the two-batch submit/wait pipeline, K7 model rows + dwell, and the
allocation-free formatter.
"""
import ctypes as C
import dataclasses
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from poregen_b200 import _capi as capi, synth  # noqa: E402
from poregen_b200.batch import GmoveOptions  # noqa: E402
from poregen_b200.gmove import GmoveEngine  # noqa: E402

N_READS = int(os.environ.get("SAN_READS", "12"))
SAMPLES = int(os.environ.get("SAN_SAMPLES", "3000"))


def opts_k(k, **kw):
    n = 4 ** k
    base = dict(kmer_size=k, sample_limit=20, file_limit=n, index_end=n, scaling=0)
    base.update(kw)
    return GmoveOptions(**base).resolve_range()


def batch_of(mode, rna=False, indel=0.0, seed=1):
    cfg = synth.SynthConfig(n_reads=N_READS, median_samples=SAMPLES, mode=mode, rna_orientation=rna, p_del=indel, p_ins=indel, seed=seed)
    return synth.make_batch(cfg)


def main():
    done = []
    for name, mode, rna, indel, o in [
        ("moves_char k5 scaling", capi.PGM_MODE_MOVES_CHAR, False, 0.0, opts_k(5, scaling=1)),
        ("moves_i8 k5", capi.PGM_MODE_MOVES_I8, False, 0.0, opts_k(5)),
        ("moves_char k9 (two-digit sort)", capi.PGM_MODE_MOVES_CHAR, False, 0.0, opts_k(9, sample_limit=3)),
        ("ss dna k6 indels", capi.PGM_MODE_SS, False, 0.05, opts_k(6, signal_print_margin=2)),
        ("ss rna k5 indels scaling", capi.PGM_MODE_SS, True, 0.05, opts_k(5, rna=True, scaling=1)),
        ("moves_char k5 text", capi.PGM_MODE_MOVES_CHAR, False, 0.0, opts_k(5, output=capi.PGM_OUT_TEXT)),
        ("ss rna k5 text -d", capi.PGM_MODE_SS, True, 0.02, opts_k(5, rna=True, output=capi.PGM_OUT_TEXT, delimit_files=True)),
    ]:
        b = batch_of(mode, rna, indel)
        with GmoveEngine(o, device=0) as eng:
            r1 = eng.process(b)
            r2 = eng.process(batch_of(mode, rna, indel, seed=2))          # totals carried: the quota path with priors
        done.append((name, int(r1.ev_len.size if r1.ev_len is not None else 0), int(r2.slot_sel.sum())))

    # svb-zd upload + decode on the device
    o = opts_k(5, scaling=1)
    ab = synth.align_raw(batch_of(capi.PGM_MODE_MOVES_CHAR), 8)
    blob, off = synth.svb_streams(ab)
    with GmoveEngine(o, device=0) as eng:
        eng.upload_svb(ab, blob, off)
        eng.count(); eng.emit()
        r = eng.download()
    done.append(("svb-zd upload", int(blob.size), int(r.slot_sel.sum())))

    # submit / wait: two batches in flight on sibling streams
    o = opts_k(5, output=capi.PGM_OUT_TEXT)
    with GmoveEngine(o, device=0) as eng:
        ids = [eng.submit(batch_of(capi.PGM_MODE_MOVES_CHAR, seed=s)) for s in (3, 4)]
        rs = [eng.wait(i) for i in ids]
        ids = [eng.submit(batch_of(capi.PGM_MODE_MOVES_CHAR, seed=5))]
        rs += [eng.wait(i) for i in ids]
    done.append(("submit/wait x3", len(rs), int(rs[-1].slot_sel.sum())))

    # K7: model rows and dwell medians over two batches
    o = opts_k(5, sample_limit=40)
    with GmoveEngine(o, device=0) as eng:
        eng.model_begin()
        for s in (6, 7):
            eng.process(batch_of(capi.PGM_MODE_MOVES_CHAR, seed=s))
        rows = eng.model_finalize(3.1)
        dw = eng.model_dwell()
    done.append(("model rows + dwell", int(rows.size), int((dw >= 0).sum())))

    # allocation-free formatter, two calls on the shared static scratch
    import torch
    lib = capi.load()
    for n in (20000, 700):
        x = torch.from_numpy(np.random.default_rng(n).normal(90, 30, n)).cuda()
        text = torch.zeros(n * 24, dtype=torch.uint8, device="cuda")
        ln = torch.zeros(1, dtype=torch.int64, device="cuda")
        rc = lib.pgm_format_noalloc(0, torch.cuda.current_stream().cuda_stream, x.data_ptr(), n, text.data_ptr(), text.numel(), ln.data_ptr())
        assert rc == 0
        torch.cuda.synchronize()
        done.append((f"format_noalloc n={n}", n, int(ln.item())))
    for d in done:
        print("ran:", *d)
    print("sanitize_run: all cases ran")


if __name__ == "__main__":
    main()
