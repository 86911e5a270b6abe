"""Digest helpers shared by the golden tests (same digest as tests/golden/make_golden.py)."""
import hashlib
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def load_golden():
    return json.load(open(os.path.join(HERE, "golden", "reference_outputs.json")))


def digest(files: dict) -> dict:
    d = {"freq": files["freq.txt"].decode(), "dump": {}}
    for k in sorted(files):
        if k != "freq.txt" and files[k]:
            d["dump"][k] = hashlib.sha256(files[k]).hexdigest()
    d["n_files"] = len(files) - 1
    return d


def format_result_files(opts, res) -> dict:
    """dump/<kmer> and freq.txt from a result that carries doubles only: the host-side writer's
    job, gmove.cpp:674-678 ("%.8f," between samples, "%.8f;" after the last one of an event)."""
    names = opts.slot_kmers()
    out = {}
    ev_off = np.concatenate(([0], np.cumsum(res.slot_sel.astype(np.int64))))
    s_off = np.concatenate(([0], np.cumsum(res.ev_len.astype(np.int64))))
    for s, name in enumerate(names):
        parts = []
        for e in range(ev_off[s], ev_off[s + 1]):
            vals = res.samples[s_off[e]:s_off[e + 1]]
            parts.append(",".join("%.8f" % v for v in vals) + ";")
        out[name] = "".join(parts).encode()
    out["freq.txt"] = "".join(f"{n}\t{int(res.slot_total[s])}\n" for s, n in enumerate(names)).encode()
    return out


def assert_matches_golden(got: dict, want: dict, label: str):
    assert got["n_files"] == want["n_files"], f"{label}: number of k-mer files"
    assert got["freq"] == want["freq"], f"{label}: freq.txt differs from the reference"
    assert got["dump"].keys() == want["dump"].keys(), f"{label}: set of non-empty dump files differs"
    bad = [k for k in want["dump"] if got["dump"][k] != want["dump"][k]]
    assert not bad, f"{label}: dump/{bad[:5]} differ from the reference"
