"""Generates the committed golden vectors by RUNNING THE UNMODIFIED REFERENCE (oracle/_ref/poregen_ref,
built by oracle/build_ref.sh from /root/reference) in the build container. Not run on the GPU box.

  python tests/golden/make_golden.py

Outputs (committed):
  tests/golden/single_read.npz   the reference's only gmove fixture (test/data/raw/gmove/single_read)
                                 parsed into arrays: raw signal, calibration, move string, sequence,
                                 ss string, PAF/SAM scalars. Data, not code.
  tests/golden/reference_outputs.json
                                 per case: freq.txt verbatim + sha256 of every non-empty dump/<kmer>
                                 as written by the reference, for (a) the fixture in its TSV / PAF /
                                 SAM forms and (b) every seeded synthetic case of tests/cases.py.
"""
import hashlib
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as ol  # noqa: E402
from cases import parity_cases  # noqa: E402
from fixture import FIXTURE_CASES, load_fixture_from_reference  # noqa: E402
from poregen_b200 import synth  # noqa: E402

REF_FIX = "/root/reference/test/data/raw/gmove/single_read"


def digest(files: dict) -> dict:
    d = {"freq": files["freq.txt"].decode(), "dump": {}}
    for k in sorted(files):
        if k != "freq.txt" and files[k]:
            d["dump"][k] = hashlib.sha256(files[k]).hexdigest()
    d["n_files"] = len(files) - 1
    return d


def main():
    out = {"fixture": {}, "synthetic": {}}
    fx = load_fixture_from_reference(REF_FIX)
    np.savez_compressed(os.path.join(HERE, "single_read.npz"), **fx)
    for name, form, opts, kfile in FIXTURE_CASES():
        d = tempfile.mkdtemp()
        try:
            align = {"tsv": "guppy_move", "paf": "guppy_move.paf", "sam": "guppy_move.sam"}[form]
            files = ol.run_reference(opts, os.path.join(REF_FIX, "reads.slow5"), os.path.join(REF_FIX, align), os.path.join(d, "o"),
                                     fastq=os.path.join(REF_FIX, "read_0.fastq") if form == "paf" else None,
                                     kmer_file=os.path.join(REF_FIX, kfile) if kfile else None)
            out["fixture"][name] = digest(files)
        finally:
            shutil.rmtree(d, ignore_errors=True)
        print("fixture", name, len(out["fixture"][name]["dump"]), "non-empty files")
    for label, cfg, opts in parity_cases():
        batch = synth.make_batch(cfg)
        d = tempfile.mkdtemp()
        try:
            synth.write_blow5(batch, os.path.join(d, "reads.blow5"))
            fl = synth.write_alignment(batch, d)
            kf = None
            if opts.kmer_list is not None:
                kf = os.path.join(d, "kmers.txt")
                open(kf, "w").write("".join(k + "\n" for k in opts.kmer_list))
            files = ol.run_reference(opts, os.path.join(d, "reads.blow5"), fl["align"], os.path.join(d, "o"), fastq=fl["fastq"], kmer_file=kf)
            out["synthetic"][label] = digest(files)
        finally:
            shutil.rmtree(d, ignore_errors=True)
        print("synthetic", label, len(out["synthetic"][label]["dump"]), "non-empty files")
    with open(os.path.join(HERE, "reference_outputs.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
