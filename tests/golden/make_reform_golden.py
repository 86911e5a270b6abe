"""Regenerates tests/golden/reform_outputs.json: SHA-256 of what the UNMODIFIED reference `reform`
(oracle/_ref/reform_ref, built by oracle/build_ref.sh from /root/reference/src/reform.cpp) prints for the
synthetic SAM files of tests/reform_cases.py. Run in the build container (needs oracle/_ref)."""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from poregen_b200 import synth  # noqa: E402
from reform_cases import CASES  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "reform_ref")


def main():
    out = {}
    for label, cfg, args in CASES:
        with tempfile.TemporaryDirectory() as d:
            fl = synth.write_alignment(synth.make_batch(cfg), d)
            p = subprocess.run([REF, "reform"] + args + [fl["align"]], capture_output=True, check=True)
            out[label] = {"sha256": hashlib.sha256(p.stdout).hexdigest(), "bytes": len(p.stdout)}
    json.dump(out, open(os.path.join(HERE, "reform_outputs.json"), "w"), indent=1, sort_keys=True)
    print("wrote", len(out), "cases")


if __name__ == "__main__":
    main()
