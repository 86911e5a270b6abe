"""Times `poregen gmove` on a C2 prefix written to disk (BLOW5 + PAF + FASTQ), with the CLI's own phase split.
usage: python profiles/micro/cli_time.py [n_reads]"""
import os, subprocess, sys, tempfile, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from poregen_b200 import synth, workloads
import oracle_lib as ol
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
cfg, opts = workloads.c2(n_reads=n)
b = synth.make_batch(cfg)
d = tempfile.mkdtemp(prefix="pgm_clitime_")
synth.write_blow5(b, os.path.join(d, "reads.blow5"))
fl = synth.write_alignment(b, d)
cli = os.path.join(ROOT, "poregen_b200", "bin", "poregen")
for i in range(3):
    out = os.path.join(d, "o%d" % i)
    cmd = [cli, "gmove"] + ol.cli_args(opts) + ["--fastq", fl["fastq"], os.path.join(d, "reads.blow5"), fl["align"], out]
    t0 = time.perf_counter()
    p = subprocess.run(cmd, capture_output=True, env=dict(os.environ, PGM_CLI_STATS="1"))
    dt = time.perf_counter() - t0
    print("run %d: %.3f s rc=%d" % (i, dt, p.returncode), [l for l in p.stderr.decode().splitlines() if "[gmove]" in l or "Real time" in l])
