import os, sys; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from poregen_b200 import synth, workloads
from poregen_b200.gmove import GmoveEngine
cfg, opts = workloads.c3(n_reads=int(sys.argv[1]) if len(sys.argv) > 1 else 6000)
b = synth.make_batch(cfg)
with GmoveEngine(opts) as e:
    for i in range(2):
        e.upload(b); e.count(); e.emit(); e.sync()
