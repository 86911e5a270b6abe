import os, sys; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from poregen_b200 import synth, workloads
from poregen_b200.gmove import GmoveEngine
cfg,opts=workloads.c2(n_reads=10000)
b=synth.make_batch(cfg)
with GmoveEngine(opts) as e:
    for i in range(2):
        e.upload(b); e.count(); e.emit(); e.sync()
    if "svb" in sys.argv[1:]:
        ab = synth.align_raw(b, 8)
        blob, off = synth.svb_streams(ab)
        for i in range(2):
            e.upload_svb(ab, blob, off); e.sync()
