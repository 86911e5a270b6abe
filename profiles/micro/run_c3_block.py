"""One C3 block (3,125 reads of the bench's generator, k = 9, empty totals: every passing event is selected) through count + emit,
a few times: the command ncu profiles (profiles/ncu_r02_*.md). usage: python profiles/micro/run_c3_block.py [reads] [reps] [text]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import __graft_entry__ as g
g.build()
from poregen_b200 import _capi as capi, synth, workloads
from poregen_b200.gmove import GmoveEngine
reads = int(sys.argv[1]) if len(sys.argv) > 1 else 3125
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg, opts = workloads.c3(n_reads=100000)
if len(sys.argv) > 3 and sys.argv[3] == "text":
    opts.output = capi.PGM_OUT_TEXT_ONLY
b = synth.make_moves_fast(cfg, 0, reads)
with GmoveEngine(opts) as e:
    e.upload(b)
    for _ in range(reps):
        e.reset(); e.count(); e.emit()
    s = e.summary()
    print("examined", s.n_events_examined, "pass", s.n_events_pass, "selected", s.n_sel_events, "samples", s.n_sel_samples)
