"""Stage times of one C3 block (3,125 reads, k = 9, empty totals) with the library's own event timers: the quick loop for kernel work.
usage: python profiles/micro/time_c3_block.py [reads] [reps]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from poregen_b200 import _capi as capi, synth, workloads
from poregen_b200.gmove import GmoveEngine
reads = int(sys.argv[1]) if len(sys.argv) > 1 else 3125
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
cfg, opts = workloads.c3(n_reads=100000)
b = synth.make_moves_fast(cfg, 0, reads)
with GmoveEngine(opts) as e:
    e.upload(b)
    for _ in range(3):
        e.reset(); e.count(); e.emit()
    e.sync()
    e.profile_enable(True); e.profile_read()
    for _ in range(reps):
        e.reset(); e.count(); e.emit()
    st = e.profile_read()
    e.profile_enable(False)
    s = e.summary()
tot = sum(v[0] for v in st.values()) / reps
print(os.environ.get("TAG", ""), "total %.3f ms |" % tot, " ".join("%s=%.0f" % (k, v[0] / reps * 1e3) for k, v in st.items() if v[0] > 0), "us | sel", s.n_sel_events)
