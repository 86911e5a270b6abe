"""Resident step time of the other BASELINE configurations on one GPU (C3: one GPU's eighth of the 100k reads; C4 DNA), next to C2
which bench.py reports. usage: python profiles/micro/time_configs.py > gpurun_out/configs.json"""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from poregen_b200 import synth, workloads
from poregen_b200.gmove import GmoveEngine

out = {}
for name, (cfg, opts) in (("c3_share_12500_reads", workloads.c3(n_reads=12500)), ("c4_dna", workloads.c4())):
    b = synth.make_batch(cfg)
    n = int(b.reads["raw_len"].astype("int64").sum())
    with GmoveEngine(opts) as e:
        e.upload(b)
        for _ in range(3):
            e.reset(); e.count(); e.emit()
        e.sync()
        e.profile_enable(True); e.profile_read()
        for _ in range(5):
            e.reset(); e.count(); e.emit()
        st = e.profile_read()
        e.profile_enable(False)
        e.sync()
        t0 = time.perf_counter()
        for _ in range(5):
            e.reset(); e.count(); e.emit()
        e.sync()
        ms = (time.perf_counter() - t0) / 5 * 1e3
        r = e.download()
    out[name] = {"samples": n, "ms_per_step": ms, "samples_per_s": n / (ms * 1e-3), "events_selected": int(r.ev_len.size),
                 "samples_selected": int(r.samples.size), "algorithmic_bytes": b.algorithmic_bytes(e.n_slots if hasattr(e, "n_slots") else 0, int(r.ev_len.size), int(r.samples.size)),
                 "stages_ms": {k: v[0] / 5 for k, v in st.items() if v[0] > 0}}
    out[name]["gbs"] = out[name]["algorithmic_bytes"] / (ms * 1e-3) / 1e9
print(json.dumps(out, indent=1))
