#!/usr/bin/env python
"""bench.py — gmove hot-path benchmark (BASELINE.json metric: signal samples/s and events/s).

A "step" is one full pass of the gmove record loop over the whole C2 workload (configs[1]:
10,000 synthetic RNA004 reads x ~10k samples, ss records, -k 5 --rna --scaling 1
--sample_limit 5000 --min_dur 20 --max_dur 40), starting from empty per-k-mer totals.

  value : samples/s, whole job, inputs resident in HBM, timed with CUDA events on the stream the
          kernels run on, max over ranks.
  e2e   : the same through the host-buffer C-ABI path (pinned host arrays in, host result out;
          H2D and D2H inside the timed region).
  N > 1 : weak scaling. Each rank owns a contiguous range of records (its own C2-sized shard);
          the only exchange is the all-gather of the per-k-mer count vector + exclusive scan over
          ranks that fixes every rank's share of sample_limit.

`--impl reference` times the UNMODIFIED reference (oracle/_ref/poregen_ref, single thread: it has
no other mode) on a bounded prefix of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "gmove signal samples/sec (events/sec alongside)"
UNIT = "samples/s"


def workload_desc(n_reads):
    return {"workload": "C2 synthetic RNA004 ss records: gmove -k 5 --rna --scaling 1 --sample_limit 5000 --min_dur 20 --max_dur 40 --file_limit 5000",
            "reads_per_gpu": n_reads, "median_samples_per_read": 10000,
            "l2": "inputs (200 MB int16 per step) exceed the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------------------------------------
def run_reference_arm(args):
    """Times oracle/_ref/poregen_ref on a bounded prefix of C2 (file form). Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from poregen_b200 import synth, workloads
    import oracle_lib as ol
    kind = "reference"
    binary = ol.REF_BIN
    if not os.path.exists(binary):
        kind = "port"
    n_reads = args.ref_reads
    cfg, opts = workloads.c2(n_reads=n_reads)
    batch = synth.make_batch(cfg)
    n_samples = int(batch.reads["raw_len"].astype("int64").sum())
    d = tempfile.mkdtemp(prefix="pgm_ref_")
    times = []
    examined = None
    try:
        if kind == "reference":
            synth.write_blow5(batch, os.path.join(d, "reads.blow5"))
            fl = synth.write_alignment(batch, d)
            for i in range(args.warmup + args.steps):
                out = os.path.join(d, f"out{i}")
                t0 = time.perf_counter()
                ol.run_reference(opts, os.path.join(d, "reads.blow5"), fl["align"], out, fastq=fl["fastq"])
                dt = time.perf_counter() - t0
                shutil.rmtree(out, ignore_errors=True)
                if i >= args.warmup:
                    times.append(dt)
        else:
            for i in range(args.warmup + args.steps):
                t0 = time.perf_counter()
                ol.oracle_run(opts, batch)
                dt = time.perf_counter() - t0
                if i >= args.warmup:
                    times.append(dt)
        examined = int(ol.oracle_run(opts, batch).n_events_examined)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    t = sum(times) / len(times)
    val = n_samples / t
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_desc(n_reads),
            "events_per_sec": examined / t,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": kind,
                             "sample": f"first {n_reads} reads of the C2 generator ({n_samples} samples) per step; "
                                       "reference binary built -O2, end to end incl. BLOW5 read and dump/ text output"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks line of /opt/skills/guides/B200_PROFILING.md, sampled during the timed region."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
class _DevArr:
    def __init__(self, ptr, n, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


def bind_near_gpu(local):
    """Pin this process to the cores of the NUMA node its GPU hangs off, BEFORE any pinned buffer is allocated
    (first touch then places the staging memory next to the GPU's PCIe root). With eight ranks streaming
    0.5 GB per step each through host memory, remote placement is what the end-to-end number would measure.
    Returns a description for the JSON line; silently does nothing where sysfs does not say."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if node < 0 or not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"gpu": bdf, "numa_node": node, "cpus": len(cpus)}
    except Exception:
        return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(stage):
    """dram bytes per launch of the stage's kernel from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(stage)
        except Exception:
            return None
    return None


def cpu_baseline_leg(n_reads):
    """Reference (or oracle port) on a bounded prefix of C2, ~10-30 s of one host core."""
    from poregen_b200 import synth, workloads
    import oracle_lib as ol
    cfg, opts = workloads.c2(n_reads=n_reads)
    batch = synth.make_batch(cfg)
    n_samples = int(batch.reads["raw_len"].astype("int64").sum())
    if os.path.exists(ol.REF_BIN):
        d = tempfile.mkdtemp(prefix="pgm_cpu_")
        try:
            synth.write_blow5(batch, os.path.join(d, "reads.blow5"))
            fl = synth.write_alignment(batch, d)
            t0 = time.perf_counter()
            ref_files = ol.run_reference(opts, os.path.join(d, "reads.blow5"), fl["align"], os.path.join(d, "out"), fastq=fl["fastq"])
            dt = time.perf_counter() - t0
            cli = _cli_leg(opts, d, fl, ref_files, n_samples)
        finally:
            shutil.rmtree(d, ignore_errors=True)
        kind = "reference"
        what = "unmodified reference binary (-O2), end to end incl. BLOW5 read and dump/ text output"
    else:
        t0 = time.perf_counter()
        ol.oracle_run(opts, batch)
        dt = time.perf_counter() - t0
        kind = "port"
        what = "CPU oracle port (oracle/gmove_oracle.c), in memory, incl. text formatting"
    out = {"value": n_samples / dt, "unit": UNIT, "cores": 1, "kind": kind,
           "sample": f"first {n_reads} reads of the C2 generator ({n_samples} samples, {dt:.1f} s): {what}"}
    if kind == "reference" and cli:
        out["same_files_through_cli"] = cli
    return out


def _cli_leg(opts, d, fl, ref_files, n_samples):
    """The C++ command line of this build on the SAME files the reference just processed (file to file:
    BLOW5 decode on host threads, device batches, dump/ written). Also checks the bytes."""
    import oracle_lib as ol
    cli = os.path.join(ROOT, "poregen_b200", "bin", "poregen")
    if not os.path.exists(cli):
        return None
    cmd = [cli, "gmove"] + ol.cli_args(opts) + ["--fastq", fl["fastq"], os.path.join(d, "reads.blow5"), fl["align"], os.path.join(d, "cli_out")]
    best = None
    for i in range(2):   # the first run pays CUDA context creation and the SLOW5 index
        shutil.rmtree(os.path.join(d, "cli_out"), ignore_errors=True)
        t0 = time.perf_counter()
        p = subprocess.run(cmd, capture_output=True)
        dt = time.perf_counter() - t0
        if p.returncode != 0:
            return {"error": p.stderr.decode()[-300:]}
        best = dt if best is None else min(best, dt)
    same = True
    for name, data in ref_files.items():
        path = os.path.join(d, "cli_out", name if name == "freq.txt" else os.path.join("dump", name))
        if not os.path.exists(path) or open(path, "rb").read() != data:
            same = False
            break
    return {"value": n_samples / best, "unit": UNIT, "seconds": best, "identical_to_reference": same,
            "note": "process start to exit, incl. CUDA context creation (~0.3-0.5 s)"}


def run_own_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    from poregen_b200 import _capi as capi, synth, workloads
    from poregen_b200.batch import ReadBatch
    from poregen_b200.gmove import GmoveEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the gmove hot path has no CPU fallback")
    torch.cuda.set_device(local)
    numa = bind_near_gpu(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        g.build()
    if world > 1:
        dist.barrier()
    capi.load()

    cfg, opts = workloads.c2(n_reads=args.reads, seed=0xC2 + rank)
    hb = synth.make_batch(cfg)
    # pinned host copies: the e2e path reads its inputs from pinned memory
    pin = lambda a: torch.from_numpy(np.array(a, copy=True).view(np.uint8).reshape(-1)).pin_memory()
    p_raw, p_align, p_bases, p_reads = pin(hb.raw), pin(hb.align), pin(hb.bases), pin(hb.reads)
    batch = ReadBatch(hb.mode, p_raw.numpy().view(np.int16), p_align.numpy(), p_bases.numpy(),
                      p_reads.numpy().view(np.dtype(capi.READ_DTYPE)), hb.read_ids)
    n_samples = int(batch.reads["raw_len"].astype(np.int64).sum())
    # a real (non-default) stream: handle 0 would make the library create a private stream that
    # torch's events could not bracket
    stream = torch.cuda.Stream(device=local)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng = GmoveEngine(opts, device=local, stream=stream.cuda_stream)
    NS = eng.n_slots
    limit = opts.sample_limit

    counts_t = None
    if world > 1:
        counts_t = torch.as_tensor(_DevArr(eng.counts_ptr(), NS), device=f"cuda:{local}")
        all_counts = torch.empty((world, NS), dtype=torch.int32, device=f"cuda:{local}")

    def step_resident():
        eng.reset()
        eng.count()
        if world > 1:
            dist.all_gather_into_tensor(all_counts.view(-1), counts_t)
            eng.set_gathered(all_counts.data_ptr(), world, rank)   # lower-rank / all-rank sums inside k_quota
        eng.emit()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident (kernel-only) timing
    eng.upload(batch)
    for _ in range(args.warmup):
        step_resident()
    barrier()
    res0 = eng.download()               # result of the last warm-up step: counts for the byte model
    eng.upload(batch)
    barrier()
    # Stage table: an UNTIMED pass of the same K steps with every stage bracketed by an event pair (the pairs cost
    # ~40 us of stream time per step, so they stay out of the timed region).
    eng.profile_enable(True)
    eng.profile_read()
    for _ in range(args.steps):
        step_resident()
    stages = eng.profile_read()
    dom = max(stages, key=lambda k: stages[k][0])
    # Timed region: only the dominant kernel keeps its event pair (roofline.achieved comes from these K launches).
    eng.profile_enable(True, only=dom)
    eng.profile_read()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_resident()
    e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = eng.launches - l0
    dom_timed = eng.profile_read()[dom]
    eng.profile_enable(False)
    clocks = sampler.stop() if rank == 0 else None
    t_res = torch.tensor([ms_total], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t_res, op=dist.ReduceOp.MAX)
    ms_step = float(t_res.item()) / args.steps

    # ---- K7 (model rows of the whole batch; scripts/poregen.sh:54-85), timed once around pgm_model_finalize
    eng.model_begin()
    eng.reset()
    eng.upload(batch); eng.count(); eng.emit()
    eng.download()
    torch.cuda.synchronize()
    tm0 = time.perf_counter()
    rows = eng.model_finalize(1e300)
    model_ms = (time.perf_counter() - tm0) * 1e3
    model_info = {"ms": model_ms, "values": int(rows["n"].sum()), "rows": int(rows.size),
                  "note": "pgm_model_finalize on this rank's C2 batch: median by int64 radix selection + exact 128-bit variance per k-mer, incl. the rows' download"}

    # ---- end to end: pinned host buffers in, the dump TEXT (what the reference fprintf's, formatted on the
    # device) out to host memory, every step. A second context with PGM_OUT_TEXT does this leg.
    import copy
    import ctypes as C
    topts = copy.copy(opts)
    topts.output = capi.PGM_OUT_TEXT_ONLY   # what the command line uses: the text, per-slot offsets and counts
    eng.close()

    # Batches in flight: while one batch's text travels device -> host, the next one's inputs travel host -> device
    # and its kernels run (PCIe is full duplex). Each slot is a context with its own stream; every step still does
    # its own upload, its own count/[all-gather]/emit and its own download into host memory inside the timed region.
    class Slot:
        svb = None      # (aligned batch, stream bytes, stream offsets): upload the compressed signal instead of raw[]

        def __init__(self):
            self.stream = torch.cuda.Stream(device=local)
            self.eng = GmoveEngine(topts, device=local, stream=self.stream.cuda_stream)
            if world > 1:
                self.counts_t = torch.as_tensor(_DevArr(self.eng.counts_ptr(), NS), device=f"cuda:{local}")
                self.all_counts = torch.empty((world, NS), dtype=torch.int32, device=f"cuda:{local}")

        def submit(self):
            e = self.eng
            with torch.cuda.stream(self.stream):
                e.reset()
                if Slot.svb is not None:
                    e.upload_svb(*Slot.svb)
                else:
                    e.upload(batch)
                e.count()
                if world > 1:
                    dist.all_gather_into_tensor(self.all_counts.view(-1), self.counts_t)
                    e.set_gathered(self.all_counts.data_ptr(), world, rank)
                e.emit()

        def collect(self):
            r = capi.PgmResult()
            self.eng._chk(self.eng.lib.pgm_download(self.eng._ctx, C.byref(r)))
            return r

    depth = max(1, args.e2e_depth)
    slots = [Slot() for _ in range(depth)]
    eng = slots[0].eng

    def run_e2e(k):
        r = None
        for i in range(min(depth, k)):
            slots[i % depth].submit()
        for i in range(k):
            sl = slots[i % depth]
            r = sl.collect()
            if i + depth < k:
                sl.submit()
        return r

    r = run_e2e(max(depth, args.warmup))
    barrier()
    t0 = time.perf_counter()
    r = run_e2e(args.steps)
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    t_e2e = torch.tensor([e2e_s], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_s = float(t_e2e.item())
    h2d = int(batch.raw.nbytes + batch.align.nbytes + batch.bases.nbytes + batch.reads.nbytes)
    text_bytes = int(C.cast(r.slot_text_off, C.POINTER(C.c_uint64))[NS])
    d2h = int(text_bytes + NS * 16 + batch.n_reads + 48)
    examined = int(r.n_events_examined)

    # ---- the same step fed with the records' svb-zd streams (what a BLOW5 file stores) instead of int16 samples:
    # k_svb_decode produces raw[] in HBM. Single-GPU runs only (the encoder here is numpy).
    svb_info = None
    if world == 1 and not args.no_svb:
        ab = synth.align_raw(hb, 8)
        sb, so = synth.svb_streams(ab)
        p_sb, p_so = pin(sb), pin(so)
        ab = ReadBatch(ab.mode, ab.raw, batch.align, batch.bases, pin(ab.reads).numpy().view(np.dtype(capi.READ_DTYPE)), ab.read_ids)
        svb_args = (ab, p_sb.numpy(), p_so.numpy().view(np.uint64))
        e = slots[0].eng
        with torch.cuda.stream(slots[0].stream):
            for _ in range(3):
                e.upload_svb(*svb_args)
            e.sync()
            e.profile_enable(True); e.profile_read()
            for _ in range(args.steps):
                e.upload_svb(*svb_args)
            st_svb = e.profile_read()["svb"]
            e.profile_enable(False)
        live = np.zeros(ab.raw.size + 1, np.int32)          # the padding between records is not written by the decoder
        np.add.at(live, ab.reads["raw_off"].astype(np.int64), 1)
        np.add.at(live, ab.reads["raw_off"].astype(np.int64) + ab.reads["raw_len"].astype(np.int64), -1)
        live = np.cumsum(live[:-1]) > 0
        assert np.array_equal(e.device_raw(ab.raw.size)[live], ab.raw[live]), "k_svb_decode: samples differ from the encoder's input"
        Slot.svb = svb_args
        rs = run_e2e(max(depth, args.warmup))
        barrier()
        t0 = time.perf_counter()
        rs = run_e2e(args.steps)
        barrier()
        svb_s = (time.perf_counter() - t0) / args.steps
        Slot.svb = None
        assert int(rs.n_sel_samples) == int(r.n_sel_samples) and int(C.cast(rs.slot_text_off, C.POINTER(C.c_uint64))[NS]) == text_bytes
        k_ms = st_svb[0] / max(1, st_svb[1])
        alg_svb = int(sb.nbytes + 2 * n_samples)
        pk, _ = measured_peak()
        svb_info = {"kernel": "k_svb_decode", "ms_per_launch": k_ms, "stream_bytes": int(sb.nbytes), "bytes_per_sample": sb.nbytes / n_samples,
                    "algorithmic_bytes": alg_svb, "gbs": alg_svb / (k_ms * 1e-3) / 1e9, "frac": alg_svb / (k_ms * 1e-3) / 1e9 / pk,
                    "e2e": {"value": n_samples / svb_s, "unit": UNIT, "ms_per_step": svb_s * 1e3,
                            "h2d_bytes_per_step": int(sb.nbytes + so.nbytes + batch.align.nbytes + batch.bases.nbytes + batch.reads.nbytes),
                            "d2h_bytes_per_step": d2h},
                    "note": "same C2 step with the records' slow5 svb-zd streams uploaded instead of int16 samples; decoded in HBM"}

    tot = torch.tensor([n_samples, examined, int(r.n_sel_events), int(r.n_sel_samples)], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(tot)
    tot_samples, tot_events, tot_sel_ev, tot_sel_samp = [float(x) for x in tot.tolist()]

    if rank == 0:
        # ---- byte model per stage (this rank), DESIGN.md "Kernels and their algorithmic bytes"
        rd = batch.reads
        T = int(rd["align_len"].astype(np.int64).sum()); NB = int(rd["seq_len"].astype(np.int64).sum()); R = batch.n_reads
        E_pass, E_sel, S_sel = int(res0.n_events_pass), int(res0.ev_len.size), int(res0.samples.size)
        act = np.unique(res0.ev_read)
        S_act = int((rd["raw_len"][act].astype(np.int64) - rd["trim"][act].astype(np.int64)).sum())
        alg = {"prep": 120 * R, "decode": T + NB + 88 * R + 16 * E_pass, "place": 32 * E_pass + 8 * R,
               "sort": 16 * E_pass + 12 * E_sel, "select": 16 * NS, "stats": (2 * S_act + 48 * R) if opts.scaling else 0,
               "gather": 2 * S_sel + 8 * S_sel + 20 * E_sel,
               "scan": 12 * E_sel, "misc": 8 * NS, "text": 0}
        alg_step = 2 * n_samples + T + NB + 24 * R + 8 * S_sel + 12 * E_sel + 4 * NS
        peak, peak_src = measured_peak()
        stage_rows = {}
        for name, (ms, ln) in stages.items():
            per = ms / args.steps
            stage_rows[name] = {"ms_per_step": per, "launches_per_step": ln / args.steps,
                                "algorithmic_bytes": alg.get(name, 0),
                                "gbs": (alg.get(name, 0) / (per * 1e-3) / 1e9) if per > 0 else None}
        # the dominant kernel's duration as measured INSIDE the timed region (its own event pair only)
        dom_per_launch_ms = dom_timed[0] / max(1, dom_timed[1])
        achieved = stage_rows[dom]["algorithmic_bytes"] / max(1.0, stage_rows[dom]["launches_per_step"]) / (dom_per_launch_ms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": tot_samples / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": dict(workload_desc(args.reads), parallelism=f"read-range x{world}"),
                "events_per_sec": tot_events / (ms_step * 1e-3),
                "events_selected_per_step": tot_sel_ev, "samples_selected_per_step": tot_sel_samp,
                "e2e": {"value": tot_samples / e2e_s, "unit": UNIT, "ms_per_step": e2e_s * 1e3, "events_per_sec": tot_events / e2e_s,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "output": "dump text (%.8f, formatted on the device), per-k-mer text offsets and counts, per-record flags",
                        "batches_in_flight": depth, "host_binding": numa},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": ncu_traffic(dom), "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": stage_rows[dom]["algorithmic_bytes"] / max(1.0, stage_rows[dom]["launches_per_step"]),
                             "ms_per_launch": dom_per_launch_ms},
                "roofline_step": {"algorithmic_bytes": alg_step, "gbs": alg_step / (ms_step * 1e-3) / 1e9,
                                  "frac": alg_step / (ms_step * 1e-3) / 1e9 / peak},
                "stages": stage_rows, "stages_note": "per-stage times from an untimed pass of the same steps with all stage timers on; the timed region brackets only the roofline kernel",
                "model_aggregation": model_info, "clocks": clocks}
        if svb_info:
            line["svb_signal"] = svb_info
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_leg(args.cpu_reads)
        emit_line(line)
    for sl in slots:
        sl.eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_FD = None


def protect_stdout():
    """The contract is ONE JSON line on stdout. Libraries (NCCL's version banner, build scripts) write to
    fd 1 behind Python's back, so fd 1 is pointed at stderr for the whole run and the line goes to a
    duplicate of the original stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--reads", type=int, default=10000, help="records per GPU (C2 = 10000)")
    ap.add_argument("--ref-reads", type=int, default=1500, help="--impl reference: records per step")
    ap.add_argument("--cpu-reads", type=int, default=3000, help="cpu_baseline leg: records in the sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-svb", action="store_true", help="skip the compressed-signal (svb-zd) leg")
    ap.add_argument("--e2e-depth", type=int, default=2, help="e2e leg: batches in flight (1 = strictly serial steps)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "own":
        args.warmup = 3
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_own_arm(args)


if __name__ == "__main__":
    main()
