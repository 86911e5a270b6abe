#!/usr/bin/env python
"""bench.py — gmove hot-path benchmark (BASELINE.json metric: signal samples/s and events/s).

Workload: C3 = BASELINE.json configs[2], the configuration the metric is quoted on and the largest that
fits one GPU: 100,000 synthetic DNA R10.4.1 move-table reads at 5 kHz (~52,000 samples each, 5.2e9 samples,
10.4 GB of int16), `gmove -k 9 --sample_limit 1000 --file_limit 262144` (262,144 k-mer files, no scaling).

A "step" is ONE gmove job over all 100,000 reads, from empty per-k-mer totals: the alignment records are streamed
in blocks of --block-reads consecutive records (what a host that reads a file hands over; 12,500 resident, 3,125 end to end), block after block,
with the per-k-mer totals (kmer_frequency_map) carried, so the result is the reference's single loop.

  value : samples/s, whole job, every block resident in HBM before the timed region; CUDA events on the stream
          the kernels run on; max over ranks.
  e2e   : the same job through the host-buffer C ABI: every block uploaded from pinned host memory and the dump
          TEXT (what the reference fprintf's) of every block downloaded to host memory inside the timed region.
  N > 1 : STRONG scaling of the same job. Blocks are dealt to the ranks round-robin: round j = blocks
          [jN, jN+N), rank g takes block jN+g - contiguous record ranges inside every round. Per round one NCCL
          all-gather of the per-k-mer counts + the exclusive scan over ranks (fused in k_quota) fixes every
          rank's share of sample_limit; totals are carried from round to round.
  c2    : resident figure of configs[1] (round 1's workload) for comparison, N=1 only.

`--impl reference` times the UNMODIFIED reference (oracle/_ref/poregen_ref, single thread: it has no other
mode) on a bounded prefix of the same generator's reads.
"""
from __future__ import annotations

import argparse
import copy
import ctypes as C
import json
import os
import resource
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
C3_READS = 100000


def workload_desc(reads, block, world):
    return {"workload": "C3 synthetic DNA R10.4.1 5 kHz move-table reads: gmove -k 9 --sample_limit 1000 --file_limit 262144 "
                        "(262,144 k-mer buckets, no scaling)",
            "reads": reads, "median_samples_per_read": 50000, "block_reads": block,
            "parallelism": f"read-range blocks dealt round-robin x{world}, one count all-gather per round",
            "l2": "inputs (10.4 GB int16 per step) exceed the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the unmodified reference binary on a prefix of the same reads
# ------------------------------------------------------------------------------------------------
def _raise_nofile():
    soft, hard = resource.getrlimit(resource.RLIMIT_NOFILE)
    want = hard if hard != resource.RLIM_INFINITY else 1 << 20
    for target in (want, 1 << 20, 300000):
        try:
            resource.setrlimit(resource.RLIMIT_NOFILE, (min(target, want) if hard != resource.RLIM_INFINITY else target, hard))
            break
        except (ValueError, OSError):
            continue
    return resource.getrlimit(resource.RLIMIT_NOFILE)[0]


def reference_sample(n_reads):
    """(cfg, opts, batch, scaled) for the first n_reads reads of C3. --sample_limit is scaled with the prefix
    (1000 * n_reads / 100000) so the same share of the events (~65 %) is written as in the whole job: with the
    full limit a 1 % prefix never reaches the quota and the reference would format MORE text per sample than it
    does on the real job."""
    import dataclasses
    from poregen_b200 import synth, workloads
    cfg, opts = workloads.c3(n_reads=C3_READS)
    batch = synth.make_moves_fast(cfg, 0, n_reads)
    limit = max(1, round(opts.sample_limit * n_reads / C3_READS))
    return cfg, dataclasses.replace(opts, sample_limit=limit), batch


class RefFiles:
    """BLOW5 + move-table TSV of a batch in a temp dir; runs the reference (windowed over the k-mer index range
    when the fd limit is below 4^k: one FILE* per k-mer, reference src/gmove.cpp:460-473)."""

    def __init__(self, opts, batch):
        from poregen_b200 import synth
        self.opts, self.batch = opts, batch
        # tmpfs when it has room (the fastest file system for the reference's 262,144 fopen()s), else the default temp dir
        base = None
        try:
            st = os.statvfs("/dev/shm")
            if st.f_bavail * st.f_frsize > (8 << 30):
                base = "/dev/shm"
        except OSError:
            pass
        self.dir = tempfile.mkdtemp(prefix="pgm_ref_", dir=base)
        self.blow5 = os.path.join(self.dir, "reads.blow5")
        synth.write_blow5(batch, self.blow5)
        self.files = synth.write_alignment(batch, self.dir)
        nofile = _raise_nofile()
        n = opts.index_end - opts.index_start + 1
        per = max(1, min(n, nofile - 64))
        self.windows = [(a, min(a + per - 1, opts.index_end)) for a in range(opts.index_start, opts.index_end + 1, per)]
        self.n = 0

    def run_reference(self, binary):
        """One reference job (all windows); returns (seconds, [out dirs])."""
        import dataclasses
        import oracle_lib as ol
        outs, total = [], 0.0
        for (a, b) in self.windows:
            o = dataclasses.replace(self.opts, index_start=a, index_end=b)
            out = os.path.join(self.dir, f"ref{self.n}")
            self.n += 1
            cmd = [binary, "gmove"] + ol.cli_args(o) + [self.blow5, self.files["align"], out]
            t0 = time.perf_counter()
            p = subprocess.run(cmd, capture_output=True)
            total += time.perf_counter() - t0
            if p.returncode != 0:
                raise RuntimeError(f"reference failed rc={p.returncode}: {p.stderr.decode()[-500:]}")
            outs.append(out)
        return total, outs

    def close(self):
        shutil.rmtree(self.dir, ignore_errors=True)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as g
    g.build()
    import oracle_lib as ol
    kind = "reference" if os.path.exists(ol.REF_BIN) else "port"
    nofile = _raise_nofile()
    n_reads = args.ref_reads
    if kind == "reference" and nofile < 262144 + 64:
        n_reads = max(50, n_reads // (262144 // max(1, nofile - 64) + 1))   # every window is a full pass over the reads
    cfg, opts, batch = reference_sample(n_reads)
    n_samples = int(batch.reads["raw_len"].astype("int64").sum())
    times = []
    if kind == "reference":
        rf = RefFiles(opts, batch)
        try:
            ref_warm = min(args.warmup, 1)      # one pass warms the page cache and the SLOW5 index; a step costs tens of seconds
            for i in range(ref_warm + args.steps):
                dt, outs = rf.run_reference(ol.REF_BIN)
                for o in outs:
                    shutil.rmtree(o, ignore_errors=True)
                if i >= ref_warm:
                    times.append(dt)
            windows = len(rf.windows)
        finally:
            rf.close()
        what = (f"unmodified reference binary (-O2, 1 thread: it has no other mode), file to file (BLOW5 read, TSV parse, "
                f"262,144 dump files written) in {windows} k-mer index window(s) (RLIMIT_NOFILE {nofile})")
    else:
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            ol.oracle_run(opts, batch)
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
        what = "CPU oracle port (oracle/gmove_oracle.c), in memory, incl. text formatting"
    examined = int(ol.oracle_run(opts, batch).n_events_examined)
    t = sum(times) / len(times)
    val = n_samples / t
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_desc(C3_READS, args.block_reads or min(12500, max(256, -(-C3_READS // (4 * args.gpus)))), args.gpus),
            "events_per_sec": examined / t,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": kind,
                             "sample": f"first {n_reads} of the 100,000 C3 reads ({n_samples} samples) per step, --sample_limit "
                                       f"{opts.sample_limit} (1000 scaled with the prefix so the same share of events is written); {what}"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)


def cpu_baseline_leg(n_reads):
    """The reference on a bounded prefix (~10-30 s of one host core), and this build's command line on the same files."""
    import oracle_lib as ol
    nofile = _raise_nofile()
    if nofile < 262144 + 64:
        n_reads = max(50, n_reads // (262144 // max(1, nofile - 64) + 1))
    cfg, opts, batch = reference_sample(n_reads)
    n_samples = int(batch.reads["raw_len"].astype("int64").sum())
    if not os.path.exists(ol.REF_BIN):
        t0 = time.perf_counter()
        ol.oracle_run(opts, batch)
        dt = time.perf_counter() - t0
        return {"value": n_samples / dt, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"first {n_reads} C3 reads ({n_samples} samples, {dt:.1f} s), --sample_limit {opts.sample_limit}: CPU oracle port, in memory"}
    rf = RefFiles(opts, batch)
    try:
        dt, outs = rf.run_reference(ol.REF_BIN)
        out = {"value": n_samples / dt, "unit": UNIT, "cores": 1, "kind": "reference", "seconds": dt,
               "sample": f"first {n_reads} of the 100,000 C3 reads ({n_samples} samples, {dt:.1f} s), --sample_limit {opts.sample_limit} "
                         f"(1000 scaled with the prefix): unmodified reference binary (-O2), file to file, {len(rf.windows)} index window(s)"}
        cli = _cli_leg(rf, outs, n_samples)
        if cli:
            out["same_files_through_cli"] = cli
    finally:
        rf.close()
    return out


def _cli_leg(rf, ref_outs, n_samples):
    """`poregen gmove` of this build on the SAME files the reference just processed (file to file: BLOW5 decode on
    host threads, device batches, dump/ written), its phase breakdown (PGM_CLI_STATS), and a byte comparison."""
    import oracle_lib as ol
    cli = os.path.join(ROOT, "poregen_b200", "bin", "poregen")
    if not os.path.exists(cli):
        return None
    out_dir = os.path.join(rf.dir, "cli_out")
    cmd = [cli, "gmove"] + ol.cli_args(rf.opts) + [rf.blow5, rf.files["align"], out_dir]
    best, stats = None, None
    env = dict(os.environ, PGM_CLI_STATS="1")
    for i in range(2):   # the first run pays the SLOW5 index and a cold CUDA context
        shutil.rmtree(out_dir, ignore_errors=True)
        t0 = time.perf_counter()
        p = subprocess.run(cmd, capture_output=True, env=env)
        dt = time.perf_counter() - t0
        if p.returncode != 0:
            return {"error": p.stderr.decode()[-300:]}
        if best is None or dt < best:
            best = dt
            stats = None
            for ln in p.stderr.decode().splitlines():      # "[gmove] N records, ...; start 0.1s create 0.2s parse ... close 0.0s" (gmove_cli.cpp, PGM_CLI_STATS)
                if ln.startswith("[gmove]") and "; start " in ln:
                    import re
                    stats = {k + "_s": float(v) for k, v in re.findall(r"(start|create|parse|read|device|write|close) ([0-9.]+)s", ln.split(";", 1)[1])}
    # byte comparison: freq.txt of every window concatenated == ours; every non-empty dump file identical
    same = True
    try:
        ours = open(os.path.join(out_dir, "freq.txt"), "rb").read()
        ref = b"".join(open(os.path.join(o, "freq.txt"), "rb").read() for o in ref_outs)
        same = ours == ref
        checked = 0
        for o in ref_outs:
            d = os.path.join(o, "dump")
            for name in sorted(os.listdir(d))[::97]:      # every 97th file of 262,144: ~2,700 files
                if open(os.path.join(d, name), "rb").read() != open(os.path.join(out_dir, "dump", name), "rb").read():
                    same = False
                checked += 1
    except OSError:
        same, checked = False, 0
    return {"value": n_samples / best, "unit": UNIT, "seconds": best, "identical_to_reference": same, "dump_files_compared": checked,
            "phases": stats,
            "phases_note": "start = option parsing + k-mer list + opening 4^k dump files (overlaps CUDA start-up); create = pgm_create; parse = alignment "
                           "text; read = waiting for SLOW5 records (decoded on host threads); device = pgm_submit/pgm_wait; write = fwrite of dump text; close = fclose + freq.txt",
            "note": "process start to exit, incl. CUDA context creation"}


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


class _DevArr:
    def __init__(self, ptr, n, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


def bind_near_gpu(local):
    """Pin this process to the cores of the NUMA node its GPU hangs off, BEFORE any pinned buffer is allocated
    (first touch then places the staging memory next to the GPU's PCIe root). Returns a description for the JSON
    line, or the reason it did nothing."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        if not os.path.exists(base + "/numa_node"):
            return {"bound": False, "why": f"{base}/numa_node absent"}
        node = int(open(base + "/numa_node").read())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if node < 0:
            return {"bound": False, "why": "numa_node = -1 (single-node host or no affinity exposed)", "gpu": bdf}
        if not cpus:
            return {"bound": False, "why": "no local cpu in this process's affinity mask", "gpu": bdf, "numa_node": node}
        os.sched_setaffinity(0, cpus)
        return {"bound": True, "gpu": bdf, "numa_node": node, "cpus": len(cpus)}
    except Exception as e:   # noqa: BLE001
        return {"bound": False, "why": repr(e)[:120]}


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
            return json.load(open(p)).get("c3", {}).get(stage)
        except Exception:
            return None
    return None


def stage_bytes(mode_scaling, T, NB, R, NS, S_act, E_pass, E_sel, S_sel):
    """Algorithmic bytes per stage."""
    return {"prep": 120 * R, "decode": T + NB + 88 * R + 16 * E_pass, "place": 32 * E_pass + 8 * R,
            "sort": 16 * E_pass + 12 * E_sel, "select": 16 * NS, "stats": (2 * S_act + 48 * R) if mode_scaling else 0,
            "gather": 2 * S_sel + 8 * S_sel + 20 * E_sel, "scan": 12 * E_sel, "misc": 8 * NS, "text": 0}


# ------------------------------------------------------------------------------------------------
def pcie_calibration(local, nbytes=1 << 30):
    """Plain pinned cudaMemcpyAsync in both directions at once on this rank's GPU (all ranks run it together):
    what the e2e step can reach at best. Returns (h2d GB/s, d2h GB/s) measured concurrently."""
    import torch
    dev = torch.device("cuda", local)
    h_in = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_out = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    d_out = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    res = []
    for rep in range(3):
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s1):
            a0.record(); d_in.copy_(h_in, non_blocking=True); a1.record()
        with torch.cuda.stream(s2):
            b0.record(); h_out.copy_(d_out, non_blocking=True); b1.record()
        torch.cuda.synchronize()
        res.append((nbytes / a0.elapsed_time(a1) / 1e6, nbytes / b0.elapsed_time(b1) / 1e6))
    return max(r[0] for r in res), max(r[1] for r in res)


def run_own_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    from poregen_b200 import _capi as capi, synth, workloads
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
    dev = torch.device("cuda", local)

    cfg, opts = workloads.c3(n_reads=args.reads)
    # resident block: 12,500 records on one GPU (larger blocks amortise the per-block table scans); with several GPUs at least four
    # rounds per rank: the job's first blocks select every event and its last ones few, and a rank's share evens out over the rounds
    # (one round per rank at N=8 left rank 0 with the heaviest block: 5.4 ms against 4.8 with four rounds). Two rounds in flight per
    # rank (sibling contexts) were measured too: N=2 16.99 against 16.56 ms, and every kernel's bracket then holds the other lane's kernels.
    B = args.block_reads or min(12500, max(256, -(-args.reads // (4 * world))))
    n_blocks = (args.reads + B - 1) // B
    rounds = (n_blocks + world - 1) // world
    my_blocks = [j * world + rank for j in range(rounds)]            # may run past n_blocks in the last round: an empty block
    t_gen0 = time.perf_counter()
    keep_alive = []

    def pinned(nbytes):
        t = torch.empty(max(1, nbytes), dtype=torch.uint8, pin_memory=True)
        keep_alive.append(t)
        return t.numpy()[:nbytes]

    # The e2e job is the same job dealt in its own, finer blocks (the same round-robin rule): --e2e-block-reads, by default at most
    # 3,125 records and at least 8 rounds per rank - with few rounds the two-deep pipeline is mostly fill and drain (the first
    # upload and the last download overlap nothing). Pinned host memory. (The e2e figure moves by 5 % from box to
    # box of the pool - 805 or 850 ms per job for the same library - always at 0.93-0.98 of the PCIe rate measured in the same run.)
    e2e_B = args.e2e_block_reads or min(3125, max(256, -(-args.reads // (8 * world))))
    e2e_nb = (args.reads + e2e_B - 1) // e2e_B
    e2e_rounds = (e2e_nb + world - 1) // world
    e2e_host_blocks = []
    for j in range(0 if args.no_e2e else e2e_rounds):
        lo, hi = (j * world + rank) * e2e_B, min(args.reads, (j * world + rank + 1) * e2e_B)
        if lo >= hi:
            e2e_host_blocks.append(None)
            continue
        hb = synth.make_moves_fast(cfg, lo, hi - lo, alloc=pinned)
        p_reads = pinned(hb.reads.nbytes)
        p_reads[:] = hb.reads.view(np.uint8).reshape(-1)
        hb.reads = p_reads.view(np.dtype(capi.READ_DTYPE))
        e2e_host_blocks.append(hb)
    host_blocks, dev_blocks, n_samples_mine = [], [], 0
    res_T = res_NB = res_R = res_nblk = 0              # this rank's share of the resident job, for the byte model
    for b in my_blocks:
        lo, hi = b * B, min(args.reads, (b + 1) * B)
        if lo >= hi:
            host_blocks.append(None); dev_blocks.append(None)
            continue
        hb = synth.make_moves_fast(cfg, lo, hi - lo)       # pageable: these host copies only travel to HBM once, here
        host_blocks.append(None)
        n_samples_mine += int(hb.reads["raw_len"].astype(np.int64).sum())
        res_T += int(hb.reads["align_len"].astype(np.int64).sum()); res_NB += int(hb.reads["seq_len"].astype(np.int64).sum())
        res_R += hb.n_reads; res_nblk += 1
        d = {k: torch.from_numpy(np.ascontiguousarray(getattr(hb, k)).view(np.uint8).reshape(-1)).to(dev) for k in ("raw", "align", "bases", "reads")}
        cb = hb.to_c()
        cb.raw, cb.align, cb.bases, cb.reads = d["raw"].data_ptr(), d["align"].data_ptr(), d["bases"].data_ptr(), d["reads"].data_ptr()
        dev_blocks.append((cb, d))
    gen_s = time.perf_counter() - t_gen0

    stream = torch.cuda.Stream(device=local)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng = GmoveEngine(opts, device=local, stream=stream.cuda_stream)
    NS = eng.n_slots
    counts_t = all_counts = None
    if world > 1:
        counts_t = torch.as_tensor(_DevArr(eng.counts_ptr(), NS), device=dev)
        all_counts = torch.empty((world, NS), dtype=torch.int32, device=dev)
    empty_batch = capi.PgmBatch()    # a rank without a block in the last round still joins the all-gather

    def job_resident(summaries=None):
        eng.reset()
        for j in range(rounds):
            blk = dev_blocks[j]
            eng.attach_device(blk[0] if blk else empty_batch)
            eng.count()
            if world > 1:
                dist.all_gather_into_tensor(all_counts.view(-1), counts_t)
                eng.set_gathered(all_counts.data_ptr(), world, rank)
            eng.emit()
            if summaries is not None:
                s = eng.summary()
                summaries.append((int(s.n_events_examined), int(s.n_events_pass), int(s.n_sel_events), int(s.n_sel_samples)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident timing
    for _ in range(args.warmup - 1):
        job_resident()
    summ = []
    job_resident(summ)                      # last warm-up pass also collects the per-block counters for the byte model
    barrier()
    eng.profile_enable(True)
    eng.profile_read()
    for _ in range(args.steps):             # UNTIMED pass with every stage bracketed -> stage table
        job_resident()
    stages = eng.profile_read()
    dom = max(stages, key=lambda k: stages[k][0])
    eng.profile_enable(True, only=dom)      # timed region: only the dominant kernel keeps its event pair
    eng.profile_read()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        job_resident()
    e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = eng.launches - l0
    dom_timed = eng.profile_read()[dom]
    eng.profile_enable(False)
    clocks = sampler.stop() if rank == 0 else None
    t_res = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_res, op=dist.ReduceOp.MAX)
    ms_step = float(t_res.item()) / args.steps

    # ---- K7 (BASELINE.json configs[4]: per-k-mer median / stddev of everything the job dumped; scripts/poregen.sh:54-85):
    # one more pass of the job with the model log on (every block's selected values stay in HBM as the integers their
    # "%.8f" text spells), then the aggregation itself is timed: one GPU: pgm_model_finalize; several: all-to-all of the
    # logged values by k-mer range over NVLink (NCCL), pgm_model_finalize on the owners, all-gather of the rows.
    model_info = None
    if not args.no_model:
        from poregen_b200.gmove import model_exchange
        eng.model_begin()
        eng.reset()
        for j in range(rounds):
            blk = dev_blocks[j]
            eng.attach_device(blk[0] if blk else empty_batch)
            eng.count()
            if world > 1:
                dist.all_gather_into_tensor(all_counts.view(-1), counts_t)
                eng.set_gathered(all_counts.data_ptr(), world, rank)
            eng.emit()
            eng.model_log()
        mt = {}
        if world > 1:   # NCCL opens its peer-to-peer channels on the first all-to-all: not part of the aggregation
            w_in = torch.zeros(world * 1024, dtype=torch.int64, device=dev)
            w_out = torch.empty_like(w_in)
            dist.all_to_all_single(w_out, w_in)
            torch.cuda.synchronize(dev)
        barrier()
        tm0 = time.perf_counter()
        rows = eng.model_finalize(3.1) if world == 1 else model_exchange(eng, opts, world, rank, 3.1, dev, timings=mt)
        barrier()
        model_s = time.perf_counter() - tm0
        tmx = torch.tensor([model_s], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tmx, op=dist.ReduceOp.MAX)
        model_info = {"ms": float(tmx.item()) * 1e3, "values": int(rows["n"].sum()), "rows": int(rows.size),
                      "rows_with_stddev": int(np.isfinite(rows["stddev"]).sum()), "rank0_breakdown": mt or None,
                      "how": "pgm_model_finalize (median by int64 radix selection + exact 128-bit variance per k-mer)" if world == 1 else
                             "NCCL all-to-all of the logged values by k-mer range, pgm_model_finalize on the owners, all-gather of the rows"}
    eng.close()

    # ---- end to end: every block from pinned host memory in, its dump TEXT out to host memory. Two blocks are in flight:
    # while block j's text travels device -> host, block j+1 is uploaded and its kernels run (PCIe is full duplex); the
    # per-k-mer totals are carried on the device from block to block. One GPU: pgm_submit / pgm_wait. Several: two sibling
    # contexts per rank (shared totals), the split-phase calls with the all-gather of the counts between count and emit.
    rounds = e2e_rounds
    host_blocks = e2e_host_blocks
    topts = copy.copy(opts)
    topts.output = capi.PGM_OUT_TEXT_ONLY
    teng = GmoveEngine(topts, device=local, stream=stream.cuda_stream)
    text_bytes = [0]

    def text_len(r):
        return int(C.cast(r.slot_text_off, C.POINTER(C.c_uint64))[NS])

    if world == 1:
        def job_e2e():
            teng.reset()
            ids, tb = [], 0
            for j in range(rounds):
                ids.append(teng.submit(host_blocks[j]))
                if len(ids) == 2:
                    tb += text_len(teng.wait_raw(ids.pop(0)))
            while ids:
                tb += text_len(teng.wait_raw(ids.pop(0)))
            text_bytes[0] = tb
        sib = None
    else:
        stream_b = torch.cuda.Stream(device=local)
        sib = teng.sibling(stream_b.cuda_stream)
        lanes = [(teng, stream), (sib, stream_b)]
        lane_counts = [torch.as_tensor(_DevArr(e.counts_ptr(), NS), device=dev) for e, _ in lanes]
        lane_all = [torch.empty((world, NS), dtype=torch.int32, device=dev) for _ in lanes]

        def job_e2e():
            with torch.cuda.stream(stream):
                teng.reset()
            pending, tb = [], 0

            def collect(e):
                r = capi.PgmResult()
                e._chk(e.lib.pgm_download(e._ctx, C.byref(r)))
                return text_len(r)
            trace = [] if (rank == 0 and os.environ.get("PGM_BENCH_TRACE")) else None
            for j in range(rounds):
                e, st_ = lanes[j % 2]
                t_a = time.perf_counter()
                if len(pending) == 2:
                    tb += collect(pending.pop(0))
                t_b = time.perf_counter()
                hb = host_blocks[j]
                with torch.cuda.stream(st_):
                    if hb is not None:
                        e.upload(hb)
                    else:
                        e.attach_device(empty_batch)
                    e.count()
                    dist.all_gather_into_tensor(lane_all[j % 2].view(-1), lane_counts[j % 2])
                    e.set_gathered(lane_all[j % 2].data_ptr(), world, rank)
                    e.emit()
                pending.append(e)
                if trace is not None:
                    trace.append((round((t_b - t_a) * 1e3, 1), round((time.perf_counter() - t_b) * 1e3, 1)))
            t_a = time.perf_counter()
            while pending:
                tb += collect(pending.pop(0))
            if trace is not None:
                print(f"[e2e trace rank 0] per round (collect ms, launch ms): {trace}; drain {round((time.perf_counter() - t_a) * 1e3, 1)} ms", file=sys.stderr)
            text_bytes[0] = tb

    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    if args.no_e2e:             # (a run that only wants the resident job and K7, e.g. the 1 M-read C5 shape)
        rounds, host_blocks, e2e_steps = 0, [], 0
    job_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        job_e2e()
    barrier()
    e2e_s = (time.perf_counter() - t0) / max(1, e2e_steps)
    t_e2e = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_s = float(t_e2e.item())
    if sib is not None:
        sib.close()
    teng.close()
    h2d = sum(int(hb.raw.nbytes + hb.align.nbytes + hb.bases.nbytes + hb.reads.nbytes) for hb in host_blocks if hb is not None)
    d2h = int(text_bytes[0] + sum((NS + 1) * 8 + 2 * NS * 4 + hb.n_reads + 48 for hb in host_blocks if hb is not None))
    pc_h2d, pc_d2h = pcie_calibration(local)

    sm = np.array(summ, dtype=np.float64).sum(axis=0) if summ else np.zeros(4)
    tot = torch.tensor([n_samples_mine, sm[0], sm[1], sm[2], sm[3], h2d, d2h, pc_h2d, pc_d2h], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    tot_samples, tot_events, tot_pass, tot_sel_ev, tot_sel_samp, tot_h2d, tot_d2h, sum_h2d_gbs, sum_d2h_gbs = [float(x) for x in tot.tolist()]

    if rank == 0:
        T, NB, R, nblk = res_T, res_NB, res_R, res_nblk
        E_pass, E_sel, S_sel = int(sm[1]), int(sm[2]), int(sm[3])
        alg = stage_bytes(opts.scaling, T, NB, R, NS * nblk, 0, E_pass, E_sel, S_sel)
        alg_step = 2 * n_samples_mine + T + NB + 24 * R + 8 * S_sel + 12 * E_sel + 4 * NS * nblk
        peak, peak_src = measured_peak()
        stage_rows = {}
        for name, (ms, ln) in stages.items():
            per = ms / args.steps
            stage_rows[name] = {"ms_per_step": per, "launches_per_step": ln / args.steps, "algorithmic_bytes": alg.get(name, 0),
                                "gbs": (alg.get(name, 0) / (per * 1e-3) / 1e9) if per > 0 else None}
        dom_launches = max(1.0, stage_rows[dom]["launches_per_step"])
        dom_per_launch_ms = dom_timed[0] / max(1, dom_timed[1])
        alg_per_launch = stage_rows[dom]["algorithmic_bytes"] / dom_launches
        achieved = alg_per_launch / (dom_per_launch_ms * 1e-3) / 1e9
        pcie_need_s = max(tot_h2d / world / (sum_h2d_gbs / world * 1e9), tot_d2h / world / (sum_d2h_gbs / world * 1e9))
        line = {"metric": METRIC, "value": tot_samples / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": workload_desc(args.reads, B, world),
                "events_per_sec": tot_events / (ms_step * 1e-3),
                "events_examined_per_step": tot_events, "events_pass_per_step": tot_pass,
                "events_selected_per_step": tot_sel_ev, "samples_selected_per_step": tot_sel_samp,
                "e2e": None if args.no_e2e else {"value": tot_samples / e2e_s, "unit": UNIT, "ms_per_step": e2e_s * 1e3, "events_per_sec": tot_events / e2e_s,
                        "h2d_bytes_per_step": int(tot_h2d), "d2h_bytes_per_step": int(tot_d2h), "steps": e2e_steps,
                        "output": "dump text (%.8f, formatted on the device), per-k-mer text offsets and counts, per-record flags",
                        "batches_in_flight": 2, "how": "pgm_submit / pgm_wait" if world == 1 else "two sibling contexts per rank, split-phase calls",
                        "block_reads": e2e_B,
                        "pcie_peak_gbs": {"h2d": sum_h2d_gbs, "d2h": sum_d2h_gbs, "how": "pinned cudaMemcpyAsync, 1 GiB, both directions at once, all ranks at once (sum over ranks)"},
                        "frac_of_pcie": pcie_need_s / e2e_s, "host_binding": numa},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": ncu_traffic(dom), "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": alg_per_launch, "ms_per_launch": dom_per_launch_ms,
                             "launches_per_step": dom_launches, "note": "rank 0's share of the job; per-launch = average over the blocks of a step"},
                "roofline_step": {"algorithmic_bytes": alg_step, "gbs": alg_step / (ms_step * 1e-3) / 1e9,
                                  "frac": alg_step / (ms_step * 1e-3) / 1e9 / peak, "note": "rank 0's share / step time"},
                "stages": stage_rows, "stages_note": "per-stage times from an untimed pass of the same steps with all stage timers on; the timed region brackets only the roofline kernel",
                "model_aggregation": model_info,
                "setup": {"generate_and_upload_s": gen_s}, "clocks": clocks}
        if world == 1 and not args.no_c2:
            line["c2"] = c2_leg(args, local, stream)
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_leg(args.cpu_reads)
        emit_line(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def c2_leg(args, local, stream):
    """configs[1] (round 1's bench workload), resident: 10,000 RNA004 ss records, -k 5 --rna --scaling 1 --sample_limit 5000
    --min_dur 20 --max_dur 40. One batch per step."""
    import numpy as np
    import torch
    from poregen_b200 import synth, workloads
    from poregen_b200.gmove import GmoveEngine
    cfg, opts = workloads.c2()
    hb = synth.make_batch(cfg)
    n_samples = int(hb.reads["raw_len"].astype(np.int64).sum())
    with GmoveEngine(opts, device=local, stream=stream.cuda_stream) as eng:
        eng.upload(hb)

        def step():
            eng.reset(); eng.count(); eng.emit()
        for _ in range(3):
            step()
        res = eng.download()
        eng.upload(hb)
        eng.profile_enable(True); eng.profile_read()
        for _ in range(args.steps):
            step()
        stages = eng.profile_read()
        eng.profile_enable(False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(args.steps):
            step()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
    rd = hb.reads
    T = int(rd["align_len"].astype(np.int64).sum()); NB = int(rd["seq_len"].astype(np.int64).sum()); R = hb.n_reads
    act = np.unique(res.ev_read)
    S_act = int((rd["raw_len"][act].astype(np.int64) - rd["trim"][act].astype(np.int64)).sum())
    alg = stage_bytes(1, T, NB, R, res.n_slots, S_act, res.n_events_pass, res.ev_len.size, res.samples.size)
    alg_step = 2 * n_samples + T + NB + 24 * R + 8 * res.samples.size + 12 * res.ev_len.size + 4 * res.n_slots
    peak, _ = measured_peak()
    rows = {k: {"ms": v[0] / args.steps, "gbs": alg.get(k, 0) / (v[0] / args.steps * 1e-3) / 1e9 if v[0] > 0 else None} for k, v in stages.items() if v[0] > 0}
    return {"workload": "C2 synthetic RNA004 ss records (10,000 reads): -k 5 --rna --scaling 1 --sample_limit 5000 --min_dur 20 --max_dur 40",
            "value": n_samples / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
            "roofline_step": {"algorithmic_bytes": alg_step, "frac": alg_step / (ms * 1e-3) / 1e9 / peak}, "stages": rows}


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
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--reads", type=int, default=C3_READS, help="records of the whole job (C3 = 100000)")
    ap.add_argument("--block-reads", type=int, default=0, help="consecutive records handed over per call (per rank and round), resident job (0: 12500, at least 4 rounds per rank)")
    ap.add_argument("--e2e-block-reads", type=int, default=0, help="the same for the end-to-end job (0: at most 3125, at least 8 rounds per rank)")
    ap.add_argument("--ref-reads", type=int, default=1000, help="--impl reference: records per step (prefix of the same reads)")
    ap.add_argument("--cpu-reads", type=int, default=2000, help="cpu_baseline leg: records in the sample")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (the line then has e2e: null)")
    ap.add_argument("--e2e-steps", type=int, default=3, help="steps of the end-to-end leg (each moves ~55 GB over PCIe)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-c2", action="store_true", help="skip the C2 comparison leg")
    ap.add_argument("--no-model", action="store_true", help="skip the K7 model aggregation leg")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "own":
        args.warmup = 3
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_own_arm(args)


if __name__ == "__main__":
    main()
