"""Python handle on the C ABI (include/pgm_gmove.h) — test/bench plumbing, not a second
implementation. Every method is a direct call into libpgm_gmove.so; there is no CPU path.

`GmoveEngine.process(batch)` is one pass of the reference's record loop
(process_move_table_file / _paf / _bam, reference src/gmove.cpp:539-706, 707-975, 1061-1266) over
`batch`, carrying the per-k-mer totals (kmer_frequency_map) from batch to batch.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _capi as capi
from .batch import GmoveOptions, GmoveResult, ReadBatch


class PgmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"pgm error {code}: {msg}")
        self.code = code


class GmoveEngine:
    def __init__(self, opts: GmoveOptions, device: int = 0, stream: Optional[int] = None):
        self.lib = capi.load()
        self.opts = opts
        self._copts, self._keep = opts.to_c()
        self._ctx = C.c_void_p()
        rc = self.lib.pgm_create(C.byref(self._copts), device, C.c_void_p(stream or 0), C.byref(self._ctx))
        if rc != capi.PGM_OK:
            raise PgmError(rc, self.lib.pgm_last_error(None).decode())
        self.n_slots = int(self._copts.n_slots)
        self._batch_keep = None

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            self.lib.pgm_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc: int):
        if rc != capi.PGM_OK:
            raise PgmError(rc, self.lib.pgm_last_error(self._ctx).decode())

    # -- batches in flight ---------------------------------------------------------------------
    def submit(self, batch: ReadBatch, svb=None) -> int:
        """pgm_submit: queue one whole pass of the next batch; returns its id. `svb` = (stream bytes, stream offsets)."""
        cb = batch.to_c()
        sv = None
        if svb is not None:
            cb.raw = None
            sv = capi.PgmSvb()
            sv.n_bytes, sv.bytes, sv.off = int(svb[0].size), svb[0].ctypes.data, svb[1].ctypes.data
        bid = C.c_uint64(0)
        self._inflight = getattr(self, "_inflight", {})
        self._chk(self.lib.pgm_submit(self._ctx, C.byref(cb), C.byref(sv) if sv is not None else None, C.byref(bid)))
        self._inflight[int(bid.value)] = (batch, svb)        # keeps the host buffers alive until the batch was waited for
        return int(bid.value)

    def wait_raw(self, batch_id: int) -> capi.PgmResult:
        """pgm_wait: the C result structure (pointers into the context's pinned buffers)."""
        r = capi.PgmResult()
        self._chk(self.lib.pgm_wait(self._ctx, batch_id, C.byref(r)))
        self._inflight.pop(batch_id, None)
        return r

    def wait(self, batch_id: int) -> GmoveResult:
        return GmoveResult.from_c(self.wait_raw(batch_id), want_text=self.opts.output != capi.PGM_OUT_F64)

    def sibling(self, stream: Optional[int] = None) -> "GmoveEngine":
        """pgm_create_sibling: another context of the same family (shared per-k-mer totals)."""
        k = GmoveEngine.__new__(GmoveEngine)
        k.lib, k.opts, k._copts, k._keep, k.n_slots, k._batch_keep = self.lib, self.opts, self._copts, self._keep, self.n_slots, None
        k._ctx = C.c_void_p()
        self._chk(self.lib.pgm_create_sibling(self._ctx, C.c_void_p(stream or 0), C.byref(k._ctx)))
        return k

    # -- one-call form (host buffers in, host result out) -------------------------------------
    def process(self, batch: ReadBatch) -> GmoveResult:
        cb = batch.to_c()
        r = capi.PgmResult()
        self._chk(self.lib.pgm_process(self._ctx, C.byref(cb), C.byref(r)))
        return GmoveResult.from_c(r, want_text=self.opts.output != capi.PGM_OUT_F64)

    def format_doubles(self, vals) -> bytes:
        """Device formatter test hook: the dump text of one event made of `vals`."""
        v = np.ascontiguousarray(vals, dtype=np.float64)
        n = C.c_uint64(0)
        self._chk(self.lib.pgm_debug_format(self._ctx, v.ctypes.data, v.size, None, 0, C.byref(n)))
        buf = C.create_string_buffer(int(n.value) + 1)
        self._chk(self.lib.pgm_debug_format(self._ctx, v.ctypes.data, v.size, buf, n.value, C.byref(n)))
        return buf.raw[:int(n.value)]

    # -- split-phase form -----------------------------------------------------------------------
    def upload(self, batch: ReadBatch):
        self._batch_keep = batch
        cb = batch.to_c()
        self._chk(self.lib.pgm_upload(self._ctx, C.byref(cb)))

    def upload_svb(self, batch: ReadBatch, stream_bytes: np.ndarray, stream_off: np.ndarray):
        """pgm_upload_svb: the records' svb-zd streams travel instead of raw[]; the samples are decoded in HBM."""
        self._batch_keep = (batch, stream_bytes, stream_off)
        cb = batch.to_c()
        cb.raw = None
        sv = capi.PgmSvb()
        sv.n_bytes = int(stream_bytes.size)
        sv.bytes = stream_bytes.ctypes.data
        sv.off = stream_off.ctypes.data
        assert stream_bytes.dtype == np.uint8 and stream_off.dtype == np.uint64 and stream_off.size == batch.n_reads + 1
        self._chk(self.lib.pgm_upload_svb(self._ctx, C.byref(cb), C.byref(sv)))

    def device_raw(self, n: int) -> np.ndarray:
        """Test hook (pgm_debug_raw): the loaded batch's first n samples as the device holds them."""
        out = np.empty(n, dtype=np.int16)
        self._chk(self.lib.pgm_debug_raw(self._ctx, out.ctypes.data, n))
        return out

    def attach_device(self, cbatch: capi.PgmBatch, keep=None):
        self._batch_keep = keep
        self._chk(self.lib.pgm_attach_device(self._ctx, C.byref(cbatch)))

    def count(self):
        self._chk(self.lib.pgm_count(self._ctx))

    def emit(self):
        self._chk(self.lib.pgm_emit(self._ctx))

    def download(self) -> GmoveResult:
        r = capi.PgmResult()
        self._chk(self.lib.pgm_download(self._ctx, C.byref(r)))
        return GmoveResult.from_c(r, want_text=self.opts.output != capi.PGM_OUT_F64)

    def summary(self) -> capi.PgmResult:
        """pgm_summary: the emitted batch's counters only (nothing but 48 bytes leaves the device)."""
        r = capi.PgmResult()
        self._chk(self.lib.pgm_summary(self._ctx, C.byref(r)))
        return r

    def sync(self):
        self._chk(self.lib.pgm_sync(self._ctx))

    def reset(self):
        self._chk(self.lib.pgm_reset(self._ctx))

    def counts_ptr(self) -> int:
        p = C.c_void_p()
        self._chk(self.lib.pgm_counts_device(self._ctx, C.byref(p)))
        return int(p.value)

    def totals_ptr(self) -> int:
        p = C.c_void_p()
        self._chk(self.lib.pgm_totals_device(self._ctx, C.byref(p)))
        return int(p.value)

    def set_prior(self, prior_dev_ptr: int, consumed_dev_ptr: int):
        self._chk(self.lib.pgm_set_prior(self._ctx, C.c_void_p(prior_dev_ptr), C.c_void_p(consumed_dev_ptr)))

    def set_gathered(self, all_counts_dev_ptr: int, n_ranks: int, rank: int):
        """Fused exchange: the all-gathered [n_ranks, n_slots] uint32 counts (device) and this rank."""
        self._chk(self.lib.pgm_set_gathered(self._ctx, C.c_void_p(all_counts_dev_ptr), n_ranks, rank))

    def model_begin(self):
        """Start logging the selected samples of every following batch for the model rows (K7)."""
        self._chk(self.lib.pgm_model_begin(self._ctx))

    def model_finalize(self, stddev_limit: float) -> np.ndarray:
        """Per-k-mer (median, stddev, ...) over everything logged since model_begin (scripts/poregen.sh:54-85)."""
        p = C.c_void_p()
        self._chk(self.lib.pgm_model_finalize(self._ctx, float(stddev_limit), C.byref(p)))
        n = self.n_slots
        buf = (C.c_char * (n * 40)).from_address(p.value)
        return np.frombuffer(buf, dtype=np.dtype(capi.MODEL_ROW_DTYPE), count=n).copy()

    def model_log(self):
        """pgm_model_log: put the emitted batch into the model log without downloading it."""
        self._chk(self.lib.pgm_model_log(self._ctx))

    def model_log_size(self) -> int:
        n = C.c_uint32(0)
        self._chk(self.lib.pgm_model_log_size(self._ctx, C.byref(n)))
        return int(n.value)

    def model_log_entry(self, i: int):
        """(device pointer to the batch's int64 values, slot_off uint64[n_slots+1] copy) of log entry i."""
        v, t = C.c_void_p(), C.c_void_p()
        self._chk(self.lib.pgm_model_log_entry(self._ctx, i, C.byref(v), C.byref(t)))
        buf = (C.c_char * ((self.n_slots + 1) * 8)).from_address(t.value)
        return int(v.value or 0), np.frombuffer(buf, dtype=np.uint64, count=self.n_slots + 1).copy()

    def model_log_append(self, vals_ptr: int, slot_off: np.ndarray):
        so = np.ascontiguousarray(slot_off, dtype=np.uint64)
        assert so.size == self.n_slots + 1
        self._chk(self.lib.pgm_model_log_append(self._ctx, C.c_void_p(vals_ptr), int(so[-1] - so[0]), so.ctypes.data))

    def model_log_reserve(self, n_vals_total: int):
        self._chk(self.lib.pgm_model_log_reserve(self._ctx, int(n_vals_total)))

    def model_dwell(self) -> np.ndarray:
        """Twice the per-k-mer dwell median over everything logged since model_begin (scripts/poregen.sh:33-52); -1: no event."""
        p = C.c_void_p()
        self._chk(self.lib.pgm_model_dwell(self._ctx, C.byref(p)))
        buf = (C.c_char * (self.n_slots * 8)).from_address(p.value)
        return np.frombuffer(buf, dtype=np.int64, count=self.n_slots).copy()

    def profile_enable(self, on: bool = True, only: Optional[str] = None):
        """Stage timers on/off; `only` = a stage name: bracket that stage alone (PGM_PROFILE_ONLY)."""
        mode = 1 if on else 0
        if on and only is not None:
            names = [self.lib.pgm_stage_name(i).decode() for i in range(capi.PGM_N_STAGES)]
            mode = 2 + names.index(only)
        self._chk(self.lib.pgm_profile_enable(self._ctx, mode))

    def profile_read(self) -> dict:
        """{stage name: (ms, launches)} accumulated since the last read (synchronises)."""
        ms = (C.c_float * capi.PGM_N_STAGES)()
        ln = (C.c_uint32 * capi.PGM_N_STAGES)()
        self._chk(self.lib.pgm_profile_read(self._ctx, ms, ln))
        return {self.lib.pgm_stage_name(i).decode(): (float(ms[i]), int(ln[i])) for i in range(capi.PGM_N_STAGES)}

    @property
    def stream(self) -> int:
        return int(self.lib.pgm_stream(self._ctx) or 0)

    @property
    def launches(self) -> int:
        return int(self.lib.pgm_launch_count(self._ctx))


def quota_exchange(counts_all: np.ndarray, before: np.ndarray, limit: int):
    """Host-side statement of the multi-GPU quota rule (north star).

    counts_all: [n_ranks, n_slots] passing events per rank in this round (rank order = record order),
    before:     [n_slots] totals before the round.
    Returns (prior [n_ranks, n_slots], consumed [n_slots]): prior[g] = before + sum of lower ranks'
    counts (what rank g must treat as already emitted), consumed = min(limit, before + all counts).
    """
    c = counts_all.astype(np.uint64)
    excl = np.cumsum(c, axis=0) - c
    prior = np.minimum(before.astype(np.uint64)[None, :] + excl, limit).astype(np.uint32)
    consumed = np.minimum(before.astype(np.uint64) + c.sum(axis=0), limit).astype(np.uint32)
    return prior, consumed


# ---------------------------------------------------------------------------------------------
# K7 across GPUs (SURVEY.md 8e "Output merge"; BASELINE.json configs[4]): every rank logged the values of its own
# batches; the k-mers are dealt to the ranks in contiguous ranges and the owner of a range collects, round by round and
# rank by rank (= the order the values stand in the reference's dump file), what every rank logged for it.
# ---------------------------------------------------------------------------------------------
def model_owner_ranges(n_slots: int, world: int):
    return [(n_slots * g // world, n_slots * (g + 1) // world) for g in range(world)]


def _piece_table(counts: np.ndarray, lo: int, hi: int, n_slots: int) -> np.ndarray:
    """slot_off[n_slots+1] of a piece that only holds the k-mers [lo, hi) (counts = values per k-mer of the range)."""
    t = np.zeros(n_slots + 1, np.uint64)
    t[lo + 1:hi + 1] = np.cumsum(counts.astype(np.uint64))
    t[hi + 1:] = t[hi]
    return t


def model_merge_local(engines, opts: GmoveOptions, stddev_limit: float) -> np.ndarray:
    """The exchange with every rank's context on ONE device (tests): engines[g] holds rank g's log; device-to-device
    copies stand in for the all-to-all. Returns the assembled rows of all k-mers."""
    world, NS = len(engines), engines[0].n_slots
    n_rounds = engines[0].model_log_size()
    assert all(e.model_log_size() == n_rounds for e in engines)
    entries = [[e.model_log_entry(r) for e in engines] for r in range(n_rounds)]
    rows = np.zeros(NS, np.dtype(capi.MODEL_ROW_DTYPE))
    for o, (lo, hi) in enumerate(model_owner_ranges(NS, world)):
        with GmoveEngine(opts) as agg:
            agg.model_begin()
            for r in range(n_rounds):
                for g in range(world):
                    ptr, so = entries[r][g]
                    agg.model_log_append(ptr + 8 * int(so[lo]), _piece_table(np.diff(so[lo:hi + 1]), lo, hi, NS))
            rows[lo:hi] = agg.model_finalize(stddev_limit)[lo:hi]
    return rows


def model_exchange(eng: GmoveEngine, opts: GmoveOptions, world: int, rank: int, stddev_limit: float, device, timings: Optional[dict] = None) -> np.ndarray:
    """One process per GPU: NCCL all-to-all of the logged values by k-mer range (torch.distributed on the log's device
    memory), pgm_model_finalize on the owner, all-gather of the rows. Every rank returns all rows. Collective."""
    import torch
    import torch.distributed as dist

    class _Dev:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (ptr, False), "version": 3}

    import time
    tm = {"counts_s": 0.0, "values_s": 0.0, "append_s": 0.0, "finalize_s": 0.0, "rows_s": 0.0, "bytes_sent": 0}

    def lap(key, t0):
        torch.cuda.synchronize(device)
        tm[key] += time.perf_counter() - t0
        return time.perf_counter()

    NS = eng.n_slots
    ranges = model_owner_ranges(NS, world)
    lo, hi = ranges[rank]
    n_rounds = eng.model_log_size()
    nr = torch.tensor([n_rounds], device=device)
    dist.all_reduce(nr, op=dist.ReduceOp.MAX)
    assert int(nr.item()) == n_rounds, "every rank logs one entry per round (empty batches included)"
    agg = GmoveEngine(opts, device=device.index)
    try:
        agg.model_begin()
        # phase 1: the per-k-mer counts of every owner's range, all rounds (NS int64 in, my range from every rank out per round):
        # the owner then knows how many values it will hold and reserves the log once
        t0 = time.perf_counter()
        in_split = [b - a for a, b in ranges]
        entries, got_cnts = [], []
        for r in range(n_rounds):
            ptr, so = eng.model_log_entry(r)
            entries.append((ptr, so))
            cnt = torch.from_numpy(np.diff(so).astype(np.int64)).to(device)
            got_cnt = torch.empty(world * (hi - lo), dtype=torch.int64, device=device)
            dist.all_to_all_single(got_cnt, cnt, output_split_sizes=[hi - lo] * world, input_split_sizes=in_split)
            got_cnts.append(got_cnt.view(world, hi - lo))
        cnt_hosts = [g.cpu().numpy() for g in got_cnts]
        agg.model_log_reserve(int(sum(int(c.sum()) for c in cnt_hosts)))
        t0 = lap("counts_s", t0)
        # phase 2: the values, one all-to-all per round; the received pieces are appended in (round, rank) = job order
        for r in range(n_rounds):
            ptr, so = entries[r]
            cnt_host = cnt_hosts[r]
            recv_n = [int(x) for x in cnt_host.sum(axis=1)]
            send_n = [int(so[b] - so[a]) for a, b in ranges]
            n_mine = int(so[-1])
            vals = torch.as_tensor(_Dev(ptr, n_mine), device=device) if n_mine else torch.empty(0, dtype=torch.int64, device=device)
            recv = torch.empty(int(sum(recv_n)), dtype=torch.int64, device=device)
            dist.all_to_all_single(recv, vals, output_split_sizes=recv_n, input_split_sizes=send_n)
            tm["bytes_sent"] += 8 * (n_mine - send_n[rank])
            t0 = lap("values_s", t0)
            at = 0
            for g in range(world):
                agg.model_log_append(recv.data_ptr() + 8 * at, _piece_table(cnt_host[g], lo, hi, NS))
                at += int(recv_n[g])
            t0 = lap("append_s", t0)
        t0 = time.perf_counter()
        mine = agg.model_finalize(stddev_limit)[lo:hi]
        t0 = lap("finalize_s", t0)
    finally:
        agg.close()
    # rows of every range to every rank (byte view: 40 bytes per row)
    width = max(b - a for a, b in ranges)
    buf = np.zeros(width, np.dtype(capi.MODEL_ROW_DTYPE))
    buf[:hi - lo] = mine
    t = torch.from_numpy(buf.view(np.uint8).copy()).to(device)
    allt = torch.empty(world * t.numel(), dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allt, t)
    allr = allt.cpu().numpy().view(np.dtype(capi.MODEL_ROW_DTYPE)).reshape(world, width)
    rows = np.zeros(NS, np.dtype(capi.MODEL_ROW_DTYPE))
    for g, (a, b) in enumerate(ranges):
        rows[a:b] = allr[g, :b - a]
    lap("rows_s", t0)
    if timings is not None:
        timings.update(tm)
    return rows
