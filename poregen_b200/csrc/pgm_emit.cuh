// pgm_emit.cuh — per-record preparation, quota selection, exact Med-MAD, sample gather.
#pragma once
#include "pgm_device.cuh"
#include "pgm_scan.cuh"

// ------------------------------------------------------------------------------------------------
// k_read_prep: one thread per record. Validates the record, derives the in-range code interval and
// sets the record-level skip flags that do not need the signal:
//   host PGM_READ_SKIP (ss: fetched slice < k, gmove.cpp:806-808), fastq_len < 10 in the move-table
//   paths (gmove.cpp:616-619, 1178-1181), trim >= raw_len (assert gmove.cpp:590, 1147).
// It also lays down the order in which the one-warp-per-record kernels (k_decode, k_stats) take the records:
// longest first, so that the last warps to finish hold short records (the kernels' run time is the slowest warp's,
// and a warp gets only two or three records). Records are counted into 64 length classes with an atomic rank
// inside the class; k_order turns the class counts into offsets and writes order[].
#define PREP_CLASSES 64
struct PrepOrder {
    uint32_t *cls_count;   // [PREP_CLASSES], zeroed per batch
    uint32_t *packed;      // [R] class << 26 | rank in class
    uint32_t *order;       // [R] out
};

__global__ void __launch_bounds__(PGM_BLOCK) k_read_prep(DevOpts o, DevBatch b, ReadCal *cal, uint8_t *rflags, uint32_t *err_flag, PrepOrder po,
                                                         unsigned long long *rec_src0) {
    const uint32_t r = blockIdx.x * PGM_BLOCK + threadIdx.x;
    if (r < b.n_reads) {
        const pgm_read rd = b.reads[r];
        uint32_t f = 0;
        if (rd.raw_off + rd.raw_len > b.n_raw || rd.align_off + rd.align_len > b.n_align || rd.seq_off + rd.seq_len > b.n_bases) {
            atomicOr(err_flag, 1u);
            f |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED;
        }
        if (b.mode == PGM_MODE_SS && rd.trim != 0) {    // the ss path has no trim (gmove.cpp:750-753): a caller that sets one is mistaken
            atomicOr(err_flag, 1u);
            f |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED;
        }
        rec_src0[r] = rd.raw_off + rd.trim;
        if (rd.flags & PGM_READ_SKIP) f |= PGM_RF_SKIPPED;
        const double scale = __ddiv_rn(rd.range, rd.digitisation);
        if (!isfinite(scale) || !isfinite(rd.offset)) f |= PGM_RF_BAD_CALIB | PGM_RF_SKIPPED;
        if (rd.trim >= rd.raw_len) f |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED;
        if (b.mode != PGM_MODE_SS && rd.fastq_len < 10) f |= PGM_RF_SKIPPED;
        ReadCal c;
        c.scale = scale; c.offset = rd.offset; c.med = 0.0; c.mad = 1.0; c.clo = 0; c.chi = -1; c.pad0 = c.pad1 = 0; c.pad2 = c.pad3 = 0.0;
        if (!(f & PGM_RF_BAD_CALIB)) pgm_code_interval(rd.offset, scale, o.pa_min, o.pa_max, &c.clo, &c.chi);
        cal[r] = c;
        rflags[r] = (uint8_t)f;
        // class 0 = longest (512-sample steps, everything from 32256 samples up together); skipped records last
        const uint32_t cls = f ? PREP_CLASSES - 1 : (PREP_CLASSES - 1) - min((uint32_t)PREP_CLASSES - 1, rd.raw_len >> 9);
        po.packed[r] = (cls << 26) | atomicAdd(&po.cls_count[cls], 1u);
    }
}

// k_order: class counts -> offsets (every CTA redoes the 64-element scan), then each record goes to its place.
__global__ void __launch_bounds__(PGM_BLOCK) k_order(uint32_t n_reads, PrepOrder po) {
    __shared__ uint32_t sm_off[PREP_CLASSES];
    if (threadIdx.x < 32) {
        const uint32_t lane = threadIdx.x;
        const uint32_t c0 = po.cls_count[lane], c1 = po.cls_count[lane + 32];
        uint32_t i0 = c0, i1 = c1;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t0 = __shfl_up_sync(0xffffffffu, i0, d), t1 = __shfl_up_sync(0xffffffffu, i1, d);
            if (lane >= (uint32_t)d) { i0 += t0; i1 += t1; }
        }
        const uint32_t half = __shfl_sync(0xffffffffu, i0, 31);
        sm_off[lane] = i0 - c0;
        sm_off[lane + 32] = half + i1 - c1;
    }
    __syncthreads();
    const uint32_t i = blockIdx.x * PGM_BLOCK + threadIdx.x;
    if (i < n_reads) {
        const uint32_t p = po.packed[i];
        po.order[sm_off[p >> 26] + (p & 0x03ffffffu)] = i;
    }
}

// ------------------------------------------------------------------------------------------------
// k_oor_check (PGM_MODE_MOVES_I8 only): the SAM/BAM path drops a record as soon as one sample of
// the trimmed signal is outside [pa_min, pa_max] (gmove.cpp:1149-1160). One CTA per record.
__global__ void __launch_bounds__(PGM_BLOCK) k_oor_check(DevBatch b, const ReadCal *cal, uint8_t *rflags) {
    const uint32_t r = blockIdx.x;
    if (rflags[r] != 0) return;
    const pgm_read rd = b.reads[r];
    const ReadCal c = cal[r];
    const int16_t *x = b.raw + rd.raw_off + rd.trim;
    const uint32_t n = rd.raw_len - rd.trim;
    int bad = 0;
    for (uint32_t i = threadIdx.x; i < n; i += PGM_BLOCK) {
        const int v = x[i];
        bad |= (v < c.clo) | (v > c.chi);
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) rflags[r] |= PGM_RF_SKIPPED;
}

// ------------------------------------------------------------------------------------------------
// Quota. prior[s] = events of slot s already emitted before this batch (this rank's view),
// counts[s] = passing events of this batch. keep = clamp(limit - prior, 0, counts): the first
// `keep` passing events of the slot, in (record, position) order, are the ones the reference
// would have written (gmove.cpp:650-652, 680-685).
// ------------------------------------------------------------------------------------------------
// k_quota: one launch for everything the quota needs. Per slot s: keep = clamp(limit - prior, 0, counts)
// and the totals after the batch; a packed look-back scan gives bucket_base = exscan(counts) (where each
// slot's bucket starts in the sorted order) and sel_base = exscan(keep) (where its emitted events start).
struct QuotaIO {
    uint32_t n_slots, limit;
    const uint32_t *prior;          // totals this rank must treat as already emitted
    const uint32_t *user_consumed;  // multi-GPU: totals after the round (or nullptr: prior + keep)
    const uint32_t *counts;
    const uint32_t *gathered;       // multi-GPU, fused form: counts of every rank [n_ranks][n_slots] (or nullptr)
    uint32_t n_ranks, rank;
    uint32_t *keep, *consumed, *bucket_base, *sel_base;   // consumed may alias prior
    uint32_t *snapshot;             // this batch's copy of the totals after it (siblings share `consumed` and move on)
};

__global__ void __launch_bounds__(PGM_BLOCK) k_quota(QuotaIO io, LbBuf st) {
    lb_scan<unsigned long long>(
        io.n_slots, st,
        [&](uint64_t s) {
            const uint32_t c = io.counts[s];
            uint32_t p = min(io.prior[s], io.limit), after;
            if (io.gathered) {
                // prior = before + counts of the lower ranks, totals after the round = before + all ranks
                unsigned long long lower = p, all = p;
                for (uint32_t h = 0; h < io.n_ranks; h++) {
                    const uint32_t v = io.gathered[(size_t)h * io.n_slots + s];
                    all += v;
                    if (h < io.rank) lower += v;
                }
                p = (uint32_t)min(lower, (unsigned long long)io.limit);
                after = (uint32_t)min(all, (unsigned long long)io.limit);
            }
            const uint32_t k = min(c, io.limit - p);
            io.keep[s] = k;
            if (!io.gathered) after = io.user_consumed ? io.user_consumed[s] : p + k;
            io.consumed[s] = after;
            io.snapshot[s] = after;
            return (unsigned long long)c | ((unsigned long long)k << 31);   // two 31-bit sums in one 62-bit scan value
        },
        [&](uint64_t s, unsigned long long ex, unsigned long long) {
            io.bucket_base[s] = (uint32_t)(ex & 0x7fffffffull);
            io.sel_base[s] = (uint32_t)(ex >> 31);
        },
        [&](unsigned long long tot) {
            io.bucket_base[io.n_slots] = (uint32_t)(tot & 0x7fffffffull);
            io.sel_base[io.n_slots] = (uint32_t)(tot >> 31);
        });
}

struct BatchScalars {  // device -> host summary of one batch
    unsigned long long examined;
    unsigned long long n_sel_samples;
    uint32_t n_pass;
    uint32_t n_sel;
    uint32_t err_flag;
    uint32_t pad;
    unsigned long long n_text;   // PGM_OUT_TEXT: characters of the dump text
};

// k_scan_sel: sample offsets of the selected events (exclusive scan of their lengths, 64-bit, decoupled look-back over tiles
// of LB_TILE events) and the batch summary, in one launch. The rows are read and the offsets written warp-contiguous
// (a thread's sixteen consecutive events take the detour through shared memory): the kernel moves 16 bytes in and 8 bytes out
// per event and nothing else.
#define SCANSEL_PAD(j) ((j) + ((j) >> 4))      // one pad word per sixteen: the per-thread (stride 16) and the per-warp (stride 1) walks both stay conflict-free
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_sel(const uint4 *sel, unsigned long long *sel_off, const uint32_t *n_sel_ptr,
                                                        const uint32_t *n_pass_ptr, BatchScalars *sc, LbBuf st) {
    __shared__ uint32_t sm_v[LB_TILE + LB_TILE / 16];
    __shared__ unsigned long long sm_scan[PGM_WARPS + 1];
    __shared__ unsigned long long sm_prefix;
    __shared__ uint32_t sm_tile;
    const uint32_t n = *n_sel_ptr;
    if (threadIdx.x == 0) sm_tile = atomicAdd(st.ticket, 1u);
    __syncthreads();
    const uint32_t tile = sm_tile;
    const uint32_t ntiles = (n + LB_TILE - 1) / LB_TILE;
    if (tile >= ntiles) {
        if (n == 0 && tile == 0 && threadIdx.x == 0) { sel_off[0] = 0; sc->n_sel_samples = 0; sc->n_sel = 0; sc->n_pass = *n_pass_ptr; }
        return;
    }
    const uint32_t base = tile * LB_TILE;
#pragma unroll
    for (int i = 0; i < LB_ITEMS; i++) {
        const uint32_t j = threadIdx.x + PGM_BLOCK * i;
        sm_v[SCANSEL_PAD(j)] = (base + j < n) ? (__ldg(sel + base + j).z & 0x00ffffffu) : 0u;
    }
    __syncthreads();
    uint32_t v[LB_ITEMS];
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < LB_ITEMS; i++) { v[i] = sm_v[SCANSEL_PAD(threadIdx.x * LB_ITEMS + i)]; s += v[i]; }
    unsigned long long tot;
    const unsigned long long ex = block_excl_scan_add_t<unsigned long long>(s, sm_scan, &tot);
    if (threadIdx.x < 32) {
        const unsigned long long excl = lb_lookback(st, tile, tot);
        if (threadIdx.x == 0) sm_prefix = excl;
    }
    if (tot > 0xffffffffull && threadIdx.x == 0) atomicOr(&sc->err_flag, 4u);     // offsets inside a tile travel as 32 bits
    uint32_t run = (uint32_t)ex;
#pragma unroll
    for (int i = 0; i < LB_ITEMS; i++) { sm_v[SCANSEL_PAD(threadIdx.x * LB_ITEMS + i)] = run; run += v[i]; }
    __syncthreads();
    const unsigned long long prefix = sm_prefix;
#pragma unroll
    for (int i = 0; i < LB_ITEMS; i++) {
        const uint32_t j = threadIdx.x + PGM_BLOCK * i;
        if (base + j < n) sel_off[base + j] = prefix + sm_v[SCANSEL_PAD(j)];
    }
    if (tile == ntiles - 1 && threadIdx.x == 0) {
        sel_off[n] = prefix + tot;
        sc->n_sel_samples = prefix + tot; sc->n_sel = n; sc->n_pass = *n_pass_ptr;
    }
}

