// pgm_emit.cuh — per-record preparation, quota selection, exact Med-MAD, sample gather.
#pragma once
#include "pgm_device.cuh"

// ------------------------------------------------------------------------------------------------
// k_read_prep: one thread per record. Validates the record, derives the in-range code interval and
// sets the record-level skip flags that do not need the signal:
//   host PGM_READ_SKIP (ss: fetched slice < k, gmove.cpp:806-808), fastq_len < 10 in the move-table
//   paths (gmove.cpp:616-619, 1178-1181), trim >= raw_len (assert gmove.cpp:590, 1147).
__global__ void __launch_bounds__(PGM_BLOCK) k_read_prep(DevOpts o, DevBatch b, ReadCal *cal, uint8_t *rflags, uint32_t *err_flag) {
    const uint32_t r = blockIdx.x * PGM_BLOCK + threadIdx.x;
    if (r >= b.n_reads) return;
    const pgm_read rd = b.reads[r];
    uint32_t f = 0;
    if (rd.raw_off + rd.raw_len > b.n_raw || rd.align_off + rd.align_len > b.n_align || rd.seq_off + rd.seq_len > b.n_bases) {
        atomicOr(err_flag, 1u);
        f |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED;
    }
    if (rd.flags & PGM_READ_SKIP) f |= PGM_RF_SKIPPED;
    const double scale = __ddiv_rn(rd.range, rd.digitisation);
    if (!isfinite(scale) || !isfinite(rd.offset)) f |= PGM_RF_BAD_CALIB | PGM_RF_SKIPPED;
    if (rd.trim >= rd.raw_len) f |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED;
    if (b.mode != PGM_MODE_SS && rd.fastq_len < 10) f |= PGM_RF_SKIPPED;
    ReadCal c;
    c.scale = scale; c.med = 0.0; c.mad = 1.0; c.clo = 0; c.chi = -1;
    if (!(f & PGM_RF_BAD_CALIB)) pgm_code_interval(rd.offset, scale, o.pa_min, o.pa_max, &c.clo, &c.chi);
    cal[r] = c;
    rflags[r] = (uint8_t)f;
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
__global__ void __launch_bounds__(PGM_BLOCK) k_select_prep(uint32_t n_slots, uint32_t limit, const uint32_t *prior, const uint32_t *counts,
                                                           uint32_t *keep, uint32_t *totals_out) {
    const uint32_t s = blockIdx.x * PGM_BLOCK + threadIdx.x;
    if (s >= n_slots) return;
    const uint32_t p = min(prior[s], limit);
    const uint32_t k = min(counts[s], limit - p);
    keep[s] = k;
    if (totals_out) totals_out[s] = p + k;
}

struct SelectIO {
    const uint32_t *n_events;     // device scalar: number of passing events E
    const uint32_t *skey, *sval;  // sorted (slot, event index)
    const uint32_t *bucket_base;  // [n_slots+1] exclusive scan of counts
    const uint32_t *keep;         // [n_slots]
    const uint32_t *sel_base;     // [n_slots+1] exclusive scan of keep
    const uint32_t *ev_start, *ev_len, *ev_read;  // [E] record-then-position order
    uint32_t *sel_start, *sel_len, *sel_read;     // [E_sel] grouped by slot
    uint8_t *read_has_sel;        // [R]
};

__global__ void __launch_bounds__(PGM_BLOCK) k_select(SelectIO io) {
    const uint32_t n = *io.n_events;
    for (uint32_t p = blockIdx.x * PGM_BLOCK + threadIdx.x; p < n; p += gridDim.x * PGM_BLOCK) {
        const uint32_t slot = io.skey[p];
        const uint32_t rank = p - io.bucket_base[slot];
        if (rank < io.keep[slot]) {
            const uint32_t id = io.sval[p];
            const uint32_t q = io.sel_base[slot] + rank;
            const uint32_t rr = io.ev_read[id];
            io.sel_start[q] = io.ev_start[id];
            io.sel_len[q] = io.ev_len[id];
            io.sel_read[q] = rr;
            io.read_has_sel[rr] = 1;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_stats: exact Med-MAD per record (reference gmove.cpp:142-184, 600-609), without sorting doubles.
//
// pA(code) is monotone in the int16 code, so the multiset of values {x_i} (with every out-of-range
// sample replaced by 0.0) is fully described by a histogram over the in-range codes
// plus one "zero mass" count. The upper median ksmall(n, n/2) and the n/2-th smallest
// |x - med| are then weighted order statistics over <= nb distinct values, each evaluated with the
// reference's own double expression, so ties and rounding are identical.
//
// Persistent CTAs pull records from an atomic queue; records with no selected event are skipped
// (their scale is never used). Histogram in shared memory when the in-range code span fits,
// otherwise in a per-CTA 65536-bin global scratch.
#define STATS_SMEM_BINS 8192

struct StatsCtx {
    uint32_t *H;       // histogram -> inclusive cumulative counts in VALUE order
    uint32_t nb;       // bins
    int clo;           // code of bin 0
    bool inc;          // value order == bin order
    double off, scale;
};
__device__ __forceinline__ uint32_t st_bin(const StatsCtx &c, uint32_t v) { return c.inc ? v : c.nb - 1 - v; }
__device__ __forceinline__ double st_val(const StatsCtx &c, uint32_t v) { return pgm_pa(c.clo + (int)st_bin(c, v), c.off, c.scale); }
__device__ __forceinline__ uint32_t st_cumI(const StatsCtx &c, uint32_t v) { return c.H[st_bin(c, v)]; }
__device__ __forceinline__ uint32_t st_cumE(const StatsCtx &c, uint32_t v) { return v ? c.H[st_bin(c, v - 1)] : 0u; }
// first v in [lo, hi) for which pred(v) holds (pred monotone false->true), else hi
template <typename P>
__device__ __forceinline__ uint32_t st_first(uint32_t lo, uint32_t hi, P pred) {
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (pred(mid)) hi = mid; else lo = mid + 1;
    }
    return lo;
}

__global__ void __launch_bounds__(PGM_BLOCK) k_stats(DevOpts o, DevBatch b, const uint8_t *rflags, const uint8_t *read_has_sel,
                                                     ReadCal *cal, uint32_t *queue, uint32_t *gscratch) {
    __shared__ uint32_t sm_hist[STATS_SMEM_BINS];
    __shared__ uint32_t sm_u[PGM_WARPS + 1];
    __shared__ uint32_t sm_r;
    __shared__ double sm_res;
    const uint32_t tid = threadIdx.x;
    for (;;) {
        __syncthreads();
        if (tid == 0) sm_r = atomicAdd(queue, 1u);
        __syncthreads();
        const uint32_t r = sm_r;
        if (r >= b.n_reads) break;
        if (rflags[r] != 0 || !read_has_sel[r]) continue;
        const pgm_read rd = b.reads[r];
        const ReadCal rc = cal[r];
        const uint64_t first = rd.raw_off + rd.trim;
        const uint32_t n = rd.raw_len - rd.trim;
        StatsCtx c;
        c.nb = (rc.chi >= rc.clo) ? (uint32_t)(rc.chi - rc.clo + 1) : 0u;
        c.clo = rc.clo; c.off = rd.offset; c.scale = rc.scale; c.inc = !(rc.scale < 0.0);
        c.H = (c.nb <= STATS_SMEM_BINS) ? sm_hist : gscratch + (size_t)blockIdx.x * 65536;
        for (uint32_t i = tid; i < c.nb; i += PGM_BLOCK) c.H[i] = 0;
        __syncthreads();
        // ---- histogram of in-range codes: 8 samples per 16-byte load on the aligned body
        {
            const bool use_sm = (c.nb <= STATS_SMEM_BINS);
#define ST_ADD(u_) do { if (use_sm) atomicAdd(&sm_hist[u_], 1u); else atomicAdd(&c.H[u_], 1u); } while (0)
            const int16_t *raw = b.raw;
            const uint64_t end = first + n;
            const uint64_t body0 = (first + 7) & ~(uint64_t)7, body1 = end & ~(uint64_t)7;
            if (body0 < body1) {
                for (uint64_t i = first + tid; i < body0; i += PGM_BLOCK) {
                    const uint32_t u = (uint32_t)((int)raw[i] - c.clo);
                    if (u < c.nb) ST_ADD(u);
                }
                const int4 *v4 = reinterpret_cast<const int4 *>(raw + body0);
                const uint64_t ng = (body1 - body0) >> 3;
                for (uint64_t g = tid; g < ng; g += PGM_BLOCK) {
                    const int4 q = __ldg(v4 + g);
                    const int w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t u0 = (uint32_t)((int)(int16_t)(w[j] & 0xffff) - c.clo);
                        const uint32_t u1 = (uint32_t)((w[j] >> 16) - c.clo);
                        if (u0 < c.nb) ST_ADD(u0);
                        if (u1 < c.nb) ST_ADD(u1);
                    }
                }
                for (uint64_t i = body1 + tid; i < end; i += PGM_BLOCK) {
                    const uint32_t u = (uint32_t)((int)raw[i] - c.clo);
                    if (u < c.nb) ST_ADD(u);
                }
            } else {
                for (uint64_t i = first + tid; i < end; i += PGM_BLOCK) {
                    const uint32_t u = (uint32_t)((int)raw[i] - c.clo);
                    if (u < c.nb) ST_ADD(u);
                }
            }
#undef ST_ADD
        }
        __syncthreads();
        // ---- histogram -> inclusive cumulative counts, in value order, in place
        uint32_t carry = 0;
        for (uint32_t base = 0; base < c.nb; base += PGM_BLOCK * 4) {
            const uint32_t v0 = base + tid * 4;
            uint32_t h[4], s = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) { h[j] = (v0 + j < c.nb) ? c.H[st_bin(c, v0 + j)] : 0u; s += h[j]; }
            uint32_t tot;
            uint32_t ex = carry + block_excl_scan_add(s, sm_u, &tot);
#pragma unroll
            for (int j = 0; j < 4; j++) { ex += h[j]; if (v0 + j < c.nb) c.H[st_bin(c, v0 + j)] = ex; }
            carry += tot;
        }
        __syncthreads();
        const uint32_t n_in = carry;
        const uint32_t nz = n - n_in;          // samples replaced by 0.0
        const uint32_t k = n / 2;              // ksmall(n, n/2)
        // ---- upper median (every thread computes it redundantly: uniform control flow)
        double med;
        {
            const uint32_t v0 = st_first(0u, c.nb, [&](uint32_t v) { return st_val(c, v) >= 0.0; });
            const uint32_t a = st_cumE(c, v0);  // in-range samples with value < 0.0
            if (k >= a && k < a + nz) med = 0.0;
            else {
                const uint32_t kk = (k < a) ? k : k - nz;
                const uint32_t v = st_first(0u, c.nb, [&](uint32_t vv) { return st_cumI(c, vv) > kk; });
                med = st_val(c, v);
            }
        }
        // ---- n/2-th smallest |x - med|: rank every distinct value in the order (dev, zero-mass < left < right)
        double mad_raw = 0.0;
        if (n > 1) {
            const uint32_t vm = st_first(0u, c.nb, [&](uint32_t v) { return st_val(c, v) >= med; });
            const uint32_t cm = st_cumE(c, vm);
            const double d0 = fabs(__dsub_rn(0.0, med));
            auto dev = [&](uint32_t v) { return fabs(__dsub_rn(st_val(c, v), med)); };
            // right side [vm, nb): dev non-decreasing in v.  left side [0, vm): dev non-increasing in v.
            auto cntR_lt = [&](double t) { return st_cumE(c, st_first(vm, c.nb, [&](uint32_t v) { return dev(v) >= t; })) - cm; };
            auto cntL_lt = [&](double t) { return cm - st_cumE(c, st_first(0u, vm, [&](uint32_t v) { return dev(v) < t; })); };
            auto cntL_le = [&](double t) { return cm - st_cumE(c, st_first(0u, vm, [&](uint32_t v) { return dev(v) <= t; })); };
            if (tid == 0) {
                sm_res = 0.0;
                if (nz) {
                    const uint32_t rs = cntL_lt(d0) + cntR_lt(d0);
                    if (rs <= k && k < rs + nz) sm_res = d0;
                }
            }
            __syncthreads();
            for (uint32_t v = tid; v < c.nb; v += PGM_BLOCK) {
                const uint32_t ci = st_cumI(c, v), ce = st_cumE(c, v);
                const uint32_t w = ci - ce;
                if (!w) continue;
                const double d = dev(v);
                uint32_t rs = (d0 <= d) ? nz : 0u;
                if (v < vm) rs += (cm - ci) + cntR_lt(d);
                else rs += (ce - cm) + cntL_le(d);
                if (rs <= k && k < rs + w) sm_res = d;
            }
            __syncthreads();
            mad_raw = sm_res;
        }
        if (tid == 0) {
            double mad = __dmul_rn(mad_raw, 1.4826);       // gmove.cpp:162,183
            mad = (mad > 1.0) ? mad : 1.0;                 // gmove.cpp:609
            cal[r].med = med;
            cal[r].mad = mad;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_gather: one warp per selected event; writes the doubles the reference prints (gmove.cpp:664-678)
// at the event's place in the slot-grouped output. raw -> pA (0.0 when out of range)
// -> optional (x - med) / mad (gmove.cpp:611-613).
struct GatherIO {
    const uint32_t *n_sel;        // device scalar E_sel
    const uint32_t *sel_start, *sel_len, *sel_read;
    const uint64_t *sel_off;      // [E_sel+1] exclusive scan of sel_len
    const ReadCal *cal;
    double *samples;
    uint64_t samples_cap;
};

__global__ void __launch_bounds__(PGM_BLOCK) k_gather(DevOpts o, DevBatch b, GatherIO io) {
    const uint32_t n = *io.n_sel;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wpg = gridDim.x * PGM_WARPS;
    for (uint32_t q = blockIdx.x * PGM_WARPS + (threadIdx.x >> 5); q < n; q += wpg) {
        const uint32_t r = io.sel_read[q];
        const uint32_t len = io.sel_len[q];
        const uint64_t off = io.sel_off[q];
        if (off + len > io.samples_cap) continue;
        const pgm_read *rd = b.reads + r;
        const ReadCal rc = io.cal[r];
        const double offset = rd->offset;
        const int16_t *x = b.raw + rd->raw_off + rd->trim + io.sel_start[q];
        double *out = io.samples + off;
        for (uint32_t t = lane; t < len; t += 32) {
            const int cval = x[t];
            double v = (cval >= rc.clo && cval <= rc.chi) ? pgm_pa(cval, offset, rc.scale) : 0.0;
            if (o.scaling) v = __ddiv_rn(__dsub_rn(v, rc.med), rc.mad);
            out[t] = v;
        }
    }
}
