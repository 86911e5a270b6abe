// pgm_signal.cuh — the signal side of gmove: exact Med-MAD and the gather of selected samples,
// fused into one kernel per record so the raw signal is read from HBM once.
//
// Med-MAD without sorting doubles (reference gmove.cpp:142-184, 600-609): pA(code) is monotone in the
// int16 code, so the multiset {x_i} (every out-of-range sample replaced by 0.0) is a
// histogram over the in-range codes plus one "zero mass". The upper median ksmall(n, n/2)
//  and the n/2-th smallest |x - med| are weighted order statistics over the distinct
// values, each evaluated with the reference's own double expression, so ties and rounding agree.
#pragma once
#include "pgm_device.cuh"

#define SIG_SMEM_BINS 2048      // in-range code span handled in shared memory by k_signal
#define STATS_SMEM_BINS 8192    // k_stats_wide: shared memory up to here, then global scratch

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

// Histogram of the in-range codes of raw[first, first+n) into H[0, nb) (zeroed by the caller).
// SMEM selects shared-memory atomics. 8 samples per 16-byte load on the aligned body.
template <bool SMEM>
__device__ __forceinline__ void stats_histogram(const int16_t *raw, uint64_t first, uint32_t n, int clo, uint32_t nb, uint32_t *H) {
    const uint32_t tid = threadIdx.x;
    const uint64_t end = first + n;
    const uint64_t body0 = (first + 7) & ~(uint64_t)7, body1 = end & ~(uint64_t)7;
    auto add = [&](int code) {
        const uint32_t u = (uint32_t)(code - clo);
        if (u < nb) atomicAdd(&H[u], 1u);
    };
    if (body0 < body1) {
        for (uint64_t i = first + tid; i < body0; i += PGM_BLOCK) add(raw[i]);
        const int4 *v4 = reinterpret_cast<const int4 *>(raw + body0);
        const uint32_t ng = (uint32_t)((body1 - body0) >> 3);
        for (uint32_t g = tid; g < ng; g += PGM_BLOCK) {
            const int4 q = __ldg(v4 + g);
            add((int)(int16_t)(q.x & 0xffff)); add(q.x >> 16);
            add((int)(int16_t)(q.y & 0xffff)); add(q.y >> 16);
            add((int)(int16_t)(q.z & 0xffff)); add(q.z >> 16);
            add((int)(int16_t)(q.w & 0xffff)); add(q.w >> 16);
        }
        for (uint64_t i = body1 + tid; i < end; i += PGM_BLOCK) add(raw[i]);
    } else {
        for (uint64_t i = first + tid; i < end; i += PGM_BLOCK) add(raw[i]);
    }
    (void)SMEM;
}

// Shared-memory variant for k_signal: branch-free. Out-of-range codes are clamped into one extra bin
// H[nb] (never read by the order statistics; the zero mass is n - sum of the others), and the
// increment is a `red.shared` on a 32-bit shared address computed once.
__device__ __forceinline__ void red_shared_inc(uint32_t saddr) { asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(saddr) : "memory"); }

__device__ __forceinline__ void stats_histogram_shared(const int16_t *raw, uint64_t first, uint32_t n, int clo, uint32_t nb, uint32_t *H) {
    const uint32_t tid = threadIdx.x;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(H);
    const uint64_t end = first + n;
    const uint64_t body0 = (first + 7) & ~(uint64_t)7, body1 = end & ~(uint64_t)7;
    auto add = [&](int code) { red_shared_inc(sbase + 4u * min((uint32_t)(code - clo), nb)); };
    if (body0 < body1) {
        const uint32_t head = (uint32_t)(body0 - first), tail = (uint32_t)(end - body1);
        if (tid < head) add(raw[first + tid]);
        if (tid < tail) add(raw[body1 + tid]);
        const int4 *v4 = reinterpret_cast<const int4 *>(raw + body0);
        const uint32_t ng = (uint32_t)((body1 - body0) >> 3);
        for (uint32_t g = tid; g < ng; g += PGM_BLOCK) {
            const int4 q = __ldg(v4 + g);
            add((int)(int16_t)(q.x & 0xffff)); add(q.x >> 16);
            add((int)(int16_t)(q.y & 0xffff)); add(q.y >> 16);
            add((int)(int16_t)(q.z & 0xffff)); add(q.z >> 16);
            add((int)(int16_t)(q.w & 0xffff)); add(q.w >> 16);
        }
    } else {
        for (uint32_t i = tid; i < n; i += PGM_BLOCK) add(raw[first + i]);
    }
}

// Shared scratch of stats_medmad.
struct StatsShared {
    uint32_t scan[PGM_WARPS + 1];
    uint32_t idx;      // result of a parallel index search
    uint32_t found;    // MAD candidate verified
    double res;        // MAD value
};

// cumulative count of in-range samples with value index <= x (x may be negative or >= nb)
__device__ __forceinline__ uint32_t st_cumAt(const StatsCtx &c, int64_t x, uint32_t n_in) {
    if (x < 0) return 0u;
    if (x >= (int64_t)c.nb) return n_in;
    return c.H[st_bin(c, (uint32_t)x)];
}

// From the histogram in c.H (all threads of the block call this together): turns it into cumulative
// counts, then computes med and the clamped mad exactly as the reference does. Index searches are
// done in parallel over the bins with integer compares; the reference's double expression is only
// evaluated for the handful of bins that can hold the answer.
__device__ __forceinline__ void stats_medmad(const StatsCtx &c, uint32_t n, StatsShared *sh, double *med_out, double *mad_out) {
    const uint32_t tid = threadIdx.x;
    // ---- histogram -> inclusive cumulative counts, in value order, in place
    uint32_t carry = 0;
    for (uint32_t base = 0; base < c.nb; base += PGM_BLOCK * 4) {
        const uint32_t v0 = base + tid * 4;
        uint32_t h[4], s = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) { h[j] = (v0 + j < c.nb) ? c.H[st_bin(c, v0 + j)] : 0u; s += h[j]; }
        uint32_t tot;
        uint32_t ex = carry + block_excl_scan_add(s, sh->scan, &tot);
#pragma unroll
        for (int j = 0; j < 4; j++) { ex += h[j]; if (v0 + j < c.nb) c.H[st_bin(c, v0 + j)] = ex; }
        carry += tot;
    }
    const uint32_t n_in = carry;
    const uint32_t nz = n - n_in;          // samples replaced by 0.0
    const uint32_t k = n / 2;              // ksmall(n, n/2)
    // ---- where 0.0 sorts among the in-range values: vz = first v with value >= 0.0
    if (tid == 0) {
        uint32_t vz;
        if (c.nb == 0 || st_val(c, 0) >= 0.0) vz = 0;                       // the usual case: pa_min > 0
        else if (st_val(c, c.nb - 1) < 0.0) vz = c.nb;
        else vz = st_first(0u, c.nb, [&](uint32_t v) { return st_val(c, v) >= 0.0; });
        sh->idx = vz;
        sh->found = 0;
    }
    __syncthreads();
    const uint32_t vz = sh->idx;
    const uint32_t a = st_cumE(c, vz);     // in-range samples with value < 0.0
    const bool med_is_zero = (k >= a && k < a + nz);
    const uint32_t kk = (k < a) ? k : k - nz;
    __syncthreads();
    // ---- upper median: the bin whose cumulative range holds rank kk
    if (!med_is_zero) {
        for (uint32_t v = tid; v < c.nb; v += PGM_BLOCK) {
            const uint32_t ci = st_cumI(c, v);
            if (kk < ci && st_cumE(c, v) <= kk) sh->idx = v;
        }
    }
    __syncthreads();
    const uint32_t vmed = sh->idx;
    const double med = med_is_zero ? 0.0 : st_val(c, vmed);
    double mad_raw = 0.0;
    if (n > 1) {
        // vm = first v with value >= med
        uint32_t vm = vmed;               // med_is_zero: vmed == vz, which is that index by definition
        if (!med_is_zero) while (vm > 0 && st_val(c, vm - 1) >= med) vm--;   // only moves for equal-valued neighbours
        const uint32_t cm = st_cumE(c, vm);
        const double d0 = fabs(__dsub_rn(0.0, med));
        auto dev = [&](uint32_t v) { return fabs(__dsub_rn(st_val(c, v), med)); };
        // right side [vm, nb): dev non-decreasing in v.  left side [0, vm): dev non-increasing in v.
        auto cntR_lt = [&](double t) { return st_cumE(c, st_first(vm, c.nb, [&](uint32_t v) { return dev(v) >= t; })) - cm; };
        auto cntL_lt = [&](double t) { return cm - st_cumE(c, st_first(0u, vm, [&](uint32_t v) { return dev(v) < t; })); };
        auto cntL_le = [&](double t) { return cm - st_cumE(c, st_first(0u, vm, [&](uint32_t v) { return dev(v) <= t; })); };
        // exact test: is the n/2-th smallest deviation the one of bin v? order on ties: zero mass < left < right
        auto try_bin = [&](uint32_t v) {
            const uint32_t ci = st_cumI(c, v), ce = st_cumE(c, v);
            const uint32_t w = ci - ce;
            if (!w) return;
            const double d = dev(v);
            uint32_t rs = (d0 <= d) ? nz : 0u;
            if (v < vm) rs += (cm - ci) + cntR_lt(d);
            else rs += (ce - cm) + cntL_le(d);
            if (rs <= k && k < rs + w) { sh->res = d; sh->found = 1; }
        };
        auto try_zero = [&]() {
            if (!nz) return;
            const uint32_t rs = cntL_lt(d0) + cntR_lt(d0);
            if (rs <= k && k < rs + nz) { sh->res = d0; sh->found = 1; }
        };
        __syncthreads();   // everyone has read vmed out of sh->idx
        // integer pre-search: smallest radius rho with #{samples within rho bins of vm} >= k+1
        // (deviation grows with the bin distance; which side wins at equal distance is settled exactly below)
        for (uint32_t rho = tid; rho <= c.nb; rho += PGM_BLOCK) {
            const uint32_t cur = st_cumAt(c, (int64_t)vm + rho, n_in) - st_cumAt(c, (int64_t)vm - rho - 1, n_in);
            const uint32_t prev = rho ? st_cumAt(c, (int64_t)vm + rho - 1, n_in) - st_cumAt(c, (int64_t)vm - rho, n_in) : 0u;
            if (prev < k + 1 && k + 1 <= cur) sh->idx = rho;
        }
        if (tid == 0 && n_in < k + 1) sh->idx = c.nb;   // only the zero mass can complete the rank
        __syncthreads();
        const int64_t rho = sh->idx;
        // candidates: bins at distance rho-2..rho+2 on both sides (11 lanes) + the zero mass
        if (tid < 10) {
            const int64_t dd = rho + (int64_t)(tid % 5) - 2;
            const int64_t v = (tid < 5) ? (int64_t)vm + dd : (int64_t)vm - dd;
            if (dd >= 0 && v >= 0 && v < (int64_t)c.nb) try_bin((uint32_t)v);
        } else if (tid == 31) try_zero();
        __syncthreads();
        if (!sh->found) {   // unusual calibration or a heavy zero mass: rank every bin
            if (tid == 0) try_zero();
            for (uint32_t v = tid; v < c.nb; v += PGM_BLOCK) try_bin(v);
            __syncthreads();
        }
        mad_raw = sh->res;
    }
    double mad = __dmul_rn(mad_raw, 1.4826);       // gmove.cpp:162,183
    mad = (mad > 1.0) ? mad : 1.0;                 // gmove.cpp:609
    *med_out = med;
    *mad_out = mad;
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// k_stats_wide: fallback for records whose in-range code span exceeds SIG_SMEM_BINS (unusual pa_min /
// pa_max or calibration). Persistent CTAs pull records from an atomic queue; histogram in shared memory
// up to STATS_SMEM_BINS bins, else in a per-CTA 65536-bin global scratch. Writes med/mad to cal[].
__global__ void __launch_bounds__(PGM_BLOCK) k_stats_wide(DevOpts o, DevBatch b, const uint8_t *rflags, const uint8_t *read_has_sel,
                                                          ReadCal *cal, uint32_t *queue, uint32_t *gscratch) {
    __shared__ uint32_t sm_hist[STATS_SMEM_BINS];
    __shared__ StatsShared sh;
    __shared__ uint32_t sm_r;
    const uint32_t tid = threadIdx.x;
    for (;;) {
        __syncthreads();
        if (tid == 0) sm_r = atomicAdd(queue, 1u);
        __syncthreads();
        const uint32_t r = sm_r;
        if (r >= b.n_reads) break;
        if (rflags[r] != 0 || !read_has_sel[r]) continue;
        const ReadCal rc = cal[r];
        StatsCtx c;
        c.nb = (rc.chi >= rc.clo) ? (uint32_t)(rc.chi - rc.clo + 1) : 0u;
        if (c.nb <= SIG_SMEM_BINS) continue;   // k_signal does these itself
        const pgm_read rd = b.reads[r];
        const uint64_t first = rd.raw_off + rd.trim;
        const uint32_t n = rd.raw_len - rd.trim;
        c.clo = rc.clo; c.off = rd.offset; c.scale = rc.scale; c.inc = !(rc.scale < 0.0);
        const bool use_sm = c.nb <= STATS_SMEM_BINS;
        c.H = use_sm ? sm_hist : gscratch + (size_t)blockIdx.x * 65536;
        for (uint32_t i = tid; i < c.nb; i += PGM_BLOCK) c.H[i] = 0;
        __syncthreads();
        if (use_sm) stats_histogram<true>(b.raw, first, n, c.clo, c.nb, sm_hist);
        else stats_histogram<false>(b.raw, first, n, c.clo, c.nb, c.H);
        __syncthreads();
        double med, mad;
        stats_medmad(c, n, &sh, &med, &mad);
        if (tid == 0) { cal[r].med = med; cal[r].mad = mad; }
    }
}

// ------------------------------------------------------------------------------------------------
// k_signal: one CTA per record that owns at least one selected event.
//   1. (scaling) histogram of the record's in-range codes in shared memory -> exact med / mad
//   2. gather: for every selected event of the record, raw -> pA (0.0 when out of range)
//      -> (x - med) / mad (gmove.cpp:611-613) -> the event's place in the slot-grouped output
//      (the doubles gmove.cpp:664-678 prints). The second touch of the record's signal hits L2.
// The record's passing events are the contiguous id range [ev_base[r], ev_base[r+1]); ev_out[id]
// is the event's index in the selected order (or PGM_INVALID), sel_off its sample offset.
struct SignalIO {
    const uint8_t *rflags, *read_has_sel;
    ReadCal *cal;
    const uint32_t *ev_base;                      // [R+1]
    const uint32_t *ev_start, *ev_len, *ev_out;   // [E]
    const uint64_t *sel_off;                      // [E_sel+1]
    double *samples;
    uint64_t samples_cap;
};

__global__ void __launch_bounds__(PGM_BLOCK) k_signal(DevOpts o, DevBatch b, SignalIO io) {
    __shared__ uint32_t sm_hist[SIG_SMEM_BINS + 1];
    __shared__ StatsShared sh;
    const uint32_t r = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    if (io.rflags[r] != 0 || !io.read_has_sel[r]) return;
    const pgm_read rd = b.reads[r];
    const ReadCal rc = io.cal[r];
    const uint64_t first = rd.raw_off + rd.trim;
    const uint32_t n = rd.raw_len - rd.trim;
    double med = rc.med, mad = rc.mad;   // wide records: filled by k_stats_wide
    if (o.scaling) {
        StatsCtx c;
        c.nb = (rc.chi >= rc.clo) ? (uint32_t)(rc.chi - rc.clo + 1) : 0u;
        if (c.nb <= SIG_SMEM_BINS) {
            c.clo = rc.clo; c.off = rd.offset; c.scale = rc.scale; c.inc = !(rc.scale < 0.0);
            c.H = sm_hist;
            for (uint32_t i = tid; i <= c.nb; i += PGM_BLOCK) sm_hist[i] = 0;
            __syncthreads();
            stats_histogram_shared(b.raw, first, n, c.clo, c.nb, sm_hist);
            __syncthreads();
            stats_medmad(c, n, &sh, &med, &mad);
        }
    }
    // (x - med) / mad with one reciprocal per record: y = RN(1/mad); q0 = RN(a*y); two residual
    // corrections r = fma(-mad, q, a), q = fma(r, y, q). After the first the quotient is faithful,
    // and Markstein's theorem makes the second one the correctly rounded a/mad (= the reference's
    // IEEE division) unless mad's significand is all ones; that case takes the plain division.
    const double rcp = __drcp_rn(mad);
    const bool fast_div = ((__double_as_longlong(mad) & 0x000FFFFFFFFFFFFFll) != 0x000FFFFFFFFFFFFFll);
    // ---- gather: lanes over 32 events' metadata, then the warp walks each selected event
    const uint32_t lane = tid & 31, w = tid >> 5;
    const uint32_t e0 = io.ev_base[r], e1 = io.ev_base[r + 1];
    const int16_t *x = b.raw + first;
    const double offset = rd.offset, scale = rc.scale;
    const int clo = rc.clo, chi = rc.chi;
    for (uint32_t gb = e0 + w * 32; gb < e1; gb += PGM_WARPS * 32) {
        const uint32_t id = gb + lane;
        uint32_t q = PGM_INVALID, st = 0, ln = 0;
        uint64_t off = 0;
        if (id < e1) {
            q = io.ev_out[id];
            if (q != PGM_INVALID) { st = io.ev_start[id]; ln = io.ev_len[id]; off = io.sel_off[q]; }
        }
        uint32_t live = __ballot_sync(0xffffffffu, q != PGM_INVALID);
        while (live) {
            const int src = __ffs(live) - 1;
            live &= live - 1;
            const uint32_t s0 = __shfl_sync(0xffffffffu, st, src);
            const uint32_t l0 = __shfl_sync(0xffffffffu, ln, src);
            const uint64_t o0 = __shfl_sync(0xffffffffu, off, src);
            if (o0 + l0 > io.samples_cap) continue;
            double *out = io.samples + o0;
            for (uint32_t t = lane; t < l0; t += 32) {
                const int cval = x[s0 + t];
                double v = (cval >= clo && cval <= chi) ? pgm_pa(cval, offset, scale) : 0.0;
                if (o.scaling) {
                    const double a = __dsub_rn(v, med);
                    if (fast_div) {
                        double q = __dmul_rn(a, rcp);
                        q = __fma_rn(__fma_rn(-mad, q, a), rcp, q);
                        v = __fma_rn(__fma_rn(-mad, q, a), rcp, q);
                    } else v = __ddiv_rn(a, mad);
                }
                out[t] = v;
            }
        }
    }
}
