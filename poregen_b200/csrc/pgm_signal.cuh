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

// From the histogram in c.H (all threads of the block call this together): turns it into cumulative
// counts, then computes med and the clamped mad exactly as the reference does. sm_u: PGM_WARPS+1
// words, sm_res: one double, both shared. Returns the values to every thread.
__device__ __forceinline__ void stats_medmad(const StatsCtx &c, uint32_t n, uint32_t *sm_u, double *sm_res, double *med_out, double *mad_out) {
    const uint32_t tid = threadIdx.x;
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
            *sm_res = 0.0;
            if (nz) {
                const uint32_t rs = cntL_lt(d0) + cntR_lt(d0);
                if (rs <= k && k < rs + nz) *sm_res = d0;
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
            if (rs <= k && k < rs + w) *sm_res = d;
        }
        __syncthreads();
        mad_raw = *sm_res;
    }
    double mad = __dmul_rn(mad_raw, 1.4826);       // gmove.cpp:162,183
    mad = (mad > 1.0) ? mad : 1.0;                 // gmove.cpp:609
    *med_out = med;
    *mad_out = mad;
}

// ------------------------------------------------------------------------------------------------
// k_stats_wide: fallback for records whose in-range code span exceeds SIG_SMEM_BINS (unusual pa_min /
// pa_max or calibration). Persistent CTAs pull records from an atomic queue; histogram in shared memory
// up to STATS_SMEM_BINS bins, else in a per-CTA 65536-bin global scratch. Writes med/mad to cal[].
__global__ void __launch_bounds__(PGM_BLOCK) k_stats_wide(DevOpts o, DevBatch b, const uint8_t *rflags, const uint8_t *read_has_sel,
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
        stats_medmad(c, n, sm_u, &sm_res, &med, &mad);
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
    __shared__ uint32_t sm_hist[SIG_SMEM_BINS];
    __shared__ uint32_t sm_u[PGM_WARPS + 1];
    __shared__ double sm_res;
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
            for (uint32_t i = tid; i < c.nb; i += PGM_BLOCK) sm_hist[i] = 0;
            __syncthreads();
            stats_histogram<true>(b.raw, first, n, c.clo, c.nb, sm_hist);
            __syncthreads();
            stats_medmad(c, n, sm_u, &sm_res, &med, &mad);
        }
    }
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
                if (o.scaling) v = __ddiv_rn(__dsub_rn(v, med), mad);
                out[t] = v;
            }
        }
    }
}
