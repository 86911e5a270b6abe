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
#define SIG_THREADS 128         // k_signal CTA: many small CTAs per SM overlap each other's serial phases
#define SIG_WARPS (SIG_THREADS / 32)
#define SIG_CH 4096             // samples per TMA chunk (8 KB)
#define SIG_NST 3               // ring stages; records of up to SIG_NST*SIG_CH samples stay resident for the gather
#define SIG_EV_TILE SIG_THREADS
#define PGM_FULL 0xffffffffu

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
// cumulative count of in-range samples with value index <= x (x may be negative or >= nb)
__device__ __forceinline__ uint32_t st_cumAt(const StatsCtx &c, int64_t x, uint32_t n_in) {
    if (x < 0) return 0u;
    if (x >= (int64_t)c.nb) return n_in;
    return c.H[st_bin(c, (uint32_t)x)];
}
// first v in [lo, hi) for which pred(v) holds (pred monotone false->true), else hi. One thread.
template <typename P>
__device__ __forceinline__ uint32_t st_first(uint32_t lo, uint32_t hi, P pred) {
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (pred(mid)) hi = mid; else lo = mid + 1;
    }
    return lo;
}
// The same search by a whole warp (uniform lo/hi): 32 probes per round.
template <typename P>
__device__ __forceinline__ uint32_t warp_first(uint32_t lo, uint32_t hi, P pred) {
    const uint32_t lane = threadIdx.x & 31;
    while (hi - lo > 32) {
        const uint32_t step = (hi - lo + 31) >> 5;
        uint32_t idx = lo + (lane + 1) * step - 1;
        if (idx >= hi) idx = hi - 1;
        const uint32_t m = __ballot_sync(PGM_FULL, pred(idx));
        if (!m) return hi;
        const uint32_t L = __ffs(m) - 1;
        uint32_t nhi = lo + (L + 1) * step;
        if (nhi > hi) nhi = hi;
        lo = lo + L * step;
        hi = nhi;
    }
    const uint32_t idx = lo + lane;
    const uint32_t m = __ballot_sync(PGM_FULL, idx < hi && pred(idx));
    return m ? lo + __ffs(m) - 1 : hi;
}

__device__ __forceinline__ double warp_min_f64(double v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v = fmin(v, __shfl_xor_sync(PGM_FULL, v, d));
    return v;
}
__device__ __forceinline__ double warp_max_f64(double v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v = fmax(v, __shfl_xor_sync(PGM_FULL, v, d));
    return v;
}

// Histogram of the in-range codes of raw[first, first+n) into H[0, nb) (zeroed by the caller), with
// generic atomics (shared or global H). Used by k_stats_wide only.
__device__ __forceinline__ void stats_histogram(const int16_t *raw, uint64_t first, uint32_t n, int clo, uint32_t nb, uint32_t *H) {
    const uint32_t tid = threadIdx.x;
    const uint64_t end = first + n;
    const uint64_t body0 = (first + 7) & ~(uint64_t)7, body1 = end & ~(uint64_t)7;
    auto add = [&](int code) {
        const uint32_t u = (uint32_t)(code - clo);
        if (u < nb) atomicAdd(&H[u], 1u);
    };
    if (body0 < body1) {
        for (uint64_t i = first + tid; i < body0; i += blockDim.x) add(raw[i]);
        const int4 *v4 = reinterpret_cast<const int4 *>(raw + body0);
        const uint32_t ng = (uint32_t)((body1 - body0) >> 3);
        for (uint32_t g = tid; g < ng; g += blockDim.x) {
            const int4 q = __ldg(v4 + g);
            add((int)(int16_t)(q.x & 0xffff)); add(q.x >> 16);
            add((int)(int16_t)(q.y & 0xffff)); add(q.y >> 16);
            add((int)(int16_t)(q.z & 0xffff)); add(q.z >> 16);
            add((int)(int16_t)(q.w & 0xffff)); add(q.w >> 16);
        }
        for (uint64_t i = body1 + tid; i < end; i += blockDim.x) add(raw[i]);
    } else {
        for (uint64_t i = first + tid; i < end; i += blockDim.x) add(raw[i]);
    }
}

__device__ __forceinline__ void red_shared_inc(uint32_t saddr) { asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(saddr) : "memory"); }

// ------------------------------------------------------------------------------------------------
// warp_medmad: executed by ONE full warp. Turns the histogram in c.H into inclusive cumulative counts
// (value order, in place), then computes med and the clamped mad exactly as the reference does.
//
// MAD = the n/2-th smallest deviation (0-based) = the smallest sample deviation d whose count
// N(d) = #{i : |x_i - med| <= d} reaches n/2 + 1. Deviations are monotone in the bin distance on
// either side of the median bin, so N(d) is two cumulative-count lookups at the bins where each side
// crosses d. Fast path: an integer search finds the bin radius rho at which the count is reached;
// the deviations of a 16-bin window per side around rho are evaluated once (one lane each) and N(d)
// of the bins at distance rho-2..rho+2 (plus the zero mass) is read off with ballots. The result is
// taken only when every value outside that candidate set is provably below a failing candidate or
// not below the answer; anything else (odd calibrations, heavy zero mass) ranks every bin instead.
__device__ __forceinline__ void warp_medmad(const StatsCtx &c, uint32_t n, bool force_slow, double *med_out, double *mad_out) {
    const uint32_t lane = threadIdx.x & 31;
    // ---- histogram -> inclusive cumulative counts, in value order, in place (odd per-lane span: no bank conflicts)
    const uint32_t per = ((c.nb + 31) >> 5) | 1u;
    const uint32_t v0 = lane * per;
    uint32_t s = 0;
    for (uint32_t j = 0; j < per; j++) {
        const uint32_t v = v0 + j;
        if (v < c.nb) s += c.H[st_bin(c, v)];
    }
    uint32_t incl = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(PGM_FULL, incl, d);
        if (lane >= (uint32_t)d) incl += t;
    }
    const uint32_t n_in = __shfl_sync(PGM_FULL, incl, 31);
    uint32_t run = incl - s;
    for (uint32_t j = 0; j < per; j++) {
        const uint32_t v = v0 + j;
        if (v < c.nb) { run += c.H[st_bin(c, v)]; c.H[st_bin(c, v)] = run; }
    }
    __syncwarp();
    const uint32_t nz = n - n_in;          // samples replaced by 0.0
    const uint32_t k = n / 2;              // ksmall(n, n/2)
    // ---- where 0.0 sorts among the in-range values: vz = first v with value >= 0.0
    uint32_t vz;
    if (c.nb == 0 || st_val(c, 0) >= 0.0) vz = 0;                       // the usual case: pa_min > 0
    else if (st_val(c, c.nb - 1) < 0.0) vz = c.nb;
    else vz = warp_first(0u, c.nb, [&](uint32_t v) { return st_val(c, v) >= 0.0; });
    const uint32_t a = st_cumE(c, vz);     // in-range samples with value < 0.0
    const bool med_is_zero = (k >= a && k < a + nz);
    const uint32_t kk = (k < a) ? k : k - nz;
    // ---- upper median: the bin whose cumulative range holds rank kk
    uint32_t vmed = vz;
    if (!med_is_zero) {
        vmed = warp_first(0u, c.nb, [&](uint32_t v) { return st_cumI(c, v) > kk; });
        if (vmed >= c.nb) vmed = c.nb ? c.nb - 1 : 0;   // cannot happen (kk < n_in)
    }
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
        bool done = false;
        if (!force_slow && n_in >= k + 1) {
            const uint32_t rho = warp_first(0u, c.nb + 1, [&](uint32_t q) {
                return st_cumAt(c, (int64_t)vm + q, n_in) - st_cumAt(c, (int64_t)vm - (int64_t)q - 1, n_in) >= k + 1;
            });
            const bool isR = lane < 16;
            const uint32_t baseR = (rho > 8) ? rho - 8 : 0u, baseL = (rho > 8) ? rho - 8 : 1u;
            const uint32_t dist = (isR ? baseR : baseL) + (lane & 15);
            const int64_t vv = isR ? (int64_t)vm + dist : (int64_t)vm - (int64_t)dist;
            const bool inb = vv >= 0 && vv < (int64_t)c.nb;
            const double dv = inb ? dev((uint32_t)vv) : CUDART_INF;
            const uint32_t wv = inb ? st_cumI(c, (uint32_t)vv) - st_cumE(c, (uint32_t)vv) : 0u;
            const bool cand = inb && wv > 0 && dist + 2 >= rho && dist <= rho + 2;
            const uint32_t inbm = __ballot_sync(PGM_FULL, inb);
            const uint32_t inbR = __popc(inbm & 0xffffu), inbL = __popc(inbm >> 16);
            const bool closerR = baseR > 0 && vm < c.nb, closerL = baseL > 1 && vm >= 1;
            const bool beyondR = (uint64_t)vm + baseR + 16 < c.nb, beyondL = (uint64_t)vm >= (uint64_t)baseL + 16;
            // N(d) with its exactness flag; uniform across the warp
            auto evalN = [&](double d, bool *det) {
                const uint32_t mm = __ballot_sync(PGM_FULL, inb && dv <= d);
                const uint32_t nR = __popc(mm & 0xffffu), nL = __popc(mm >> 16);
                bool ok = true;
                uint32_t cR = 0, cL = 0;
                if (nR == 0) { if (closerR) ok = false; }
                else { cR = st_cumI(c, vm + baseR + nR - 1) - cm; if (nR == inbR && beyondR) ok = false; }
                if (nL == 0) { if (closerL) ok = false; }
                else { cL = cm - st_cumE(c, vm - (baseL + nL - 1)); if (nL == inbL && beyondL) ok = false; }
                *det = ok;
                return ((d0 <= d) ? nz : 0u) + cR + cL;
            };
            bool my_det = false;
            uint32_t my_N = 0;
            const uint32_t candm = __ballot_sync(PGM_FULL, cand);
            for (uint32_t m = candm; m; m &= m - 1) {
                const int src = __ffs(m) - 1;
                const double d = __shfl_sync(PGM_FULL, dv, src);
                bool det;
                const uint32_t N = evalN(d, &det);
                if ((int)lane == src) { my_det = det; my_N = N; }
            }
            double av = (cand && my_det && my_N >= k + 1) ? dv : CUDART_INF;     // smallest passing candidate
            double bv = (cand && my_det && my_N < k + 1) ? dv : -1.0;            // largest failing candidate
            bool det0 = true;
            uint32_t N0 = 0;
            if (nz) N0 = evalN(d0, &det0);
            double a_ = warp_min_f64(av), b_ = warp_max_f64(bv);
            if (nz && det0) { if (N0 >= k + 1) a_ = fmin(a_, d0); else b_ = fmax(b_, d0); }
            bool bad = cand && !my_det && dv > b_ && dv < a_;                    // an unranked candidate between the two
            bad |= inb && dist + 2 < rho && dv > b_;                             // closer bins must all fail
            bad |= inb && dist > rho + 2 && dv < a_;                             // farther bins must not undercut
            bool badu = (nz && !det0 && d0 > b_ && d0 < a_);
            badu |= (closerR && inbR == 0) || (closerL && inbL == 0);
            if (!__any_sync(PGM_FULL, bad) && !badu && a_ < CUDART_INF) { mad_raw = a_; done = true; }
        }
        if (!done) {   // rank every bin: is the n/2-th smallest deviation the one of bin v? order on ties: zero mass < left < right
            auto cntR_lt = [&](double t) { return st_cumE(c, st_first(vm, c.nb, [&](uint32_t v) { return dev(v) >= t; })) - cm; };
            auto cntL_lt = [&](double t) { return cm - st_cumE(c, st_first(0u, vm, [&](uint32_t v) { return dev(v) < t; })); };
            auto cntL_le = [&](double t) { return cm - st_cumE(c, st_first(0u, vm, [&](uint32_t v) { return dev(v) <= t; })); };
            double res = 0.0;
            bool found = false;
            if (lane == 0 && nz) {
                const uint32_t rs = cntL_lt(d0) + cntR_lt(d0);
                if (rs <= k && k < rs + nz) { res = d0; found = true; }
            }
            for (uint32_t v = lane; v < c.nb; v += 32) {
                const uint32_t ci = st_cumI(c, v), ce = st_cumE(c, v);
                const uint32_t w = ci - ce;
                if (!w) continue;
                const double d = dev(v);
                uint32_t rs = (d0 <= d) ? nz : 0u;
                if (v < vm) rs += (cm - ci) + cntR_lt(d);
                else rs += (ce - cm) + cntL_le(d);
                if (rs <= k && k < rs + w) { res = d; found = true; }
            }
            __syncwarp();
            const uint32_t fm = __ballot_sync(PGM_FULL, found);
            mad_raw = fm ? __shfl_sync(PGM_FULL, res, __ffs(fm) - 1) : 0.0;
        }
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
        c.H = (c.nb <= STATS_SMEM_BINS) ? sm_hist : gscratch + (size_t)blockIdx.x * 65536;
        for (uint32_t i = tid; i < c.nb; i += PGM_BLOCK) c.H[i] = 0;
        __syncthreads();
        stats_histogram(b.raw, first, n, c.clo, c.nb, c.H);
        __threadfence_block();
        __syncthreads();
        if (tid < 32) {
            double med, mad;
            warp_medmad(c, n, (o.debug & 1u) != 0, &med, &mad);
            if (tid == 0) { cal[r].med = med; cal[r].mad = mad; }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// TMA 1-D bulk copy + mbarrier (sm_90+/sm_100a PTX): one elected thread streams a record's signal
// from HBM into a shared-memory ring; the copy engine signals a transaction barrier per chunk.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------------------------------------
// k_signal: one CTA per record that owns at least one selected event.
//   1. (scaling) the record's signal is streamed by TMA bulk copies into a shared-memory ring and
//      histogrammed from there (in-range codes); warp 0 then derives the exact med / mad. Records of
//      up to SIG_NST * SIG_CH samples stay resident in the ring for step 2.
//   2. gather: for every selected event of the record, raw -> pA (0.0 when out of range)
//      -> (x - med) / mad (gmove.cpp:611-613) -> the event's place in the slot-grouped output
//      (the doubles gmove.cpp:664-678 prints). Selected events are compacted into shared memory
//      (start, length, output offset) and every warp keeps four events in flight.
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

struct SelEv { uint32_t st, ln; uint64_t off; };

struct SigConv {   // raw code -> the double the reference prints
    int clo; uint32_t nbins;
    double offset, scale, med, mad, rcp;
    bool scaling, fast_div;
    __device__ __forceinline__ double operator()(int c) const {
        double v = ((uint32_t)(c - clo) < nbins) ? pgm_pa(c, offset, scale) : 0.0;
        if (scaling) {
            // (x - med) / mad with one reciprocal per record: y = RN(1/mad); q0 = RN(a*y); two residual
            // corrections r = fma(-mad, q, a), q = fma(r, y, q). After the first the quotient is faithful,
            // and Markstein's theorem makes the second one the correctly rounded a/mad (= the reference's
            // IEEE division) unless mad's significand is all ones; that case takes the plain division.
            const double a = __dsub_rn(v, med);
            if (fast_div) {
                double q = __dmul_rn(a, rcp);
                q = __fma_rn(__fma_rn(-mad, q, a), rcp, q);
                v = __fma_rn(__fma_rn(-mad, q, a), rcp, q);
            } else v = __ddiv_rn(a, mad);
        }
        return v;
    }
};

// Every warp walks the compacted selected events four at a time (independent load -> convert -> store
// chains); x is the record's signal relative to its first (trimmed) sample, in shared or global memory.
__device__ __forceinline__ void sig_gather(const SelEv *sel, uint32_t ns, const int16_t *__restrict__ x, const SigConv &cv,
                                           double *__restrict__ samples) {
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t base = w * 4; base < ns; base += SIG_WARPS * 4) {
        SelEv m[4];
        int c[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t e = base + u;
            m[u] = sel[min(e, ns - 1)];
            if (e >= ns) m[u].ln = 0;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) c[u] = (lane < m[u].ln) ? (int)x[m[u].st + lane] : 0;
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (lane < m[u].ln) samples[m[u].off + lane] = cv(c[u]);
        if (max(max(m[0].ln, m[1].ln), max(m[2].ln, m[3].ln)) > 32) {
#pragma unroll
            for (int u = 0; u < 4; u++)
                for (uint32_t t = lane + 32; t < m[u].ln; t += 32) samples[m[u].off + t] = cv((int)x[m[u].st + t]);
        }
    }
}

__global__ void __launch_bounds__(SIG_THREADS) k_signal(DevOpts o, DevBatch b, SignalIO io) {
    __shared__ __align__(128) int16_t sm_ring[SIG_NST * SIG_CH];
    __shared__ __align__(8) uint64_t sm_full[SIG_NST];
    __shared__ __align__(16) uint32_t sm_hist[SIG_SMEM_BINS + 4];   // reused for the selected-event list after the statistics
    __shared__ double sm_mm[2];
    __shared__ uint32_t sm_nsel;
    SelEv *sm_sel = reinterpret_cast<SelEv *>(sm_hist);
    static_assert(SIG_EV_TILE * sizeof(SelEv) <= sizeof(uint32_t) * (SIG_SMEM_BINS + 4), "event list must fit the histogram area");
    const uint32_t r = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    if (io.rflags[r] != 0 || !io.read_has_sel[r]) return;
    const pgm_read *rp = b.reads + r;
    const uint64_t first = rp->raw_off + rp->trim;
    const uint32_t n = rp->raw_len - rp->trim;
    const double offset = rp->offset;
    const ReadCal rc = io.cal[r];
    const uint32_t e0 = io.ev_base[r], e1 = io.ev_base[r + 1];
    // metadata of the first tile of events: requested now, used after the statistics
    uint32_t pq = PGM_INVALID, pst = 0, pln = 0;
    uint64_t poff = 0;
    if (e0 + tid < e1) {
        pq = io.ev_out[e0 + tid];
        if (pq != PGM_INVALID) { pst = io.ev_start[e0 + tid]; pln = io.ev_len[e0 + tid]; poff = io.sel_off[pq]; }
    }
    if (tid == 0) sm_nsel = 0;
    SigConv cv;
    cv.clo = rc.clo; cv.nbins = (rc.chi >= rc.clo) ? (uint32_t)(rc.chi - rc.clo + 1) : 0u;
    cv.offset = offset; cv.scale = rc.scale; cv.med = rc.med; cv.mad = rc.mad;   // wide records: filled by k_stats_wide
    cv.scaling = o.scaling != 0;
    bool resident = false;
    uint32_t head = 0;   // samples between the 16-byte aligned start of the ring image and the record's first sample
    if (o.scaling && cv.nbins <= SIG_SMEM_BINS) {
        const uint32_t nb = cv.nbins;
        const int clo = rc.clo;
        const uint64_t endp = first + n;
        const uint64_t a0 = first & ~(uint64_t)7;
        uint64_t a1 = (endp + 7) & ~(uint64_t)7;
        if (a1 > (b.n_raw & ~(uint64_t)7)) a1 = b.n_raw & ~(uint64_t)7;   // never read past the array: the last few samples go scalar
        if (a1 < a0) a1 = a0;
        const uint32_t nal = (uint32_t)(a1 - a0);
        const uint32_t nch = (nal + SIG_CH - 1) / SIG_CH;
        head = (uint32_t)(first - a0);
        const uint32_t rel_end = head + n;       // record = [head, rel_end) relative to a0
        if (tid == 0) {
            for (int s = 0; s < SIG_NST; s++) mbar_init(&sm_full[s], 1);
            mbar_fence_init();
        }
        for (uint32_t i = tid; i <= nb; i += SIG_THREADS) sm_hist[i] = 0;
        __syncthreads();
        auto issue = [&](uint32_t j) {
            const uint32_t s = j % SIG_NST;
            const uint32_t len = min((uint32_t)SIG_CH, nal - j * SIG_CH);
            mbar_expect_tx(&sm_full[s], len * 2);
            tma_bulk_g2s(sm_ring + s * SIG_CH, b.raw + a0 + (uint64_t)j * SIG_CH, len * 2, &sm_full[s]);
        };
        if (tid == 0)
            for (uint32_t j = 0; j < nch && j < SIG_NST; j++) issue(j);
        const uint32_t sbase = smem_u32(sm_hist);
        // out-of-range codes are clamped into the extra bin H[nb] (never read by the order statistics)
        auto add = [&](int code) { red_shared_inc(sbase + 4u * min((uint32_t)(code - clo), nb)); };
        {   // samples the aligned image does not cover (only at the very end of raw[])
            const uint64_t t0 = (a1 > first) ? a1 : first;
            for (uint64_t i = t0 + tid; i < endp; i += SIG_THREADS) add(b.raw[i]);
        }
        for (uint32_t j = 0; j < nch; j++) {
            const uint32_t s = j % SIG_NST;
            while (!mbar_try_wait(&sm_full[s], (j / SIG_NST) & 1)) {}
            const uint32_t rel0 = j * SIG_CH;
            const uint32_t ng = min((uint32_t)SIG_CH, nal - rel0) >> 3;
            const int4 *v4 = reinterpret_cast<const int4 *>(sm_ring + s * SIG_CH);
            for (uint32_t g = tid; g < ng; g += SIG_THREADS) {
                const int4 q = v4[g];
                const uint32_t i0 = rel0 + 8u * g;
                const int v[8] = {(int)(int16_t)(q.x & 0xffff), q.x >> 16, (int)(int16_t)(q.y & 0xffff), q.y >> 16,
                                  (int)(int16_t)(q.z & 0xffff), q.z >> 16, (int)(int16_t)(q.w & 0xffff), q.w >> 16};
                if (i0 >= head && i0 + 8 <= rel_end) {
#pragma unroll
                    for (int t = 0; t < 8; t++) add(v[t]);
                } else {
#pragma unroll
                    for (int t = 0; t < 8; t++)
                        if (i0 + t >= head && i0 + t < rel_end) add(v[t]);
                }
            }
            if (j + SIG_NST < nch) {   // ring slot is reused: everybody is done reading it
                __syncthreads();
                if (tid == 0) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue(j + SIG_NST);
                }
            }
        }
        resident = (nch <= SIG_NST) && (a1 >= endp);
        __syncthreads();
        if (tid < 32) {
            StatsCtx c;
            c.nb = nb; c.clo = clo; c.off = offset; c.scale = rc.scale; c.inc = !(rc.scale < 0.0);
            c.H = sm_hist;
            double med, mad;
            warp_medmad(c, n, (o.debug & 1u) != 0, &med, &mad);
            if (tid == 0) { sm_mm[0] = med; sm_mm[1] = mad; }
        }
        __syncthreads();
        cv.med = sm_mm[0]; cv.mad = sm_mm[1];
    }
    cv.rcp = __drcp_rn(cv.mad);
    cv.fast_div = ((__double_as_longlong(cv.mad) & 0x000FFFFFFFFFFFFFll) != 0x000FFFFFFFFFFFFFll);
    const int16_t *xg = b.raw + first;
    const int16_t *xs = sm_ring + head;
    for (uint32_t t0 = e0; t0 < e1; t0 += SIG_EV_TILE) {
        if (t0 != e0) {
            __syncthreads();   // previous tile gathered
            if (tid == 0) sm_nsel = 0;
            pq = PGM_INVALID;
            if (t0 + tid < e1) {
                pq = io.ev_out[t0 + tid];
                if (pq != PGM_INVALID) { pst = io.ev_start[t0 + tid]; pln = io.ev_len[t0 + tid]; poff = io.sel_off[pq]; }
            }
            __syncthreads();
        }
        if (pq != PGM_INVALID && poff + pln <= io.samples_cap) {
            const uint32_t p = atomicAdd(&sm_nsel, 1u);
            SelEv e; e.st = pst; e.ln = pln; e.off = poff;
            sm_sel[p] = e;
        }
        __syncthreads();
        const uint32_t ns = sm_nsel;
        if (resident) sig_gather(sm_sel, ns, xs, cv, io.samples);
        else sig_gather(sm_sel, ns, xg, cv, io.samples);
    }
}
