// pgm_model.cuh — K7: the per-k-mer model row of reference scripts/poregen.sh:54-85
// (calculate_mean_stddev_all): over all values of dump/<KMER> except the first one (`tail -n +2`),
// `datamash median` (mean of the two middle values for an even count) and `datamash sstdev`
// (sample standard deviation, n-1), the latter clamped to a limit.
//
// The script works on the TEXT of the dump, so its inputs are the "%.8f" decimals, not the doubles:
// every value is N * 1e-8 with N the integer the formatter prints (pgm_text.cuh). That makes the row an
// exact integer problem: the median is a selection on int64 keys, and
//     sum (x - mean)^2 = (m * sum N^2 - (sum N)^2) / m * 1e-16
// is evaluated exactly in 128-bit integers, with one rounding when it becomes a double.
// Values accumulate across batches in a per-context log (the dump files accumulate across the
// reference's record loop in the same way); pgm_model_finalize groups them by slot and runs k_model_rows.
#pragma once
#include "pgm_text.cuh"

#define PGM_MODEL_BAD_KEY 0x7fffffffffffffffll   // not representable: |x| * 1e8 >= 2^63, inf, nan

// doubles -> the integers their "%.8f" text spells (sign applied). Grid-stride.
__global__ void __launch_bounds__(PGM_BLOCK) k_model_scale(const double *x, const unsigned long long *n_ptr, long long *out, uint32_t *bad) {
    const unsigned long long n = *n_ptr;
    for (unsigned long long i = (unsigned long long)blockIdx.x * PGM_BLOCK + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * PGM_BLOCK) {
        const Fmt8 f = fmt8_prepare(x[i]);
        long long v = PGM_MODEL_BAD_KEY;
        if (f.big == 0 && f.ip < 92233720368ull) {
            v = (long long)(f.ip * 100000000ull + f.fp);
            if (f.neg) v = -v;
        } else atomicOr(bad, 1u);
        out[i] = v;
    }
}

// Copies every slot's range of one batch behind what earlier batches put into the grouped array.
// One CTA per slot (grid-stride over slots).
__global__ void __launch_bounds__(PGM_BLOCK) k_model_group(const long long *vals, const unsigned long long *slot_off, const unsigned long long *dst_off,
                                                           uint32_t n_slots, long long *grouped) {
    for (uint32_t s = blockIdx.x; s < n_slots; s += gridDim.x) {
        const unsigned long long a = slot_off[s], n = slot_off[s + 1] - a, d = dst_off[s];
        for (unsigned long long i = threadIdx.x; i < n; i += PGM_BLOCK) grouped[d + i] = vals[a + i];
    }
}

// Dwell medians (reference scripts/poregen.sh:33-52, calculate_dwell_times_medians): per dump file, the comma count of every
// ';'-separated field — event length - 1, plus a 0 for the empty field behind the last ';' — and `datamash median` over them.
// The counts of one batch's slot range go behind two leading zeros (the first is the value k_model_rows drops, the second the
// empty field) and what earlier batches put there; k_model_rows then selects the median exactly like for the level values.
__global__ void __launch_bounds__(PGM_BLOCK) k_dwell_group(const uint32_t *sel_len, const uint32_t *ev_off, const unsigned long long *dst_off,
                                                           uint32_t n_slots, long long *grouped) {
    for (uint32_t s = blockIdx.x; s < n_slots; s += gridDim.x) {
        const uint32_t a = ev_off[s], n = ev_off[s + 1] - a;
        const unsigned long long d = dst_off[s];
        for (uint32_t i = threadIdx.x; i < n; i += PGM_BLOCK) grouped[d + i] = (long long)sel_len[a + i] - 1;
    }
}
__global__ void __launch_bounds__(PGM_BLOCK) k_dwell_heads(const unsigned long long *off, uint32_t n_slots, long long *grouped) {
    const uint32_t s = blockIdx.x * PGM_BLOCK + threadIdx.x;
    if (s < n_slots) { grouped[off[s]] = 0; grouped[off[s] + 1] = 0; }
}

struct ModelRowDev {     // mirrors pgm_model_row (include/pgm_gmove.h)
    double median, stddev;
    long long median_num;          // median = median_num / 2e8, exactly
    unsigned long long n;          // values used (the slot's values minus the first)
    uint32_t flags, pad;
};
#define PGM_MODEL_F_EMPTY 1u       // fewer than two values in the file: datamash prints nothing
#define PGM_MODEL_F_NO_SD 2u       // one value after the drop: sstdev is nan
#define PGM_MODEL_F_RANGE 4u       // a value of the slot is not representable as int64 * 1e-8
#define PGM_MODEL_F_CLAMPED 8u     // stddev was above the limit

__device__ __forceinline__ unsigned long long model_key(long long v) { return (unsigned long long)v ^ 0x8000000000000000ull; }

// k-th smallest (0-based) of keys[0, m) by MSD radix selection, 8 bits per pass, whole CTA.
// Returns the key; *less = number of keys strictly below it. `hist` needs 256 words, `sh` 4 words.
__device__ unsigned long long block_select(const long long *v, unsigned long long m, unsigned long long k, uint32_t *hist, unsigned long long *sh,
                                           unsigned long long *less) {
    unsigned long long prefix = 0, pmask = 0, below = 0;
    for (int shift = 56; shift >= 0; shift -= 8) {
        for (uint32_t i = threadIdx.x; i < 256; i += PGM_BLOCK) hist[i] = 0;
        __syncthreads();
        for (unsigned long long i = threadIdx.x; i < m; i += PGM_BLOCK) {
            const unsigned long long key = model_key(v[i]);
            if ((key & pmask) == prefix) atomicAdd(&hist[(uint32_t)(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long run = below;
            uint32_t b = 0;
            for (; b < 255; b++) {
                if (run + hist[b] > k) break;
                run += hist[b];
            }
            sh[0] = b; sh[1] = run;
        }
        __syncthreads();
        prefix |= sh[0] << shift;
        pmask |= 0xffull << shift;
        below = sh[1];
        __syncthreads();
    }
    *less = below;
    return prefix;
}

__device__ __forceinline__ void u128_add(unsigned long long &hi, unsigned long long &lo, unsigned long long ah, unsigned long long al) {
    const unsigned long long l = lo + al;
    hi += ah + (l < lo ? 1ull : 0ull);
    lo = l;
}

// One CTA per slot. grouped[off[s] .. off[s+1]) are the slot's values in file order.
__global__ void __launch_bounds__(PGM_BLOCK) k_model_rows(const long long *grouped, const unsigned long long *off, uint32_t n_slots, double limit,
                                                          ModelRowDev *rows) {
    __shared__ uint32_t hist[256];
    __shared__ unsigned long long sh[4];
    __shared__ unsigned long long red[PGM_WARPS][5];
    for (uint32_t s = blockIdx.x; s < n_slots; s += gridDim.x) {
        __syncthreads();
        const unsigned long long a = off[s], cnt = off[s + 1] - a;
        ModelRowDev row;
        row.median = CUDART_NAN; row.stddev = CUDART_NAN; row.median_num = 0; row.n = cnt ? cnt - 1 : 0; row.flags = 0; row.pad = 0;
        if (cnt < 2) {
            row.flags = PGM_MODEL_F_EMPTY;
            if (threadIdx.x == 0) rows[s] = row;
            continue;
        }
        const long long *v = grouped + a + 1;      // `tail -n +2`
        const unsigned long long m = cnt - 1;
        // ---- exact sums: S1 = sum N (signed 128), S2 = sum N^2 (unsigned 128); any unrepresentable value?
        unsigned long long s1h = 0, s1l = 0, s2h = 0, s2l = 0, bad = 0;
        for (unsigned long long i = threadIdx.x; i < m; i += PGM_BLOCK) {
            const long long x = v[i];
            if (x == PGM_MODEL_BAD_KEY) { bad = 1; continue; }
            u128_add(s1h, s1l, x < 0 ? ~0ull : 0ull, (unsigned long long)x);
            const unsigned long long ax = (unsigned long long)(x < 0 ? -x : x);
            u128_add(s2h, s2l, __umul64hi(ax, ax), ax * ax);
        }
        // warp then block reduction of the five words (128-bit adds)
        for (int d = 16; d; d >>= 1) {
            const unsigned long long oh = __shfl_xor_sync(0xffffffffu, s1h, d), ol = __shfl_xor_sync(0xffffffffu, s1l, d);
            const unsigned long long ph = __shfl_xor_sync(0xffffffffu, s2h, d), pl = __shfl_xor_sync(0xffffffffu, s2l, d);
            u128_add(s1h, s1l, oh, ol);
            u128_add(s2h, s2l, ph, pl);
            bad |= __shfl_xor_sync(0xffffffffu, bad, d);
        }
        if ((threadIdx.x & 31) == 0) {
            unsigned long long *r = red[threadIdx.x >> 5];
            r[0] = s1h; r[1] = s1l; r[2] = s2h; r[3] = s2l; r[4] = bad;
        }
        __syncthreads();
        s1h = s1l = s2h = s2l = bad = 0;
        for (int w = 0; w < PGM_WARPS; w++) {
            u128_add(s1h, s1l, red[w][0], red[w][1]);
            u128_add(s2h, s2l, red[w][2], red[w][3]);
            bad |= red[w][4];
        }
        if (bad) {
            row.flags = PGM_MODEL_F_RANGE;
            if (threadIdx.x == 0) rows[s] = row;
            continue;
        }
        // ---- median: ranks (m-1)/2 and m/2 (equal for an odd count)
        const unsigned long long k = m / 2;
        unsigned long long less;
        const unsigned long long key_hi = block_select(v, m, k, hist, sh, &less);
        long long hi_v = (long long)(key_hi ^ 0x8000000000000000ull), lo_v = hi_v;
        if ((m & 1) == 0 && less == k) {
            // rank k-1 is the largest value below the k-th
            long long best = (long long)0x8000000000000000ull;
            for (unsigned long long i = threadIdx.x; i < m; i += PGM_BLOCK) {
                const long long x = v[i];
                if (x < hi_v && x > best) best = x;
            }
            for (int d = 16; d; d >>= 1) {
                const long long o = __shfl_xor_sync(0xffffffffu, best, d);
                best = o > best ? o : best;
            }
            __syncthreads();
            if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][0] = (unsigned long long)best;
            __syncthreads();
            best = (long long)red[0][0];
            for (int w = 1; w < PGM_WARPS; w++) {
                const long long o = (long long)red[w][0];
                best = o > best ? o : best;
            }
            lo_v = best;
        }
        if (threadIdx.x == 0) {
            row.median_num = lo_v + hi_v;
            row.median = (double)row.median_num / 2e8;
            if (m < 2) row.flags |= PGM_MODEL_F_NO_SD;
            else {
                // D = m * S2 - S1^2 >= 0, exactly
                const __int128 S1 = (__int128)(((unsigned __int128)s1h << 64) | s1l);
                const unsigned __int128 S2 = ((unsigned __int128)s2h << 64) | s2l;
                const unsigned __int128 a1 = (unsigned __int128)(S1 < 0 ? -S1 : S1);
                // m * S2 and S1^2 can exceed 128 bits only for astronomically large inputs; both are computed in
                // double then, which costs the exactness but not the order of magnitude
                const double S2d = (double)S2, a1d = (double)a1, md = (double)m;
                double Dd;
                if (S2d * md < 1.5e38 && a1d * a1d < 1.5e38) {
                    const unsigned __int128 D = (unsigned __int128)m * S2 - a1 * a1;
                    Dd = (double)D;
                } else Dd = md * S2d - a1d * a1d;
                const double var = Dd / (md * (md - 1.0));
                double sd = sqrt(var) * 1e-8;
                if (sd > limit) { sd = limit; row.flags |= PGM_MODEL_F_CLAMPED; }   // poregen.sh:77-79
                row.stddev = sd;
            }
            rows[s] = row;
        }
    }
}
