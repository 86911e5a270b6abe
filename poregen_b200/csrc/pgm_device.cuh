// pgm_device.cuh — device-side types and small helpers shared by the gmove kernels (sm_100a).
//
// All arithmetic on signal values is IEEE double with explicit round-to-nearest intrinsics
// (__dadd_rn / __dmul_rn / __ddiv_rn) so nothing is contracted into an FMA: the reference is
// plain x86-64 SSE2 code (no FMA), see SURVEY.md F2.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include "../../include/pgm_gmove.h"

#define PGM_INVALID 0xFFFFFFFFu
#define PGM_BLOCK 256
#define PGM_WARPS (PGM_BLOCK / 32)

// Options as the kernels see them.
struct DevOpts {
    uint32_t k, m, s_off, margin, limit, max_dur, min_dur, pick_margin;
    uint32_t scaling, n_slots, index_lo, mode;
    uint32_t use_lut;       // 1: slot = lut[id] (kmer_list), 0: slot = id - index_lo
    uint32_t fourth;        // 'T' or 'U': the 4th letter of the k-mer table's alphabet
    uint32_t debug;         // bit 0: Med-MAD always takes the rank-every-bin path (env PGM_DEBUG_SLOW_STATS, tests)
    uint32_t reserved;
    double pa_min, pa_max;
    const uint32_t *lut;    // [4^k] when use_lut
};

// Device view of a batch.
struct DevBatch {
    uint32_t n_reads, mode;
    uint64_t n_raw, n_align, n_bases;
    const int16_t *raw;
    const uint8_t *align;
    const uint8_t *bases;
    const pgm_read *reads;
};

// Per-read derived calibration (k_read_prep) and Med-MAD result (k_stats).
struct __align__(64) ReadCal {   // 64 bytes: what a conversion without Med-MAD needs sits in ONE 32-byte sector (k_gather reads it per event)
    double scale;       // range / digitisation
    double offset;      // pgm_read.offset (copied: the record table itself is a second random access away)
    int32_t clo, chi;   // int16 codes c with clo <= c <= chi are inside [pa_min, pa_max]; clo>chi: none
    int32_t pad0, pad1;
    double med, mad;    // only meaningful with scaling and when the read has selected events
    double pad2, pad3;
};

// ---------------------------------------------------------------- pA and the in-range interval
// TO_PICOAMPS (reference src/poregen.h:30): ((raw)+(offset))*((range)/(digitisation))
// (double)c for any int32 without the conversion unit: 2^52 + 2^31 + c is exactly representable with the mantissa
// bits 0x80000000 ^ c, and subtracting the constant back is exact. I2F.F64 shares a narrow pipe; DADD does not.
__device__ __forceinline__ double pgm_i2d(int c) {
    return __dsub_rn(__hiloint2double(0x43300000, (int)((uint32_t)c ^ 0x80000000u)), 4503601774854144.0);
}
__device__ __forceinline__ double pgm_pa(int c, double offset, double scale) {
    return __dmul_rn(__dadd_rn(pgm_i2d(c), offset), scale);
}

// pA(c) is monotone in c (each fp op is monotone), non-decreasing for scale>=0 and non-increasing
// for scale<0, so {c : !(pA<pa_min || pA>pa_max)} (reference gmove.cpp:594) is an interval of codes.
// Returns it as [clo, chi] (empty when clo > chi).
__device__ inline void pgm_code_interval(double offset, double scale, double pa_min, double pa_max, int *clo, int *chi) {
    const bool inc = !(scale < 0.0);
    // lower bound test: pA >= pa_min ; upper bound test: pA <= pa_max
    // inc: codes >= a satisfy lower, codes <= b satisfy upper.  dec: codes <= a' satisfy lower, codes >= b' satisfy upper.
    int lo = -32768, hi = 32768;  // first code (in code order) for which the "becomes true" predicate holds
    // P1(c): inc ? pA(c) >= pa_min : pA(c) <= pa_max   (monotone false->true as c grows)
    while (lo < hi) {
        int mid = lo + ((hi - lo) >> 1);
        double v = pgm_pa(mid, offset, scale);
        bool p = inc ? (v >= pa_min) : (v <= pa_max);
        if (p) hi = mid; else lo = mid + 1;
    }
    const int first = lo;
    // P2(c): inc ? pA(c) <= pa_max : pA(c) >= pa_min   (monotone true->false as c grows); find last true
    lo = -32769; hi = 32767;
    while (lo < hi) {
        int mid = lo + ((hi - lo + 1) >> 1);
        double v = pgm_pa(mid, offset, scale);
        bool p = inc ? (v <= pa_max) : (v >= pa_min);
        if (p) lo = mid; else hi = mid - 1;
    }
    *clo = first;
    *chi = lo;
}

// ---------------------------------------------------------------- k-mer codes
// 2-bit code of a base as the k-mer table sees it, or 0xFF. `rna_rec`: record-level T->U
// replacement (reference gmove.cpp:815-817).
__device__ __forceinline__ uint32_t pgm_base_code(uint32_t c, uint32_t fourth, bool rna_rec) {
    if (rna_rec && c == 'T') c = 'U';
    if (c == 'A') return 0;
    if (c == 'C') return 1;
    if (c == 'G') return 2;
    if (c == fourth) return 3;
    return 0xFF;
}

__device__ __forceinline__ uint32_t pgm_slot_of(const DevOpts &o, uint32_t id) {
    if (o.use_lut) return o.lut[id];
    uint32_t s = id - o.index_lo;
    return (s < o.n_slots) ? s : PGM_INVALID;
}

// ---------------------------------------------------------------- block-wide scans (PGM_BLOCK threads)
__device__ __forceinline__ uint32_t warp_incl_scan_add(uint32_t v) {
    const unsigned lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= (unsigned)d) v += t;
    }
    return v;
}
__device__ __forceinline__ int warp_incl_scan_max(int v) {
    const unsigned lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= (unsigned)d) v = max(v, t);
    }
    return v;
}

// Exclusive sum over the block; *total = block sum. `sm` needs PGM_WARPS+1 words. Contains barriers.
__device__ __forceinline__ uint32_t block_excl_scan_add(uint32_t v, uint32_t *sm, uint32_t *total) {
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan_add(v);
    __syncthreads();
    if (lane == 31) sm[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t x = (lane < PGM_WARPS) ? sm[lane] : 0;
        uint32_t xi = warp_incl_scan_add(x);
        if (lane < PGM_WARPS) sm[lane] = xi - x;
        if (lane == PGM_WARPS - 1) sm[PGM_WARPS] = xi;
    }
    __syncthreads();
    *total = sm[PGM_WARPS];
    return sm[w] + inc - v;
}

// Exclusive max over the block (identity -1); *total = block max. `sm` needs PGM_WARPS+1 ints.
__device__ __forceinline__ int block_excl_scan_max(int v, int *sm, int *total) {
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = warp_incl_scan_max(v);
    int exc = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) exc = -1;
    __syncthreads();
    if (lane == 31) sm[w] = inc;
    __syncthreads();
    if (w == 0) {
        int x = (lane < PGM_WARPS) ? sm[lane] : -1;
        int xi = warp_incl_scan_max(x);
        int xe = __shfl_up_sync(0xffffffffu, xi, 1);
        if (lane == 0) xe = -1;
        if (lane < PGM_WARPS) sm[lane] = xe;
        if (lane == PGM_WARPS - 1) sm[PGM_WARPS] = xi;
    }
    __syncthreads();
    *total = sm[PGM_WARPS];
    return max(sm[w], exc);
}
