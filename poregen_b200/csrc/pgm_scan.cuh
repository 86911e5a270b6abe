// pgm_scan.cuh — device-wide exclusive prefix sums (three-kernel reduce / top / down-sweep).
// Used for: per-record event totals -> event offsets, per-slot counts -> bucket offsets,
// per-event lengths -> sample offsets (64-bit), radix-sort digit histograms.
#pragma once
#include "pgm_device.cuh"

#define SCAN_ITEMS 16
#define SCAN_TILE (PGM_BLOCK * SCAN_ITEMS)

template <typename T>
__device__ __forceinline__ T warp_incl_scan_add_t(T v) {
    const unsigned lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= (unsigned)d) v += t;
    }
    return v;
}

// exclusive block scan for any additive T; sm: PGM_WARPS+1 elements
template <typename T>
__device__ __forceinline__ T block_excl_scan_add_t(T v, T *sm, T *total) {
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    T inc = warp_incl_scan_add_t<T>(v);
    __syncthreads();
    if (lane == 31) sm[w] = inc;
    __syncthreads();
    if (w == 0) {
        T x = (lane < PGM_WARPS) ? sm[lane] : (T)0;
        T xi = warp_incl_scan_add_t<T>(x);
        if (lane < PGM_WARPS) sm[lane] = xi - x;
        if (lane == PGM_WARPS - 1) sm[PGM_WARPS] = xi;
    }
    __syncthreads();
    *total = sm[PGM_WARPS];
    return sm[w] + inc - v;
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_reduce(const TIn *in, uint64_t n, const uint32_t *n_ptr, TOut *block_sums) {
    __shared__ TOut sm[PGM_WARPS + 1];
    if (n_ptr) n = *n_ptr;  // element count known only on the device; the grid covers the capacity
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    TOut s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        const uint64_t idx = base + (uint64_t)i * PGM_BLOCK + threadIdx.x;
        if (idx < n) s += (TOut)in[idx];
    }
    TOut tot;
    block_excl_scan_add_t<TOut>(s, sm, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// single CTA: in-place exclusive scan of block_sums[nb]; total -> *total_out (may be null)
template <typename TOut>
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_top(TOut *block_sums, uint32_t nb, TOut *total_out) {
    __shared__ TOut sm[PGM_WARPS + 1];
    TOut carry = 0;
    for (uint32_t base = 0; base < nb; base += PGM_BLOCK) {
        const uint32_t idx = base + threadIdx.x;
        TOut v = (idx < nb) ? block_sums[idx] : (TOut)0;
        TOut tot;
        TOut ex = block_excl_scan_add_t<TOut>(v, sm, &tot);
        if (idx < nb) block_sums[idx] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

// out[i] = sum_{j<i} in[j]; out has n+1 entries when write_total (out[n] = total)
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_down(const TIn *in, uint64_t n, const uint32_t *n_ptr, const TOut *block_sums, TOut *out,
                                                         int write_total) {
    __shared__ TOut sm[PGM_WARPS + 1];
    if (n_ptr) n = *n_ptr;
    if (n == 0 && write_total && blockIdx.x == 0 && threadIdx.x == 0) out[0] = 0;
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
    TOut v[SCAN_ITEMS];
    TOut s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        const uint64_t idx = base + i;
        v[i] = (idx < n) ? (TOut)in[idx] : (TOut)0;
        s += v[i];
    }
    TOut tot;
    TOut ex = block_excl_scan_add_t<TOut>(s, sm, &tot) + block_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        const uint64_t idx = base + i;
        if (idx < n) out[idx] = ex;
        ex += v[i];
        if (write_total && idx + 1 == n) out[n] = ex;
    }
}
