// pgm_scan.cuh — device-wide exclusive prefix sums in one launch (decoupled look-back).
// Used for: per-record event totals -> event offsets, per-slot counts -> bucket offsets and quota,
// per-event lengths -> sample / text offsets (64-bit), radix-sort digit histograms.
#pragma once
#include "pgm_device.cuh"


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

// ------------------------------------------------------------------------------------------------
// Single-pass exclusive scan with decoupled look-back.
// Tiles are handed out by an atomic ticket, so a tile only ever waits on tiles that already run.
// One 64-bit status word per tile, value << 2 | state (1: tile aggregate, 2: inclusive prefix), read
// and written whole, so no fence is needed between "flag" and "value". Values are limited to 62 bits.
// The status words and the ticket are zeroed before the launch (they live in the per-batch zero region).
#define LB_ITEMS 16
#define LB_TILE (PGM_BLOCK * LB_ITEMS)

struct LbBuf {
    unsigned long long *status;  // [tiles]
    uint32_t *ticket;
};

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p) { return *reinterpret_cast<const volatile unsigned long long *>(p); }
__device__ __forceinline__ void st_volatile_u64(unsigned long long *p, unsigned long long v) { *reinterpret_cast<volatile unsigned long long *>(p) = v; }

template <typename T>
__device__ __forceinline__ T warp_sum_t(T v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// The look-back step on its own, for kernels that bring their own tile aggregate (k_gather): called by the 32 lanes of
// ONE warp with uniform arguments. Publishes the aggregate of `tile`, sums the tiles in front of it (waiting for each
// one's aggregate or inclusive prefix), publishes the inclusive prefix, returns the exclusive one in every lane.
__device__ __forceinline__ unsigned long long lb_lookback(LbBuf st, uint32_t tile, unsigned long long tot) {
    const uint32_t lane = threadIdx.x & 31;
    unsigned long long excl = 0;
    if (tile == 0) {
        if (lane == 0) st_volatile_u64(st.status, (tot << 2) | 2ull);
        return 0ull;
    }
    if (lane == 0) st_volatile_u64(st.status + tile, (tot << 2) | 1ull);
    int64_t pos = (int64_t)tile - 1;
    for (;;) {
        const int64_t idx = pos - (int64_t)lane;
        unsigned long long w = 2ull;   // before tile 0: prefix 0
        if (idx >= 0) {
            do { w = ld_volatile_u64(st.status + idx); } while ((w & 3ull) == 0);
        }
        const uint32_t m2 = __ballot_sync(0xffffffffu, (w & 3ull) == 2ull);
        const unsigned long long val = w >> 2;
        if (m2) {
            const uint32_t L = __ffs(m2) - 1;
            excl += warp_sum_t<unsigned long long>(lane <= L ? val : 0ull);
            break;
        }
        excl += warp_sum_t<unsigned long long>(val);
        pos -= 32;
    }
    if (lane == 0) st_volatile_u64(st.status + tile, ((excl + tot) << 2) | 2ull);
    return excl;
}

// load(idx) -> T, store(idx, exclusive_prefix, own_value), fin(total) (one thread of the last tile).
template <typename T, typename LoadF, typename StoreF, typename FinalF>
__device__ __forceinline__ void lb_scan(uint64_t n, LbBuf st, LoadF load, StoreF store, FinalF fin) {
    __shared__ T sm_scan[PGM_WARPS + 1];
    __shared__ T sm_prefix;
    __shared__ uint32_t sm_tile;
    if (threadIdx.x == 0) sm_tile = atomicAdd(st.ticket, 1u);
    __syncthreads();
    const uint32_t tile = sm_tile;
    const uint64_t ntiles = (n + LB_TILE - 1) / LB_TILE;
    if (tile >= ntiles) {
        if (n == 0 && tile == 0 && threadIdx.x == 0) fin((T)0);
        return;
    }
    const uint64_t base = (uint64_t)tile * LB_TILE + (uint64_t)threadIdx.x * LB_ITEMS;
    T v[LB_ITEMS];
    T s = 0;
#pragma unroll
    for (int i = 0; i < LB_ITEMS; i++) {
        const uint64_t idx = base + i;
        v[i] = (idx < n) ? load(idx) : (T)0;
        s += v[i];
    }
    T tot;
    const T ex = block_excl_scan_add_t<T>(s, sm_scan, &tot);
    if (threadIdx.x < 32) {
        const unsigned long long excl = lb_lookback(st, tile, (unsigned long long)tot);
        if (threadIdx.x == 0) sm_prefix = (T)excl;
    }
    __syncthreads();
    T run = sm_prefix + ex;
#pragma unroll
    for (int i = 0; i < LB_ITEMS; i++) {
        const uint64_t idx = base + i;
        if (idx < n) store(idx, run, v[i]);
        run += v[i];
    }
    if (tile == ntiles - 1 && threadIdx.x == 0) fin(sm_prefix + tot);
}
