// pgm_sort.cuh — stable LSD radix sort of (slot key, event index) pairs.
//
// This is the "stable counting-sort bucketing of surviving events by k-mer id" of the north star:
// the passing events are produced in (record, position) order, so a STABLE sort by slot leaves
// every bucket in exactly the order the reference would have appended to dump/<kmer>
// (reference gmove.cpp:650-652, 680-685), and rank-in-bucket < remaining quota == "emitted".
//
// Per digit pass: per-tile digit histogram -> device-wide exclusive scan (digit-major) -> stable
// scatter. Inside a tile, stability comes from processing items in (warp, round, lane) order with
// __match_any_sync peer ranks and per-warp digit counters.
#pragma once
#include "pgm_device.cuh"
#include "pgm_scan.cuh"

#define SORT_ITEMS 16
#define SORT_TILE (PGM_BLOCK * SORT_ITEMS)
#define SORT_MAX_BITS 10
#define SORT_MAX_BINS (1 << SORT_MAX_BITS)

__global__ void __launch_bounds__(PGM_BLOCK) k_sort_hist(const uint32_t *keys, const uint32_t *n_ptr, uint32_t shift, uint32_t bits,
                                                         uint32_t *hist, uint32_t ntiles) {
    __shared__ uint32_t sm[SORT_MAX_BINS];
    const uint32_t n = *n_ptr;
    const uint32_t bins = 1u << bits, mask = bins - 1;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) sm[i] = 0;
        __syncthreads();
        const uint32_t base = tile * SORT_TILE;
        if (base < n) {
#pragma unroll
            for (int i = 0; i < SORT_ITEMS; i++) {
                const uint32_t idx = base + i * PGM_BLOCK + threadIdx.x;
                if (idx < n) atomicAdd(&sm[(keys[idx] >> shift) & mask], 1u);
            }
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) hist[(size_t)i * ntiles + tile] = sm[i];
    }
}

// Event rows carried along with the keys (two-pass sorts): the first pass reads them in event order and writes them in
// pass order, so the last pass (k_rank) reads them sequentially instead of gathering three 4-byte words per event from
// arrays far larger than L2 (C3: 7.9 GB of DRAM reads for 25 M events).
struct SortPayload {
    const uint32_t *s_in, *l_in, *r_in;
    uint32_t *s_out, *l_out, *r_out;      // null: keys (and vals) only
};

// vals_in == nullptr: the value of item i is i (first pass). vals_out == nullptr: the values are not written.
__global__ void __launch_bounds__(PGM_BLOCK) k_sort_scatter(const uint32_t *keys_in, const uint32_t *vals_in, uint32_t *keys_out,
                                                            uint32_t *vals_out, const uint32_t *n_ptr, uint32_t shift, uint32_t bits,
                                                            const uint32_t *hist_scanned, uint32_t ntiles, SortPayload pl) {
    __shared__ uint32_t wcount[PGM_WARPS][SORT_MAX_BINS];
    __shared__ uint32_t goff[SORT_MAX_BINS];
    const uint32_t n = *n_ptr;
    const uint32_t bins = 1u << bits, mask = bins - 1;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t used_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used_tiles; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < PGM_WARPS * SORT_MAX_BINS; i += PGM_BLOCK) (&wcount[0][0])[i] = 0;
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) goff[i] = hist_scanned[(size_t)i * ntiles + tile];
        __syncthreads();
        const uint32_t wbase = tile * SORT_TILE + w * (32 * SORT_ITEMS);
        uint32_t key[SORT_ITEMS], val[SORT_ITEMS], lrank[SORT_ITEMS];
#pragma unroll
        for (int rd = 0; rd < SORT_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            const bool valid = idx < n;
            key[rd] = valid ? keys_in[idx] : 0;
            val[rd] = valid ? (vals_in ? vals_in[idx] : idx) : 0;
            const uint32_t d = valid ? ((key[rd] >> shift) & mask) : bins;  // invalid lanes form their own group
            const uint32_t peers = __match_any_sync(0xffffffffu, d);
            const uint32_t leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if (valid && lane == leader) {
                old = wcount[w][d];
                wcount[w][d] = old + __popc(peers);
            }
            old = __shfl_sync(0xffffffffu, old, leader);
            lrank[rd] = old + __popc(peers & ((1u << lane) - 1));
            __syncwarp();
        }
        __syncthreads();
        // exclusive scan over warps, per digit
        for (uint32_t b = threadIdx.x; b < bins; b += PGM_BLOCK) {
            uint32_t run = 0;
#pragma unroll
            for (int ww = 0; ww < PGM_WARPS; ww++) {
                const uint32_t t = wcount[ww][b];
                wcount[ww][b] = run;
                run += t;
            }
        }
        __syncthreads();
#pragma unroll
        for (int rd = 0; rd < SORT_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            if (idx < n) {
                const uint32_t d = (key[rd] >> shift) & mask;
                const uint32_t pos = goff[d] + wcount[w][d] + lrank[rd];
                keys_out[pos] = key[rd];
                if (vals_out) vals_out[pos] = val[rd];
                if (pl.s_out) { pl.s_out[pos] = pl.s_in[val[rd]]; pl.l_out[pos] = pl.l_in[val[rd]]; pl.r_out[pos] = pl.r_in[val[rd]]; }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// One-launch scan of the digit-major tile histogram hist[bin * ntiles_cap + tile] over the tiles in
// use (n_events known on the device only), in place.
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_hist(uint32_t *hist, const uint32_t *n_ptr, uint32_t bins, uint32_t ntiles_cap,
                                                         LbBuf st) {
    const uint32_t n_ev = *n_ptr;
    const uint32_t ntu = (n_ev + SORT_TILE - 1) / SORT_TILE;
    const uint64_t n = (uint64_t)bins * ntu;
    auto addr = [&](uint64_t e) { return (size_t)(e / ntu) * ntiles_cap + (size_t)(e % ntu); };
    lb_scan<uint32_t>(n, st, [&](uint64_t e) { return hist[addr(e)]; },
                      [&](uint64_t e, uint32_t ex, uint32_t) { hist[addr(e)] = ex; }, [](uint32_t) {});
}

// ------------------------------------------------------------------------------------------------
// k_rank: the LAST radix pass fused with the quota selection. For every passing event it computes
// the position the stable sort would give it (so its rank inside its slot), decides "emitted" =
// rank < keep[slot] (reference gmove.cpp:650-652, 680-685) and writes the event straight to its
// place in the slot-grouped output; no sorted intermediate is materialised.
// vals_in == nullptr: single pass, the value of item i is i.
struct RankIO {
    const uint32_t *n_events;
    const uint32_t *keys_in, *vals_in;      // full slot keys (in the order of the previous pass) and original event ids
    const uint32_t *hist_scanned;           // this pass's scanned tile histogram
    uint32_t ntiles_cap, shift, bits;
    const uint32_t *bucket_base, *keep, *sel_base;
    const uint32_t *ev_start, *ev_len, *ev_read;  // [E] record-then-position order
    const uint32_t *p_start, *p_len, *p_read;     // the same rows in the order of the previous pass (SortPayload), or null
    uint32_t *sel_start, *sel_len, *sel_read;     // [E_sel] grouped by slot
    unsigned long long *sel_src;                  // [E_sel] index in raw[] of the event's first sample
    const pgm_read *reads;
    uint8_t *read_has_sel;                        // [R]
};

// 512-thread CTAs: a tile's 4096 events are 8 rounds per warp instead of 16 (the rounds of a warp are serial: every one
// reads the counters the previous one wrote), and the 830 tiles of C2 make 2.8 waves of half-length tiles instead of 1.4 of full ones.
#define RANK_BLOCK 512
#define RANK_WARPS (RANK_BLOCK / 32)
#define RANK_ITEMS (SORT_TILE / RANK_BLOCK)
#define RANK_SMEM ((RANK_WARPS + 1) * SORT_MAX_BINS * 4)
__global__ void __launch_bounds__(RANK_BLOCK, 2) k_rank(RankIO io) {
    extern __shared__ __align__(16) uint32_t sm_rank[];
    uint32_t(*wcount)[SORT_MAX_BINS] = reinterpret_cast<uint32_t(*)[SORT_MAX_BINS]>(sm_rank);
    uint32_t *goff = sm_rank + RANK_WARPS * SORT_MAX_BINS;
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.bits, mask = bins - 1, shift = io.shift;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t used_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used_tiles; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < RANK_WARPS * bins; i += RANK_BLOCK) wcount[i / bins][i % bins] = 0;
        for (uint32_t i = threadIdx.x; i < bins; i += RANK_BLOCK) goff[i] = io.hist_scanned[(size_t)i * io.ntiles_cap + tile];
        __syncthreads();
        const uint32_t wbase = tile * SORT_TILE + w * (32 * RANK_ITEMS);
        uint32_t key[RANK_ITEMS], lrank[RANK_ITEMS], peers[RANK_ITEMS];
        // the matches of all rounds first (they depend on nothing but the keys), then the serial counter updates
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            key[rd] = (idx < n) ? io.keys_in[idx] : 0;
        }
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const bool valid = wbase + rd * 32 + lane < n;
            const uint32_t d = valid ? ((key[rd] >> shift) & mask) : bins;  // invalid lanes form their own group
            peers[rd] = __match_any_sync(0xffffffffu, d);
        }
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const bool valid = wbase + rd * 32 + lane < n;
            const uint32_t d = (key[rd] >> shift) & mask;
            const uint32_t leader = __ffs(peers[rd]) - 1;
            uint32_t old = 0;
            if (valid && lane == leader) {
                old = wcount[w][d];
                wcount[w][d] = old + __popc(peers[rd]);
            }
            old = __shfl_sync(0xffffffffu, old, leader);
            lrank[rd] = old + __popc(peers[rd] & ((1u << lane) - 1));
            __syncwarp();
        }
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < bins; b += RANK_BLOCK) {   // exclusive scan over warps, per digit
            uint32_t run = 0;
#pragma unroll
            for (int ww = 0; ww < RANK_WARPS; ww++) {
                const uint32_t t = wcount[ww][b];
                wcount[ww][b] = run;
                run += t;
            }
        }
        __syncthreads();
        // where every event of the tile goes (or that it is over its quota), then the selected events' rows: all loads of
        // a level are issued before the first one is used (id -> event row -> record row is a three-deep chain)
        uint32_t qd[RANK_ITEMS], idv[RANK_ITEMS];
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            qd[rd] = PGM_INVALID; idv[rd] = 0;
            if (idx < n) {
                const uint32_t slot = key[rd];
                const uint32_t d = (slot >> shift) & mask;
                const uint32_t pos = goff[d] + wcount[w][d] + lrank[rd];
                const uint32_t rank = pos - io.bucket_base[slot];
                idv[rd] = io.vals_in ? io.vals_in[idx] : idx;
                if (rank < io.keep[slot]) qd[rd] = io.sel_base[slot] + rank;
            }
        }
        uint32_t rrv[RANK_ITEMS], stv[RANK_ITEMS], lnv[RANK_ITEMS];
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            rrv[rd] = stv[rd] = lnv[rd] = 0;
            if (qd[rd] != PGM_INVALID) {
                if (io.p_read) {     // rows arrived with the keys, in this pass's order
                    const uint32_t idx = wbase + rd * 32 + lane;
                    rrv[rd] = io.p_read[idx]; stv[rd] = io.p_start[idx]; lnv[rd] = io.p_len[idx];
                } else { rrv[rd] = io.ev_read[idv[rd]]; stv[rd] = io.ev_start[idv[rd]]; lnv[rd] = io.ev_len[idv[rd]]; }
            }
        }
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            if (qd[rd] != PGM_INVALID) {
                const uint32_t q = qd[rd], rr = rrv[rd];
                io.sel_start[q] = stv[rd];
                io.sel_src[q] = io.reads[rr].raw_off + io.reads[rr].trim + stv[rd];
                io.sel_len[q] = lnv[rd];
                io.sel_read[q] = rr;
                io.read_has_sel[rr] = 1;
            }
        }
    }
}
