// pgm_sort.cuh — stable bucketing of the passing events by slot (k-mer) and the quota selection.
//
// This is the "stable counting-sort bucketing of surviving events by k-mer id" of the north star: the passing
// events are produced in (record, position) order, so a STABLE split by slot leaves every bucket in exactly the
// order the reference would have appended to dump/<kmer> (reference gmove.cpp:650-652, 680-685), and
// rank-in-bucket < remaining quota == "emitted".
//
// An event travels as ONE 16-byte row {slot, start, len, record}: a scatter is one 16-byte store per event, and two
// events of a bucket complete a 32-byte sector (the 4-byte structure-of-arrays stores of round 1 left partial sectors
// that DRAM had to read back: profiles/ncu_r01_count_sort.md).
//
// The decoders leave every record's passing events in a per-record temporary (they cannot know the record's place in
// the batch before every earlier record is done). Nothing copies them into a dense array: the kernels here walk the
// dense (record, position) order in tiles of SORT_TILE events and fetch the rows of a tile straight from the
// temporaries (EvSource: dense index -> record by a search in ev_base[], restricted to the few records of the tile).
//
//   tables up to 2^SORT_MAX_BITS slots (k <= 5):   k_tile_hist -> k_scan_hist -> k_rank                (one digit)
//   larger tables (k = 9: 262,144 slots):          k_tile_hist -> k_scan_hist -> k_msd_scatter -> k_msd_rank
// Two digits, most significant first: k_msd_scatter splits the batch by the high digit (stable); every high-digit
// bucket then holds a CONTIGUOUS range of slots, and k_msd_rank gives a bucket to one CTA, which walks it in order
// with one running counter per slot in shared memory - no second histogram, no second scan, the quota data of the
// bucket's slots (keep, sel_base) loaded once, coalesced, and all writes of a CTA inside one bucket's output range.
// Inside a tile, stability comes from processing items in (warp, round, lane) order with __match_any_sync peer ranks
// and per-warp digit counters.
#pragma once
#include "pgm_device.cuh"
#include "pgm_scan.cuh"

#define SORT_ITEMS 16
#define SORT_TILE (PGM_BLOCK * SORT_ITEMS)
#define SORT_MAX_BITS 10
#define SORT_MAX_BINS (1 << SORT_MAX_BITS)
#define MSD_MAX_LO_BITS 11      // k_msd_rank: slots per high-digit bucket (shared-memory counters)

// ------------------------------------------------------------------------------------------------
// Where the passing events are: record r's events, in position order, at rows_tmp[tmp_base(r) + j], j < n_pass[r];
// dense index of the first one = ev_base[r] (exclusive scan of n_pass, k_scan_records).
struct EvSource {
    const uint4 *rows_tmp;
    const uint32_t *ev_base;     // [R+1]
    const pgm_read *reads;
    uint32_t n_reads;
    uint32_t shift;              // temporaries are indexed by align_off >> shift (ss: two characters per event at least)
};

// smallest r in [lo, hi] with ev_base[r+1] > i (the record that owns dense index i; hi is known to satisfy it)
__device__ __forceinline__ uint32_t ev_find_rec(const uint32_t *ev_base, uint32_t lo, uint32_t hi, uint32_t i) {
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (ev_base[mid + 1] > i) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// The same search by a whole warp (uniform arguments): 32 probes per round.
__device__ __forceinline__ uint32_t ev_find_rec_warp(const uint32_t *ev_base, uint32_t lo, uint32_t hi, uint32_t i) {
    const uint32_t lane = threadIdx.x & 31;
    while (hi - lo >= 32) {
        const uint32_t step = (hi - lo + 32) >> 5;            // probes lo + step-1, lo + 2*step-1, ...
        uint32_t idx = lo + (lane + 1) * step - 1;
        if (idx > hi) idx = hi;
        const uint32_t m = __ballot_sync(0xffffffffu, ev_base[idx + 1] > i);   // monotone false -> true (true at hi)
        const uint32_t L = __ffs(m) - 1;
        uint32_t nhi = lo + (L + 1) * step - 1;
        if (nhi > hi) nhi = hi;
        lo = lo + L * step;
        hi = nhi;
    }
    const uint32_t idx = min(lo + lane, hi);
    const uint32_t m = __ballot_sync(0xffffffffu, ev_base[idx + 1] > i);
    return lo + __ffs(m) - 1;
}

// Records of the tile [t0, t1) (t0 < t1 <= n_ev) -> sm_rec[0] = first, sm_rec[1] = last. Called by all threads; contains a barrier.
__device__ __forceinline__ void ev_tile_records(const EvSource &s, uint32_t t0, uint32_t t1, uint32_t *sm_rec) {
    if (threadIdx.x < 32) {
        const uint32_t r_lo = ev_find_rec_warp(s.ev_base, 0u, s.n_reads - 1, t0);
        const uint32_t r_hi = ev_find_rec_warp(s.ev_base, r_lo, s.n_reads - 1, t1 - 1);
        if (threadIdx.x == 0) { sm_rec[0] = r_lo; sm_rec[1] = r_hi; }
    }
    __syncthreads();
}

__device__ __forceinline__ uint4 ev_fetch(const EvSource &s, uint32_t i, uint32_t r_lo, uint32_t r_hi) {
    const uint32_t r = (r_lo == r_hi) ? r_lo : ev_find_rec(s.ev_base, r_lo, r_hi, i);
    const size_t t = (size_t)(s.reads[r].align_off >> s.shift) + (i - s.ev_base[r]);
    return __ldg(s.rows_tmp + t);
}

// slot of dense event i only (the histogram pass needs nothing else)
__device__ __forceinline__ uint32_t ev_fetch_slot(const EvSource &s, uint32_t i, uint32_t r_lo, uint32_t r_hi) {
    const uint32_t r = (r_lo == r_hi) ? r_lo : ev_find_rec(s.ev_base, r_lo, r_hi, i);
    const size_t t = (size_t)(s.reads[r].align_off >> s.shift) + (i - s.ev_base[r]);
    return __ldg(reinterpret_cast<const uint32_t *>(s.rows_tmp + t));
}

// ------------------------------------------------------------------------------------------------
// k_tile_hist: per-tile histogram of one digit of the slot, hist[digit * ntiles_cap + tile]. Every used tile writes
// all its bins (nothing is zeroed beforehand, nothing is atomic in global memory).
__global__ void __launch_bounds__(PGM_BLOCK) k_tile_hist(EvSource src, const uint32_t *n_ptr, uint32_t shift, uint32_t bits, uint32_t *hist,
                                                         uint32_t ntiles_cap) {
    __shared__ uint32_t sm[1 << MSD_MAX_LO_BITS];
    __shared__ uint32_t sm_rec[2];
    const uint32_t n = *n_ptr;
    const uint32_t bins = 1u << bits, mask = bins - 1;
    const uint32_t used = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) sm[i] = 0;
        const uint32_t base = tile * SORT_TILE, end = min(n, base + SORT_TILE);
        ev_tile_records(src, base, end, sm_rec);
        const uint32_t r_lo = sm_rec[0], r_hi = sm_rec[1];
        uint32_t slot[SORT_ITEMS];
#pragma unroll
        for (int i = 0; i < SORT_ITEMS; i++) {
            const uint32_t idx = base + i * PGM_BLOCK + threadIdx.x;
            slot[i] = (idx < end) ? ev_fetch_slot(src, idx, r_lo, r_hi) : 0u;
        }
#pragma unroll
        for (int i = 0; i < SORT_ITEMS; i++) {
            const uint32_t idx = base + i * PGM_BLOCK + threadIdx.x;
            if (idx < end) atomicAdd(&sm[(slot[i] >> shift) & mask], 1u);
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) hist[(size_t)i * ntiles_cap + tile] = sm[i];
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
// Stable rank of every item of a tile inside its digit: the tile's items are taken in (warp, round, lane) order,
// 512-thread CTAs (a tile's 4096 events are 8 rounds per warp; the rounds of a warp are serial: every one reads the
// counters the previous one wrote). On return wcount[w][d] holds the items of digit d in warps < w (exclusive
// over warps) and lrank[rd] the item's rank among the items of its digit in its own warp.
#define RANK_BLOCK 512
#define RANK_WARPS (RANK_BLOCK / 32)
#define RANK_ITEMS (SORT_TILE / RANK_BLOCK)

template <typename Counter>
__device__ __forceinline__ void tile_rank(const uint32_t (&dig)[RANK_ITEMS], uint32_t nvalid_warp, uint32_t bins, Counter *wcount /* [RANK_WARPS][bins] */,
                                          uint32_t (&lrank)[RANK_ITEMS]) {
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    Counter *mine = wcount + (size_t)w * bins;
    uint32_t peers[RANK_ITEMS];
    // the matches of all rounds first (they depend on nothing but the digits), then the serial counter updates
#pragma unroll
    for (int rd = 0; rd < RANK_ITEMS; rd++) {
        const bool valid = (uint32_t)(rd * 32) + lane < nvalid_warp;
        peers[rd] = __match_any_sync(0xffffffffu, valid ? dig[rd] : 0xffffffffu);   // invalid lanes form their own group
    }
#pragma unroll
    for (int rd = 0; rd < RANK_ITEMS; rd++) {
        const bool valid = (uint32_t)(rd * 32) + lane < nvalid_warp;
        const uint32_t leader = __ffs(peers[rd]) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = mine[dig[rd]];
            mine[dig[rd]] = (Counter)(old + __popc(peers[rd]));
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        lrank[rd] = old + __popc(peers[rd] & ((1u << lane) - 1));
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// k_rank (tables of one digit): the digit pass fused with the quota selection. For every passing event it computes
// the position the stable sort would give it (so its rank inside its slot), decides "emitted" =
// rank < keep[slot] (reference gmove.cpp:650-652, 680-685) and writes the event's row straight to its
// place in the slot-grouped output; no sorted intermediate is materialised.
struct RankIO {
    const uint32_t *n_events;
    EvSource src;
    const uint32_t *hist_scanned;           // scanned tile histogram of the digit
    uint32_t ntiles_cap, bits;
    const uint32_t *bucket_base, *keep, *sel_base;
    uint4 *sel;                             // [E_sel] selected events grouped by slot, as SelRow
    const unsigned long long *rec_src0;     // [R] index in raw[] of the record's sample 0 (k_read_prep)
    uint8_t *read_has_sel;                  // [R] or null (only the Med-MAD pass needs it)
};

// A selected event as the gather wants it: {src lo, src hi, len, record}, src = index in raw[] of its first sample.
__device__ __forceinline__ uint4 sel_row(const uint4 &row, const unsigned long long *rec_src0) {
    const unsigned long long src = rec_src0[row.w] + row.y;
    return make_uint4((uint32_t)src, (uint32_t)(src >> 32), row.z, row.w);
}

#define RANK_SMEM ((RANK_WARPS + 1) * SORT_MAX_BINS * 4)
__global__ void __launch_bounds__(RANK_BLOCK, 2) k_rank(RankIO io) {
    extern __shared__ __align__(16) uint32_t sm_rank[];
    __shared__ uint32_t sm_rec[2];
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.bits, mask = bins - 1;
    uint32_t *wcount = sm_rank;                         // [RANK_WARPS][bins]
    uint32_t *goff = sm_rank + RANK_WARPS * bins;       // [bins]
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t used_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used_tiles; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < RANK_WARPS * bins; i += RANK_BLOCK) wcount[i] = 0;
        for (uint32_t i = threadIdx.x; i < bins; i += RANK_BLOCK) goff[i] = io.hist_scanned[(size_t)i * io.ntiles_cap + tile];
        const uint32_t base = tile * SORT_TILE, end = min(n, base + SORT_TILE);
        ev_tile_records(io.src, base, end, sm_rec);
        const uint32_t r_lo = sm_rec[0], r_hi = sm_rec[1];
        const uint32_t wbase = base + w * (32 * RANK_ITEMS);
        const uint32_t nvw = (end > wbase) ? min(end - wbase, (uint32_t)(32 * RANK_ITEMS)) : 0u;
        uint4 row[RANK_ITEMS];
        uint32_t dig[RANK_ITEMS], lrank[RANK_ITEMS];
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            row[rd] = (idx < end) ? ev_fetch(io.src, idx, r_lo, r_hi) : make_uint4(0, 0, 0, 0);
            dig[rd] = row[rd].x & mask;
        }
        tile_rank<uint32_t>(dig, nvw, bins, wcount, lrank);
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < bins; b += RANK_BLOCK) {   // exclusive scan over warps, per digit
            uint32_t run = 0;
#pragma unroll
            for (int ww = 0; ww < RANK_WARPS; ww++) {
                const uint32_t t = wcount[ww * bins + b];
                wcount[ww * bins + b] = run;
                run += t;
            }
        }
        __syncthreads();
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            if (idx < end) {
                const uint32_t slot = row[rd].x;
                const uint32_t pos = goff[dig[rd]] + wcount[w * bins + dig[rd]] + lrank[rd];
                const uint32_t rank = pos - io.bucket_base[slot];
                if (rank < io.keep[slot]) {
                    io.sel[io.sel_base[slot] + rank] = sel_row(row[rd], io.rec_src0);
                    if (io.read_has_sel) io.read_has_sel[row[rd].w] = 1;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_msd_scatter (tables of two digits, first pass): stable split of the batch by the HIGH digit of the slot.
// rows_out[pos], pos = (events of lower high digits) + (events of the same high digit in earlier tiles) + rank in tile.
struct MsdScatterIO {
    const uint32_t *n_events;
    EvSource src;
    const uint32_t *hist_scanned;
    uint32_t ntiles_cap, shift, bits;
    uint4 *rows_out;
};

__global__ void __launch_bounds__(RANK_BLOCK, 2) k_msd_scatter(MsdScatterIO io) {
    extern __shared__ __align__(16) uint32_t sm_rank[];
    __shared__ uint32_t sm_rec[2];
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.bits;
    uint32_t *wcount = sm_rank;
    uint32_t *goff = sm_rank + RANK_WARPS * bins;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t used_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used_tiles; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < RANK_WARPS * bins; i += RANK_BLOCK) wcount[i] = 0;
        for (uint32_t i = threadIdx.x; i < bins; i += RANK_BLOCK) goff[i] = io.hist_scanned[(size_t)i * io.ntiles_cap + tile];
        const uint32_t base = tile * SORT_TILE, end = min(n, base + SORT_TILE);
        ev_tile_records(io.src, base, end, sm_rec);
        const uint32_t r_lo = sm_rec[0], r_hi = sm_rec[1];
        const uint32_t wbase = base + w * (32 * RANK_ITEMS);
        const uint32_t nvw = (end > wbase) ? min(end - wbase, (uint32_t)(32 * RANK_ITEMS)) : 0u;
        uint4 row[RANK_ITEMS];
        uint32_t dig[RANK_ITEMS], lrank[RANK_ITEMS];
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            row[rd] = (idx < end) ? ev_fetch(io.src, idx, r_lo, r_hi) : make_uint4(0, 0, 0, 0);
            dig[rd] = (row[rd].x >> io.shift) & (bins - 1);
        }
        tile_rank<uint32_t>(dig, nvw, bins, wcount, lrank);
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < bins; b += RANK_BLOCK) {
            uint32_t run = goff[b];
#pragma unroll
            for (int ww = 0; ww < RANK_WARPS; ww++) {
                const uint32_t t = wcount[ww * bins + b];
                wcount[ww * bins + b] = run;
                run += t;
            }
        }
        __syncthreads();
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            if (idx < end) io.rows_out[wcount[w * bins + dig[rd]] + lrank[rd]] = row[rd];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_msd_rank (second pass): one CTA per high-digit bucket at a time. The bucket is the contiguous range
// [start(h), start(h+1)) of rows_in (start = the scanned histogram's entry of tile 0) and holds the slots
// h << lo_bits ... in their (record, position) order. The CTA walks it tile by tile with one running counter per
// slot: rank = events of the slot seen so far; rank < keep[slot] -> the row goes to sel[sel_base[slot] + rank].
struct MsdRankIO {
    const uint32_t *n_events;
    const uint4 *rows_in;
    const uint32_t *hist_scanned;           // first pass's scanned histogram: [hi_bins][ntiles_cap]
    uint32_t ntiles_cap, hi_bits, lo_bits, n_slots;
    const uint32_t *keep, *sel_base;
    uint4 *sel;
    const unsigned long long *rec_src0;
    uint8_t *read_has_sel;                  // or null
    uint32_t *ticket;                       // bucket hand-out (zeroed per batch)
};

__global__ void __launch_bounds__(RANK_BLOCK, 2) k_msd_rank(MsdRankIO io) {
    extern __shared__ __align__(16) uint32_t sm_rank[];
    __shared__ uint32_t sm_bucket;
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.lo_bits, mask = bins - 1;
    uint32_t *wcount = sm_rank;                          // [RANK_WARPS][bins]
    uint32_t *cnt = sm_rank + RANK_WARPS * bins;         // [bins] events of the slot so far
    uint32_t *keep_s = cnt + bins, *selb_s = keep_s + bins;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t n_buckets = (io.n_slots + bins - 1) >> io.lo_bits;
    if (n == 0) return;                                  // nothing was scanned: the histogram holds no bucket starts
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sm_bucket = atomicAdd(io.ticket, 1u);
        __syncthreads();
        const uint32_t h = sm_bucket;
        if (h >= n_buckets) break;
        const uint32_t start = io.hist_scanned[(size_t)h * io.ntiles_cap];
        const uint32_t stop = (h + 1 < (1u << io.hi_bits)) ? io.hist_scanned[(size_t)(h + 1) * io.ntiles_cap] : n;
        if (start >= stop) continue;
        const uint32_t slot0 = h << io.lo_bits;
        for (uint32_t i = threadIdx.x; i < bins; i += RANK_BLOCK) {
            const uint32_t s = slot0 + i;
            cnt[i] = 0;
            keep_s[i] = (s < io.n_slots) ? io.keep[s] : 0u;
            selb_s[i] = (s < io.n_slots) ? io.sel_base[s] : 0u;
        }
        for (uint32_t base = start; base < stop; base += SORT_TILE) {
            __syncthreads();
            for (uint32_t i = threadIdx.x; i < RANK_WARPS * bins; i += RANK_BLOCK) wcount[i] = 0;
            __syncthreads();
            const uint32_t end = min(stop, base + SORT_TILE);
            const uint32_t wbase = base + w * (32 * RANK_ITEMS);
            const uint32_t nvw = (end > wbase) ? min(end - wbase, (uint32_t)(32 * RANK_ITEMS)) : 0u;
            uint4 row[RANK_ITEMS];
            uint32_t dig[RANK_ITEMS], lrank[RANK_ITEMS];
#pragma unroll
            for (int rd = 0; rd < RANK_ITEMS; rd++) {
                const uint32_t idx = wbase + rd * 32 + lane;
                row[rd] = (idx < end) ? __ldg(io.rows_in + idx) : make_uint4(0, 0, 0, 0);
                dig[rd] = row[rd].x & mask;
            }
            tile_rank<uint32_t>(dig, nvw, bins, wcount, lrank);
            __syncthreads();
            for (uint32_t b = threadIdx.x; b < bins; b += RANK_BLOCK) {   // running count of the slot, then exclusive over warps
                uint32_t run = cnt[b];
#pragma unroll
                for (int ww = 0; ww < RANK_WARPS; ww++) {
                    const uint32_t t = wcount[ww * bins + b];
                    wcount[ww * bins + b] = run;
                    run += t;
                }
                cnt[b] = run;
            }
            __syncthreads();
#pragma unroll
            for (int rd = 0; rd < RANK_ITEMS; rd++) {
                const uint32_t idx = wbase + rd * 32 + lane;
                if (idx < end) {
                    const uint32_t d = dig[rd];
                    const uint32_t rank = wcount[w * bins + d] + lrank[rd];
                    if (rank < keep_s[d]) {
                        io.sel[selb_s[d] + rank] = sel_row(row[rd], io.rec_src0);
                        if (io.read_has_sel) io.read_has_sel[row[rd].w] = 1;
                    }
                }
            }
        }
    }
}
#define MSD_RANK_SMEM(lo_bits) ((size_t)(RANK_WARPS + 3) * (1u << (lo_bits)) * 4)

// Selected rows -> the structure-of-arrays form of pgm_result (ev_start / ev_len / ev_read), for callers that download them.
__global__ void __launch_bounds__(PGM_BLOCK) k_unpack_sel(const uint4 *sel, const uint32_t *n_sel_ptr, const unsigned long long *rec_src0, uint32_t *ev_start,
                                                          uint32_t *ev_len, uint32_t *ev_read) {
    const uint32_t n = *n_sel_ptr;
    for (uint32_t i = blockIdx.x * PGM_BLOCK + threadIdx.x; i < n; i += gridDim.x * PGM_BLOCK) {
        const uint4 r = sel[i];
        const unsigned long long src = ((unsigned long long)r.y << 32) | r.x;
        ev_start[i] = (uint32_t)(src - rec_src0[r.w]); ev_len[i] = r.z; ev_read[i] = r.w;
    }
}
