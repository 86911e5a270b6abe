// pgm_sort.cuh — stable bucketing of the passing events by slot (k-mer) and the quota selection.
//
// This is the "stable counting-sort bucketing of surviving events by k-mer id" of the north star: the passing
// events are produced in (record, position) order, so a STABLE split by slot leaves every bucket in exactly the
// order the reference would have appended to dump/<kmer> (reference gmove.cpp:650-652, 680-685), and
// rank-in-bucket < remaining quota == "emitted".
//
// An event travels as ONE 16-byte row (EvRow below): a scatter is one 16-byte store per event, and two events of a
// bucket complete a 32-byte sector (the 4-byte structure-of-arrays stores of round 1 left partial sectors that DRAM had
// to read back: profiles/ncu_r01_count_sort.md). The row already holds what the gather needs (the event's first sample
// as an index into raw[], its length, its record), so the selected rows ARE the selection: nothing is looked up again.
//
// The decoders leave every record's passing events in a per-record temporary (they cannot know the record's place in
// the batch before every earlier record is done). Nothing copies them into a dense array: the kernels here walk the
// dense (record, position) order in tiles of SORT_TILE events and fetch the rows of a tile straight from the
// temporaries (EvSource: dense index -> record; the handful of records a tile touches sit in shared memory).
//
//   tables up to 2^SORT_MAX_BITS slots (k <= 5):   k_tile_hist -> k_scan_hist -> k_rank                (one digit)
//   larger tables (k = 9: 262,144 slots):          k_tile_hist -> k_scan_hist -> k_msd_scatter -> k_msd_rank
// Two digits, most significant first: k_msd_scatter splits the batch by the high digit (stable); every high-digit
// bucket then holds a CONTIGUOUS range of slots, and k_msd_rank gives a bucket to one CTA, which walks it in order
// with one running counter per slot in shared memory - no second histogram, no second scan, the quota data of the
// bucket's slots (keep, sel_base) loaded once, coalesced, and all writes of a CTA inside one bucket's output range.
// Inside a tile, stability comes from processing items in (warp, round, lane) order with peer masks (the lanes of a round that hold
// the same digit: ballots over the digit's bits - MATCH.ANY, -DSORT_MATCH_ANY, is 25 % slower for the whole sort stage)
// and per-warp digit counters.
#pragma once
#include "pgm_device.cuh"
#include "pgm_scan.cuh"

#define SORT_ITEMS 16
#define SORT_TILE (PGM_BLOCK * SORT_ITEMS)
#define SORT_MAX_BITS 10
#define SORT_MAX_BINS (1 << SORT_MAX_BITS)
#define MSD_MAX_LO_BITS 10      // slots per high-digit bucket (shared-memory counters beside the 64 KB tile): tables up to 2^20 slots (k = 10)

// ------------------------------------------------------------------------------------------------
// EvRow, the 16 bytes of an event:  x = slot,  y = src bits 0-31,  z = src bits 32-39 << 24 | len,  w = record.
// src = index in raw[] of the event's first sample (after --margin), len = its samples (< 2^24: longer events make the
// decoders flag the batch), record = index of its alignment record in the batch.
#define EV_MAX_LEN 0x00ffffffu
__device__ __forceinline__ uint4 ev_row(uint32_t slot, unsigned long long src, uint32_t len, uint32_t rec) {
    return make_uint4(slot, (uint32_t)src, ((uint32_t)(src >> 32) << 24) | len, rec);
}
__device__ __forceinline__ unsigned long long ev_row_src(const uint4 &r) { return ((unsigned long long)(r.z >> 24) << 32) | r.y; }
__device__ __forceinline__ uint32_t ev_row_len(const uint4 &r) { return r.z & EV_MAX_LEN; }

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

// The records of one tile, in shared memory: a tile of long records touches one or two of them, so dense index ->
// temporary address is a compare against EV_SEGS cached boundaries; tiles that touch more records (short records, or
// batches where the quota leaves a few events per record) search ev_base[] between the tile's first and last record.
#define EV_SEGS 4
struct EvTile {
    uint32_t r_lo, r_hi;
    uint32_t seg_start[EV_SEGS];             // dense index of the first event of record r_lo + k (0xffffffff: not cached)
    long long seg_delta[EV_SEGS];            // temporary address - dense index, for record r_lo + k
};

// Fills *t for the tile [t0, t1) (t0 < t1 <= n_ev). Called by all threads of the CTA; contains a barrier.
__device__ __forceinline__ void ev_tile_setup(const EvSource &s, uint32_t t0, uint32_t t1, EvTile *t) {
    if (threadIdx.x < 32) {
        const uint32_t r_lo = ev_find_rec_warp(s.ev_base, 0u, s.n_reads - 1, t0);
        const uint32_t r_hi = ev_find_rec_warp(s.ev_base, r_lo, s.n_reads - 1, t1 - 1);
        if (threadIdx.x == 0) { t->r_lo = r_lo; t->r_hi = r_hi; }
        if (threadIdx.x < EV_SEGS) {
            const uint32_t r = r_lo + threadIdx.x;
            uint32_t st = 0xffffffffu;
            long long dl = 0;
            if (r <= r_hi) {
                const uint32_t eb = s.ev_base[r];
                st = threadIdx.x ? eb : 0u;          // everything below the second record's start belongs to the first
                dl = (long long)(s.reads[r].align_off >> s.shift) - (long long)eb;
            }
            t->seg_start[threadIdx.x] = st; t->seg_delta[threadIdx.x] = dl;
        }
    }
    __syncthreads();
}

// temporary address of dense event i of the tile
__device__ __forceinline__ size_t ev_addr(const EvSource &s, const EvTile &t, uint32_t i) {
    if (t.r_hi - t.r_lo < EV_SEGS) {
        uint32_t k = 0;
#pragma unroll
        for (int q = 1; q < EV_SEGS; q++) k += (i >= t.seg_start[q]) ? 1u : 0u;     // a record without events shares its successor's start: the successor wins
        return (size_t)((long long)i + t.seg_delta[k]);
    }
    const uint32_t r = ev_find_rec(s.ev_base, t.r_lo, t.r_hi, i);
    return (size_t)(s.reads[r].align_off >> s.shift) + (i - s.ev_base[r]);
}
__device__ __forceinline__ uint4 ev_fetch(const EvSource &s, const EvTile &t, uint32_t i) { return __ldg(s.rows_tmp + ev_addr(s, t, i)); }
// slot of dense event i only (the histogram pass needs nothing else)
__device__ __forceinline__ uint32_t ev_fetch_slot(const EvSource &s, const EvTile &t, uint32_t i) {
    return __ldg(reinterpret_cast<const uint32_t *>(s.rows_tmp + ev_addr(s, t, i)));
}

// ------------------------------------------------------------------------------------------------
// k_tile_hist: per-tile histogram of one digit of the slot, tile-major: hist[tile * bins + digit]. Every used tile
// writes all its bins (nothing is zeroed beforehand, nothing is atomic in global memory).
__global__ void __launch_bounds__(PGM_BLOCK) k_tile_hist(EvSource src, const uint32_t *n_ptr, uint32_t shift, uint32_t bits, uint32_t *hist) {
    __shared__ uint32_t sm[1 << MSD_MAX_LO_BITS];
    __shared__ EvTile sm_tile;
    const uint32_t n = *n_ptr;
    const uint32_t bins = 1u << bits, mask = bins - 1;
    const uint32_t used = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) sm[i] = 0;
        const uint32_t base = tile * SORT_TILE, end = min(n, base + SORT_TILE);
        ev_tile_setup(src, base, end, &sm_tile);
        uint32_t slot[SORT_ITEMS];
#pragma unroll
        for (int i = 0; i < SORT_ITEMS; i++) {
            const uint32_t idx = base + i * PGM_BLOCK + threadIdx.x;
            slot[i] = (idx < end) ? ev_fetch_slot(src, sm_tile, idx) : 0u;
        }
#pragma unroll
        for (int i = 0; i < SORT_ITEMS; i++) {
            const uint32_t idx = base + i * PGM_BLOCK + threadIdx.x;
            if (idx < end) atomicAdd(&sm[(slot[i] >> shift) & mask], 1u);
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < bins; i += PGM_BLOCK) hist[(size_t)tile * bins + i] = sm[i];
    }
}

// ------------------------------------------------------------------------------------------------
// One-launch scan of the tile histogram in digit-major ORDER (all tiles of digit 0, then digit 1, ...) over the tiles in
// use (n_events known on the device only), in place. The storage is tile-major (hist[tile * bins + digit]): the tile
// kernels read and write their bins coalesced; this kernel moves 4 bytes per 32-byte sector, on a table of a few megabytes.
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_hist(uint32_t *hist, const uint32_t *n_ptr, uint32_t bins, LbBuf st) {
    const uint32_t n_ev = *n_ptr;
    const uint32_t ntu = (n_ev + SORT_TILE - 1) / SORT_TILE;
    const uint64_t n = (uint64_t)bins * ntu;
    auto addr = [&](uint64_t e) { return (size_t)(e % ntu) * bins + (size_t)(e / ntu); };
    lb_scan<uint32_t>(n, st, [&](uint64_t e) { return hist[addr(e)]; },
                      [&](uint64_t e, uint32_t ex, uint32_t) { hist[addr(e)] = ex; }, [](uint32_t) {});
}

// ------------------------------------------------------------------------------------------------
// Stable rank of every item of a tile inside its digit: the tile's items are taken in (warp, round, lane) order: warp w
// owns the 32 * ITEMS consecutive items from w * 32 * ITEMS on, ITEMS rounds of 32 (the rounds of a warp are serial: every
// one reads the counters the previous one wrote). On return wcount[w][d] holds the warp's items of digit d and lrank[rd]
// the item's rank among the items of its digit in its own warp.
template <int ITEMS>
__device__ __forceinline__ void tile_rank(const uint32_t (&dig)[ITEMS], uint32_t nvalid_warp, uint32_t bins, uint32_t *wcount, uint32_t (&lrank)[ITEMS]) {
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t *mine = wcount + (size_t)w * bins;
    uint32_t peers[ITEMS];
    // the matches of all rounds first (they depend on nothing but the digits), then the serial counter updates
#pragma unroll
    for (int rd = 0; rd < ITEMS; rd++) {
        const bool valid = (uint32_t)(rd * 32) + lane < nvalid_warp;
#ifdef SORT_MATCH_ANY
        peers[rd] = __match_any_sync(0xffffffffu, valid ? dig[rd] : 0xffffffffu);   // invalid lanes form their own group
#else
        // the lanes with my digit, bit by bit with ballots (one VOTE and two logic operations per bit: MATCH.ANY stood for a third
        // of the sort kernels' stall samples, profiles/ncu_r02_c3block.md)
        uint32_t m = __ballot_sync(0xffffffffu, valid);
        if (!valid) m = ~m;
        for (uint32_t bit = 1; bit < bins; bit <<= 1) {
            const uint32_t v = __ballot_sync(0xffffffffu, (dig[rd] & bit) != 0);
            m &= (dig[rd] & bit) ? v : ~v;
        }
        peers[rd] = m;
#endif
    }
#pragma unroll
    for (int rd = 0; rd < ITEMS; rd++) {
        const bool valid = (uint32_t)(rd * 32) + lane < nvalid_warp;
        const uint32_t leader = __ffs(peers[rd]) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = mine[dig[rd]];
            mine[dig[rd]] = old + __popc(peers[rd]);
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
// 512-thread CTAs: a tile's 4096 events are 8 rounds per warp instead of 16, and the 830 tiles of C2 make 2.8 waves of
// half-length tiles instead of 1.4 of full ones.
#define RANK_BLOCK 512
#define RANK_WARPS (RANK_BLOCK / 32)
#define RANK_ITEMS (SORT_TILE / RANK_BLOCK)
struct RankIO {
    const uint32_t *n_events;
    EvSource src;
    const uint32_t *hist_scanned;           // scanned tile histogram of the digit
    uint32_t bits;
    const uint32_t *bucket_base, *keep, *sel_base;
    uint4 *sel;                             // [E_sel] selected events (EvRow) grouped by slot
    uint8_t *read_has_sel;                  // [R] or null (only the Med-MAD pass needs it)
};

#define RANK_SMEM ((RANK_WARPS + 1) * SORT_MAX_BINS * 4)
__global__ void __launch_bounds__(RANK_BLOCK, 2) k_rank(RankIO io) {
    extern __shared__ __align__(128) uint32_t sm_rank[];
    __shared__ EvTile sm_tile;
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.bits, mask = bins - 1;
    uint32_t *wcount = sm_rank;                         // [RANK_WARPS][bins]
    uint32_t *goff = sm_rank + RANK_WARPS * bins;       // [bins]
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t used_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    for (uint32_t tile = blockIdx.x; tile < used_tiles; tile += gridDim.x) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < RANK_WARPS * bins; i += RANK_BLOCK) wcount[i] = 0;
        for (uint32_t i = threadIdx.x; i < bins; i += RANK_BLOCK) goff[i] = io.hist_scanned[(size_t)tile * bins + i];
        const uint32_t base = tile * SORT_TILE, end = min(n, base + SORT_TILE);
        ev_tile_setup(io.src, base, end, &sm_tile);
        const uint32_t wbase = base + w * (32 * RANK_ITEMS);
        const uint32_t nvw = (end > wbase) ? min(end - wbase, (uint32_t)(32 * RANK_ITEMS)) : 0u;
        uint4 row[RANK_ITEMS];
        uint32_t dig[RANK_ITEMS], lrank[RANK_ITEMS];
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            row[rd] = (idx < end) ? ev_fetch(io.src, sm_tile, idx) : make_uint4(0, 0, 0, 0);
            dig[rd] = row[rd].x & mask;
        }
        tile_rank<RANK_ITEMS>(dig, nvw, bins, wcount, lrank);
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
                    io.sel[io.sel_base[slot] + rank] = row[rd];
                    if (io.read_has_sel) io.read_has_sel[row[rd].w] = 1;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// The two-digit kernels. k_msd_scatter works on the histogram's tiles (SORT_TILE events, 512 threads x 8 rounds like
// k_rank). k_msd_rank is free to pick its own tile (MSD_BLOCK threads x 8 rounds); measured on a C3 block (13 M events,
// whole sort stage): 512 threads 515 us, 128 threads 528 us, 256 threads 573 us.
#ifndef MSD_BLOCK
#define MSD_BLOCK 512
#endif
#ifndef MSD_MIN_CTAS
#define MSD_MIN_CTAS 2
#endif
#define MSD_WARPS (MSD_BLOCK / 32)
#define MSD_ITEMS 8
#define MSD_TILE (MSD_BLOCK * MSD_ITEMS)

// k_msd_scatter (first pass): stable split of the batch by the HIGH digit of the slot.
// rows_out[pos], pos = (events of lower high digits) + (events of the same high digit in earlier tiles) + rank in tile.
// Bulk copies (TMA, cp.async.bulk -> SASS UBLKCP) with an mbarrier (SYNCS) as the completion signal: one thread asks for a whole
// tile of rows, the copy engine writes it into shared memory, the CTA waits on the barrier's phase.
__device__ __forceinline__ uint32_t sort_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sort_smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sort_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(sort_smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_load(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {   // 16-byte aligned, bytes % 16 == 0
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(sort_smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(sort_smem_u32(bar)) : "memory");
}

struct MsdScatterIO {
    const uint32_t *n_events;
    EvSource src;
    const uint32_t *hist_scanned;
    uint32_t shift, bits;
    uint4 *rows_out;
};

__global__ void __launch_bounds__(RANK_BLOCK, 2) k_msd_scatter(MsdScatterIO io) {
    extern __shared__ __align__(128) uint32_t sm_rank[];
    __shared__ EvTile sm_tile;
    __shared__ __align__(8) uint64_t sm_bar;
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.bits;
    uint4 *tile_rows = reinterpret_cast<uint4 *>(sm_rank);               // [SORT_TILE] the tile's rows, written by the copy engine
    uint32_t *wcount = sm_rank + SORT_TILE * 4;          // [RANK_WARPS][bins]
    uint32_t *goff = wcount + RANK_WARPS * bins;          // [bins] first output position of the tile's rows of the digit
    uint32_t *tile_off = goff + bins;                     // [bins] first place of the digit in the tile's digit order
    __shared__ uint32_t sm_wsum[RANK_WARPS];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t used_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    if (blockIdx.x >= used_tiles) return;
    if (threadIdx.x == 0) mbar_init(&sm_bar, 1);
    uint32_t phase = 0;
    // A tile of long records is one to EV_SEGS contiguous stretches of the temporaries (sm_tile's segment table): one thread
    // asks the copy engine for them (bulk copies, one mbarrier for the lot) and the tile arrives in dense order in shared
    // memory while the CTA clears its counters and loads the tile's bases. Tiles that touch more records are fetched row by
    // row (ev_fetch). (The copy of the NEXT tile under this tile's ranking was measured too: 513 -> 485 us per block; the
    // buffer serves the coalesced write-out instead, which is worth more: 461, and 4xx with both kernels.)
    auto bulk_tile = [&]() { return sm_tile.r_hi - sm_tile.r_lo < EV_SEGS; };
    auto request = [&](uint32_t base, uint32_t end) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the generic-proxy reads of the buffer come first
        mbar_expect_tx(&sm_bar, (end - base) * 16u);
#pragma unroll
        for (int k = 0; k < EV_SEGS; k++) {
            const uint32_t lo = max(base, sm_tile.seg_start[k]);
            const uint32_t hi = min(end, (k + 1 < EV_SEGS) ? sm_tile.seg_start[k + 1] : 0xffffffffu);
            if (lo < hi) bulk_load(tile_rows + (lo - base), io.src.rows_tmp + (size_t)((long long)lo + sm_tile.seg_delta[k]), (hi - lo) * 16u, &sm_bar);
        }
    };
    for (uint32_t tile = blockIdx.x; tile < used_tiles; tile += gridDim.x) {
        __syncthreads();                                  // the previous tile's write-out has left the buffer and the counters
        const uint32_t base = tile * SORT_TILE, end = min(n, base + SORT_TILE);
        ev_tile_setup(io.src, base, end, &sm_tile);       // (its barrier also publishes the mbarrier's initialisation)
        const bool bulk = bulk_tile();
        if (threadIdx.x == 0 && bulk) request(base, end);
        for (uint32_t i = threadIdx.x; i < RANK_WARPS * bins; i += RANK_BLOCK) wcount[i] = 0;
        for (uint32_t i = threadIdx.x; i < bins; i += RANK_BLOCK) goff[i] = io.hist_scanned[(size_t)tile * bins + i];
        const uint32_t wbase = base + w * (32 * RANK_ITEMS);
        const uint32_t nvw = (end > wbase) ? min(end - wbase, (uint32_t)(32 * RANK_ITEMS)) : 0u;
        uint4 row[RANK_ITEMS];
        uint32_t dig[RANK_ITEMS], lrank[RANK_ITEMS];
        if (bulk) { mbar_wait(&sm_bar, phase); phase ^= 1; }
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            row[rd] = (idx >= end) ? make_uint4(0, 0, 0, 0) : bulk ? tile_rows[idx - base] : ev_fetch(io.src, sm_tile, idx);
            dig[rd] = (row[rd].x >> io.shift) & (bins - 1);
        }
        __syncthreads();                                  // counters cleared, every row in registers: the buffer is free for the sorted copy
        tile_rank<RANK_ITEMS>(dig, nvw, bins, wcount, lrank);
        __syncthreads();
        {   // per digit: the tile's first output position (-> exclusive over warps) and its first place in the tile's digit order
            const uint32_t bpt = (bins + RANK_BLOCK - 1) / RANK_BLOCK;
            uint32_t tsum = 0;
            for (uint32_t kk = 0; kk < bpt; kk++) {
                const uint32_t b = threadIdx.x * bpt + kk;
                if (b < bins) {
                    uint32_t run = goff[b];
#pragma unroll
                    for (int ww = 0; ww < RANK_WARPS; ww++) {
                        const uint32_t t = wcount[ww * bins + b];
                        wcount[ww * bins + b] = run;
                        run += t;
                    }
                    tile_off[b] = tsum;
                    tsum += run - goff[b];
                }
            }
            const uint32_t inc = warp_incl_scan_add(tsum);
            if (lane == 31) sm_wsum[w] = inc;
            __syncthreads();
            uint32_t before = inc - tsum;
            for (uint32_t ww = 0; ww < w; ww++) before += sm_wsum[ww];
            for (uint32_t kk = 0; kk < bpt; kk++) {
                const uint32_t b = threadIdx.x * bpt + kk;
                if (b < bins) tile_off[b] += before;
            }
        }
        __syncthreads();
        // the tile in digit order in shared memory, then out: the rows of one digit go to consecutive places of rows_out
#pragma unroll
        for (int rd = 0; rd < RANK_ITEMS; rd++) {
            const uint32_t idx = wbase + rd * 32 + lane;
            if (idx < end) {
                const uint32_t d = dig[rd];
                tile_rows[tile_off[d] + (wcount[w * bins + d] + lrank[rd] - goff[d])] = row[rd];
            }
        }
        __syncthreads();
        for (uint32_t p = threadIdx.x; p < end - base; p += RANK_BLOCK) {
            const uint4 r = tile_rows[p];
            const uint32_t d = (r.x >> io.shift) & (bins - 1);
            io.rows_out[goff[d] + (p - tile_off[d])] = r;
        }
    }
}
#define MSD_SCATTER_SMEM(bits) ((size_t)SORT_TILE * 16 + (size_t)(RANK_WARPS + 2) * (1u << (bits)) * 4)

// ------------------------------------------------------------------------------------------------
// k_msd_rank (second pass): one CTA per high-digit bucket at a time. The bucket is the contiguous range
// [start(h), start(h+1)) of rows_in (start = the scanned histogram's entry of tile 0) and holds the slots
// h << lo_bits ... in their (record, position) order. The CTA walks it tile by tile with one running counter per
// slot: rank = events of the slot seen so far; rank < keep[slot] -> the row goes to sel[sel_base[slot] + rank].
struct MsdRankIO {
    const uint32_t *n_events;
    const uint4 *rows_in;
    const uint32_t *hist_scanned;           // first pass's scanned histogram (tile-major: tile 0's row holds the bucket starts)
    uint32_t hi_bits, lo_bits, n_slots;
    const uint32_t *keep, *sel_base;
    uint4 *sel;
    uint8_t *read_has_sel;                  // or null
    uint32_t *ticket;                       // bucket hand-out (zeroed per batch)
};

__global__ void __launch_bounds__(MSD_BLOCK, MSD_MIN_CTAS) k_msd_rank(MsdRankIO io) {
    extern __shared__ __align__(128) uint32_t sm_rank[];
    __shared__ uint32_t sm_bucket;
    __shared__ __align__(8) uint64_t sm_bar;
    const uint32_t n = *io.n_events;
    const uint32_t bins = 1u << io.lo_bits, mask = bins - 1;
    uint4 *tile_rows = reinterpret_cast<uint4 *>(sm_rank);                 // [MSD_TILE] the tile's rows, written by the copy engine
    uint32_t *wcount = sm_rank + MSD_TILE * 4;           // [MSD_WARPS][bins]
    uint32_t *cnt = wcount + MSD_WARPS * bins;           // [bins] events of the slot so far
    uint32_t *keep_s = cnt + bins, *selb_s = keep_s + bins;
    uint32_t *cold = selb_s + bins, *tile_off = cold + bins;   // [bins] the slot's count before this tile; its first place in the tile's slot order
    __shared__ uint32_t sm_wsum[MSD_WARPS];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t n_buckets = (io.n_slots + bins - 1) >> io.lo_bits;
    if (n == 0) return;                                  // nothing was scanned: the histogram holds no bucket starts
    if (threadIdx.x == 0) mbar_init(&sm_bar, 1);
    uint32_t phase = 0;
    // one thread asks the copy engine for the rows [base, end) of the bucket; the tile buffer must be free (all rows of the
    // previous tile are in registers: the callers stand behind a block barrier)
    auto request = [&](uint32_t base, uint32_t end) {
        const uint32_t bytes = (end - base) * 16u;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the generic-proxy reads of the buffer come first
        mbar_expect_tx(&sm_bar, bytes);
        bulk_load(tile_rows, io.rows_in + base, bytes, &sm_bar);
    };
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sm_bucket = atomicAdd(io.ticket, 1u);
        __syncthreads();
        const uint32_t h = sm_bucket;
        if (h >= n_buckets) break;
        const uint32_t start = io.hist_scanned[h];
        const uint32_t stop = (h + 1 < (1u << io.hi_bits)) ? io.hist_scanned[h + 1] : n;
        if (start >= stop) continue;
        // the bucket's first tile is asked for while its slot tables load (the barriers above: the previous bucket's last tile
        // has left the buffer, and the mbarrier's initialisation is visible)
        if (threadIdx.x == 0) request(start, min(stop, start + MSD_TILE));
        const uint32_t slot0 = h << io.lo_bits;
        for (uint32_t i = threadIdx.x; i < bins; i += MSD_BLOCK) {
            const uint32_t s = slot0 + i;
            cnt[i] = 0;
            keep_s[i] = (s < io.n_slots) ? io.keep[s] : 0u;
            selb_s[i] = (s < io.n_slots) ? io.sel_base[s] : 0u;
        }
        for (uint32_t base = start; base < stop; base += MSD_TILE) {
            __syncthreads();                              // the previous tile's write-out has left the buffer and the counters
            const uint32_t end = min(stop, base + MSD_TILE);
            if (threadIdx.x == 0 && base != start) request(base, end);
            for (uint32_t i = threadIdx.x; i < MSD_WARPS * bins; i += MSD_BLOCK) wcount[i] = 0;
            const uint32_t wbase = base + w * (32 * MSD_ITEMS);
            const uint32_t nvw = (end > wbase) ? min(end - wbase, (uint32_t)(32 * MSD_ITEMS)) : 0u;
            uint4 row[MSD_ITEMS];
            uint32_t dig[MSD_ITEMS], lrank[MSD_ITEMS];
            mbar_wait(&sm_bar, phase);
            phase ^= 1;
#pragma unroll
            for (int rd = 0; rd < MSD_ITEMS; rd++) {
                const uint32_t idx = wbase + rd * 32 + lane;
                row[rd] = (idx < end) ? tile_rows[idx - base] : make_uint4(0, 0, 0, 0);
                dig[rd] = row[rd].x & mask;
            }
            __syncthreads();                              // counters cleared, every row in registers: the buffer is free for the sorted copy
            tile_rank<MSD_ITEMS>(dig, nvw, bins, wcount, lrank);
            __syncthreads();
            // Per slot: running count (-> exclusive over warps), and where the slot's events of THIS tile start in the tile's
            // slot-sorted order (exclusive scan over the slots; a thread owns bpt consecutive slots).
            {
                const uint32_t bpt = (bins + MSD_BLOCK - 1) / MSD_BLOCK;
                uint32_t tsum = 0;
                for (uint32_t kk = 0; kk < bpt; kk++) {
                    const uint32_t b = threadIdx.x * bpt + kk;
                    if (b < bins) {
                        uint32_t run = cnt[b];
                        cold[b] = run;
#pragma unroll
                        for (int ww = 0; ww < MSD_WARPS; ww++) {
                            const uint32_t t = wcount[ww * bins + b];
                            wcount[ww * bins + b] = run;
                            run += t;
                        }
                        tile_off[b] = tsum;
                        tsum += run - cnt[b];
                        cnt[b] = run;
                    }
                }
                const uint32_t inc = warp_incl_scan_add(tsum);
                if (lane == 31) sm_wsum[w] = inc;
                __syncthreads();
                uint32_t before = inc - tsum;             // exclusive prefix of this thread's slots inside the tile
                for (uint32_t ww = 0; ww < w; ww++) before += sm_wsum[ww];
                for (uint32_t kk = 0; kk < bpt; kk++) {
                    const uint32_t b = threadIdx.x * bpt + kk;
                    if (b < bins) tile_off[b] += before;
                }
            }
            __syncthreads();
            // The tile in slot order, in shared memory: the rows of a slot stand together, and a slot's rows go to CONSECUTIVE
            // places of sel[]; written out by consecutive threads they leave as whole 32/64/128-byte pieces instead of one
            // 16-byte store per event scattered over 512 places (the scattered form cost as much as everything else in the kernel).
#pragma unroll
            for (int rd = 0; rd < MSD_ITEMS; rd++) {
                const uint32_t idx = wbase + rd * 32 + lane;
                if (idx < end) {
                    const uint32_t d = dig[rd];
                    const uint32_t rank = wcount[w * bins + d] + lrank[rd];
                    tile_rows[tile_off[d] + (rank - cold[d])] = row[rd];
                }
            }
            __syncthreads();
            for (uint32_t p = threadIdx.x; p < end - base; p += MSD_BLOCK) {
                const uint4 r = tile_rows[p];
                const uint32_t d = r.x & mask;
                const uint32_t rank = cold[d] + (p - tile_off[d]);
                if (rank < keep_s[d]) {
                    io.sel[selb_s[d] + rank] = r;
                    if (io.read_has_sel) io.read_has_sel[r.w] = 1;
                }
            }
        }
    }
}
#define MSD_RANK_SMEM(lo_bits) ((size_t)MSD_TILE * 16 + (size_t)(MSD_WARPS + 5) * (1u << (lo_bits)) * 4)

// Selected rows -> the structure-of-arrays form of pgm_result (ev_start / ev_len / ev_read), for callers that download them.
// ev_start is relative to the record's first sample after trim (rec_src0, k_read_prep).
__global__ void __launch_bounds__(PGM_BLOCK) k_unpack_sel(const uint4 *sel, const uint32_t *n_sel_ptr, const unsigned long long *rec_src0, uint32_t *ev_start,
                                                          uint32_t *ev_len, uint32_t *ev_read) {
    const uint32_t n = *n_sel_ptr;
    for (uint32_t i = blockIdx.x * PGM_BLOCK + threadIdx.x; i < n; i += gridDim.x * PGM_BLOCK) {
        const uint4 r = sel[i];
        ev_start[i] = (uint32_t)(ev_row_src(r) - rec_src0[r.w]); ev_len[i] = ev_row_len(r); ev_read[i] = r.w;
    }
}
