// pgm_decode.cuh — alignment decode: move tables and ss:Z strings -> candidate events -> filters.
//
// One WARP per alignment record, one launch per batch. Records are handed out by an atomic ticket in
// alignment-file order. Per record:
//   1. tokenise: the alignment bytes are read as aligned 16-byte words (32 lanes = 512 bytes per step);
//      "boundary" bytes (a set move, or an ss operator) are found with byte-parallel word arithmetic,
//      one lane per boundary parses the run length, and warp prefix scans turn the boundaries into
//      event indices and raw-signal ranges;
//   2. evaluate: the shared back end (pgm_eval_event) applies the reference's filters in the reference's
//      order; passing events are counted per slot and written compactly, in position order, to a
//      per-record temporary;
//   3. k_scan_records: a scan over the per-record totals gives the record's offset in the batch-wide
//      (record, position) order; the sort kernels (pgm_sort.cuh) read the rows from the temporaries in that order.
#pragma once
#include "pgm_device.cuh"
#include "pgm_scan.cuh"
#include "pgm_sort.cuh"

#define DEC_WARPS PGM_WARPS
#define DEC_WSM 2368   // bytes of shared staging per warp
#define DEC_SLOT_SMEM 1024   // k-mer tables up to this many slots keep quota state and counts in shared memory

struct DecodeIO {
    // quota state
    const uint32_t *consumed;  // [n_slots] kmer_frequency_map at batch start
    uint32_t *counts;          // [n_slots] passing events of this batch
    uint32_t *n_pass;          // [R] passing events per record
    uint32_t *ev_base;         // [R+1] exclusive prefix of n_pass
    uint32_t ev_cap;           // capacity of the event arrays
    uint8_t *rflags;           // [R] PGM_RF_*
    unsigned long long *examined;
    uint32_t *err_flag;        // bit 1: an event longer than EV_MAX_LEN samples (its row cannot hold it)
    // per-record temporaries: record r's passing events, in position order, as rows (EvRow, pgm_sort.cuh) at
    // rows_tmp[(align_off >> tmp_shift) + j] (EvSource, pgm_sort.cuh)
    uint4 *rows_tmp;
    uint32_t tmp_shift;
    // ss scratch, indexed by the record's alignment offset (+ event / indel index)
    uint32_t *sc_start, *sc_end;   // event bounds
    uint8_t *sc_code;              // 2-bit base code per event
    int32_t *sc_indel;             // indel positions
    uint32_t *ticket;          // record hand-out
    const uint32_t *order;     // [R] ticket -> record, longest records first (k_read_prep)
};

// Reference gmove.cpp:641-662 (and 916-937, 1204-1225): duration filters, k-mer lookup, quota
// check, margin widening and end clamp. Returns true when the event would be emitted provided the
// quota still has room when its turn comes.
__device__ __forceinline__ bool pgm_eval_event(const DevOpts &o, const uint32_t *consumed, uint32_t start, uint32_t end,
                                               uint32_t id, bool id_ok, uint64_t n_sig, uint32_t *slot, uint32_t *s2, uint32_t *len) {
    const uint32_t dur = end - start;
    if (dur > o.max_dur) return false;
    if (dur < o.min_dur) return false;
    if (!id_ok) return false;
    const uint32_t sl = pgm_slot_of(o, id);
    if (sl == PGM_INVALID) return false;
    if (consumed[sl] >= o.limit) return false;
    const uint32_t a = (start < o.margin) ? 0u : start - o.margin;                 // U2
    const uint32_t e1 = end + o.margin;
    const uint32_t b = ((uint64_t)e1 > n_sig) ? (uint32_t)n_sig : e1;
    if (a >= b) return false;                                                       // U3
    *slot = sl; *s2 = a; *len = b - a;
    return true;
}

__device__ __forceinline__ uint32_t warp_excl_scan_add(uint32_t v, uint32_t *total) {
    const uint32_t inc = warp_incl_scan_add(v);
    *total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}
__device__ __forceinline__ int warp_excl_scan_max(int v, int *total) {
    const int inc = warp_incl_scan_max(v);
    *total = __shfl_sync(0xffffffffu, inc, 31);
    int ex = __shfl_up_sync(0xffffffffu, inc, 1);
    if ((threadIdx.x & 31) == 0) ex = -1;
    return ex;
}

// Byte-parallel helpers on 32-bit words: 0x80 in every byte of y that is zero (exact, no borrow
// artefacts), and the four flag bits 7,15,23,31 packed into bits 0..3.
__device__ __forceinline__ uint32_t byte_is_zero(uint32_t y) { return ~(((y & 0x7f7f7f7fu) + 0x7f7f7f7fu) | y | 0x7f7f7f7fu); }
__device__ __forceinline__ uint32_t byte_nibble(uint32_t m) { return (((m >> 7) * 0x00204081u) >> 21) & 0xfu; }

// The 16 alignment bytes a lane owns in the current 512-byte step.
struct Lane16 {
    uint32_t w[4];
    __device__ __forceinline__ uint32_t byte(int i) const { return (w[i >> 2] >> (8 * (i & 3))) & 0xffu; }
};
// Aligned 16-byte load of the word at `off` from the aligned base; words without any byte of the
// record are not touched (a word holding at least one valid byte lies inside the allocation).
__device__ __forceinline__ Lane16 load_lane16(const uint8_t *abase, uint32_t off, uint32_t span) {
    Lane16 x;
    if (off < span) {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(abase + off));
        x.w[0] = q.x; x.w[1] = q.y; x.w[2] = q.z; x.w[3] = q.w;
    } else { x.w[0] = x.w[1] = x.w[2] = x.w[3] = 0; }
    return x;
}

// ------------------------------------------------------------------------------------------------
// Move tables (TSV '0'/'1' bytes, or mv:B:c int8): reference gmove.cpp:620-690 and 1182-1252.
// Event e (0-based over ALL set moves) lies between the e-th and (e+1)-th set move; it is examined
// when e >= m and labelled with the k-mer at base s_off + (e - m). The run after the last set move
// is never an event (the loop bound reads the terminator, gmove.cpp:632). Decoder: k_decode_moves below.

// ------------------------------------------------------------------------------------------------
// ss:Z strings: reference gmove.cpp:831-955. Grammar (<uint>(','|'I'|'D'))*.
//   ','  a base mapped to n samples: event j (the j-th ','), raw range [i_raw, i_raw+n)
//   'I'  n samples with no base: i_raw += n, indel position = number of ',' so far
//   'D'  n bases with no samples: i_k += n, indel position = number of ',' so far
// The warp scan carries (raw advance, ',' count, deleted bases, indel count).
__device__ __forceinline__ bool ss_is_op(uint32_t c) { return c == ',' || c == 'I' || c == 'D'; }
__device__ __forceinline__ bool ss_is_digit(uint32_t c) { return c - '0' <= 9u; }

// indel_pos as the reference builds it (gmove.cpp:822-826, 843-845, 877-885), entries 0..n:
//   P[0] = -st_k ; DNA: P[t] = pos[t-1] ; RNA: P[t] = L - pos[n-t]      (non-decreasing)
// the last entry P[n+1] = end_k + kmer_pick_margin is passed separately.
__device__ __forceinline__ int ss_pget(const int32_t *pos, uint32_t n, uint32_t t, int L, int st_k, bool rna) {
    if (t == 0) return -st_k;
    return rna ? L - pos[n - t] : pos[t - 1];
}

// pick_this_kmer (gmove.cpp:204-211): exists t>=1 with left+k+margin <= P[t] && P[t-1] <= left-margin.
// P[0..n] is sorted, so it is enough to look at the largest admissible t (U6 for t==0).
__device__ __forceinline__ bool ss_pick(const int32_t *pos, uint32_t n, int last, int left, int k, int margin, int L, int st_k, bool rna) {
    const int A = (int)((uint32_t)left + (uint32_t)k + (uint32_t)margin);
    const int B = (int)((uint32_t)left - (uint32_t)margin);
    if (n == 0) return -st_k <= B && last >= A;        // no indel in the record (record-uniform): P = {-st_k}, then `last`
    // u = number of t in [0, n] with P[t] <= B
    uint32_t lo = 0, hi = n + 1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (ss_pget(pos, n, mid, L, st_k, rna) <= B) lo = mid + 1; else hi = mid;
    }
    const uint32_t u = lo;
    const uint32_t ts = min(u, n);
    if (ts >= 1 && ss_pget(pos, n, ts, L, st_k, rna) >= A) return true;
    if (u == n + 1 && last >= A) return true;
    return false;
}

__device__ __forceinline__ uint32_t decode_ss_record(const DevOpts &o, const DevBatch &b, const DecodeIO &io, const pgm_read &rd, uint32_t r,
                                                     uint8_t *wsm, uint32_t *exam, bool *bad_out) {
    const uint32_t lane = threadIdx.x & 31;
    const uint8_t *ss = b.align + rd.align_off;
    const uint8_t *seq = b.bases + rd.seq_off;
    const uint32_t T = rd.align_len;
    const bool rna = (rd.flags & PGM_READ_RNA) != 0;
    const uint64_t n_sig = rd.raw_len;                  // the ss path never trims (gmove.cpp:750-753)
    uint32_t *sc_start = io.sc_start + rd.align_off, *sc_end = io.sc_end + rd.align_off;
    uint4 *t_rows = io.rows_tmp + (rd.align_off >> io.tmp_shift);   // at most align_len / 2 events: every ',' token has a digit in front
    uint8_t *sc_code = io.sc_code + rd.align_off;
    int32_t *sc_indel = io.sc_indel + rd.align_off;
    const uint32_t head = (uint32_t)((uintptr_t)ss & 15);
    const uint8_t *abase = ss - head;
    const uint32_t span = head + T;

    // ---- 1. tokenise: 512 bytes per step are staged in the warp's shared buffer (with the previous 16
    // bytes as a halo); operator positions come from byte-parallel word tricks, then ONE LANE PER TOKEN
    // parses its run length backwards and three warp scans place the token. The next step's bytes are
    // requested before this step is parsed, and the base look-ups of up to 8 token rounds are in flight
    // together.
    uint8_t *wbuf = wsm;                                            // [16 + 512]
    uint16_t *wpos = reinterpret_cast<uint16_t *>(wsm + 528);       // [512]
    uint32_t c_raw = rd.query_start, c_com = 0, c_del = 0, c_ind = 0;   // carries
    bool bad = false;
    __syncwarp();   // the previous record's tiles are done with the staging area
    if (lane < 4) reinterpret_cast<uint32_t *>(wbuf)[lane] = 0;
    Lane16 xn = load_lane16(abase, 16 * lane, span);
    for (uint32_t cb = 0; cb < span; cb += 512) {
        const uint32_t off = cb + 16 * lane;
        const Lane16 x = xn;
        xn = load_lane16(abase, off + 512, span);
        // bytes of the lane that belong to the string: bit i <=> head <= off + i < span
        const uint32_t lo = (off < head) ? min(16u, head - off) : 0u;
        const uint32_t hi = (off >= span) ? 0u : min(16u, span - off);
        const uint32_t vm = (hi > lo) ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
        uint32_t opm = 0, dgm = 0;
#pragma unroll
        for (int wi = 0; wi < 4; wi++) {
            const uint32_t w = x.w[wi];
            const uint32_t e = byte_is_zero(w ^ 0x2c2c2c2cu) | byte_is_zero(w ^ 0x49494949u) | byte_is_zero(w ^ 0x44444444u);   // ',' 'I' 'D'
            const uint32_t wl = w & 0x7f7f7f7fu;
            const uint32_t dg = (wl + 0x50505050u) & ~(wl + 0x46464646u) & ~w & 0x80808080u;                 // '0'..'9'
            opm |= byte_nibble(e) << (4 * wi);
            dgm |= byte_nibble(dg) << (4 * wi);
        }
        opm &= vm; dgm &= vm;
        if ((opm | dgm) != vm) bad = true;                                   // gmove.cpp:867
        uint32_t ntok;
        uint32_t tb = warp_excl_scan_add(__popc(opm), &ntok);
        *reinterpret_cast<uint4 *>(wbuf + 16 + 16 * lane) = make_uint4(x.w[0], x.w[1], x.w[2], x.w[3]);
        for (uint32_t m = opm; m; m &= m - 1) wpos[tb++] = (uint16_t)(16 * lane + __ffs(m) - 1);
        __syncwarp();
        const int sb = (cb == 0) ? (int)head : -16;                          // first byte a run length may use (chunk coordinates)
        for (uint32_t g0 = 0; g0 < ntok; g0 += 256) {
            uint32_t base_ch[8], ev_idx[8];
            // run lengths of all 8 rounds first, branch-free: the four bytes in front of the operator come as one
            // unaligned word from the staged copy (numbers up to 9999; longer ones finish in the rare loop), so the
            // dependent shared-memory loads of the rounds overlap instead of queueing behind each other's loops
            uint32_t chs[8], nums[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const uint32_t t = g0 + 32 * u + lane;
                chs[u] = 0; nums[u] = 0;
                if (t < ntok) {
                    const int pos = wpos[t];
                    const uint32_t bidx = (uint32_t)(16 + pos - 4);
                    const uint32_t *wb = reinterpret_cast<const uint32_t *>(wbuf) + (bidx >> 2);
                    const uint32_t v = __funnelshift_r(wb[0], wb[1], (bidx & 3u) * 8u);   // byte 3 = the byte before the operator
                    chs[u] = wbuf[16 + pos];
                    const uint32_t d3 = (v >> 24) - '0', d2 = ((v >> 16) & 0xffu) - '0', d1 = ((v >> 8) & 0xffu) - '0', d0 = (v & 0xffu) - '0';
                    const bool k3 = d3 <= 9u && pos - 1 >= sb, k2 = k3 && d2 <= 9u && pos - 2 >= sb;
                    const bool k1 = k2 && d1 <= 9u && pos - 3 >= sb, k0 = k1 && d0 <= 9u && pos - 4 >= sb;
                    uint32_t num = (k3 ? d3 : 0u) + (k2 ? 10u * d2 : 0u) + (k1 ? 100u * d1 : 0u) + (k0 ? 1000u * d0 : 0u);
                    if (!k3) bad = true;                                     // no digit: gmove.cpp:834
                    if (k0) {                                                // five digits or more
                        uint32_t nd = 4, mult = 10000;
                        for (int q = pos - 5; q >= sb && nd < 10; q--) {
                            const uint32_t dgt = (uint32_t)wbuf[16 + q] - '0';
                            if (dgt > 9) break;
                            if (nd < 9) num += dgt * mult;
                            mult *= 10;
                            nd++;
                        }
                        if (nd > 9) bad = true;                              // U7
                    }
                    nums[u] = num;
                }
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const uint32_t t0 = g0 + 32 * u;
                ev_idx[u] = PGM_INVALID;
                base_ch[u] = 0;
                if (t0 >= ntok) continue;
                const uint32_t ch = chs[u], num = nums[u];
                const uint32_t a = (ch == 'D') ? 0u : num, d = (ch == 'D') ? num : 0u;
                // raw advance: a warp scan. Deleted bases: a scan only where the round has a 'D'. The ',' and indel
                // counts are flags: two ballots instead of a scan.
                uint32_t ta, td = 0;
                const uint32_t xa = c_raw + warp_excl_scan_add(a, &ta);
                uint32_t xd = c_del;
                if (__any_sync(0xffffffffu, d != 0)) xd += warp_excl_scan_add(d, &td);
                const uint32_t lt = (1u << lane) - 1u;
                const uint32_t mcom = __ballot_sync(0xffffffffu, ch == ','), mind = __ballot_sync(0xffffffffu, ch != 0 && ch != ',');
                const uint32_t xc = c_com + __popc(mcom & lt), xg = c_ind + __popc(mind & lt);
                if (ch == ',') {
                    const uint64_t ik = (uint64_t)xc + xd;
                    const int64_t si = rna ? (int64_t)rd.seq_len - (int64_t)ik - 1 : (int64_t)ik;
                    ev_idx[u] = xc;
                    if (si >= 0 && si < (int64_t)rd.seq_len) base_ch[u] = __ldg(seq + si);   // consumed after the 8 rounds
                    sc_start[xc] = xa; sc_end[xc] = xa + num;
                } else if (ch) sc_indel[xg] = (int32_t)xc;
                c_raw += ta; c_del += td; c_com += __popc(mcom); c_ind += __popc(mind);
            }
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (ev_idx[u] != PGM_INVALID) sc_code[ev_idx[u]] = (uint8_t)pgm_base_code(base_ch[u], o.fourth, rna);
        }
        __syncwarp();
        if (lane == 31) *reinterpret_cast<uint4 *>(wbuf) = make_uint4(x.w[0], x.w[1], x.w[2], x.w[3]);   // halo of the next step
    }
    if (__any_sync(0xffffffffu, bad)) { *bad_out = true; return 0; }   // U7
    __syncwarp();   // scratch writes of this warp are visible to it from here on
    // ---- 2. evaluate: one k-mer start per lane and round, passing events compacted in place. The scratch
    // of 256 events at a time is staged in shared memory, so a round waits on shared, not on L2.
    uint32_t *t_st = reinterpret_cast<uint32_t *>(wsm);            // [256] start of event i + m
    uint32_t *t_en = t_st + 256;                                    // [256] end of event i + m
    uint8_t *t_cd = wsm + 2048;                                     // [256 + 16] base codes from event i on
    const uint32_t L = c_com, n_indel = c_ind;
    const int k = (int)o.k;
    const int last = (int)((uint32_t)rd.end_k + o.pick_margin);             // gmove.cpp:885
    const uint32_t n_i = ((int)L >= k) ? L - o.k + 1 : 0;                   // U5
    uint32_t pbase = 0, my_exam = 0;
    for (uint32_t tile = 0; tile < n_i; tile += 256) {
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const uint32_t q = 32 * u + lane;
            const uint64_t e = (uint64_t)tile + q + o.m;
            t_st[q] = (e < L) ? sc_start[e] : 0u;
            t_en[q] = (e < L) ? sc_end[e] : 0u;
        }
#pragma unroll
        for (int u = 0; u < 9; u++) {
            const uint32_t q = 32 * u + lane;
            if (q < 256 + 16) t_cd[q] = (tile + q < L) ? sc_code[tile + q] : (uint8_t)0xFF;
        }
        __syncwarp();
        const uint32_t tend = min(tile + 256, n_i);
        for (uint32_t base = tile; base < tend; base += 32) {
            const uint32_t i = base + lane, q = i - tile;
            const uint64_t j = (uint64_t)i + o.m;
            bool ok = false;
            uint32_t slot = 0, s2 = 0, len = 0;
            if (i < n_i && j < L) {                                              // gmove.cpp:892 (-1 cell), U5
                my_exam++;
                const int left = rna ? (int)L - (int)i - k : (int)i;            // gmove.cpp:899,906
                if (ss_pick(sc_indel, n_indel, last, left, k, (int)o.pick_margin, (int)L, rd.st_k, rna)) {
                    uint32_t id = 0;
                    bool id_ok = true;
                    for (int t = 0; t < k; t++) {
                        const uint32_t c = t_cd[rna ? (q + k - 1 - t) : (q + t)];
                        id_ok &= (c != 0xFF);
                        id = (id << 2) | (c & 3);
                    }
                    ok = pgm_eval_event(o, io.consumed, t_st[q], t_en[q], id, id_ok, n_sig, &slot, &s2, &len);
                }
            }
            const uint32_t bm = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const uint32_t p = pbase + __popc(bm & ((1u << lane) - 1));
                atomicAdd(io.counts + slot, 1u);
                if (len > EV_MAX_LEN) atomicOr(io.err_flag, 2u);
                t_rows[p] = ev_row(slot, rd.raw_off + s2, len, r);            // the ss path never trims (gmove.cpp:750-753)
            }
            pbase += __popc(bm);
        }
    }
    __syncwarp();
    *exam += my_exam;
    return pbase;
}

// ------------------------------------------------------------------------------------------------
// k_decode_moves: the move-table decoder (PGM_MODE_MOVES_CHAR / _I8), one CTA per record at a time.
// Reference gmove.cpp:620-690, 1182-1252; the work is laid out for long records (C3: 8,700 moves, 4,200 events each):
//   * a step is DM_CHUNK move bytes, eight per thread (one aligned 8-byte load, the next step's already on its way);
//     a block scan numbers the set moves and every thread drops the positions of its own into a shared list;
//   * the events of the step are then dealt out EVENLY: thread t takes the c = ceil(n / 256) consecutive events from
//     t * c on (a thread's own eight bytes hold anything from 0 to 8 set moves: walking those would run every warp
//     for the worst lane - profiles/ncu_r02_c3block.md);
//   * the bases the step's events need are one contiguous window of the sequence: fetched as 16-byte words at the
//     top of the step, staged as 2-bit codes in shared memory; a thread's events are consecutive, so its k-mer id
//     ROLLS (one code per event);
//   * every event is evaluated ONCE; the quota look-ups of a thread's events are issued together; passing events stay
//     in registers until a second block scan gives their place in the record's temporary (position order).
#define DM_THREADS 256
#define DM_WARPS (DM_THREADS / 32)
#define DM_PER 8
#define DM_CHUNK (DM_THREADS * DM_PER)
#ifndef DM_MIN_CTAS
#define DM_MIN_CTAS 4
#endif

__global__ void __launch_bounds__(DM_THREADS, DM_MIN_CTAS) k_decode_moves(DevOpts o, DevBatch b, DecodeIO gio) {
    __shared__ __align__(16) uint8_t sm_codes[DM_CHUNK + 64];
    __shared__ uint16_t sm_pos[DM_CHUNK];            // step-relative byte index of every set move of the step
    __shared__ uint32_t sm_cnt_w[DM_WARPS], sm_pass_w[DM_WARPS];
    __shared__ uint32_t sm_rec;
    __shared__ uint32_t sm_cons[DEC_SLOT_SMEM], sm_cnt[DEC_SLOT_SMEM];
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    DecodeIO io = gio;
    const bool local_slots = o.n_slots <= DEC_SLOT_SMEM;
    if (local_slots) {
        for (uint32_t s = tid; s < o.n_slots; s += DM_THREADS) { sm_cons[s] = gio.consumed[s]; sm_cnt[s] = 0; }
        io.consumed = sm_cons; io.counts = sm_cnt;
    }
    const uint32_t one8 = (b.mode == PGM_MODE_MOVES_CHAR) ? 0x31313131u : 0x01010101u;
    const uint32_t idmask = (o.k >= 16) ? 0xffffffffu : ((1u << (2 * o.k)) - 1u);
    const uint8_t *bases_end = b.bases + b.n_bases;
    uint32_t my_exam = 0;
    for (;;) {
        __syncthreads();
        if (tid == 0) {
            const uint32_t t = atomicAdd(io.ticket, 1u);
            sm_rec = (t < b.n_reads) ? io.order[t] : 0xffffffffu;
        }
        __syncthreads();
        const uint32_t r = sm_rec;
        if (r >= b.n_reads) break;
        if (io.rflags[r] != 0) { if (tid == 0) io.n_pass[r] = 0; continue; }
        const pgm_read rd = b.reads[r];
        const uint8_t *mv = b.align + rd.align_off;
        const uint8_t *seq = b.bases + rd.seq_off;
        const uint32_t M = rd.align_len;
        const uint64_t n_sig = (uint64_t)rd.raw_len - rd.trim;
        const uint32_t head = (uint32_t)((uintptr_t)mv & 7);
        const uint8_t *abase = mv - head;
        const uint32_t span = head + M;
        uint4 *t_rows = io.rows_tmp + (rd.align_off >> io.tmp_shift);
        const unsigned long long src0 = rd.raw_off + rd.trim;       // raw[] index of the record's sample 0
        uint32_t ones_carry = 0, pbase = 0;
        int last_carry = -1;
        uint2 xn = make_uint2(0u, 0u);
        if (DM_PER * tid < span) xn = __ldg(reinterpret_cast<const uint2 *>(abase + DM_PER * tid));   // a word with one valid byte lies inside the allocation
        for (uint32_t cb = 0; cb < span; cb += DM_CHUNK) {
            const uint32_t off = cb + DM_PER * tid;
            const uint2 x = xn;
            if (off + DM_CHUNK < span) xn = __ldg(reinterpret_cast<const uint2 *>(abase + off + DM_CHUNK));
            // ---- the bases of the step's events. The set move with `rank` set moves of the record in front of it closes
            // event rank-1, examined when rank > m, labelled with the k-mer at base s_off + rank - m - 1. The step's first set
            // move has rank ones_carry: its window starts at base w0 and is at most DM_CHUNK + k - 1 bases long. Thread t
            // fetches the aligned 16-byte word t of it now and turns it into codes while the set moves are being counted.
            const int64_t w0 = (int64_t)o.s_off + (int64_t)ones_carry - (int64_t)o.m - 1;
            const int64_t wlo = max(w0, (int64_t)0);
            const uint8_t *wbase = seq + (wlo & ~(int64_t)15) - (((uintptr_t)seq) & 15);     // 16-byte aligned address at or below seq + wlo
            const int shw = (int)(w0 - (int64_t)(wbase - seq));                              // codes index of window position 0 (may be negative: bases in front of the sequence)
            if (16 * tid < DM_CHUNK + 48) {
                const uint8_t *wword = wbase + 16 * tid;
                const int64_t sp0 = (int64_t)(wword - seq);                                   // sequence position of the word's byte 0
                const int lo_b = (int)min(max(wlo - sp0, (int64_t)0), (int64_t)16), hi_b = (int)min(max((int64_t)rd.seq_len - sp0, (int64_t)0), (int64_t)16);
                uint32_t cw[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};     // 0xFF = no base here (U4: beyond the sequence = no match)
                if (lo_b < hi_b) {
                    uint32_t wv[4];
                    if (wword >= b.bases && wword + 16 <= bases_end) {
                        const uint4 bw = __ldg(reinterpret_cast<const uint4 *>(wword));
                        wv[0] = bw.x; wv[1] = bw.y; wv[2] = bw.z; wv[3] = bw.w;
                    } else {                                                               // the array's first / last partial word: byte by byte
                        wv[0] = wv[1] = wv[2] = wv[3] = 0;
                        for (int q = lo_b; q < hi_b; q++) wv[q >> 2] |= (uint32_t)__ldg(wword + q) << (8 * (q & 3));
                    }
                    const uint32_t m16 = ((1u << hi_b) - 1u) & ~((1u << lo_b) - 1u);
                    const uint32_t f4 = o.fourth * 0x01010101u;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t w = wv[j];
                        const uint32_t eC = byte_is_zero(w ^ 0x43434343u), eG = byte_is_zero(w ^ 0x47474747u), eF = byte_is_zero(w ^ f4);
                        uint32_t valid = byte_is_zero(w ^ 0x41414141u) | eC | eG | eF;                     // 0x80 per byte that is A, C, G or the fourth letter
                        valid &= ((((m16 >> (4 * j)) & 0xfu) * 0x00204081u) & 0x01010101u) << 7;              // and lies inside the sequence
                        const uint32_t code = ((eC | eF) >> 7) | ((eG | eF) >> 6);
                        cw[j] = (code & (valid >> 7) * 3u) | (((~valid & 0x80808080u) >> 7) * 0xffu);
                    }
                }
                *reinterpret_cast<uint4 *>(sm_codes + 16 * tid) = make_uint4(cw[0], cw[1], cw[2], cw[3]);
            }
            const uint32_t lo = (off < head) ? min((uint32_t)DM_PER, head - off) : 0u;
            const uint32_t hi = (off >= span) ? 0u : min((uint32_t)DM_PER, span - off);
            uint32_t bits = byte_nibble(byte_is_zero(x.x ^ one8)) | (byte_nibble(byte_is_zero(x.y ^ one8)) << 4);
            bits &= (hi > lo) ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
            const uint32_t cnt = __popc(bits);
            const uint32_t inc_cnt = warp_incl_scan_add(cnt);
            if (lane == 31) sm_cnt_w[wid] = inc_cnt;
            __syncthreads();                                                            // A
            uint32_t wbase_cnt = 0, tot_cnt = 0;
#pragma unroll
            for (int w = 0; w < DM_WARPS; w++) {
                const uint32_t c = sm_cnt_w[w];
                if ((uint32_t)w < wid) wbase_cnt += c;
                tot_cnt += c;
            }
            {   // the positions of the thread's set moves, in order, into the step's list
                uint32_t q = wbase_cnt + inc_cnt - cnt;
                for (uint32_t m = bits; m; m &= m - 1) sm_pos[q++] = (uint16_t)(DM_PER * tid + __ffs(m) - 1);
            }
            __syncthreads();                                                            // B
            // ---- the thread's events: c consecutive ones from tid * c on
            const uint32_t c_ev = (tot_cnt + DM_THREADS - 1) / DM_THREADS;
            const uint32_t e0 = tid * c_ev;
            const uint32_t c_mine = (e0 < tot_cnt) ? min(c_ev, tot_cnt - e0) : 0u;
            uint32_t ev_slot[DM_PER], ev_s2[DM_PER], ev_len[DM_PER];
            uint32_t cand = 0;                       // bit j: the thread's j-th event passed everything but the quota
            if (c_mine) {
                uint32_t id = 0, run = 0;            // run = consecutive valid codes ending at the newest one (saturates at k)
                for (uint32_t t = 0; t + 1 < o.k; t++) {
                    const uint32_t c = sm_codes[max((int)(e0 + t) + shw, 0)];
                    id = ((id << 2) | (c & 3u)) & idmask;
                    run = (c == 0xFF) ? 0u : min(run + 1u, o.k);
                }
                int prev = e0 ? (int)cb + (int)sm_pos[e0 - 1] - (int)head : last_carry;
#pragma unroll
                for (int j = 0; j < DM_PER; j++) {
                    ev_slot[j] = 0; ev_s2[j] = 0; ev_len[j] = 0;
                    if ((uint32_t)j < c_mine) {
                        const uint32_t e = e0 + j;
                        const uint32_t c = sm_codes[max((int)(e + o.k - 1) + shw, 0)];
                        id = ((id << 2) | (c & 3u)) & idmask;
                        run = (c == 0xFF) ? 0u : min(run + 1u, o.k);
                        const uint32_t rank = ones_carry + e;
                        const int p = (int)cb + (int)sm_pos[e] - (int)head;
                        if (rank > o.m) {
                            my_exam++;
                            const uint32_t start = (uint32_t)prev * rd.stride, end = (uint32_t)p * rd.stride;   // modulo 2^32, like the reference's uint32_t
                            const uint32_t dur = end - start;
                            const uint32_t sl = (run >= o.k) ? pgm_slot_of(o, id) : PGM_INVALID;
                            const uint32_t a = (start < o.margin) ? 0u : start - o.margin;                 // U2
                            const uint32_t e1 = end + o.margin;
                            const uint32_t e2 = ((uint64_t)e1 > n_sig) ? (uint32_t)n_sig : e1;
                            if (!(dur > o.max_dur) && !(dur < o.min_dur) && sl != PGM_INVALID && a < e2) {   // gmove.cpp:641-662, U3
                                cand |= 1u << j;
                                ev_slot[j] = sl; ev_s2[j] = a; ev_len[j] = e2 - a;
                            }
                        }
                        prev = p;
                    }
                }
            }
            const int last_here = tot_cnt ? (int)cb + (int)sm_pos[tot_cnt - 1] - (int)head : -1;
            // quota look-ups of all candidates together (gmove.cpp:650-652), then the survivors
            uint32_t cons[DM_PER];
#pragma unroll
            for (int j = 0; j < DM_PER; j++) cons[j] = ((cand >> j) & 1u) ? io.consumed[ev_slot[j]] : 0xffffffffu;
            uint32_t okmask = 0;
#pragma unroll
            for (int j = 0; j < DM_PER; j++)
                if (((cand >> j) & 1u) && cons[j] < o.limit) { okmask |= 1u << j; atomicAdd(io.counts + ev_slot[j], 1u); }
            const uint32_t np = __popc(okmask);
            const uint32_t inc_p = warp_incl_scan_add(np);
            if (lane == 31) sm_pass_w[wid] = inc_p;
            __syncthreads();                                                            // C
            uint32_t wbase_p = 0, tot_p = 0;
#pragma unroll
            for (int w = 0; w < DM_WARPS; w++) {
                const uint32_t c = sm_pass_w[w];
                if ((uint32_t)w < wid) wbase_p += c;
                tot_p += c;
            }
            uint32_t wr = pbase + wbase_p + inc_p - np;
#pragma unroll
            for (int j = 0; j < DM_PER; j++)
                if ((okmask >> j) & 1u) {
                    if (ev_len[j] > EV_MAX_LEN) atomicOr(io.err_flag, 2u);
                    t_rows[wr++] = ev_row(ev_slot[j], src0 + ev_s2[j], ev_len[j], r);
                }
            pbase += tot_p;
            ones_carry += tot_cnt;
            last_carry = max(last_carry, last_here);
        }
        if (tid == 0) {
            const bool bad = ones_carry <= o.m;                                        // U1
            io.n_pass[r] = bad ? 0u : pbase;
            if (bad) io.rflags[r] |= (PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED);
        }
    }
    my_exam = warp_sum_t<uint32_t>(my_exam);
    if (lane == 0 && my_exam) atomicAdd(io.examined, (unsigned long long)my_exam);
    if (local_slots) {
        __syncthreads();
        for (uint32_t s = tid; s < o.n_slots; s += DM_THREADS) {
            const uint32_t n = sm_cnt[s];
            if (n) atomicAdd(gio.counts + s, n);
        }
    }
}

// ------------------------------------------------------------------------------------------------
#ifndef DEC_MIN_CTAS
#define DEC_MIN_CTAS 4
#endif
// k_decode: the ss:Z decoder, one warp per record.
__global__ void __launch_bounds__(PGM_BLOCK, DEC_MIN_CTAS) k_decode(DevOpts o, DevBatch b, DecodeIO gio) {
    __shared__ __align__(16) uint8_t sm_w[DEC_WARPS][DEC_WSM];   // per-warp staging (tokeniser, then evaluation tiles)
    // Small k-mer tables (k <= 5 at full range): the quota state every event looks up and the per-slot counts every passing
    // event increments live in shared memory for the CTA's lifetime; the counts go to the global vector once, at the end.
    __shared__ uint32_t sm_cons[DEC_SLOT_SMEM], sm_cnt[DEC_SLOT_SMEM];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    DecodeIO io = gio;
    const bool local_slots = o.n_slots <= DEC_SLOT_SMEM;
    if (local_slots) {
        for (uint32_t s = threadIdx.x; s < o.n_slots; s += PGM_BLOCK) { sm_cons[s] = gio.consumed[s]; sm_cnt[s] = 0; }
        __syncthreads();
        io.consumed = sm_cons; io.counts = sm_cnt;
    }
    uint32_t exam = 0;
    for (;;) {
        uint32_t r = 0;
        if (lane == 0) {
            r = atomicAdd(io.ticket, 1u);
            r = (r < b.n_reads) ? io.order[r] : 0xffffffffu;
        }
        r = __shfl_sync(0xffffffffu, r, 0);
        if (r >= b.n_reads) break;
        uint32_t n_pass = 0;
        bool bad = false;
        if (io.rflags[r] == 0) {
            const pgm_read rd = b.reads[r];
            n_pass = decode_ss_record(o, b, io, rd, r, sm_w[wid], &exam, &bad);
            if (bad) {
                n_pass = 0;
                if (lane == 0) io.rflags[r] |= (PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED);
            }
        }
        if (lane == 0) io.n_pass[r] = n_pass;
    }
    exam = warp_sum_t<uint32_t>(exam);
    if (lane == 0 && exam) atomicAdd(io.examined, (unsigned long long)exam);
    if (local_slots) {
        __syncthreads();
        for (uint32_t s = threadIdx.x; s < o.n_slots; s += PGM_BLOCK) {
            const uint32_t n = sm_cnt[s];
            if (n) atomicAdd(gio.counts + s, n);
        }
    }
}

// ev_base = exclusive scan of the per-record totals (one launch, look-back over tiles of 4096 records)
__global__ void __launch_bounds__(PGM_BLOCK) k_scan_records(const uint32_t *n_pass, uint32_t *ev_base, uint32_t R, LbBuf st) {
    lb_scan<uint32_t>(R, st, [&](uint64_t r) { return n_pass[r]; }, [&](uint64_t r, uint32_t ex, uint32_t) { ev_base[r] = ex; },
                      [&](uint32_t tot) { ev_base[R] = tot; });
}

