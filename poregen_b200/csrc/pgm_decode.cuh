// pgm_decode.cuh — alignment decode: move tables and ss:Z strings -> candidate events -> filters.
//
// One CTA per alignment record. Both encodings go through the same machinery: a block-wide
// prefix scan over the alignment bytes turns "boundary" bytes (a set move, or an ss operator
// with its run length) into event indices and raw-signal ranges; the shared back end
// (pgm_eval_event) applies the reference's filters in the reference's order.
//
// Runs twice per batch: COUNT pass (per-slot counts + per-read totals) and EMIT pass (writes the
// passing events compactly, in (record, position) order, at the offsets the scan over the
// per-read totals produced). The ss token scan is only done in the COUNT pass; its per-event
// scratch (start, end, base code, indel positions) is reused by the EMIT pass.
#pragma once
#include "pgm_device.cuh"

#define DEC_ITEMS 16  // alignment bytes per thread per chunk
#define KM_ITEMS 4    // ss: k-mer start positions per thread per chunk

struct DecodeIO {
    // quota state
    const uint32_t *consumed;  // [n_slots] kmer_frequency_map at batch start
    uint32_t *counts;          // [n_slots] passing events of this batch (COUNT)
    uint32_t *n_pass;          // [R] passing events per record (COUNT)
    const uint32_t *ev_base;   // [R+1] exclusive scan of n_pass (EMIT)
    uint32_t ev_cap;           // capacity of the event arrays
    uint32_t *key, *ev_start, *ev_len, *ev_read;  // [E] passing events, record-then-position order (EMIT)
    uint8_t *rflags;           // [R] PGM_RF_*
    unsigned long long *examined;
    // ss scratch, indexed by align_off + event index / indel index
    uint32_t *sc_start, *sc_end;
    uint8_t *sc_code;
    int32_t *sc_indel;
    uint32_t *ss_L, *ss_nindel;  // [R]
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

// ------------------------------------------------------------------------------------------------
// Move tables (TSV '0'/'1' bytes, or mv:B:c int8): reference gmove.cpp:620-690 and 1182-1252.
// Event e (0-based over ALL set moves) lies between the e-th and (e+1)-th set move; it is examined
// when e >= m and labelled with the k-mer at base s_off + (e - m). The run after the last set move
// is never an event (the loop bound reads the terminator, gmove.cpp:632).
template <bool EMIT>
__global__ void __launch_bounds__(PGM_BLOCK) k_decode_moves(DevOpts o, DevBatch b, DecodeIO io) {
    __shared__ uint32_t sm_u[PGM_WARPS + 1];
    __shared__ int sm_i[PGM_WARPS + 1];
    const uint32_t r = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    if (io.rflags[r] != 0) {
        if (!EMIT && tid == 0) io.n_pass[r] = 0;
        return;
    }
    const pgm_read rd = b.reads[r];
    const uint8_t *mv = b.align + rd.align_off;
    const uint8_t *seq = b.bases + rd.seq_off;
    const uint32_t M = rd.align_len;
    const uint8_t one = (b.mode == PGM_MODE_MOVES_CHAR) ? (uint8_t)'1' : (uint8_t)1;
    const uint64_t n_sig = (uint64_t)rd.raw_len - rd.trim;
    uint32_t ones_carry = 0;
    int last_carry = -1;
    uint32_t out_run = EMIT ? io.ev_base[r] : 0;  // next output index (EMIT)
    uint32_t pass_total = 0, exam_total = 0;

    for (uint32_t base = 0; base < M; base += PGM_BLOCK * DEC_ITEMS) {
        const uint32_t p0 = base + tid * DEC_ITEMS;
        uint32_t bits = 0;
        if (p0 < M) {
            const uint32_t lim = min((uint32_t)DEC_ITEMS, M - p0);
            for (uint32_t i = 0; i < lim; i++) bits |= (uint32_t)(__ldg(mv + p0 + i) == one) << i;
        }
        const uint32_t cnt = __popc(bits);
        const int lastpos = bits ? (int)(p0 + 31 - __clz(bits)) : -1;
        uint32_t tot_cnt; int tot_last;
        const uint32_t ex_cnt = block_excl_scan_add(cnt, sm_u, &tot_cnt);
        const int ex_last = block_excl_scan_max(lastpos, sm_i, &tot_last);

        // pass A: evaluate (count), pass B (EMIT only): evaluate again and write in order
        uint32_t my_pass = 0;
        for (int phase = 0; phase < (EMIT ? 2 : 1); phase++) {
            uint32_t rank = ones_carry + ex_cnt;
            int prev = max(last_carry, ex_last);
            uint32_t w = bits;
            uint32_t wr = 0;
            if (EMIT && phase == 1) {
                uint32_t tot_pass;
                const uint32_t ex_pass = block_excl_scan_add(my_pass, sm_u, &tot_pass);
                wr = out_run + ex_pass;
                out_run += tot_pass;
            }
            while (w) {
                const uint32_t i = __ffs(w) - 1;
                w &= w - 1;
                const uint32_t p = p0 + i;
                if (rank > o.m) {  // this set move closes event (rank-1); needs rank-1 >= m
                    const uint64_t ei = (uint64_t)rank - o.m - 1;
                    const uint64_t seqpos = (uint64_t)o.s_off + ei;
                    const uint32_t start = (uint32_t)((uint64_t)(uint32_t)prev * rd.stride);
                    const uint32_t end = (uint32_t)((uint64_t)p * rd.stride);
                    bool id_ok = seqpos + o.k <= rd.seq_len;                                   // U4
                    uint32_t id = 0;
                    if (id_ok) {
                        for (uint32_t t = 0; t < o.k; t++) {
                            const uint32_t c = pgm_base_code(__ldg(seq + seqpos + t), o.fourth, false);
                            id_ok &= (c != 0xFF);
                            id = (id << 2) | (c & 3);
                        }
                    }
                    uint32_t slot, s2, len;
                    const bool ok = pgm_eval_event(o, io.consumed, start, end, id, id_ok, n_sig, &slot, &s2, &len);
                    if (phase == 0) {
                        exam_total++;
                        if (ok) {
                            my_pass++;
                            if (!EMIT) atomicAdd(io.counts + slot, 1u);
                        }
                    } else if (ok) {
                        if (wr < io.ev_cap) {
                            io.key[wr] = slot; io.ev_start[wr] = s2; io.ev_len[wr] = len; io.ev_read[wr] = r;
                        }
                        wr++;
                    }
                }
                prev = (int)p;
                rank++;
            }
        }
        pass_total += my_pass;
        ones_carry += tot_cnt;
        last_carry = max(last_carry, tot_last);
    }
    if (!EMIT) {
        uint32_t tp, te;
        block_excl_scan_add(pass_total, sm_u, &tp);
        block_excl_scan_add(exam_total, sm_u, &te);
        if (tid == 0) {
            io.n_pass[r] = tp;
            if (te) atomicAdd(io.examined, (unsigned long long)te);
            if (ones_carry <= o.m) io.rflags[r] |= (PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED);       // U1
        }
    }
}

// ------------------------------------------------------------------------------------------------
// ss:Z strings: reference gmove.cpp:831-955. Grammar (<uint>(','|'I'|'D'))*.
//   ','  a base mapped to n samples: event j (the j-th ','), raw range [i_raw, i_raw+n)
//   'I'  n samples with no base: i_raw += n, indel position = number of ',' so far
//   'D'  n bases with no samples: i_k += n, indel position = number of ',' so far
// The exclusive block scan carries (raw advance, ',' count, deleted bases, indel count).
__device__ __forceinline__ bool ss_is_op(uint32_t c) { return c == ',' || c == 'I' || c == 'D'; }
__device__ __forceinline__ bool ss_is_digit(uint32_t c) { return c >= '0' && c <= '9'; }

// Parses the run length that ends just before position p. Returns false for U7 (no digits, or
// more than 9 digits).
__device__ __forceinline__ bool ss_number_before(const uint8_t *ss, uint32_t p, uint32_t *val) {
    uint32_t v = 0, mult = 1, nd = 0;
    int q = (int)p - 1;
    while (q >= 0) {
        const uint32_t c = __ldg(ss + q);
        if (!ss_is_digit(c)) break;
        if (nd == 9) return false;
        v += (c - '0') * mult;
        mult *= 10;
        nd++;
        q--;
    }
    *val = v;
    return nd > 0;
}

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

template <bool EMIT>
__global__ void __launch_bounds__(PGM_BLOCK) k_decode_ss(DevOpts o, DevBatch b, DecodeIO io) {
    __shared__ uint32_t sm_u[PGM_WARPS + 1];
    __shared__ uint32_t sm_bad;
    const uint32_t r = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    if (io.rflags[r] != 0) {
        if (!EMIT && tid == 0) io.n_pass[r] = 0;
        return;
    }
    const pgm_read rd = b.reads[r];
    const uint8_t *ss = b.align + rd.align_off;
    const uint8_t *seq = b.bases + rd.seq_off;
    const uint32_t T = rd.align_len;
    const bool rna = (rd.flags & PGM_READ_RNA) != 0;
    const uint64_t n_sig = rd.raw_len;                  // the ss path never trims (gmove.cpp:750-753)
    uint32_t *sc_start = io.sc_start + rd.align_off;
    uint32_t *sc_end = io.sc_end + rd.align_off;
    uint8_t *sc_code = io.sc_code + rd.align_off;
    int32_t *sc_indel = io.sc_indel + rd.align_off;

    uint32_t L, n_indel;
    if (!EMIT) {
        if (tid == 0) sm_bad = 0;
        uint32_t c_raw = rd.query_start, c_com = 0, c_del = 0, c_ind = 0;  // carries
        __syncthreads();
        for (uint32_t base = 0; base < T; base += PGM_BLOCK * DEC_ITEMS) {
            const uint32_t p0 = base + tid * DEC_ITEMS;
            const uint32_t lim = (p0 < T) ? min((uint32_t)DEC_ITEMS, T - p0) : 0;
            uint32_t a = 0, c = 0, d = 0, g = 0;
            bool bad = false;
            for (uint32_t i = 0; i < lim; i++) {
                const uint32_t ch = __ldg(ss + p0 + i);
                if (ss_is_op(ch)) {
                    uint32_t num;
                    if (!ss_number_before(ss, p0 + i, &num)) { bad = true; num = 0; }
                    if (ch == ',') { a += num; c++; }
                    else if (ch == 'I') { a += num; g++; }
                    else { d += num; g++; }
                } else if (!ss_is_digit(ch)) bad = true;
            }
            if (bad) atomicOr(&sm_bad, 1u);
            uint32_t ta, tc, td, tg;
            uint32_t xa = c_raw + block_excl_scan_add(a, sm_u, &ta);
            uint32_t xc = c_com + block_excl_scan_add(c, sm_u, &tc);
            uint32_t xd = c_del + block_excl_scan_add(d, sm_u, &td);
            uint32_t xg = c_ind + block_excl_scan_add(g, sm_u, &tg);
            for (uint32_t i = 0; i < lim; i++) {
                const uint32_t ch = __ldg(ss + p0 + i);
                if (!ss_is_op(ch)) continue;
                uint32_t num = 0;
                ss_number_before(ss, p0 + i, &num);
                if (ch == ',') {
                    const uint64_t ik = (uint64_t)xc + xd;
                    const int64_t si = rna ? (int64_t)rd.seq_len - (int64_t)ik - 1 : (int64_t)ik;
                    uint32_t code = 0xFF;
                    if (si >= 0 && si < (int64_t)rd.seq_len) code = pgm_base_code(__ldg(seq + si), o.fourth, rna);
                    sc_start[xc] = xa; sc_end[xc] = xa + num; sc_code[xc] = (uint8_t)code;
                    xa += num; xc++;
                } else if (ch == 'I') { sc_indel[xg] = (int32_t)xc; xa += num; xg++; }
                else { sc_indel[xg] = (int32_t)xc; xd += num; xg++; }
            }
            c_raw += ta; c_com += tc; c_del += td; c_ind += tg;
        }
        __syncthreads();  // scratch writes of this CTA are visible to it from here on
        L = c_com; n_indel = c_ind;
        if (sm_bad) {     // U7
            if (tid == 0) { io.rflags[r] |= (PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED); io.n_pass[r] = 0; io.ss_L[r] = 0; io.ss_nindel[r] = 0; }
            return;
        }
        if (tid == 0) { io.ss_L[r] = L; io.ss_nindel[r] = n_indel; }
    } else {
        L = io.ss_L[r]; n_indel = io.ss_nindel[r];
    }

    uint32_t pass_total = 0, exam_total = 0;
    uint32_t out_run = EMIT ? io.ev_base[r] : 0;
    const int k = (int)o.k;
    const int last = (int)((uint32_t)rd.end_k + o.pick_margin);             // gmove.cpp:885
    const uint32_t n_i = ((int)L >= k) ? L - o.k + 1 : 0;                   // U5
    for (uint32_t base = 0; base < n_i; base += PGM_BLOCK * KM_ITEMS) {
        const uint32_t i0 = base + tid * KM_ITEMS;
        uint32_t my_pass = 0;
        for (int phase = 0; phase < (EMIT ? 2 : 1); phase++) {
            uint32_t wr = 0;
            if (EMIT && phase == 1) {
                uint32_t tot_pass;
                const uint32_t ex = block_excl_scan_add(my_pass, sm_u, &tot_pass);
                wr = out_run + ex;
                out_run += tot_pass;
            }
            for (uint32_t u = 0; u < KM_ITEMS; u++) {
                const uint32_t i = i0 + u;
                if (i >= n_i) break;
                const uint64_t j = (uint64_t)i + o.m;
                if (j >= L) continue;                                        // gmove.cpp:892 (-1 cell), U5
                const int left = rna ? (int)L - (int)i - k : (int)i;        // gmove.cpp:899,906
                if (phase == 0) exam_total++;
                if (!ss_pick(sc_indel, n_indel, last, left, k, (int)o.pick_margin, (int)L, rd.st_k, rna)) continue;
                uint32_t id = 0; bool id_ok = true;
                for (int t = 0; t < k; t++) {
                    const uint32_t c = sc_code[rna ? (i + k - 1 - t) : (i + t)];
                    id_ok &= (c != 0xFF);
                    id = (id << 2) | (c & 3);
                }
                uint32_t slot, s2, len;
                const bool ok = pgm_eval_event(o, io.consumed, sc_start[j], sc_end[j], id, id_ok, n_sig, &slot, &s2, &len);
                if (!ok) continue;
                if (phase == 0) {
                    my_pass++;
                    if (!EMIT) atomicAdd(io.counts + slot, 1u);
                } else {
                    if (wr < io.ev_cap) { io.key[wr] = slot; io.ev_start[wr] = s2; io.ev_len[wr] = len; io.ev_read[wr] = r; }
                    wr++;
                }
            }
        }
        pass_total += my_pass;
    }
    if (!EMIT) {
        uint32_t tp, te;
        block_excl_scan_add(pass_total, sm_u, &tp);
        block_excl_scan_add(exam_total, sm_u, &te);
        if (tid == 0) {
            io.n_pass[r] = tp;
            if (te) atomicAdd(io.examined, (unsigned long long)te);
        }
    }
}
