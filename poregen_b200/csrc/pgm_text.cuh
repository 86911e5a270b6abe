// pgm_text.cuh — the dump text on the device: printf("%.8f,") ... printf("%.8f;") of reference
// gmove.cpp:674-678 (~85 % of the reference's run time is this formatting).
//
// "%.8f" of an IEEE double is an exact integer problem: x = m * 2^e, so N = round-half-even(m * 10^8 * 2^e)
// (the rounding glibc applies in the default mode) and the text is the decimal expansion of N with a
// point before the last 8 digits. m * 10^8 < 2^80 fits 128 bits; for |x| < 2^64 / 10^8 (every pA or
// scaled value in practice) N fits 64 bits and the digits come from a handful of constant divisions.
// Everything else (up to 1.8e308: 309 integer digits) goes through a small multi-limb path, so the
// output is right for every finite double, and inf/nan print as glibc prints them.
#pragma once
#include "pgm_device.cuh"
#include "pgm_scan.cuh"

struct Fmt8 {          // one formatted sample, fast representation
    uint64_t ip;       // integer part (fast path)
    uint32_t fp;       // 8 fractional digits as an integer < 10^8
    uint32_t len;      // characters including sign and terminator
    uint8_t neg, nd;   // sign, digits of ip
    uint8_t big;       // 1: needs fmt8_big (|x| >= 2^64 / 10^8), 2: inf, 3: nan
};

__device__ __forceinline__ uint32_t dec_digits_u64(uint64_t v) {
    uint32_t n = 1;
    while (v >= 10) { v /= 10; n++; }
    return n;
}

// 128-bit helpers on (hi, lo)
__device__ __forceinline__ bool u128_bit(uint64_t hi, uint64_t lo, uint32_t b) { return b < 64 ? (lo >> b) & 1 : (hi >> (b - 64)) & 1; }
__device__ __forceinline__ bool u128_any_below(uint64_t hi, uint64_t lo, uint32_t b) {   // any bit in [0, b) set, b <= 127
    if (b == 0) return false;
    if (b <= 64) return (b == 64 ? lo : (lo & ((1ull << b) - 1))) != 0;
    return lo != 0 || (hi & ((1ull << (b - 64)) - 1)) != 0;
}
__device__ __forceinline__ void u128_shr(uint64_t hi, uint64_t lo, uint32_t s, uint64_t *rh, uint64_t *rl) {   // s in [1, 127]
    if (s >= 64) { *rh = 0; *rl = (s == 64) ? hi : hi >> (s - 64); }
    else { *rh = hi >> s; *rl = (lo >> s) | (hi << (64 - s)); }
}

// N = round-half-even(|x| * 10^8) as 128 bits for |x| < 2^52 (e < 0); returns false when x is an integer >= 2^52
// (then N = m * 2^e * 10^8 is produced by the limb path).
__device__ __forceinline__ bool scaled_u128(uint64_t m, int e, uint64_t *nh, uint64_t *nl) {
    if (e >= 0) return false;
    const uint64_t lo = m * 100000000ull, hi = __umul64hi(m, 100000000ull);
    const uint32_t s = (uint32_t)(-e);
    if (s >= 128) { *nh = 0; *nl = 0; return true; }          // m * 10^8 < 2^80 <= 2^(s-1): rounds to 0
    uint64_t h, l;
    u128_shr(hi, lo, s, &h, &l);
    const bool guard = u128_bit(hi, lo, s - 1);
    if (guard && (u128_any_below(hi, lo, s - 1) || (l & 1))) { l++; if (l == 0) h++; }
    *nh = h; *nl = l;
    return true;
}

__device__ __forceinline__ Fmt8 fmt8_prepare(double x) {
    Fmt8 f;
    const uint64_t bits = (uint64_t)__double_as_longlong(x);
    const uint32_t ex = (uint32_t)(bits >> 52) & 0x7ffu;
    const uint64_t frac = bits & 0x000FFFFFFFFFFFFFull;
    f.neg = (uint8_t)(bits >> 63); f.big = 0; f.ip = 0; f.fp = 0; f.nd = 1;
    if (ex == 0x7ff) { f.big = frac ? 3 : 2; f.len = f.neg + 3 + 1; return f; }
    const uint64_t m = ex ? (frac | 0x0010000000000000ull) : frac;
    const int e = ex ? (int)ex - 1075 : -1074;
    uint64_t nh, nl;
    if (!scaled_u128(m, e, &nh, &nl) || nh != 0) { f.big = 1; f.len = 0; return f; }
    f.ip = nl / 100000000ull;
    f.fp = (uint32_t)(nl - f.ip * 100000000ull);
    f.nd = (uint8_t)((f.ip < 10) ? 1 : dec_digits_u64(f.ip));
    f.len = f.neg + f.nd + 1 + 8 + 1;
    return f;
}

// Multi-limb path: decimal digits of |x| * 10^8 rounded (x < 2^52 but N >= 2^64) or of the integer m * 2^e.
// WRITE = false: returns the length only. dst receives sign, digits, '.', 8 digits, terminator.
template <bool WRITE>
__device__ __noinline__ uint32_t fmt8_big(double x, char term, char *dst) {
    const uint64_t bits = (uint64_t)__double_as_longlong(x);
    const uint32_t ex = (uint32_t)(bits >> 52) & 0x7ffu;
    const uint64_t frac = bits & 0x000FFFFFFFFFFFFFull;
    const bool neg = bits >> 63;
    const uint64_t m = ex ? (frac | 0x0010000000000000ull) : frac;
    const int e = ex ? (int)ex - 1075 : -1074;
    uint32_t L[36];                 // little-endian 32-bit limbs of the integer to print
    int nl = 0;
    bool scaled;                    // the integer already carries the 8 fractional digits
    uint64_t nh, nlo;
    if (scaled_u128(m, e, &nh, &nlo)) {
        scaled = true;
        L[0] = (uint32_t)nlo; L[1] = (uint32_t)(nlo >> 32); L[2] = (uint32_t)nh; L[3] = (uint32_t)(nh >> 32);
        nl = 4;
    } else {
        scaled = false;
        for (int i = 0; i < 36; i++) L[i] = 0;
        const int w = e >> 5, b = e & 31;   // m << e
        const uint64_t lo = m << b;
        const uint64_t hi = b ? (m >> (64 - b)) : 0;
        L[w] = (uint32_t)lo; L[w + 1] = (uint32_t)(lo >> 32); L[w + 2] = (uint32_t)hi;
        nl = w + 3;
    }
    while (nl > 0 && L[nl - 1] == 0) nl--;
    // base 10^9 chunks, least significant first
    uint32_t chunk[36];
    int nc = 0;
    while (nl > 0) {
        uint64_t rem = 0;
        for (int i = nl - 1; i >= 0; i--) {
            const uint64_t cur = (rem << 32) | L[i];
            L[i] = (uint32_t)(cur / 1000000000ull);
            rem = cur - (uint64_t)L[i] * 1000000000ull;
        }
        chunk[nc++] = (uint32_t)rem;
        while (nl > 0 && L[nl - 1] == 0) nl--;
    }
    if (nc == 0) { chunk[0] = 0; nc = 1; }
    uint32_t top = chunk[nc - 1], topd = 1;
    while (top >= 10) { top /= 10; topd++; }
    const uint32_t D = (uint32_t)(nc - 1) * 9 + topd;      // decimal digits of the integer
    // digit i (0 = most significant) of the integer
    auto digit = [&](uint32_t i) {
        const uint32_t fromlsd = D - 1 - i;
        uint32_t c = chunk[fromlsd / 9];
        for (uint32_t k = fromlsd % 9; k; k--) c /= 10;
        return (char)('0' + c % 10);
    };
    uint32_t len = neg ? 1u : 0u;
    if (WRITE && neg) dst[0] = '-';
    if (scaled) {            // D > 8 here (N >= 2^64)
        const uint32_t ni = D - 8;
        if (WRITE) {
            for (uint32_t i = 0; i < ni; i++) dst[len + i] = digit(i);
            dst[len + ni] = '.';
            for (uint32_t i = 0; i < 8; i++) dst[len + ni + 1 + i] = digit(ni + i);
        }
        len += ni + 9;
    } else {
        if (WRITE) {
            for (uint32_t i = 0; i < D; i++) dst[len + i] = digit(i);
            dst[len + D] = '.';
            for (uint32_t i = 0; i < 8; i++) dst[len + D + 1 + i] = '0';
        }
        len += D + 9;
    }
    if (WRITE) dst[len] = term;
    return len + 1;
}

__device__ __forceinline__ uint32_t fmt8_len(double x) {
    const Fmt8 f = fmt8_prepare(x);
    return f.big == 1 ? fmt8_big<false>(x, ',', nullptr) : f.len;
}

// Writes the prepared sample (f.len characters) at dst.
__device__ __forceinline__ void fmt8_write(const Fmt8 &f, double x, char term, char *dst) {
    if (f.big == 1) { fmt8_big<true>(x, term, dst); return; }
    uint32_t p = 0;
    if (f.neg) dst[p++] = '-';
    if (f.big) {
        const char *s = (f.big == 2) ? "inf" : "nan";
        dst[p] = s[0]; dst[p + 1] = s[1]; dst[p + 2] = s[2]; dst[p + 3] = term;
        return;
    }
    if (f.ip < 0x100000000ull) {
        uint32_t v = (uint32_t)f.ip;
        for (int i = (int)f.nd - 1; i >= 0; i--) { const uint32_t q = v / 10; dst[p + i] = (char)('0' + (v - q * 10)); v = q; }
    } else {
        uint64_t v = f.ip;
        for (int i = (int)f.nd - 1; i >= 0; i--) { const uint64_t q = v / 10; dst[p + i] = (char)('0' + (uint32_t)(v - q * 10)); v = q; }
    }
    p += f.nd;
    dst[p] = '.';
    uint32_t v = f.fp;
#pragma unroll
    for (int i = 8; i >= 1; i--) { const uint32_t q = v / 10; dst[p + i] = (char)('0' + (v - q * 10)); v = q; }
    dst[p + 9] = term;
}

// ------------------------------------------------------------------------------------------------
// Kernels. One warp per selected event (grid-stride), lanes over its samples.
struct TextIO {
    const double *samples;
    const uint4 *sel;            // selected events (EvRow): .z & 0xffffff = samples of the event
    const unsigned long long *sel_off;
    const uint32_t *n_sel;
    uint32_t *ev_tlen;                // [E_sel] characters of the event
    unsigned long long *ev_toff;      // [E_sel+1] exclusive scan of ev_tlen
    char *text;
    unsigned long long text_cap;
};

// One warp per event, two events ahead: the row (length, offset) of event q+2 and the first 32 values of event q+1 are
// requested before event q is formatted — a warp's events are otherwise a chain of two dependent trips to memory each.
struct TextAhead {
    uint32_t q, ln;
    unsigned long long off, toff;
};
template <bool WITH_TOFF>
__device__ __forceinline__ TextAhead text_row(const TextIO &io, uint32_t q, uint32_t n) {
    TextAhead r;
    r.q = q; r.ln = 0; r.off = 0; r.toff = 0;
    if (q < n) {
        r.ln = io.sel[q].z & 0x00ffffffu; r.off = io.sel_off[q];
        if (WITH_TOFF) r.toff = io.ev_toff[q];
    }
    return r;
}

__global__ void __launch_bounds__(PGM_BLOCK) k_text_len(TextIO io) {
    const uint32_t n = *io.n_sel;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t nwarps = gridDim.x * PGM_WARPS;
    const uint32_t q0 = blockIdx.x * PGM_WARPS + (threadIdx.x >> 5);
    if (q0 >= n) return;
    TextAhead A = text_row<false>(io, q0, n), B = text_row<false>(io, q0 + nwarps, n);
    double vA = (lane < A.ln) ? io.samples[A.off + lane] : 0.0;
    while (A.q < n) {
        const double vB = (lane < B.ln) ? io.samples[B.off + lane] : 0.0;
        const TextAhead C = text_row<false>(io, B.q + nwarps, n);
        const double *x = io.samples + A.off;
        uint32_t s = (lane < A.ln) ? fmt8_len(vA) : 0u;
        for (uint32_t t = lane + 32; t < A.ln; t += 32) s += fmt8_len(x[t]);
        s = warp_sum_t<uint32_t>(s);
        if (lane == 0) io.ev_tlen[A.q] = s;
        A = B; vA = vB; B = C;
    }
}

__global__ void __launch_bounds__(PGM_BLOCK) k_scan_text(const uint32_t *ev_tlen, unsigned long long *ev_toff, const uint32_t *n_sel_ptr,
                                                         unsigned long long *n_text_out, LbBuf st) {
    const uint32_t n = *n_sel_ptr;
    lb_scan<unsigned long long>(
        n, st, [&](uint64_t i) { return (unsigned long long)ev_tlen[i]; },
        [&](uint64_t i, unsigned long long ex, unsigned long long) { ev_toff[i] = ex; },
        [&](unsigned long long tot) { ev_toff[n] = tot; *n_text_out = tot; });
}

__global__ void __launch_bounds__(PGM_BLOCK) k_text_write(TextIO io) {
    const uint32_t n = *io.n_sel;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t nwarps = gridDim.x * PGM_WARPS;
    const uint32_t q0 = blockIdx.x * PGM_WARPS + (threadIdx.x >> 5);
    if (q0 >= n) return;
    TextAhead A = text_row<true>(io, q0, n), B = text_row<true>(io, q0 + nwarps, n);
    double vA = (lane < A.ln) ? io.samples[A.off + lane] : 0.0;
    while (A.q < n) {
        const double vB = (lane < B.ln) ? io.samples[B.off + lane] : 0.0;
        const TextAhead C = text_row<true>(io, B.q + nwarps, n);
        const uint32_t ln = A.ln;
        const double *x = io.samples + A.off;
        unsigned long long pos = A.toff;
        for (uint32_t t0 = 0; t0 < ln; t0 += 32) {
            const uint32_t t = t0 + lane;
            Fmt8 f;
            double v = 0.0;
            uint32_t len = 0;
            if (t < ln) {
                v = t0 ? x[t] : vA;
                f = fmt8_prepare(v);
                len = f.big == 1 ? fmt8_big<false>(v, ',', nullptr) : f.len;
            }
            uint32_t incl = len;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= (uint32_t)d) incl += u;
            }
            const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
            const unsigned long long at = pos + (incl - len);
            if (t < ln && at + len <= io.text_cap) fmt8_write(f, v, (t + 1 == ln) ? ';' : ',', io.text + at);   // gmove.cpp:674-678
            pos += tot;
        }
        A = B; vA = vB; B = C;
    }
}

// slot_text_off[s] = first character of slot s = text offset of its first selected event
__global__ void __launch_bounds__(PGM_BLOCK) k_slot_text_off(const uint32_t *sel_base, const unsigned long long *ev_toff, uint32_t n_slots,
                                                             unsigned long long *out) {
    const uint32_t s = blockIdx.x * PGM_BLOCK + threadIdx.x;
    if (s <= n_slots) out[s] = ev_toff[sel_base[s]];
}

// Test hook: n doubles -> one event's text (',' between, ';' at the end), single warp.
__global__ void k_text_debug(const double *x, uint32_t n, char *text, unsigned long long *len_out) {
    const uint32_t lane = threadIdx.x & 31;
    unsigned long long pos = 0;
    for (uint32_t t0 = 0; t0 < n; t0 += 32) {
        const uint32_t t = t0 + lane;
        Fmt8 f;
        double v = 0.0;
        uint32_t len = 0;
        if (t < n) { v = x[t]; f = fmt8_prepare(v); len = f.big == 1 ? fmt8_big<false>(v, ',', nullptr) : f.len; }
        uint32_t incl = len;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += u;
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        if (t < n && text) fmt8_write(f, v, (t + 1 == n) ? ';' : ',', text + pos + (incl - len));
        pos += tot;
    }
    if (lane == 0) *len_out = pos;
}
