/* synth_fast.c — counter-based generator of synthetic move-table reads (the C3 / C5 shape of SURVEY.md §8d:
 * DNA, 5 kHz, stride 6, ~12.5 samples per base) for bench.py and the full-size tests.
 *
 * Input manufacturing only: nothing here is on the gmove path. Every value is a pure integer function of
 * (seed, global read index, base or sample index), so any range of reads can be produced on any rank or
 * thread and is the same bytes everywhere (the reference arm's prefix is a prefix of the GPU arm's reads).
 * The numpy generator (synth.make_batch) stays the one for the small parity cases; at 100 ns per sample it
 * cannot make 5e9 samples inside a bench run.
 *
 * Per read r (global index), with h(a, i) = mix64(a + i * K):
 *   base i    : code = h(Ab, i) & 3, dwell d_i = 1 + #{j : u_i < T[j]} strides (T = geometric tail table from
 *               the caller), u_i = bits 32..63 of the same word
 *   moves     : '1' followed by d_i - 1 '0's per base (mv:B:c style int8 0/1 when moves_char == 0)
 *   samples   : base i owns d_i * stride samples, value = L[6-mer at i] - offset_code[r] + noise, noise = the sum
 *               of four 16-bit uniforms, centred, times noise_num / 2^20; with probability
 *               spike_thr / 2^32 the sample is a spike (lo_code or hi_code: outside [pa_min, pa_max])
 *   tail      : tail_r < stride filler samples at the filler level; raw_len = sum d_i * stride + tail_r
 */
#include <stdint.h>
#include <string.h>

static inline uint64_t mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}
#define K_STEP 0x9E3779B97F4A7C15ull

typedef struct {
    uint64_t seed;
    uint32_t stride;        /* samples per move */
    uint32_t n_thr;         /* entries in dwell_thr */
    const uint32_t *dwell_thr; /* dwell = 1 + #{j < n_thr : u < dwell_thr[j]} */
    const int32_t *level;   /* [4096] level of every 6-mer, in codes (before the read's offset) */
    int32_t filler, lo_code, hi_code; /* codes before the read's offset */
    uint32_t spike_thr;     /* P(spike) * 2^32 */
    int32_t noise_num;      /* noise codes = (sum4 - 131070) * noise_num >> 20 */
    uint32_t moves_char;    /* 1: '0'/'1' bytes, 0: 0/1 */
} synth_params;

static inline uint64_t read_key(const synth_params *p, uint64_t r, uint64_t salt) { return mix64(p->seed ^ mix64(r * K_STEP + salt)); }

static inline uint32_t dwell_of(const synth_params *p, uint64_t w) {
    const uint32_t u = (uint32_t)(w >> 32);
    uint32_t d = 1;
    for (uint32_t j = 0; j < p->n_thr; j++) d += (u < p->dwell_thr[j]);
    return d;
}

/* Pass 1: moves per read (= sum of dwells) for reads [read_lo, read_lo + n); n_bases[i] given. */
void pgm_synth_moves_len(const synth_params *p, uint64_t read_lo, uint32_t n, const uint32_t *n_bases, uint64_t *moves_out) {
#pragma omp parallel for schedule(dynamic, 16)
    for (uint32_t q = 0; q < n; q++) {
        const uint64_t ab = read_key(p, read_lo + q, 1);
        uint64_t m = 0;
        for (uint32_t i = 0; i < n_bases[q]; i++) m += dwell_of(p, mix64(ab + (uint64_t)i * K_STEP));
        moves_out[q] = m;
    }
}

/* Pass 2: fill bases / moves / samples of reads [read_lo, read_lo + n) at the caller's offsets.
 * raw_cap[q] = samples reserved for the read (>= raw_len, the rest is filled with the filler level). */
void pgm_synth_moves_fill(const synth_params *p, uint64_t read_lo, uint32_t n, const uint32_t *n_bases, const int32_t *offset_code,
                          const uint64_t *raw_off, const uint64_t *raw_cap, const uint64_t *align_off,
                          const uint64_t *seq_off, int16_t *raw, uint8_t *align, uint8_t *bases) {
    static const char alpha[4] = {'A', 'C', 'G', 'T'};
    const uint8_t one = p->moves_char ? '1' : 1, zero = p->moves_char ? '0' : 0;
#pragma omp parallel for schedule(dynamic, 4)
    for (uint32_t q = 0; q < n; q++) {
        const uint64_t ab = read_key(p, read_lo + q, 1), as = read_key(p, read_lo + q, 2);
        const uint32_t nb = n_bases[q];
        uint8_t *bq = bases + seq_off[q];
        uint8_t *mq = align + align_off[q];
        int16_t *rq = raw + raw_off[q];
        const int32_t off = offset_code[q];
        /* bases first: the level of base i looks five bases ahead (wrapping inside the read) */
        for (uint32_t i = 0; i < nb; i++) bq[i] = (uint8_t)(mix64(ab + (uint64_t)i * K_STEP) & 3);
        uint64_t mpos = 0, spos = 0;
        uint32_t kid = 0;
        for (uint32_t t = 0; t < 6; t++) kid = (kid << 2) | bq[t % nb];
        for (uint32_t i = 0; i < nb; i++) {
            const uint64_t w = mix64(ab + (uint64_t)i * K_STEP);
            const uint32_t d = dwell_of(p, w);
            mq[mpos++] = one;
            for (uint32_t j = 1; j < d; j++) mq[mpos++] = zero;
            const int32_t lev = p->level[kid & 4095] - off;
            const uint32_t ns = d * p->stride;
            for (uint32_t j = 0; j < ns; j++, spos++) {
                const uint64_t z = mix64(as + spos * K_STEP);
                const int32_t s4 = (int32_t)((z & 0xffff) + ((z >> 16) & 0xffff) + ((z >> 32) & 0xffff) + ((z >> 48) & 0xffff)) - 131070;
                int32_t v = lev + (int32_t)(((int64_t)s4 * p->noise_num) >> 20);
                const uint32_t u = (uint32_t)((z * 0xD6E8FEB86659FD93ull) >> 32);
                if (u < p->spike_thr) v = ((u & 1) ? p->hi_code : p->lo_code) - off;
                rq[spos] = (int16_t)(v < -32768 ? -32768 : v > 32767 ? 32767 : v);
            }
            kid = (kid << 2) | bq[(i + 6) % nb];
        }
        for (; spos < raw_cap[q]; spos++) {   /* the tail samples and the padding up to the next read */
            const int32_t v = p->filler - off;
            rq[spos] = (int16_t)(v < -32768 ? -32768 : v > 32767 ? 32767 : v);
        }
        for (uint32_t i = 0; i < nb; i++) bq[i] = (uint8_t)alpha[bq[i]];
    }
}
