/* oracle/gmove_oracle.c — TEST INFRASTRUCTURE ONLY. See gmove_oracle.h.
 *
 * Record-at-a-time, single-threaded, IEEE double, exactly like the reference
 * (/root/reference/src/gmove.cpp; all citations below are to that file unless stated).
 * Compile WITHOUT -ffast-math / -march=native and with -ffp-contract=off so that
 * (raw+offset)*(range/digitisation) and (x-med)/mad round like the reference's x86-64 build.
 *
 * Where the reference has undefined behaviour the restatement (and the CUDA path) defines it,
 * and says so:
 *   U1 fewer than m+1 set moves (623-629 runs off the string; 1185-1192 never terminates)
 *      -> record flagged PGM_RF_BAD_ALIGN, nothing emitted, no ':'.
 *   U2 --margin larger than the event start (653: unsigned `<0` never true, 656 wraps)
 *      -> start clamps to 0 (the evident intent of 653-654).
 *   U3 event whose clamped range is empty or reversed (664 builds a vector from a reversed
 *      range, 674 underflows) -> event is dropped like a duration-filter miss.
 *   U4 k-mer start beyond the sequence end (634 substr throws std::out_of_range) -> no match.
 *   U5 ss: refined length < k (891 unsigned wrap) -> no events; event index past the decoded
 *      events (892 reads beyond the -1 padding when past `cap`) -> skipped like a -1 cell.
 *   U6 pick_this_kmer reads indel_pos[-1] at i==0 (206) when the left test passes, which needs a
 *      negative margin -> that term is false.
 *   U7 ss with a bad character / missing digits / >9 digits (834,838,867 exit(1); buff[11]
 *      overflow) -> PGM_RF_BAD_ALIGN, nothing emitted (the CLI turns it into the same exit).
 *   U8 ±0.0 ties inside the order statistic: the value returned is the same, its sign of zero is
 *      whichever the selection leaves; only reachable with pa_min <= 0.
 */
#include "gmove_oracle.h"
#include <math.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---------- growable arrays ---------- */
typedef struct { uint32_t *len, *read, *start; size_t n, cap; double *s; size_t ns, caps; } slot_buf;

static void sb_push_event(slot_buf *b, uint32_t len, uint32_t read, uint32_t start, const double *x) {
    if (b->n == b->cap) {
        b->cap = b->cap ? b->cap * 2 : 8;
        b->len = (uint32_t *)realloc(b->len, b->cap * sizeof(uint32_t));
        b->read = (uint32_t *)realloc(b->read, b->cap * sizeof(uint32_t));
        b->start = (uint32_t *)realloc(b->start, b->cap * sizeof(uint32_t));
    }
    b->len[b->n] = len; b->read[b->n] = read; b->start[b->n] = start; b->n++;
    if (b->ns + len > b->caps) {
        while (b->ns + len > b->caps) b->caps = b->caps ? b->caps * 2 : 64;
        b->s = (double *)realloc(b->s, b->caps * sizeof(double));
    }
    memcpy(b->s + b->ns, x, (size_t)len * sizeof(double));
    b->ns += len;
}

/* ---------- order statistic (ksort.h:233-259 ks_ksmall: k-th smallest, 0-based) ---------- */
static double kth_smallest(double *a, size_t n, size_t k) {
    ptrdiff_t lo = 0, hi = (ptrdiff_t)n - 1, kk = (ptrdiff_t)k;
    while (lo < hi) {
        double p = a[lo + (hi - lo) / 2];
        ptrdiff_t i = lo, j = hi;
        while (i <= j) {
            while (a[i] < p) i++;
            while (p < a[j]) j--;
            if (i <= j) { double t = a[i]; a[i] = a[j]; a[j] = t; i++; j--; }
        }
        if (kk <= j) hi = j; else if (kk >= i) lo = i; else return a[k];
    }
    return a[k];
}

/* calc_median 142-150: UPPER median ksmall(n, n/2) on a copy */
double oracle_median(const double *x, size_t n) {
    double *c = (double *)malloc((n ? n : 1) * sizeof(double));
    memcpy(c, x, n * sizeof(double));
    double m = kth_smallest(c, n, n / 2);
    free(c);
    return m;
}

/* calc_madf 161-184: 1.4826 * median(|x - med|); n==1 -> 0.0 */
double oracle_mad(const double *x, size_t n, double med) {
    if (n == 1) return 0.0;
    double *d = (double *)malloc((n ? n : 1) * sizeof(double));
    for (size_t i = 0; i < n; i++) d[i] = fabs(x[i] - med);
    double m = kth_smallest(d, n, n / 2);
    free(d);
    return m * 1.4826;
}

/* pick_this_kmer 204-211 (U6) */
int oracle_pick_kmer(const int *indel_pos, size_t n, int left, int k, int margin) {
    for (size_t i = 1; i < n; i++) {
        if ((int)((uint32_t)left + (uint32_t)k + (uint32_t)margin) <= indel_pos[i] &&
            indel_pos[i - 1] <= (int)((uint32_t)left - (uint32_t)margin)) return 1;
    }
    return 0;
}

/* ---------- k-mer table: string -> slot (std::map<std::string,...>::find, 647, 922, 1210) ---------- */
typedef struct { const char *s; uint32_t slot; } kent;
static uint32_t g_k;
static int kent_cmp(const void *a, const void *b) { return memcmp(((const kent *)a)->s, ((const kent *)b)->s, g_k); }

typedef struct { kent *e; char *store; uint32_t n, k; } ktable;

static int ktable_build(ktable *t, const pgm_opts *o) {
    t->n = o->n_slots; t->k = o->kmer_size;
    t->e = (kent *)malloc((size_t)(t->n ? t->n : 1) * sizeof(kent));
    t->store = (char *)malloc((size_t)(t->n ? t->n : 1) * t->k);
    if (o->kmer_list) {
        memcpy(t->store, o->kmer_list, (size_t)t->n * t->k);
    } else {
        /* generate_kmers poregen.cpp:248-267: lexicographic over A<C<G<T|U; index == base-4 value */
        const char *al = o->flag_rna ? "ACGU" : "ACGT";
        for (uint32_t s = 0; s < t->n; s++) {
            uint64_t id = (uint64_t)o->index_lo + s;
            for (uint32_t j = 0; j < t->k; j++) t->store[(size_t)s * t->k + (t->k - 1 - j)] = al[(id >> (2 * j)) & 3];
        }
    }
    for (uint32_t s = 0; s < t->n; s++) { t->e[s].s = t->store + (size_t)s * t->k; t->e[s].slot = s; }
    g_k = t->k;
    qsort(t->e, t->n, sizeof(kent), kent_cmp);
    return 0;
}
static int64_t ktable_find(const ktable *t, const char *kmer) {
    size_t lo = 0, hi = t->n;
    while (lo < hi) {
        size_t mid = lo + (hi - lo) / 2;
        int c = memcmp(t->e[mid].s, kmer, t->k);
        if (c == 0) return t->e[mid].slot;
        if (c < 0) lo = mid + 1; else hi = mid;
    }
    return -1;
}
static void ktable_free(ktable *t) { free(t->e); free(t->store); }

/* ---------- run state ---------- */
typedef struct {
    const pgm_opts *o;
    const pgm_batch *b;
    ktable kt;
    slot_buf *slots;
    uint32_t *totals;
    uint64_t examined, passed;
} run_t;

/* The block shared by all three reference loops: duration filters, map lookup, quota, margin
 * widening, emission and counting (641-685, 916-950, 1204-1247). `kmer` may be NULL (U4).
 * n_sig is the length the end clamp uses (trimmed length in moves modes, full length for ss). */
static void consider_event(run_t *R, uint32_t read_idx, const double *x, uint64_t n_sig,
                           uint32_t start, uint32_t end, const char *kmer) {
    const pgm_opts *o = R->o;
    R->examined++;
    if ((uint32_t)(end - start) > o->max_dur) return;             /* 641, 916, 1204 */
    if ((uint32_t)(end - start) < o->min_dur) return;             /* 644, 919, 1207 */
    if (!kmer) return;
    int64_t slot = ktable_find(&R->kt, kmer);                       /* 647, 922, 1210 */
    if (slot < 0) return;
    if (R->totals[slot] >= o->sample_limit) return;                /* 650, 925, 1213 (== in the reference) */
    uint32_t s2 = (start < o->signal_print_margin) ? 0 : start - o->signal_print_margin; /* 653-657, U2 */
    uint32_t e2 = ((uint64_t)(uint32_t)(end + o->signal_print_margin) > n_sig) ? (uint32_t)n_sig
                                                                                : end + o->signal_print_margin; /* 658-662 */
    if (s2 >= e2) return;                                           /* U3 */
    R->passed++;
    sb_push_event(&R->slots[slot], e2 - s2, read_idx, s2, x + s2);  /* 664-678 */
    R->totals[slot]++;                                              /* 680 */
}

/* pA conversion + optional Med-MAD (592-614, 754-776, 1150-1176).
 * Returns 1 when MOVES_I8 must drop the read (1152-1160). x has raw_len cells, zero-filled;
 * only the first n are written (588, 751, 1145). */
static int convert_and_scale(const pgm_opts *o, const pgm_read *rd, const int16_t *raw, uint64_t first, uint64_t n,
                             int drop_on_oor, double *x) {
    for (uint64_t i = 0; i < n; i++) {
        double pA = ((double)raw[first + i] + rd->offset) * (rd->range / rd->digitisation); /* TO_PICOAMPS poregen.h:30 */
        if (pA < o->pa_min || pA > o->pa_max) {
            if (drop_on_oor) return 1;
            continue; /* stays 0.0 (594-596) */
        }
        x[i] = pA;
    }
    if (o->scaling == 1 && n > 0) {
        double med = oracle_median(x, n);                            /* 601 */
        double mad = oracle_mad(x, n, med);                          /* 605 */
        mad = (mad > 1.0) ? mad : 1.0;                               /* 609 */
        for (uint64_t i = 0; i < n; i++) x[i] = (x[i] - med) / mad;  /* 611-613 */
    }
    return 0;
}

static int calib_ok(const pgm_read *rd) {
    double s = rd->range / rd->digitisation;
    return isfinite(s) && isfinite(rd->offset);
}

/* process_move_table_file 557-700 and process_move_table_bam 1080-1261, one record */
static void do_moves_record(run_t *R, uint32_t ri, uint8_t *flag) {
    const pgm_opts *o = R->o; const pgm_batch *b = R->b; const pgm_read *rd = &b->reads[ri];
    const uint8_t one = (b->mode == PGM_MODE_MOVES_CHAR) ? (uint8_t)'1' : (uint8_t)1;
    if (rd->flags & PGM_READ_SKIP) { *flag |= PGM_RF_SKIPPED; return; }
    if (!calib_ok(rd)) { *flag |= PGM_RF_BAD_CALIB | PGM_RF_SKIPPED; return; }
    if (rd->trim >= rd->raw_len) { *flag |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED; return; } /* assert 590, 1147 */
    uint64_t n = (uint64_t)rd->raw_len - rd->trim;                  /* 591 */
    double *x = (double *)calloc(rd->raw_len, sizeof(double));      /* 588 */
    int drop = convert_and_scale(o, rd, b->raw, rd->raw_off + rd->trim, n, b->mode == PGM_MODE_MOVES_I8, x);
    if (drop) { *flag |= PGM_RF_SKIPPED; free(x); return; }        /* 1158-1160 */
    if (rd->fastq_len < 10) { *flag |= PGM_RF_SKIPPED; free(x); return; } /* 616-619, 1178-1181 */
    const uint8_t *mv = b->align + rd->align_off;
    const char *seq = (const char *)b->bases + rd->seq_off;
    uint64_t move_len = rd->align_len;
    /* 620-630: position of the (m+1)-th set move */
    uint64_t cnt = 0, idx = 0, start_idx = 0; int found = 0;
    for (idx = 0; idx < move_len; idx++) {
        if (mv[idx] == one) { cnt++; start_idx = idx; if (cnt == (uint64_t)o->sig_move_offset + 1) { found = 1; break; } }
    }
    if (!found) { *flag |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED; free(x); return; } /* U1 */
    uint64_t seq_start = o->kmer_start_offset;                       /* 631 */
    for (idx = start_idx + 1; idx < move_len; idx++) {              /* 632: the cell at move_len is NUL / ERANGE 0 */
        if (mv[idx] != one) continue;
        const char *kmer = (seq_start + o->kmer_size <= rd->seq_len) ? seq + seq_start : NULL; /* 634; short substr never matches; U4 */
        uint32_t rs = (uint32_t)(start_idx * rd->stride);            /* 636 */
        uint32_t re = (uint32_t)(idx * rd->stride);                  /* 637 */
        start_idx = idx; seq_start++;
        consider_event(R, ri, x, n, rs, re, kmer);
        /* 687-689 `break` once the next k-mer would be short: later events could only see short
         * k-mers, which never match, so it has no observable effect; not restated. */
    }
    free(x);
}

/* process_move_table_paf 732-969, one record */
static void do_ss_record(run_t *R, uint32_t ri, uint8_t *flag) {
    const pgm_opts *o = R->o; const pgm_batch *b = R->b; const pgm_read *rd = &b->reads[ri];
    if (rd->flags & PGM_READ_SKIP) { *flag |= PGM_RF_SKIPPED; return; }   /* 806-808 `continue` */
    if (!calib_ok(rd)) { *flag |= PGM_RF_BAD_CALIB | PGM_RF_SKIPPED; return; }
    if (rd->raw_len == 0) { *flag |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED; return; }      /* assert 752 */
    const int rna = (rd->flags & PGM_READ_RNA) ? 1 : 0;                  /* 793 */
    const uint64_t n = rd->raw_len;                                       /* 750: the FULL signal, no trim */
    const char *ss = (const char *)b->align + rd->align_off;
    const uint32_t ss_len = rd->align_len;
    const int fetched_len = (int)rd->seq_len;
    char *seq = (char *)malloc((size_t)rd->seq_len + 1);
    memcpy(seq, b->bases + rd->seq_off, rd->seq_len);
    if (rna) for (uint32_t i = 0; i < rd->seq_len; i++) if (seq[i] == 'T') seq[i] = 'U'; /* 815-817 */

    /* 831-871 ss walk */
    uint32_t cap_ev = ss_len / 2 + 2;
    uint32_t *ev_start = (uint32_t *)malloc(cap_ev * sizeof(uint32_t));
    uint32_t *ev_end = (uint32_t *)malloc(cap_ev * sizeof(uint32_t));
    char *refined = (char *)malloc(cap_ev + 1);
    int *indel = (int *)malloc(((size_t)cap_ev + 3) * sizeof(int));
    size_t n_indel = 0;
    if (!rna) indel[n_indel++] = -rd->st_k;                               /* 824-826 */
    uint64_t i_k = 0, i_raw = rd->query_start, num_del = 0; uint32_t n_ev = 0;
    uint64_t num = 0; int ndig = 0, bad = 0;
    for (uint32_t p = 0; p < ss_len && !bad; p++) {
        char c = ss[p];
        if (c == ',' || c == 'I' || c == 'D') {
            if (ndig <= 0 || ndig > 9) { bad = 1; break; }                /* 834, U7 */
            if (c == 'I') { i_raw += num; indel[n_indel++] = (int)(i_k - num_del); }        /* 841-843 */
            else if (c == 'D') { indel[n_indel++] = (int)(i_k - num_del); i_k += num; num_del += num; } /* 844-847 */
            else {                                                         /* 848-857 */
                int64_t si = rna ? (int64_t)fetched_len - (int64_t)i_k - 1 : (int64_t)i_k;
                refined[n_ev] = (si >= 0 && si < fetched_len) ? seq[si] : '\0'; /* out-of-slice base never matches */
                ev_start[n_ev] = (uint32_t)i_raw; i_raw += num; ev_end[n_ev] = (uint32_t)i_raw;
                i_k++; n_ev++;
            }
            num = 0; ndig = 0;
        } else if (c >= '0' && c <= '9') {
            if (ndig < 10) { num = num * 10 + (uint64_t)(c - '0'); ndig++; } /* U7 is decided at the operator */
        } else { bad = 1; break; }                                        /* 867 */
    }
    if (bad) { *flag |= PGM_RF_BAD_ALIGN | PGM_RF_SKIPPED; goto done; }
    {
        const int L = (int)n_ev;                                           /* 872 */
        if (rna) {                                                         /* 877-884 */
            for (size_t t = 0; t < n_indel; t++) indel[t] = L - indel[t];
            indel[n_indel++] = -rd->st_k;
            for (size_t a = 0, z = n_indel - 1; a < z; a++, z--) { int t = indel[a]; indel[a] = indel[z]; indel[z] = t; }
            for (int a = 0, z = L - 1; a < z; a++, z--) { char t = refined[a]; refined[a] = refined[z]; refined[z] = t; }
        }
        indel[n_indel++] = (int)((uint32_t)rd->end_k + o->kmer_pick_margin); /* 885 */

        double *x = (double *)calloc(rd->raw_len, sizeof(double));        /* 751 */
        convert_and_scale(o, rd, b->raw, rd->raw_off, n, 0, x);            /* 754-776 */
        const int k = (int)o->kmer_size;
        if (L >= k) {                                                      /* U5 */
            for (int i = 0; i <= L - k; i++) {                             /* 891 */
                uint64_t j = (uint64_t)i + o->sig_move_offset;
                if (j >= n_ev) continue;                                   /* 892-895, U5 */
                int left = rna ? L - i - k : i;                            /* 899, 906 */
                const char *kmer = refined + left;
                R->examined++;
                if (!oracle_pick_kmer(indel, n_indel, left, k, (int)o->kmer_pick_margin)) continue; /* 901, 908 */
                R->examined--; /* consider_event counts it */
                consider_event(R, ri, x, n, ev_start[j], ev_end[j], kmer);
                /* 951-953 break: i+k > fetched length cannot hold for i <= L-k <= fetched-k unless the
                 * ss has more ',' than target bases; then the remaining k-mers contain '\0' and never match. */
            }
        }
        free(x);
    }
done:
    free(seq); free(ev_start); free(ev_end); free(refined); free(indel);
}

/* Full-size checks compare the doubles, not the text: formatting 6e8 samples with snprintf would take minutes. */
static int g_want_text = 1;
void oracle_set_text(int on) { g_want_text = on; }

int oracle_gmove_batch(const pgm_opts *opts, const pgm_batch *batch, uint32_t *totals, oracle_result *out) {
    memset(out, 0, sizeof(*out));
    if (!opts || !batch || !totals || opts->kmer_size < 1 || opts->kmer_size > 15) return PGM_ERR_ARG;
    run_t R; memset(&R, 0, sizeof(R));
    R.o = opts; R.b = batch; R.totals = totals;
    ktable_build(&R.kt, opts);
    R.slots = (slot_buf *)calloc(opts->n_slots ? opts->n_slots : 1, sizeof(slot_buf));
    out->n_slots = opts->n_slots; out->n_reads = batch->n_reads;
    out->read_flags = (uint8_t *)calloc(batch->n_reads ? batch->n_reads : 1, 1);
    for (uint32_t ri = 0; ri < batch->n_reads; ri++) {
        if (batch->mode == PGM_MODE_SS) do_ss_record(&R, ri, &out->read_flags[ri]);
        else do_moves_record(&R, ri, &out->read_flags[ri]);
    }
    /* flatten, grouped by slot */
    uint64_t ne = 0, ns = 0;
    for (uint32_t s = 0; s < opts->n_slots; s++) { ne += R.slots[s].n; ns += R.slots[s].ns; }
    out->n_events_examined = R.examined; out->n_events_pass = R.passed;
    out->n_sel_events = ne; out->n_sel_samples = ns;
    out->slot_sel = (uint32_t *)calloc(opts->n_slots ? opts->n_slots : 1, sizeof(uint32_t));
    out->slot_total = (uint32_t *)calloc(opts->n_slots ? opts->n_slots : 1, sizeof(uint32_t));
    out->ev_len = (uint32_t *)malloc((ne ? ne : 1) * sizeof(uint32_t));
    out->ev_read = (uint32_t *)malloc((ne ? ne : 1) * sizeof(uint32_t));
    out->ev_start = (uint32_t *)malloc((ne ? ne : 1) * sizeof(uint32_t));
    out->samples = (double *)malloc((ns ? ns : 1) * sizeof(double));
    out->slot_text_off = (uint64_t *)calloc((size_t)opts->n_slots + 1, sizeof(uint64_t));
    /* text: "%.8f," ... "%.8f;" per event (674-678) */
    size_t tcap = g_want_text ? (size_t)ns * 14 + 16 : 16, tl = 0;
    out->text = (char *)malloc(tcap);
    uint64_t pe = 0, ps = 0;
    for (uint32_t s = 0; s < opts->n_slots; s++) {
        slot_buf *sb = &R.slots[s];
        out->slot_sel[s] = (uint32_t)sb->n; out->slot_total[s] = totals[s];
        out->slot_text_off[s] = tl;
        size_t q = 0;
        for (size_t e = 0; e < sb->n; e++) {
            out->ev_len[pe] = sb->len[e]; out->ev_read[pe] = sb->read[e]; out->ev_start[pe] = sb->start[e]; pe++;
            for (uint32_t t = 0; g_want_text && t < sb->len[e]; t++, q++) {
                if (tl + 400 > tcap) { tcap = tcap * 2 + 1024; out->text = (char *)realloc(out->text, tcap); }
                tl += (size_t)snprintf(out->text + tl, 390, (t + 1 == sb->len[e]) ? "%.8f;" : "%.8f,", sb->s[q]);
            }
        }
        memcpy(out->samples + ps, sb->s, sb->ns * sizeof(double)); ps += sb->ns;
        free(sb->len); free(sb->read); free(sb->start); free(sb->s);
    }
    out->slot_text_off[opts->n_slots] = tl;
    free(R.slots); ktable_free(&R.kt);
    return PGM_OK;
}

void oracle_free_result(oracle_result *r) {
    free(r->slot_sel); free(r->slot_total); free(r->ev_len); free(r->ev_read); free(r->ev_start);
    free(r->samples); free(r->read_flags); free(r->text); free(r->slot_text_off);
    memset(r, 0, sizeof(*r));
}

/* scripts/poregen.sh:54-85 (PARITY UNPINNED, see header) */
static int dbl_cmp(const void *a, const void *b) { double x = *(const double *)a, y = *(const double *)b; return (x > y) - (x < y); }
int oracle_model_row(const double *vals, size_t n, double limit, double *median, double *stddev) {
    if (n < 2) return PGM_ERR_ARG;            /* `tail -n +2` drops the first value (poregen.sh:73-74) */
    size_t m = n - 1;
    double *c = (double *)malloc(m * sizeof(double));
    memcpy(c, vals + 1, m * sizeof(double));
    qsort(c, m, sizeof(double), dbl_cmp);
    *median = (m & 1) ? c[m / 2] : (c[m / 2 - 1] + c[m / 2]) / 2.0;
    double mean = 0; for (size_t i = 0; i < m; i++) mean += c[i]; mean /= (double)m;
    double ss = 0; for (size_t i = 0; i < m; i++) ss += (c[i] - mean) * (c[i] - mean);
    double sd = (m > 1) ? sqrt(ss / (double)(m - 1)) : NAN;
    if (sd > limit) sd = limit;                /* poregen.sh:77 */
    *stddev = sd;
    free(c);
    return PGM_OK;
}

/* The same row computed the way the script really does it: from the TEXT of dump/<KMER>
 * (scripts/poregen.sh:73-74: `tr ';,' '\n' < file | tail -n +2 | datamash median 1` / `sstdev 1`).
 * Values are parsed with strtold, as GNU datamash does, and all arithmetic is long double (datamash's
 * field type); datamash prints with "%.14Lg" (its default numeric format), which is what *_str hold.
 * PARITY UNPINNED: GNU datamash is not installed here, its version is not pinned by the reference, and no
 * reference test holds a model row.  Returns the number of values used (after the drop), 0 when the
 * file has fewer than two values (datamash prints an empty line then). */
static int ld_cmp(const void *a, const void *b) { long double x = *(const long double *)a, y = *(const long double *)b; return (x > y) - (x < y); }
size_t oracle_model_row_text(const char *text, size_t len, double limit, double *median, double *stddev, char *median_str, char *stddev_str) {
    size_t cap = 1024, n = 0;
    long double *v = (long double *)malloc(cap * sizeof(long double));
    size_t i = 0;
    while (i < len) {
        size_t j = i;
        while (j < len && text[j] != ',' && text[j] != ';') j++;
        if (j > i) {
            char buf[512];
            size_t L = j - i < sizeof buf - 1 ? j - i : sizeof buf - 1;
            memcpy(buf, text + i, L); buf[L] = 0;
            if (n == cap) { cap *= 2; v = (long double *)realloc(v, cap * sizeof(long double)); }
            v[n++] = strtold(buf, NULL);
        }
        i = j + 1;
    }
    *median = NAN; *stddev = NAN;
    if (median_str) median_str[0] = 0;
    if (stddev_str) stddev_str[0] = 0;
    if (n < 2) { free(v); return 0; }
    size_t m = n - 1;
    long double *c = v + 1;                                   /* tail -n +2 */
    /* sstdev is a datamash run of its own (poregen.sh:74): its sums go over the values in FILE order */
    long double mean = 0; for (size_t k = 0; k < m; k++) mean += c[k]; mean /= (long double)m;
    long double ss = 0; for (size_t k = 0; k < m; k++) ss += (c[k] - mean) * (c[k] - mean);
    qsort(c, m, sizeof(long double), ld_cmp);
    long double med = (m & 1) ? c[m / 2] : (c[m / 2 - 1] + c[m / 2]) / 2.0L;
    long double sd = (m > 1) ? sqrtl(ss / (long double)(m - 1)) : (long double)NAN;
    *median = (double)med;
    if (median_str) sprintf(median_str, "%.14Lg", med);
    if (sd > (long double)limit) { sd = limit; }              /* poregen.sh:77-79 */
    *stddev = (double)sd;
    if (stddev_str) sprintf(stddev_str, "%.14Lg", sd);
    free(v);
    return m;
}

/* ---------- svb-zd signal stream ----------
 * slow5lib/src/slow5_press.c:1085-1130 (ptr_depress_svb / ptr_depress_svb_zd): u32 count, then
 * thirdparty/streamvbyte/src/streamvbyte_decode.c:57-104 (keys: 2 bits per value, (count+3)/4 bytes; data:
 * code+1 little-endian bytes per value, 36-56) and streamvbyte_zigzag.c:19-40 (zig-zag, running sum from 0,
 * stored as int16). Returns 0, or -1 where slow5lib reports an error (count differs from what the caller expects /
 * the decoder does not consume exactly nbytes-4 bytes, slow5_press.c:1097-1103). */
int oracle_svb_zd_decode(const uint8_t *p, size_t nbytes, uint32_t expect, int16_t *out) {
    if (nbytes < 4) return -1;
    uint32_t count;
    memcpy(&count, p, 4);
    if (count != expect) return -1;
    const size_t keylen = ((size_t)count + 3) / 4;
    if (4 + keylen > nbytes) return -1;
    const uint8_t *keys = p + 4, *data = keys + keylen, *end = p + nbytes;
    int32_t prev = 0;
    for (uint32_t i = 0; i < count; i++) {
        const unsigned code = (keys[i >> 2] >> (2 * (i & 3))) & 3u;
        if (data + code + 1 > end) return -1;
        uint32_t v = 0;
        memcpy(&v, data, code + 1);            /* little endian host, as the reference assumes */
        data += code + 1;
        const int32_t d = (int32_t)((v >> 1) ^ (0u - (v & 1u)));
        prev = (int32_t)((uint32_t)prev + (uint32_t)d);
        out[i] = (int16_t)prev;
    }
    return data == end ? 0 : -1;
}

/* ---------- dwell medians ----------
 * scripts/poregen.sh:33-52 calculate_dwell_times_medians on one dump file's text:
 *   awk -F';' '{for(i=1;i<=NF;i++) print gsub(",", "", $i)}' | datamash median 1
 * i.e. the comma count of every ';'-separated field of every line — the file is one line that ends with ';', so the
 * last field is empty and counts 0 — and the median (mean of the two middle values). Returns the number of values
 * (0 for an empty file: awk prints nothing); *twice_median = sum of the two middle counts (or twice the middle one).
 * PARITY UNPINNED like the model rows: datamash is not available; the median of integers has one defensible value. */
static int u64_cmp(const void *a, const void *b) {
    const uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return (x > y) - (x < y);
}
size_t oracle_dwell_median_text(const char *text, size_t len, uint64_t *twice_median) {
    *twice_median = 0;
    if (len == 0) return 0;
    size_t nf = 1;
    for (size_t i = 0; i < len; i++) nf += text[i] == ';';
    uint64_t *c = (uint64_t *)calloc(nf, sizeof(uint64_t));
    size_t f = 0;
    for (size_t i = 0; i < len; i++) {
        if (text[i] == ';') f++;
        else if (text[i] == ',') c[f]++;
    }
    qsort(c, nf, sizeof(uint64_t), u64_cmp);
    *twice_median = (nf & 1) ? 2 * c[nf / 2] : c[nf / 2 - 1] + c[nf / 2];
    free(c);
    return nf;
}
