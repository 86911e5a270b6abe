/* oracle/gmove_oracle.h — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the reference's `gmove` per-record pipeline on the same in-memory batch
 * the CUDA library takes (include/pgm_gmove.h). Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it.
 *
 * Parity pin: PINNED. tests/test_oracle_vs_ref.py feeds identical inputs to this restatement
 * (through the product's host parsers) and to the UNMODIFIED reference compiled by
 * oracle/build_ref.sh, and compares dump/ + freq.txt byte for byte; the reference outputs for
 * the shipped fixture and for seeded synthetic sets are committed under tests/golden/.
 */
#ifndef GMOVE_ORACLE_H
#define GMOVE_ORACLE_H
#include "../include/pgm_gmove.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_result {
    uint32_t n_slots, n_reads;
    uint64_t n_events_examined, n_events_pass, n_sel_events, n_sel_samples;
    uint32_t *slot_sel;   /* [n_slots] */
    uint32_t *slot_total; /* [n_slots] */
    uint32_t *ev_len, *ev_read, *ev_start; /* [n_sel_events], grouped by slot, emission order */
    double *samples;      /* [n_sel_samples] */
    uint8_t *read_flags;  /* [n_reads] */
    char *text;           /* dump text grouped by slot ("%.8f," ... "%.8f;") */
    uint64_t *slot_text_off; /* [n_slots+1] */
} oracle_result;

/* totals: in/out uint32[n_slots] = kmer_frequency_map carried across batches (zero it first).
 * Returns PGM_OK or PGM_ERR_*. */
int oracle_gmove_batch(const pgm_opts *opts, const pgm_batch *batch, uint32_t *totals, oracle_result *out);
void oracle_free_result(oracle_result *r);
void oracle_set_text(int on);   /* 0: leave text empty (slot_text_off all zero) - for full-size comparisons of the doubles */

/* Single pieces, exposed for unit tests. */
double oracle_median(const double *x, size_t n);              /* gmove.cpp:142-150 */
double oracle_mad(const double *x, size_t n, double med);     /* gmove.cpp:161-184 (before the clamp) */
int oracle_pick_kmer(const int *indel_pos, size_t n, int left, int k, int margin); /* gmove.cpp:204-211 */

/* scripts/poregen.sh:54-85 calculate_mean_stddev_all on one slot's dump text: drops the first
 * value, median (mean of the two middles), sample stddev clamped to `limit`.  PARITY UNPINNED:
 * the arithmetic lives in GNU datamash, which is neither in the reference tree nor installed. */
int oracle_model_row(const double *vals, size_t n, double limit, double *median, double *stddev);
size_t oracle_model_row_text(const char *text, size_t len, double limit, double *median, double *stddev, char *median_str, char *stddev_str);

/* scripts/poregen.sh:33-52 calculate_dwell_times_medians on one dump file's text; returns the number of values, *twice_median =
 * twice the datamash median of the per-field comma counts. PARITY UNPINNED (datamash absent). */
size_t oracle_dwell_median_text(const char *text, size_t len, uint64_t *twice_median);

/* slow5lib's svb-zd signal stream -> int16 samples (slow5_press.c:1085-1130 and the streamvbyte scalar decoder).
 * PINNED: tests/test_oracle_svb.py compares it (and the numpy encoder that makes the test inputs) with slow5lib's
 * own slow5_ptr_compress_solo / slow5_ptr_depress_solo compiled into oracle/_ref/libsvbref.so. */
int oracle_svb_zd_decode(const uint8_t *p, size_t nbytes, uint32_t expect, int16_t *out);

#ifdef __cplusplus
}
#endif
#endif
