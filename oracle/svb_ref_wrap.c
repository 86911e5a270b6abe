/* oracle/svb_ref_wrap.c — TEST INFRASTRUCTURE. Two C entry points on slow5lib's OWN signal compression (compiled
 * from the reference's submodule by build_ref.sh into oracle/_ref/libsvbref.so), so that tests can pin the oracle's
 * svb-zd decoder and the numpy encoder against it. No algorithm here. */
#include <stdint.h>
#include <stdlib.h>
#include <slow5/slow5.h>
#include <slow5/slow5_press.h>

/* slow5_press.c:1054-1083 via the public wrapper; returns a malloc'd stream of *nbytes bytes */
void *svbref_compress(const int16_t *x, size_t n_samples, size_t *nbytes) {
    return slow5_ptr_compress_solo(SLOW5_COMPRESS_SVB_ZD, x, n_samples * sizeof *x, nbytes);
}
/* slow5_press.c:1110-1130; returns malloc'd int16 samples, *out_bytes = 2 * count; NULL on error */
void *svbref_depress(const uint8_t *p, size_t nbytes, size_t *out_bytes) {
    return slow5_ptr_depress_solo(SLOW5_COMPRESS_SVB_ZD, p, nbytes, out_bytes);
}
void svbref_free(void *p) { free(p); }
