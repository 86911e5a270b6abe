/* oracle/shim/htslib/faidx.h — TEST INFRASTRUCTURE, not product code.
 *
 * Minimal stand-in for the two htslib-1.17 faidx entry points the reference's
 * gmove PAF path calls (reference src/gmove.cpp:724 fai_load, :805
 * faidx_fetch_seq). htslib is not vendored in the reference tree (its Makefile
 * downloads 1.17, Makefile:69-81) and there is no network here, so the
 * unmodified reference sources are compiled against this header instead.
 * Behaviour restated from htslib's documented contract: coordinates are
 * 0-based inclusive, clamped to the sequence, `*len=-2` + NULL when the name is
 * absent. Only an existing `<file>.fai` is read (5 or 6 columns); no indexing.
 */
#ifndef ORACLE_SHIM_FAIDX_H
#define ORACLE_SHIM_FAIDX_H
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <ctype.h>

typedef struct {
    char *name;
    long long len, offset, line_blen, line_len;
} shim_fai_ent_t;

typedef struct faidx_t {
    FILE *fp;
    shim_fai_ent_t *ent;
    size_t n;
} faidx_t;

static inline faidx_t *fai_load(const char *fn) {
    size_t l = strlen(fn);
    char *idx = (char *)malloc(l + 5);
    memcpy(idx, fn, l);
    memcpy(idx + l, ".fai", 5);
    FILE *fi = fopen(idx, "r");
    free(idx);
    if (!fi) return NULL;
    FILE *fp = fopen(fn, "r");
    if (!fp) { fclose(fi); return NULL; }
    faidx_t *f = (faidx_t *)calloc(1, sizeof(faidx_t));
    f->fp = fp;
    size_t cap = 0;
    char *line = NULL; size_t ll = 0;
    while (getline(&line, &ll, fi) != -1) {
        char nm[4096]; long long a, b, c, d;
        if (sscanf(line, "%4095s %lld %lld %lld %lld", nm, &a, &b, &c, &d) != 5) continue;
        if (f->n == cap) { cap = cap ? cap * 2 : 64; f->ent = (shim_fai_ent_t *)realloc(f->ent, cap * sizeof(shim_fai_ent_t)); }
        f->ent[f->n].name = strdup(nm);
        f->ent[f->n].len = a; f->ent[f->n].offset = b; f->ent[f->n].line_blen = c; f->ent[f->n].line_len = d;
        f->n++;
    }
    free(line);
    fclose(fi);
    return f;
}

static inline void fai_destroy(faidx_t *f) {
    if (!f) return;
    for (size_t i = 0; i < f->n; i++) free(f->ent[i].name);
    free(f->ent);
    if (f->fp) fclose(f->fp);
    free(f);
}

static inline char *faidx_fetch_seq(const faidx_t *f, const char *name, int beg, int end, int *len) {
    const shim_fai_ent_t *e = NULL;
    for (size_t i = 0; i < f->n; i++) if (strcmp(f->ent[i].name, name) == 0) { e = &f->ent[i]; break; }
    if (!e) { *len = -2; fprintf(stderr, "[E::fai_fetch] The sequence \"%s\" was not found\n", name); return NULL; }
    long long b = beg, en = end;
    if (en < b) b = en;
    if (b < 0) b = 0; else if (e->len <= b) b = e->len;
    if (en < 0) en = 0; else if (e->len <= en) en = e->len - 1;
    long long want = en + 1 - b;
    if (want < 0) want = 0;
    char *s = (char *)malloc((size_t)want + 1);
    long long got = 0;
    if (want > 0) {
        fseeko(f->fp, (off_t)(e->offset + b / e->line_blen * e->line_len + b % e->line_blen), SEEK_SET);
        int c;
        while (got < want && (c = fgetc(f->fp)) != EOF) if (isgraph(c)) s[got++] = (char)c;
    }
    s[got] = 0;
    *len = (int)got;
    return s;
}
#endif
