/* db_pack.c -- see db_pack.h.  Plain C11, OpenMP optional. */
#include "db_pack.h"
#include <stdlib.h>
#include <string.h>

void swpack_free(swpack_t *p)
{
    if (!p) return;
    free(p->order); free(p->plain_off); free(p->plain);
    free(p->scols); free(p->stream_off); free(p->stream_ncols); free(p->stream_pair0);
    memset(p, 0, sizeof(*p));
}

static int layout_streams(swpack_t *p, int32_t stream_cols, int32_t taper, int64_t *pair_col);
static int fill_streams(swpack_t *p);

int swpack_build(swpack_t *p, const int8_t *codes, const int64_t *offsets, int64_t n, int32_t stream_cols, int32_t taper)
{
    memset(p, 0, sizeof(*p));
    if (n < 0) return -1;
    p->n_targets = n;
    p->n_residues = n ? offsets[n] - offsets[0] : 0;

    /* 1. stable counting sort by length, descending */
    int64_t max_len = 0;
    for (int64_t t = 0; t < n; t++) {
        const int64_t len = offsets[t + 1] - offsets[t];
        if (len < 0 || len > 0x7fffffff) return -1;
        if (len > max_len) max_len = len;
    }
    p->max_len = (int32_t)max_len;
    int64_t *bucket = (int64_t *)calloc((size_t)max_len + 2, sizeof(int64_t));
    p->order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    p->plain_off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n + 1));
    p->plain = (int8_t *)malloc((size_t)(p->n_residues ? p->n_residues : 1));
    if (!bucket || !p->order || !p->plain_off || !p->plain) { free(bucket); swpack_free(p); return -1; }
    for (int64_t t = 0; t < n; t++) bucket[max_len - (offsets[t + 1] - offsets[t]) + 1]++;
    for (int64_t l = 0; l <= max_len; l++) bucket[l + 1] += bucket[l];
    for (int64_t t = 0; t < n; t++) p->order[bucket[max_len - (offsets[t + 1] - offsets[t])]++] = t;
    free(bucket);

    /* 2. plain residues in sorted order (the 32-bit kernels read these) */
    p->plain_off[0] = 0;
    for (int64_t s = 0; s < n; s++) {
        const int64_t t = p->order[s], len = offsets[t + 1] - offsets[t];
        p->plain_off[s + 1] = p->plain_off[s] + len;
    }
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (int64_t s = 0; s < n; s++) {
        const int64_t t = p->order[s], len = offsets[t + 1] - offsets[t];
        const int8_t *src = codes + offsets[t];
        int8_t *dst = p->plain + p->plain_off[s];
        for (int64_t i = 0; i < len; i++) { dst[i] = src[i]; bad |= (src[i] < 0 || src[i] > 31); }
    }
    if (bad) { swpack_free(p); return -1; }

    if (layout_streams(p, stream_cols, taper, NULL) != 0) return -1;
    return fill_streams(p);
}

/* Layout only: the database is described by its target lengths; `plain` and `scols` stay NULL (the caller fills
 * them on the device).  pair_col[n_pairs] receives the first stream column of every pair. */
int swpack_layout(swpack_t *p, const int32_t *lengths, int64_t n, int32_t stream_cols, int32_t taper, int64_t *pair_col)
{
    memset(p, 0, sizeof(*p));
    if (n < 0) return -1;
    p->n_targets = n;
    int64_t max_len = 0, total = 0;
    for (int64_t t = 0; t < n; t++) {
        if (lengths[t] < 0) return -1;
        if (lengths[t] > max_len) max_len = lengths[t];
        total += lengths[t];
    }
    p->n_residues = total;
    p->max_len = (int32_t)max_len;
    int64_t *bucket = (int64_t *)calloc((size_t)max_len + 2, sizeof(int64_t));
    p->order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    p->plain_off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n + 1));
    if (!bucket || !p->order || !p->plain_off) { free(bucket); swpack_free(p); return -1; }
    for (int64_t t = 0; t < n; t++) bucket[max_len - lengths[t] + 1]++;
    for (int64_t l = 0; l <= max_len; l++) bucket[l + 1] += bucket[l];
    for (int64_t t = 0; t < n; t++) p->order[bucket[max_len - lengths[t]]++] = t;
    free(bucket);
    p->plain_off[0] = 0;
    for (int64_t s = 0; s < n; s++) p->plain_off[s + 1] = p->plain_off[s] + lengths[p->order[s]];
    return layout_streams(p, stream_cols, taper, pair_col);
}

/* Length limit of a stream that starts with `remaining` of `total` columns still to place.  With `taper` the streams
 * of the last 1/16 of the database are half as long and those of the last 1/64 a quarter: a stream (times the query
 * stripes) is the unit of the kernel's work queue, and equal-sized units leave the GPU partly idle for up to one unit
 * at the end of every launch (measured: 2-6 % of a multi-stripe launch on the C2 database).  The short targets live
 * at the end of the sorted database, so the shorter streams still hold many pairs each. */
static int64_t stream_limit(int64_t stream_cols, int32_t taper, int64_t remaining, int64_t total)
{
    int64_t lim = stream_cols, frac = 16;
    for (int32_t level = 0; level < taper; level++, frac *= 4) {     /* taper = number of halvings: at 1/16, 1/64, 1/256 ... */
        if (remaining > total / frac) break;
        lim /= 2;
    }
    return lim < 64 ? 64 : lim;
}

/* 3. streams of paired targets for the systolic kernel (see db_pack.h): the stream tables from the lengths ... */
static int layout_streams(swpack_t *p, int32_t stream_cols, int32_t taper, int64_t *pair_col)
{
    const int64_t n = p->n_targets;
    const int64_t n_pairs = (n + 1) / 2;
    p->n_pairs = n_pairs;
    if (stream_cols < 32) stream_cols = 32;
    int64_t total = 0;
    for (int64_t q = 0; q < n_pairs; q++) {
        const int64_t len = p->plain_off[2 * q + 1] - p->plain_off[2 * q];
        total += len < 1 ? 1 : len;
    }
    /* pass 1: count streams.  pair length = its first (longer) member; an empty pair still takes one
     * (flagged, all-padding) column so that every pair has a flush point */
    int64_t n_streams = 0, cur = 0, done = 0, limit = stream_cols;
    for (int64_t q = 0; q < n_pairs; q++) {
        int64_t len = p->plain_off[2 * q + 1] - p->plain_off[2 * q];
        if (len < 1) len = 1;
        if (cur == 0) { n_streams++; limit = stream_limit(stream_cols, taper, total - done, total); }
        cur += len; done += len;
        if (cur >= limit) cur = 0;
    }
    p->n_streams = n_streams;
    p->stream_off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n_streams + 1));
    p->stream_ncols = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n_streams ? n_streams : 1));
    p->stream_pair0 = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n_streams ? n_streams : 1));
    if (!p->stream_off || !p->stream_ncols || !p->stream_pair0) { swpack_free(p); return -1; }
    int64_t s = -1, off = 0; cur = 0; done = 0;
    int32_t maxc = 0;
    for (int64_t q = 0; q < n_pairs; q++) {
        int64_t len = p->plain_off[2 * q + 1] - p->plain_off[2 * q];
        if (len < 1) len = 1;
        if (cur == 0) { s++; p->stream_off[s] = off; p->stream_pair0[s] = q; limit = stream_limit(stream_cols, taper, total - done, total); }
        if (pair_col) pair_col[q] = off + cur;
        cur += len; done += len;
        if (cur > 0x7fffff00) { swpack_free(p); return -1; }
        if (cur >= limit || q == n_pairs - 1) {
            p->stream_ncols[s] = (int32_t)cur;
            if (cur > maxc) maxc = (int32_t)cur;
            off += (cur + 63) & ~(int64_t)63;
            cur = 0;
        }
    }
    p->stream_off[n_streams] = off;
    p->n_long_streams = 0;
    /* long = the stream's FIRST pair alone is longer than SWPACK_LONG_STREAM_COLS (a titin-like target, a serial chain of
     * that many columns).  Pairs are sorted by length, so these streams are a prefix for any stream_cols; counting by
     * stream length instead would make every stream "long" once stream_cols >= the threshold. */
    while (p->n_long_streams < n_streams) {
        const int64_t q0 = p->stream_pair0[p->n_long_streams];
        if (p->plain_off[2 * q0 + 1] - p->plain_off[2 * q0] <= SWPACK_LONG_STREAM_COLS) break;
        p->n_long_streams++;
    }
    p->n_long_pairs = p->n_long_streams < n_streams ? p->stream_pair0[p->n_long_streams] : n_pairs;
    p->scols_len = off;
    p->max_stream_cols = maxc;
    return 0;
}

/* ... and the columns themselves from `plain` */
static int fill_streams(swpack_t *p)
{
    const int64_t n = p->n_targets, n_pairs = p->n_pairs, n_streams = p->n_streams;
    p->scols = (uint16_t *)calloc((size_t)(p->scols_len ? p->scols_len : 1), sizeof(uint16_t));
    if (!p->scols) { swpack_free(p); return -1; }
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t st = 0; st < n_streams; st++) {
        uint16_t *dst = p->scols + p->stream_off[st];
        const int64_t q_end = (st + 1 < n_streams) ? p->stream_pair0[st + 1] : n_pairs;
        for (int64_t q = p->stream_pair0[st]; q < q_end; q++) {
            const int8_t *a = p->plain + p->plain_off[2 * q];
            const int64_t la = p->plain_off[2 * q + 1] - p->plain_off[2 * q];
            const int64_t lb = (2 * q + 1 < n) ? p->plain_off[2 * q + 2] - p->plain_off[2 * q + 1] : 0;
            const int8_t *b = a + la;
            const int64_t len = la < 1 ? 1 : la;
            for (int64_t j = 0; j < len; j++) {
                const unsigned ca = j < la ? (uint8_t)a[j] : 0u, cb = j < lb ? (uint8_t)b[j] : 0u;
                dst[j] = (uint16_t)(ca | (cb << 8));
            }
            dst[0] |= 0x80u;
            dst += len;
        }
    }
    return 0;
}
