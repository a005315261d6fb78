/* db_pack.c -- see db_pack.h.  Plain C11, OpenMP optional. */
#include "db_pack.h"
#include <stdlib.h>
#include <string.h>

void swpack_free(swpack_t *p)
{
    if (!p) return;
    free(p->order); free(p->plain_off); free(p->plain);
    free(p->chunk_cols); free(p->chunk_off); free(p->cols);
    memset(p, 0, sizeof(*p));
}

int swpack_build(swpack_t *p, const int8_t *codes, const int64_t *offsets, int64_t n, int32_t long_threshold)
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

    /* 2. plain residues in sorted order (intra-sequence kernel + rescoring read these) */
    p->plain_off[0] = 0;
    int64_t n_long = 0;
    for (int64_t s = 0; s < n; s++) {
        const int64_t t = p->order[s], len = offsets[t + 1] - offsets[t];
        p->plain_off[s + 1] = p->plain_off[s] + len;
        if (len > long_threshold) n_long = s + 1;
    }
    p->n_long = n_long;
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (int64_t s = 0; s < n; s++) {
        const int64_t t = p->order[s], len = offsets[t + 1] - offsets[t];
        const int8_t *src = codes + offsets[t];
        int8_t *dst = p->plain + p->plain_off[s];
        for (int64_t i = 0; i < len; i++) { dst[i] = src[i]; bad |= (src[i] < 0 || src[i] > 31); }
    }
    if (bad) { swpack_free(p); return -1; }

    /* 3. 64-target chunks, column-interleaved */
    const int64_t n_short = n - n_long;
    const int64_t n_chunks = (n_short + 63) / 64;
    p->n_chunks = n_chunks;
    p->chunk_cols = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n_chunks ? n_chunks : 1));
    p->chunk_off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n_chunks + 1));
    if (!p->chunk_cols || !p->chunk_off) { swpack_free(p); return -1; }
    int64_t words = 0;
    int32_t max_cols = 0;
    for (int64_t c = 0; c < n_chunks; c++) {
        const int64_t s0 = n_long + c * 64;
        const int64_t len0 = p->plain_off[s0 + 1] - p->plain_off[s0];   /* longest of the chunk */
        const int32_t cols = (int32_t)((len0 + 3) & ~(int64_t)3);
        p->chunk_cols[c] = cols;
        p->chunk_off[c] = words;
        words += (int64_t)(cols / 4) * 32;
        if (cols > max_cols) max_cols = cols;
    }
    p->chunk_off[n_chunks] = words;
    p->cols_words = words;
    p->max_chunk_cols = max_cols;
    p->cols = (uint64_t *)calloc((size_t)(words ? words : 1), sizeof(uint64_t));   /* code 0 = padding */
    if (!p->cols) { swpack_free(p); return -1; }
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t c = 0; c < n_chunks; c++) {
        uint64_t *base = p->cols + p->chunk_off[c];
        for (int lane = 0; lane < 32; lane++) {
            for (int h = 0; h < 2; h++) {
                const int64_t s = n_long + c * 64 + lane * 2 + h;
                if (s >= n) continue;
                const int8_t *src = p->plain + p->plain_off[s];
                const int64_t len = p->plain_off[s + 1] - p->plain_off[s];
                for (int64_t j = 0; j < len; j++)
                    base[(j >> 2) * 32 + lane] |= (uint64_t)(uint8_t)src[j] << (16 * (j & 3) + 8 * h);
            }
        }
    }
    return 0;
}
