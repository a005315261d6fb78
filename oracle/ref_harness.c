/*
 * ref_harness.c -- ORACLE-SIDE HARNESS AROUND THE UNMODIFIED REFERENCE (test infrastructure only).
 *
 * Linked against the reference's own objects (src/alignment.c, src/alignment_scoring.c,
 * compiled where they lie under /root/reference by oracle/Makefile into oracle/_ref/).
 * It drives the reference's public library API exactly the way its driver does
 * (src/alignment_cmdline.c:429-507): groups of 16 targets, [pos][lane] interleave,
 * '*' (31) padding up to the first (longest) member's length, aligner_create +
 * alignment_fill_matrices under `omp parallel for schedule(dynamic,1)`.
 * Two deliberate differences, both to dodge reference bugs:
 *   - vector_size is always 16 and a partial last group is filled with '*'-only lanes
 *     (the reference under-allocates its row buffers for vector_size < 16);
 *   - targets are taken from memory (already encoded), not from a FASTA file.
 * Targets must be sorted by length descending (alignment_cmdline.c:439).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <omp.h>
#include "alignment.h"

/* scoring_t is built by the caller of this harness through the reference's own
 * scoring_init / scoring_add_mutation; here we accept a raw 32x32 table. */
static void fill_scoring(scoring_t *sc, const int8_t *mat, int gap_open, int gap_extend)
{
    memset(sc, 0, sizeof(*sc));
    scoring_init(sc, 1, -1, gap_open, gap_extend, true);
    memcpy(sc->swap_scores, mat, 32 * 32);
}

/* returns seconds spent inside the alignment_fill_matrices parallel region */
double refh_search16(const int8_t *q, long lq, const int8_t *residues, const int64_t *offsets, long n,
                     const int8_t *mat, int gap_open, int gap_extend, int threads, int16_t *scores)
{
    scoring_t sc;
    fill_scoring(&sc, mat, gap_open, gap_extend);
    if (threads > 0) omp_set_num_threads(threads);
    const long nb = (n + 15) / 16;
    aligner_t **al = (aligner_t **)calloc((size_t)nb, sizeof(aligner_t *));
    int8_t **bufs = (int8_t **)calloc((size_t)nb, sizeof(int8_t *));
    for (long b = 0; b < nb; b++) {
        const long first = b * 16;
        const long plen = (long)(offsets[first + 1] - offsets[first]);
        int8_t *buf = (int8_t *)aligned_alloc(32, (size_t)((plen > 0 ? plen : 1) * 16 + 31) / 32 * 32);
        memset(buf, 31, (size_t)(plen > 0 ? plen : 1) * 16);
        for (long l = 0; l < 16 && first + l < n; l++) {
            const int8_t *t = residues + offsets[first + l];
            const long lt = (long)(offsets[first + l + 1] - offsets[first + l]);
            if (lt > plen) { abort(); } /* unsorted input: the reference asserts here */
            for (long i = 0; i < lt; i++) buf[i * 16 + l] = t[i];
        }
        bufs[b] = buf;
        al[b] = aligner_create(NULL, NULL, NULL, NULL, (int8_t *)q, buf, (size_t)lq, (size_t)plen, 16, &sc);
    }
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
#pragma omp parallel for schedule(dynamic, 1)
    for (long b = 0; b < nb; b++) alignment_fill_matrices(al[b]);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    for (long b = 0; b < nb; b++) {
        for (long l = 0; l < 16 && b * 16 + l < n; l++) scores[b * 16 + l] = al[b]->max_scores[l];
        aligner_destroy(al[b]);
        free(al[b]->max_scores);
        free(al[b]);
        free(bufs[b]);
    }
    free(al); free(bufs);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

int refh_max_threads(void) { return omp_get_max_threads(); }

/* thin pass-throughs so ctypes users can exercise the reference's scalar helpers */
int refh_letters_to_index(int c) { return (int)letters_to_index((char)c); }
