/*
 * alignment.c -- aligner_t handle; alignment_fill_matrices() = one 16-lane batch on the GPU.
 * Reference counterparts: src/alignment.c:190-239 (create/update/destroy), :47-187 (fill).
 */
#include "alignment.h"
#include "swgpu.h"
#include <stdio.h>
#include <stdlib.h>

const size_t FULL_VECTOR_SIZE = 16;

void aligner_update(aligner_t *a, char *seq_a_str, char **seq_b_str_batch, char *seq_a_fasta, char **seq_b_fasta_batch,
                    int8_t *seq_a_indexes, int8_t *seq_b_batch_indexes, size_t len_a, size_t len_b, size_t vector_size,
                    const scoring_t *scoring)
{
    a->scoring = scoring;
    a->seq_a_str = seq_a_str; a->seq_b_str_batch = seq_b_str_batch;
    a->seq_a_fasta = seq_a_fasta; a->seq_b_fasta_batch = seq_b_fasta_batch;
    a->seq_a_indexes = seq_a_indexes; a->seq_b_batch_indexes = seq_b_batch_indexes;
    a->score_width = len_a + 1; a->score_height = len_b + 1;
    a->vector_size = vector_size;
}

aligner_t *aligner_create(char *seq_a_str, char **seq_b_str_batch, char *seq_a_fasta, char **seq_b_fasta_batch,
                          int8_t *seq_a_indexes, int8_t *seq_b_batch_indexes, size_t len_a, size_t len_b, size_t vector_size,
                          const scoring_t *scoring)
{
    aligner_t *a = (aligner_t *)calloc(1, sizeof(aligner_t));
    if (!a) return NULL;
    /* always 16 lanes of result storage (the reference sizes by vector_size and overruns).  ONE block:
     * like the reference (src/alignment.c:235-239) aligner_destroy() leaves max_scores to the caller, and a caller that
     * frees max_scores the way the reference's users do releases the appended int32 scores with it. */
    a->max_scores = (score_t *)calloc(16, sizeof(score_t) + sizeof(int32_t));
    if (!a->max_scores) { free(a); return NULL; }
    a->max_scores32 = (int32_t *)(void *)(a->max_scores + 16);
    aligner_update(a, seq_a_str, seq_b_str_batch, seq_a_fasta, seq_b_fasta_batch, seq_a_indexes, seq_b_batch_indexes,
                   len_a, len_b, vector_size, scoring);
    return a;
}

/* SWGPU_DEVICE=<n>: the GPU the single-batch entry point runs on (default 0) */
static int batch_device(void)
{
    static int dev = -1;
    if (dev < 0) { const char *e = getenv("SWGPU_DEVICE"); dev = (e && atoi(e) > 0) ? atoi(e) : 0; }
    return dev;
}

/* Re-entrant across distinct aligner_t, like the reference's (called from its OpenMP workers at once,
 * src/alignment_cmdline.c:504-507): every host thread has its own engine context (csrc/swgpu.cu::fill_batch16). */
void alignment_fill_matrices(aligner_t *a)
{
    const size_t len_a = a->score_width - 1, len_b = a->score_height - 1;
    int lanes = (int)(a->vector_size < 1 ? 1 : (a->vector_size > 16 ? 16 : a->vector_size));
    int32_t s[16] = {0};
    const int rc = swgpu_fill_batch16(batch_device(), a->seq_a_indexes, (int64_t)len_a, a->seq_b_batch_indexes, (int64_t)len_b, lanes,
                                      &a->scoring->swap_scores[0][0], a->scoring->gap_open, a->scoring->gap_extend, s);
    if (rc != SWGPU_OK) {
        fprintf(stderr, "Error: GPU alignment failed: %s\n", swgpu_last_error());
        exit(EXIT_FAILURE);
    }
    for (int l = 0; l < 16; l++) {
        a->max_scores32[l] = l < lanes ? s[l] : 0;
        a->max_scores[l] = (score_t)(a->max_scores32[l] > 32767 ? 32767 : a->max_scores32[l]);
    }
}

void aligner_destroy(aligner_t *a)
{
    if (!a) return;
    /* exactly what the reference frees (src/alignment.c:235-239): the three row buffers.  max_scores (with
     * max_scores32 in the same block) and the struct itself stay with the caller, who may still read the scores. */
    free(a->curr_match_scores); free(a->curr_gap_a_scores); free(a->curr_gap_b_scores);
    a->curr_match_scores = a->curr_gap_a_scores = a->curr_gap_b_scores = NULL;
}
