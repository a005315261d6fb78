/*
 * alignment.h -- batch aligner handle of the drop-in host library.
 *
 * Mirrors src/alignment.h:25-37,57-78 of the reference: aligner_t keeps the reference's fields in
 * the reference's order (print callbacks read vector_size, seq_a_fasta, seq_a_str,
 * seq_b_fasta_batch[], seq_b_str_batch[] and max_scores[], src/tools/sw_cmdline.c:45-71), and the
 * four functions keep their signatures.  One field is APPENDED:
 *     int32_t *max_scores32   exact scores.  max_scores[] is score_t = int16 and cannot hold a
 *                             score above 32 767; it carries min(score, 32767) there (the
 *                             reference prints wrapped garbage in that case).
 * alignment_fill_matrices() runs ONE 16-lane batch on the GPU through swgpu_fill_batch16(); it is
 * there for API parity -- the fast path is align_from_query_and_db() (alignment_cmdline.h), which
 * hands the whole database to the engine at once.
 */
#ifndef ALIGNMENT_HEADER_SEEN
#define ALIGNMENT_HEADER_SEEN

#include <string.h>
#include <stddef.h>
#include "alignment_scoring.h"

typedef struct
{
    const scoring_t *scoring;
    int8_t *seq_a_indexes, *seq_b_batch_indexes;   /* query codes; batch codes interleaved [pos][16] */
    char *seq_a_str, **seq_b_str_batch;
    char *seq_a_fasta, **seq_b_fasta_batch;
    size_t vector_size;                            /* filled lanes, 1..16 */
    size_t score_width, score_height;              /* len(query)+1, padded len(batch)+1 */
    score_t *curr_match_scores;                    /* unused by the GPU engine (kept for layout) */
    score_t *curr_gap_a_scores;
    score_t *curr_gap_b_scores;
    score_t *max_scores;                           /* [16], saturated at 32767 */
    int32_t *max_scores32;                         /* [16], exact (extension) */
} aligner_t;

#ifdef __cplusplus
extern "C" {
#endif

#define aligner_init(a) (memset(a, 0, sizeof(aligner_t)))

extern const size_t FULL_VECTOR_SIZE;              /* 16, src/alignment.c:21 */

void aligner_update(aligner_t *aligner,
                    char *seq_a_str, char **seq_b_str_batch,
                    char *seq_a_fasta, char **seq_b_fasta_batch,
                    int8_t *seq_a_indexes, int8_t *seq_b_batch_indexes,
                    size_t len_a, size_t len_b, size_t vector_size,
                    const scoring_t *scoring);

aligner_t *aligner_create(char *seq_a_str, char **seq_b_str_batch,
                          char *seq_a_fasta, char **seq_b_fasta_batch,
                          int8_t *seq_a_indexes, int8_t *seq_b_batch_indexes,
                          size_t len_a, size_t len_b, size_t vector_size,
                          const scoring_t *scoring);

/* GPU twin of src/alignment.c:47-187 for one batch; exits with a message if there is no CUDA device */
void alignment_fill_matrices(aligner_t *aligner);

void aligner_destroy(aligner_t *aligner);

#ifdef __cplusplus
}
#endif
#endif
