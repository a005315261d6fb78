/*
 * alignment_scoring.h -- scoring model of the drop-in host library (plain C11).
 *
 * Binary-compatible with the reference's src/alignment_scoring.h:18-61: same score_t, same
 * scoring_t field order and alignment (callers put scoring_t on their stack,
 * src/tools/sw_cmdline.c:82), same five functions.  Differences, all documented in DESIGN.md:
 *   - scoring_init() ZEROES swap_scores (the reference leaves pairs that the matrix file never
 *     mentions as stack garbage, src/alignment_scoring.c:31-50);
 *   - <x86intrin.h> is not pulled in (nothing here needs it).
 */
#ifndef ALIGNMENT_SCORING_HEADER_SEEN
#define ALIGNMENT_SCORING_HEADER_SEEN

#include <inttypes.h>
#include <stdbool.h>
#include <limits.h>
#include <stdalign.h>

typedef int16_t score_t;          /* src/alignment_scoring.h:18 */
#define SCORE_MIN INT_MIN

typedef struct
{
  score_t gap_open, gap_extend;   /* a gap of length N costs gap_open + N*gap_extend (both <= 0) */
  bool use_match_mismatch;        /* parsed and stored; the scoring path only reads swap_scores   */
  int match, mismatch;
  bool case_sensitive;
  uint32_t swap_set[32];          /* bit b of swap_set[a]: the pair (a,b) was given a score       */
  alignas(32) int8_t swap_scores[32][32];   /* [query code][target code]                           */
  int min_penalty, max_penalty;
} scoring_t;

#ifndef get_swap_bit
  #define get_swap_bit(scoring, a, b) (((scoring)->swap_set[(size_t)(a)] >> (b)) & 1U)
  #define set_swap_bit(scoring, a, b) ((scoring)->swap_set[(size_t)(a)] |= (1U << (b)))
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* src/alignment_scoring.c:31-50 */
void scoring_init(scoring_t *scoring, int match, int mismatch, int gap_open, int gap_extend, bool case_sensitive);
/* src/alignment_scoring.c:60-68; |score| must be < 128 */
void scoring_add_mutation(scoring_t *scoring, char a, char b, int score);
/* src/alignment_scoring.c:70-81: A-Z/a-z -> 1..26, '*' -> 31, anything else: message on stdout, exit(1) */
int8_t letters_to_index(char c);
/* src/alignment_scoring.c:83-92 */
char index_to_letters(int8_t c);
/* src/alignment_scoring.c:99-109: match 1, mismatch -2, gap open -4, gap extend -1 */
void scoring_system_default(scoring_t *scoring);

#ifdef __cplusplus
}
#endif
#endif
