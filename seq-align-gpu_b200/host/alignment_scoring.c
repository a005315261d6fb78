/* alignment_scoring.c -- scoring model (see include/alignment_scoring.h for the reference lines). */
#include "alignment_scoring.h"
#include <assert.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

void scoring_init(scoring_t *sc, int match, int mismatch, int gap_open, int gap_extend, bool case_sensitive)
{
    sc->gap_open = (score_t)gap_open;
    sc->gap_extend = (score_t)gap_extend;
    sc->use_match_mismatch = true;
    sc->match = match;
    sc->mismatch = mismatch;
    sc->case_sensitive = case_sensitive;
    memset(sc->swap_set, 0, sizeof(sc->swap_set));
    /* deviation from the reference (which leaves the table uninitialised): undefined pairs score 0 */
    memset(sc->swap_scores, 0, sizeof(sc->swap_scores));
    sc->min_penalty = match < mismatch ? match : mismatch;
    sc->max_penalty = match > mismatch ? match : mismatch;
}

int8_t letters_to_index(char c)
{
    if (c >= 'a' && c <= 'z') return (int8_t)(c - 'a' + 1);
    if (c >= 'A' && c <= 'Z') return (int8_t)(c - 'A' + 1);
    if (c == '*') return 31;
    /* same channel (stdout), text and exit status as the reference */
    printf("Error: %c is not a legal character for the substitution matrix!\n", c);
    exit(1);
}

char index_to_letters(int8_t c)
{
    if (c >= 1 && c <= 26) return (char)(c + 'A' - 1);
    if (c == 31) return '*';
    printf("Error: %d is not a legal index for the substitution matrix!\n", c);
    exit(1);
}

void scoring_add_mutation(scoring_t *sc, char a, char b, int score)
{
    assert(score > -128 && score < 128);
    const int ia = letters_to_index(a), ib = letters_to_index(b);
    sc->swap_scores[ia][ib] = (int8_t)score;
    set_swap_bit(sc, ia, ib);
    if (score < sc->min_penalty) sc->min_penalty = score;
    if (score > sc->max_penalty) sc->max_penalty = score;
}

void scoring_system_default(scoring_t *sc)
{
    scoring_init(sc, 1, -2, -4, -1, false);
}
