/*
 * sw_cmdline.c -- bin/smith_waterman: same flags, same stdout grammar as the reference's
 * src/tools/sw_cmdline.c:27-100 ("Entry #N:" / optional names and sequences / "score: S" / blank),
 * which test/tests.py:52 and benchmarks/benchmark.py:29 of the reference parse.
 * Scores above 32 767 are printed exactly (the reference wraps them).
 */
#include <stdio.h>
#include <stdlib.h>
#include "alignment_cmdline.h"
#include "alignment_scoring_load.h"
#include "seq_align.h"

static cmdline_t *cmd;

static void print_entry_group(aligner_t *a, size_t first_index)
{
    if (cmd->print_fasta && a->seq_a_fasta != NULL) { fputs(a->seq_a_fasta, stdout); putc('\n', stdout); }
    if (cmd->print_seq) { fputs(a->seq_a_str, stdout); putc('\n', stdout); }
    for (size_t b = 0; b < a->vector_size; b++) {
        printf("Entry #%lu:\n", (unsigned long)(first_index + b));
        if (cmd->print_fasta && a->seq_b_fasta_batch[b] != NULL) { fputs(a->seq_b_fasta_batch[b], stdout); putc('\n', stdout); }
        if (cmd->print_seq) { fputs(a->seq_b_str_batch[b], stdout); putc('\n', stdout); }
        printf("score: %i\n\n", a->max_scores32 ? a->max_scores32[b] : a->max_scores[b]);
    }
    fflush(stdout);
}

int main(int argc, char **argv)
{
    scoring_t scoring;
    scoring_system_default(&scoring);
    scoring.match = 2; scoring.mismatch = -2; scoring.gap_open = -2; scoring.gap_extend = -1;   /* SW defaults */
    cmd = cmdline_new(argc, argv, &scoring, SEQ_ALIGN_SW_CMD);
    const char *qf = cmdline_get_file1(cmd), *df = cmdline_get_file2(cmd);
    if (qf != NULL && df != NULL) align_from_query_and_db(qf, df, &scoring, &print_entry_group, !cmd->interactive);
    else { fprintf(stderr, "Error: Both query and database files must be provided\n"); fflush(stderr); }
    cmdline_free(cmd);
    return EXIT_SUCCESS;
}
