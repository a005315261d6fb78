/*
 * alignment_cmdline.h -- command line + one-to-many driver of the drop-in host library.
 *
 * Mirrors src/alignment_cmdline.h:19-60 of the reference: cmdline_t layout, the parse_entire_*
 * helpers, cmdline_new/free/set_files/get_file1/2 and align_from_query_and_db keep their
 * signatures and behaviour (flags, usage text on stderr + exit(EXIT_FAILURE), stdout grammar).
 * align_from_query_and_db() is THE accelerated boundary: where the reference builds 16-lane batches
 * and runs `omp parallel for` over alignment_fill_matrices (src/alignment_cmdline.c:422-527), this
 * one encodes the whole FASTA database, hands it to the GPU engine (swgpu.h) -- sharded over
 * SWGPU_GPUS devices if that environment variable is set -- and replays the reference's print
 * callback over groups of 16 entries in file order.
 */
#ifndef ALIGNMENT_CMDLINE_HEADER_SEEN
#define ALIGNMENT_CMDLINE_HEADER_SEEN

#include <stdarg.h>
#include <stdbool.h>
#include "alignment.h"

enum SeqAlignCmdType {SEQ_ALIGN_SW_CMD};

typedef struct
{
  char *file_path1, *file_path2;
  score_t match, mismatch, gap_open, gap_extend;
  unsigned int max_hits_per_alignment;
  bool max_hits_per_alignment_set;
  bool print_seq;
  bool print_matrices;
  bool interactive;
  bool print_fasta, print_pretty, print_colour;
} cmdline_t;

#ifdef __cplusplus
extern "C" {
#endif

char parse_entire_score_t(char *str, score_t *result);
char parse_entire_short(char *str, short *result);
char parse_entire_ushort(char *str, unsigned short *result);
char parse_entire_int(char *str, int *result);
char parse_entire_uint(char *str, unsigned int *result);

cmdline_t *cmdline_new(int argc, char **argv, scoring_t *scoring, enum SeqAlignCmdType cmd_type);
void cmdline_free(cmdline_t *cmd);
void cmdline_set_files(cmdline_t *cmd, char *p1, char *p2);
char *cmdline_get_file1(cmdline_t *cmd);
char *cmdline_get_file2(cmdline_t *cmd);

void align_from_query_and_db(const char *query_path, const char *db_path, scoring_t *scoring,
                             void (print_alignment)(aligner_t *aligner, size_t total_cnt),
                             bool use_zlib);

#ifdef __cplusplus
}
#endif
#endif
