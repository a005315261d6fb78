/*
 * alignment_scoring_load.h -- substitution-matrix file loader.
 * Replaces src/alignment_scoring_load.h:14-15 of the reference (same signature; the gzFile stays
 * owned by the caller, src/alignment_cmdline.c:229-240).
 */
#ifndef ALIGNMENT_SCORING_LOAD_HEADER_SEEN
#define ALIGNMENT_SCORING_LOAD_HEADER_SEEN

#include <zlib.h>
#include "alignment_scoring.h"

#ifdef __cplusplus
extern "C" {
#endif

void align_scoring_load_matrix(gzFile file, const char *file_path, scoring_t *scoring, char case_sensitive);

#ifdef __cplusplus
}
#endif
#endif
