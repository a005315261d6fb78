/*
 * db_pack.h -- host-side database packer (plain C).
 *
 * Takes the encoded database (what src/alignment_cmdline.c:444-446 of the reference
 * produces residue by residue) and lays it out for the GPU kernels:
 *   - stable sort by length, descending (the reference requires the FILE to be sorted,
 *     alignment_cmdline.c:431-439; here the engine sorts and keeps the file-order index);
 *   - the n_long longest targets (length > long_threshold) are left as plain byte strings
 *     for the intra-sequence kernel;
 *   - the rest is cut into chunks of 64 consecutive targets; inside a chunk lane l of a
 *     warp owns targets 2l and 2l+1, and columns are interleaved [col/4][lane] as 8-byte
 *     words holding 4 columns x (codeA | codeB << 8) -- the GPU twin of the reference's
 *     16-lane [pos][lane] interleave (alignment_cmdline.c:445); padding uses code 0
 *     instead of the reference's '*' (=31, :448-450) so padded cells can never score.
 *   - STREAMS for the systolic inter-sequence kernel (k1s): consecutive targets of the sorted
 *     list are PAIRED (2p, 2p+1 -> the two u16 halves of one packed register); pairs are
 *     concatenated into streams of about stream_cols columns.  One column = 16 bits:
 *     low byte = code of target 2p (bit 7 set on the first column of a pair), high byte = code
 *     of target 2p+1 (0 once the shorter one has ended).  A pair never straddles two streams.
 */
#ifndef SWGPU_DB_PACK_H
#define SWGPU_DB_PACK_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SWPACK_LONG_STREAM_COLS 4096

typedef struct {
    int64_t  n_targets, n_residues;
    int64_t  n_long;          /* sorted[0 .. n_long) go to the intra-sequence kernel */
    int64_t  n_chunks;        /* 64-target chunks over sorted[n_long .. n_targets)    */
    int32_t  max_len;         /* longest target                                       */
    int32_t  max_chunk_cols;  /* longest chunk (padded to a multiple of 4)            */
    int64_t *order;           /* [n] sorted position -> index in the input arrays     */
    int64_t *plain_off;       /* [n+1] offsets into plain, sorted order               */
    int8_t  *plain;           /* [n_residues] residues, sorted order                  */
    int32_t *chunk_cols;      /* [n_chunks] padded column count                       */
    int64_t *chunk_off;       /* [n_chunks] offset into cols, in 8-byte words         */
    uint64_t *cols;           /* interleaved chunk data                               */
    int64_t  cols_words;
    /* streams (systolic kernel): all targets, sorted order, paired */
    int64_t  n_pairs, n_streams;
    int32_t  max_stream_cols;
    int64_t  n_long_streams;  /* leading streams with more than SWPACK_LONG_STREAM_COLS columns */
    int64_t  n_long_pairs;    /* pairs inside those streams                                   */
    uint16_t *scols;          /* [scols_len] stream columns, each stream 64-column aligned */
    int64_t  scols_len;
    int64_t *stream_off;      /* [n_streams+1] start column of each stream in scols      */
    int32_t *stream_ncols;    /* [n_streams]   real columns                              */
    int64_t *stream_pair0;    /* [n_streams]   index of the stream's first pair          */
} swpack_t;

/* returns 0, or -1 on allocation failure / bad input (negative length, code > 31) */
int  swpack_build(swpack_t *p, const int8_t *codes, const int64_t *offsets, int64_t n_targets,
                  int32_t long_threshold, int32_t stream_cols);
void swpack_free(swpack_t *p);

#ifdef __cplusplus
}
#endif
#endif
