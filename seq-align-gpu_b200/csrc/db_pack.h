/*
 * db_pack.h -- host-side database packer (plain C).
 *
 * Takes the encoded database (what src/alignment_cmdline.c:444-446 of the reference produces
 * residue by residue) and lays it out for the GPU kernels.  GPU twin of the reference's batch
 * build (16-lane [pos][lane] interleave with '*' padding, alignment_cmdline.c:429-457):
 *   - stable sort by length, descending.  The reference requires the FILE to be sorted
 *     (alignment_cmdline.c:431-439); here the engine sorts and keeps the file-order index;
 *   - `plain`: residues of the sorted targets back to back (32-bit kernels read these);
 *   - STREAMS for the systolic 16-bit kernel (k1s): consecutive targets of the sorted list are
 *     PAIRED (2p, 2p+1 -> the two u16 halves of one packed register); pairs are concatenated
 *     into streams of about stream_cols columns.  One column = 16 bits: low byte = code of
 *     target 2p (bit 7 set on the first column of a pair), high byte = code of target 2p+1
 *     (0 once the shorter one has ended; code 0 = padding, which can never score -- the
 *     reference pads with '*' = 31, :448-450).  A pair never straddles two streams; a pair
 *     longer than stream_cols is a stream of its own.  With `taper` the streams of the last 1/16 of the columns are
 *     half as long and those of the last 1/64 a quarter (smaller work units at the end of the kernel's queue).
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
    int32_t  max_len;         /* longest target                                          */
    int64_t *order;           /* [n] sorted position -> index in the input arrays        */
    int64_t *plain_off;       /* [n+1] offsets into plain, sorted order                  */
    int8_t  *plain;           /* [n_residues] residues, sorted order                     */
    int64_t  n_pairs, n_streams;
    int32_t  max_stream_cols;
    int64_t  n_long_streams;  /* leading streams whose first pair alone has more than SWPACK_LONG_STREAM_COLS columns */
    int64_t  n_long_pairs;    /* pairs inside those streams                              */
    uint16_t *scols;          /* [scols_len] stream columns, each stream 64-column aligned */
    int64_t  scols_len;
    int64_t *stream_off;      /* [n_streams+1] start column of each stream in scols      */
    int32_t *stream_ncols;    /* [n_streams]   real columns                              */
    int64_t *stream_pair0;    /* [n_streams]   index of the stream's first pair          */
} swpack_t;

/* returns 0, or -1 on allocation failure / bad input (negative length, code outside 0..31) */
int  swpack_build(swpack_t *p, const int8_t *codes, const int64_t *offsets, int64_t n_targets, int32_t stream_cols, int32_t taper);
/* Same layout from target lengths alone (synthetic databases whose residues are generated on the device):
 * order, plain_off and the stream tables are filled, plain and scols stay NULL; pair_col[(n+1)/2] (may be NULL)
 * receives the first column of every pair in scols. */
int  swpack_layout(swpack_t *p, const int32_t *lengths, int64_t n_targets, int32_t stream_cols, int32_t taper, int64_t *pair_col);
void swpack_free(swpack_t *p);

#ifdef __cplusplus
}
#endif
#endif
