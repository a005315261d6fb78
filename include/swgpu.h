/*
 * swgpu.h -- C-ABI of the B200 Smith-Waterman database-search engine (libswgpu.so).
 *
 * Plain pointers and sizes only; no CUDA or torch types cross this boundary.
 * This is the `extern "C"` device layer the C host code (alignment_cmdline.c in
 * this repo) calls, and what a maintainer of the reference would bind instead of
 * its OpenMP loop.  Each entry point names the reference interface it replaces
 * (paths relative to the reference tree).
 *
 * Residue codes are the reference's: letters_to_index() of src/alignment_scoring.c:70-81
 * (A..Z -> 1..26, '*' -> 31).  Code 0 is reserved by the engine as padding.
 * The substitution table is the reference's scoring_t::swap_scores, int8 [32][32],
 * indexed [query code][target code] (src/alignment_scoring.h:34, src/alignment.c:33,41).
 * Gap penalties follow src/alignment_scoring.c:35-36: a gap of length N costs
 * gap_open + N*gap_extend (both normally negative).
 *
 * All functions return 0 on success and a negative SWGPU_E* code on failure;
 * swgpu_last_error() gives the message.  Nothing here ever falls back to the CPU:
 * without a CUDA device every compute entry point fails with SWGPU_ENODEV.
 */
#ifndef SWGPU_H_SEEN
#define SWGPU_H_SEEN

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SWGPU_OK        0
#define SWGPU_ENODEV   -1   /* no CUDA device / driver */
#define SWGPU_ECUDA    -2   /* a CUDA call failed */
#define SWGPU_EINVAL   -3   /* bad argument */
#define SWGPU_ENOMEM   -4

typedef struct swgpu_db swgpu_db_t;   /* packed, HBM-resident database on one GPU */

/* Packing knobs (0 = engine default). */
typedef struct {
    int32_t long_threshold;   /* > 0: targets longer than this are always scored by the 32-bit kernels */
    int32_t stream_cols;      /* length of one stream of concatenated target pairs (16-bit kernel); default: 2304, down to
                                 1152 for databases below ~110 M residues */
    int32_t no_taper;         /* != 0: every stream gets stream_cols columns (default: the streams of the last 1/16 of the
                                 database are half as long, those of the last 1/64 a quarter -- smaller work units at the
                                 end of the kernel's work queue) */
    int32_t reserved[5];
} swgpu_pack_opts_t;

/* Per-search statistics, all device-timed with CUDA events on the engine's stream. */
typedef struct {
    double  ms_total;         /* first kernel launch -> last result resident in HBM */
    double  ms_inter;         /* packed 16-bit systolic kernel K1S incl. profile build and finalize */
    double  ms_intra;         /* the separate low-latency K1S launch for long streams (runs concurrently) */
    double  ms_rescore;       /* 32-bit kernels: rescoring of saturated targets (K3 / K3X) */
    double  ms_h2d, ms_d2h;   /* query/table upload, score download (host API only) */
    int64_t cells;            /* sum over targets of Lq * Lt, real residues */
    int64_t cells_inter, cells_intra, cells_rescore;
    int64_t n_rescored;       /* targets whose 16-bit result saturated */
    int64_t n_long;           /* targets handled by the long-stream launch */
    int32_t kernel_launches;  /* engine kernels launched by this search */
    int32_t tile_rows;        /* query rows per lane (register tile) chosen for the 16-bit kernel */
    double  ms_topk;          /* last swgpu_topk(): selection + coordinate passes */
    int32_t topk_launches;    /* kernels launched by the last swgpu_topk() */
    int32_t lanes_per_pair;   /* K1S plan: lanes that share one target pair (8, 16 or 32) */
    int32_t passes;           /* K1S plan: query stripes of lanes_per_pair * tile_rows rows */
    int32_t reserved;
} swgpu_stats_t;

/* Top hit with coordinates (the reference has no counterpart). */
typedef struct {
    int32_t score;
    int32_t q_start, q_end, t_start, t_end;   /* 0-based inclusive; -1 if score == 0 */
    int64_t entry;                            /* file-order index of the target */
} swgpu_hit_t;

int         swgpu_device_count(void);
const char *swgpu_last_error(void);
const char *swgpu_version(void);

/*
 * Database packing.  Replaces the batch build of src/alignment_cmdline.c:422-457
 * (encode, 16-lane interleave, '*' padding): here targets are sorted by length
 * inside the engine (the reference demands a pre-sorted file, :431-439), paired,
 * concatenated into streams of 16-bit columns and uploaded once.
 *   codes    n_residues residue codes, targets back to back, file order
 *   offsets  n_targets+1 prefix offsets into codes
 *   entry_base  file-order index of target 0 of this shard (multi-GPU sharding)
 */
int  swgpu_db_create(swgpu_db_t **out, int device, const int8_t *codes, const int64_t *offsets,
                     int64_t n_targets, int64_t entry_base, const swgpu_pack_opts_t *opts);

/*
 * Synthetic database generated ON THE DEVICE (a UniRef50-shaped 10 G-residue database is
 * never materialised on the host or in a FASTA file).  The host supplies the target lengths only; the layout
 * (length sort, pairs, streams) is built from them and the residues are produced by a counter-based generator
 * straight into the packed buffers: residue j of the target with global entry id e is
 *     u = mix64((seed ^ e * 0x9E3779B97F4A7C15) + (j + 1) * 0xBF58476D1CE4E5B9) >> 48        (16 bits)
 *     code = codes[first i with u < cdf[i]]
 * where mix64 is the splitmix64 finaliser.  A target therefore depends on (seed, e, j) alone -- not on the shard it
 * lands in or on the number of GPUs -- and any single target can be regenerated on the host for a sampled parity
 * check (swgpu/synth.py::synth_target).  entry_ids[t] (NULL: entry_base + t) is the global id of local target t.
 * No reference counterpart (the reference reads FASTA only, src/alignment_cmdline.c:422-457).
 */
typedef struct {
    uint64_t seed;
    int32_t  n_letters;       /* 1..32 */
    int32_t  reserved;
    uint32_t cdf[32];         /* cumulative thresholds out of 65536, non-decreasing, cdf[n_letters-1] == 65536 */
    int8_t   codes[32];       /* residue code of letter i (1..31) */
} swgpu_synth_t;

int  swgpu_db_create_synthetic(swgpu_db_t **out, int device, const int32_t *lengths, const int64_t *entry_ids,
                               int64_t n_targets, int64_t entry_base, const swgpu_synth_t *gen,
                               const swgpu_pack_opts_t *opts);
/* Residues of one target (local file-order index) copied back from the packed database: *len = its length;
 * out == NULL only queries the length.  Used by the sampled parity checks of device-generated databases. */
int  swgpu_db_fetch_target(swgpu_db_t *db, int64_t local_entry, int8_t *out, int64_t cap, int64_t *len);
void swgpu_db_destroy(swgpu_db_t *db);
int64_t swgpu_db_num_targets(const swgpu_db_t *db);
int64_t swgpu_db_num_residues(const swgpu_db_t *db);
int64_t swgpu_db_device_bytes(const swgpu_db_t *db);

/*
 * One-to-many scoring: replaces the `omp parallel for` over alignment_fill_matrices
 * (src/alignment_cmdline.c:504-507, src/alignment.c:47-187) for the whole database.
 * scores[n_targets] are written in FILE order (the order Entry #N is printed in,
 * src/tools/sw_cmdline.c:55-72) as int32: exact scores, including those above the
 * reference's int16 range.  Query and table are host pointers; scores is a host pointer.
 */
int swgpu_search(swgpu_db_t *db, const int8_t *query, int64_t query_len,
                 const int8_t *table /* [32*32] */, int gap_open, int gap_extend,
                 int32_t *scores, swgpu_stats_t *stats);

/* Same search, scores left resident on the device (no D2H); used for device-timed
 * throughput.  swgpu_db_scores_d2h() copies the last result out. */
int swgpu_search_resident(swgpu_db_t *db, const int8_t *query, int64_t query_len,
                          const int8_t *table, int gap_open, int gap_extend, swgpu_stats_t *stats);
int swgpu_db_scores_d2h(swgpu_db_t *db, int32_t *scores);

/*
 * Device-side top-k of the last search (score desc, file-order entry asc) with
 * linear-space end/start coordinates recomputed on the GPU for the k hits.
 */
int swgpu_topk(swgpu_db_t *db, int k, swgpu_hit_t *hits, int *n_hits);

/*
 * Many queries against the resident database: the many-to-many front end the reference's report names as
 * the natural extension (it reads a single query, src/alignment_cmdline.c:369-380).  Up to three queries are
 * in flight on separate CUDA streams and scratch buffers, so the latency-bound tail of one query (long
 * targets, 32-bit rescoring, top-k selection and coordinates, score read-back) overlaps the bulk kernel of
 * the next.  Results are identical to n_queries calls of swgpu_search() + swgpu_topk().
 *   queries[q], query_lens[q]  host pointers
 *   scores      NULL, or n_queries host pointers (each NULL or int32[n_targets], file order); pinned buffers are
 *               written directly by the copy engine, pageable ones through an internal pinned staging buffer
 *   k, hits, n_hits   k > 0: hits[q*k .. q*k+n_hits[q]) = top hits of query q with coordinates; k == 0: none
 *   per_query   NULL or n_queries statistics (their ms_* include time spent queued behind other queries)
 *   batch       NULL or totals; ms_total = first kernel of the batch -> last result of the batch in host memory
 */
typedef struct {
    double  ms_total;
    int64_t cells;
    int64_t n_rescored;
    int32_t kernel_launches;
    int32_t slots;            /* queries kept in flight */
    int32_t reserved[4];
} swgpu_batch_stats_t;

int swgpu_search_many(swgpu_db_t *db, int32_t n_queries, const int8_t *const *queries, const int64_t *query_lens,
                      const int8_t *table, int gap_open, int gap_extend, int32_t *const *scores,
                      int32_t k, swgpu_hit_t *hits, int32_t *n_hits,
                      swgpu_stats_t *per_query, swgpu_batch_stats_t *batch);

/*
 * Alignment paths for reported hits (the caller-visible step after the path; upstream seq-align printed
 * alignments, the reference fork removed them and keeps only a score, src/alignment.c:133).  For every hit with
 * coordinates (from swgpu_topk / swgpu_search_many, any hit of THIS database) the affine global alignment of
 * query[q_start..q_end] and target[t_start..t_end] that starts and ends in a match state and scores hit.score is
 * recomputed on the GPU with direction bytes and walked back.  Tie-break: a cell prefers the diagonal, then a
 * gap in the query ('D'), then a gap in the target ('I'); a gap prefers to open rather than extend.
 * Output: run-length encoded operations in path order, ops[ops_off[k] .. ops_off[k+1]) for hit k, each
 * `length << 2 | code` with code 0 = 'M' (aligned pair, match or mismatch), 1 = 'I' (query residue against a gap),
 * 2 = 'D' (target residue against a gap).  Hits with score 0 get an empty path.  ops_off has n_hits + 1 entries.
 * The query, table and gap scores must be those of the search that produced the hits (checked: the rectangle
 * must score hit.score).
 */
int swgpu_traceback(swgpu_db_t *db, const int8_t *query, int64_t query_len, const int8_t *table,
                    int gap_open, int gap_extend, const swgpu_hit_t *hits, int32_t n_hits,
                    uint32_t *ops, int64_t ops_cap, int64_t *ops_off);

/* Statistics of the last search on this handle, with ms_topk / topk_launches of the last swgpu_topk(). */
int swgpu_db_last_stats(const swgpu_db_t *db, swgpu_stats_t *out);

/* Single-batch call with the reference's own batch layout: 16 lanes interleaved
 * [pos][lane], pad 31 -- the GPU twin of alignment_fill_matrices(aligner_t*)
 * (src/alignment.c:47).  Far too fine-grained to be fast; kept for API parity. */
int swgpu_fill_batch16(int device, const int8_t *query, int64_t query_len,
                       const int8_t *batch_interleaved, int64_t padded_len, int lanes,
                       const int8_t *table, int gap_open, int gap_extend, int32_t *scores16);

/*
 * Allocation-free scoring of a small set of targets.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  It is a variant of the 32-bit one-warp-per-target kernel (k3_intra) whose scratch
 * memory -- the per-warp carry rows between query stripes and the work-queue counter -- lives in ONE file-scope
 * `static __device__` array of SWGPU_NOALLOC_SCRATCH_BYTES (about 8 MiB), which every GPU that loads libswgpu.so holds
 * until the process exits.  The call allocates and frees nothing (no cudaMalloc / cudaFree / cudaMallocHost, no
 * allocating library routine), never synchronises, and only enqueues work on `stream`.
 *   ALL pointers are DEVICE pointers owned by the caller:
 *   d_query[query_len], d_targets (residue codes 0..31, targets back to back), d_offsets[n_targets + 1] (prefix offsets
 *   into d_targets), d_table[32*32] (int8, [query code][target code]), d_scores[n_targets] (int32, written in the
 *   order of d_offsets; exact 32-bit scores, same recurrence and results as swgpu_search()).
 *   max_target_len  the caller's bound on the longest target; rejected with SWGPU_EINVAL if it exceeds
 *                   SWGPU_NOALLOC_MAX_TARGET_LEN.  A target that is longer than the limit all the same is not scored:
 *                   its d_scores entry is set to -1.
 *   stream          a cudaStream_t passed as void* (NULL = the default stream).
 * The scratch array is SHARED by every call on that GPU and keeps the previous call's contents; the entry point clears
 * the part that must start at zero (the queue counter) on `stream` before the kernel.  Two calls that overlap on the
 * GPU (different streams, or different host threads without ordering) RACE on it: order them on one stream or with
 * events.  Gap scores must be non-positive here (this entry point carries the two-gap-state kernel only; swgpu_search()
 * also takes positive ones).
 */
#define SWGPU_NOALLOC_MAX_TARGET_LEN 4096
#define SWGPU_NOALLOC_SCRATCH_BYTES  (256 + 256 * SWGPU_NOALLOC_MAX_TARGET_LEN * 8)
int swgpu_score_targets_noalloc(int device, const int8_t *d_query, int32_t query_len,
                                const int8_t *d_targets, const int64_t *d_offsets, int32_t n_targets,
                                int32_t max_target_len, const int8_t *d_table, int gap_open, int gap_extend,
                                int32_t *d_scores, void *stream);

/* Debug mode, no counterpart in the reference: with SWGPU_GUARD=1 in the environment every device allocation of the engine carries a
 * 512-byte tail of a known pattern.  swgpu_guard_check() reads the tails of all live allocations back and returns the number of
 * damaged ones (each is described on stderr); -1 when the mode is off.  swgpu_db_destroy() runs it too and aborts on damage.
 * swgpu_guard_selftest() overruns a scratch allocation by one byte on purpose and returns 1 when the check reports exactly that. */
int swgpu_guard_check(void);
int swgpu_guard_selftest(void);

#ifdef __cplusplus
}
#endif
#endif /* SWGPU_H_SEEN */
