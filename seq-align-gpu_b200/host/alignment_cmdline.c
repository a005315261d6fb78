/*
 * alignment_cmdline.c -- command line and the one-to-many driver on top of the GPU engine.
 * Interface and observable behaviour follow the reference's src/alignment_cmdline.c (flags :177-308,
 * driver :343-539); see include/alignment_cmdline.h.  The arithmetic lives in libswgpu.so only.
 */
#define _POSIX_C_SOURCE 200809L
#include "alignment_cmdline.h"
#include "alignment_scoring_load.h"
#include "fasta_reader.h"
#include "db_loader.h"
#include <time.h>
#include "swgpu.h"

#include <limits.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>
#include <zlib.h>

/* ---------------------------------------------------------------- number parsing */

char parse_entire_short(char *str, short *result)
{
    char *end = str;
    const long v = strtol(str, &end, 10);
    if (end != str + strlen(str) || v > SHRT_MAX || v < SHRT_MIN) return 0;
    *result = (short)v; return 1;
}

char parse_entire_ushort(char *str, unsigned short *result)
{
    char *end = str;
    const long v = strtol(str, &end, 10);
    if (end != str + strlen(str) || v > USHRT_MAX) return 0;
    *result = (unsigned short)v; return 1;
}

char parse_entire_int(char *str, int *result)
{
    char *end = str;
    const long v = strtol(str, &end, 10);
    if (end != str + strlen(str) || v > INT_MAX || v < INT_MIN) return 0;
    *result = (int)v; return 1;
}

char parse_entire_uint(char *str, unsigned int *result)
{
    char *end = str;
    const unsigned long v = strtoul(str, &end, 10);
    if (end != str + strlen(str) || v > UINT_MAX) return 0;
    *result = (unsigned int)v; return 1;
}

char parse_entire_score_t(char *str, score_t *result)
{
    short v;                                   /* score_t is int16_t (include/alignment_scoring.h) */
    if (!parse_entire_short(str, &v)) return 0;
    *result = v; return 1;
}

/* ---------------------------------------------------------------- usage */

static void usage_exit(const score_t defaults[4], const char *prog, const char *fmt, ...)
    __attribute__((format(printf, 3, 4))) __attribute__((noreturn));

static void usage_exit(const score_t defaults[4], const char *prog, const char *fmt, ...)
{
    if (fmt) {
        va_list ap;
        fprintf(stderr, "Error: ");
        va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap);
        if (fmt[strlen(fmt) - 1] != '\n') fputc('\n', stderr);
    }
    fprintf(stderr, "usage: %s [OPTIONS] [seq1 seq2]\n", prog);
    fprintf(stderr,
            "  Smith-Waterman optimal local alignment (maximises score).  \n"
            "  Scores ONE query sequence against every sequence of a database file on the GPU\n"
            "  (B200, sm_100a).  Can read gzip files, FASTA and FASTQ.\n\n"
            "  OPTIONS:\n"
            "    --file <file>        Sequence file reading with gzip support\n"
            "    --files <f1> <f2>    Query file <f1> (first record) against database file <f2>\n"
            "    --stdin              Read from STDIN (same as '--file -')\n"
            "\n"
            "    --match <score>      [default: %i]\n"
            "    --mismatch <score>   [default: %i]\n"
            "    --gapopen <score>    [default: %i]\n"
            "    --gapextend <score>  [default: %i]\n"
            "\n"
            "    --substitution_matrix <file>  see details for formatting\n\n"
            "    --minscore <score>   Minimum required score\n"
            "\n"
            "    --printseq           Print sequences before local alignments\n"
            "    --printmatrices      Print dynamic programming matrices\n"
            "    --printfasta         Print fasta header lines\n"
            "    --pretty             Print with a descriptor line\n"
            "    --colour             Print with colour\n"
            "\n",
            defaults[0], defaults[1], defaults[2], defaults[3]);
    printf("\n"
           " DETAILS:\n"
           "  * Gap (of length N) penalty is: (open+N*extend)\n"
           "  * To do alignment without affine gap penalty, set '--gapopen 0'.\n"
           "  * Scoring files should be matrices, with entries separated by a single\n"
           "    character or whitespace. See files in the 'scoring' directory for examples.\n"
           "  * Environment: SWGPU_GPUS=<n> shards the database over n GPUs,\n"
           "    SWGPU_TOPK=<k> prints only the k best entries (with coordinates).\n"
           "\n");
    exit(EXIT_FAILURE);
}

void cmdline_free(cmdline_t *cmd) { free(cmd); }
void cmdline_set_files(cmdline_t *cmd, char *p1, char *p2) { cmd->file_path1 = p1; cmd->file_path2 = p2; }
char *cmdline_get_file1(cmdline_t *cmd) { return cmd->file_path1; }
char *cmdline_get_file2(cmdline_t *cmd) { return cmd->file_path2; }

#define USAGE(...) usage_exit(defaults, argv[0], __VA_ARGS__)

cmdline_t *cmdline_new(int argc, char **argv, scoring_t *scoring, enum SeqAlignCmdType cmd_type)
{
    (void)cmd_type;
    cmdline_t *cmd = (cmdline_t *)calloc(1, sizeof(cmdline_t));
    const score_t defaults[4] = {(score_t)scoring->match, (score_t)scoring->mismatch, scoring->gap_open, scoring->gap_extend};
    if (argc == 1) USAGE(NULL);
    for (int i = 1; i < argc; i++)
        if (!strcasecmp(argv[i], "--help") || !strcasecmp(argv[i], "-help") || !strcasecmp(argv[i], "-h")) USAGE(NULL);

    bool matrix_set = false, match_set = false;
    for (int i = 1; i < argc; i++) {
        char *a = argv[i];
        if (a[0] != '-') {
            if (argc - i != 2) USAGE("Unknown options: '%s'", a);
            break;
        }
        if (!strcasecmp(a, "--printseq")) cmd->print_seq = true;
        else if (!strcasecmp(a, "--printmatrices")) cmd->print_matrices = true;
        else if (!strcasecmp(a, "--printfasta")) cmd->print_fasta = true;
        else if (!strcasecmp(a, "--pretty")) cmd->print_pretty = true;
        else if (!strcasecmp(a, "--colour")) cmd->print_colour = true;
        else if (!strcasecmp(a, "--stdin")) { cmdline_set_files(cmd, "", NULL); cmd->interactive = true; }
        else if (i == argc - 1) USAGE("Unknown argument without parameter: %s", a);
        else if (!strcasecmp(a, "--scoring")) i++;                       /* accepted and ignored, like the reference */
        else if (!strcasecmp(a, "--substitution_matrix")) {
            gzFile f = gzopen(argv[i + 1], "r");
            if (!f) USAGE("Couldn't read: %s", argv[i + 1]);
            align_scoring_load_matrix(f, argv[i + 1], scoring, true);
            gzclose(f);
            matrix_set = true; i++;
        }
        else if (!strcasecmp(a, "--match")) {
            if (!parse_entire_int(argv[i + 1], &scoring->match)) USAGE("Invalid --match argument ('%s') must be an int", argv[i + 1]);
            match_set = true; i++;
        }
        else if (!strcasecmp(a, "--mismatch")) {
            if (!parse_entire_int(argv[i + 1], &scoring->mismatch)) USAGE("Invalid --mismatch argument ('%s') must be an int", argv[i + 1]);
            i++;
        }
        else if (!strcasecmp(a, "--gapopen")) {
            if (!parse_entire_score_t(argv[i + 1], &scoring->gap_open)) USAGE("Invalid --gapopen argument ('%s') must be an int", argv[i + 1]);
            i++;
        }
        else if (!strcasecmp(a, "--gapextend")) {
            if (!parse_entire_score_t(argv[i + 1], &scoring->gap_extend)) USAGE("Invalid --gapextend argument ('%s') must be an int", argv[i + 1]);
            i++;
        }
        else if (!strcasecmp(a, "--minscore")) i++;                      /* listed in the usage text; no effect on the scoring path */
        else if (!strcasecmp(a, "--file")) { cmdline_set_files(cmd, argv[i + 1], NULL); i++; }
        else if (!strcasecmp(a, "--files")) {
            if (i >= argc - 2) USAGE("--files option takes 2 arguments");
            printf("Query File=%s and Database File=%s\n", argv[i + 1], argv[i + 2]);
            if (!strcmp(argv[i + 1], "-") && !strcmp(argv[i + 2], "-")) cmdline_set_files(cmd, argv[i + 1], NULL);
            else cmdline_set_files(cmd, argv[i + 1], argv[i + 2]);
            i += 2;
        }
        else USAGE("Unknown argument '%s'", a);
    }
    if (matrix_set && !match_set) scoring->use_match_mismatch = 0;
    if (scoring->use_match_mismatch && scoring->match < scoring->mismatch) USAGE("Match value should not be less than mismatch penalty");
    if (cmd->file_path1 == NULL || cmd->file_path2 == NULL) USAGE("No input specified");
    return cmd;
}

/* ---------------------------------------------------------------- one-to-many driver */

typedef struct {
    int device;
    const int8_t *codes; const int64_t *off; int64_t n;   /* this shard */
    const int64_t *entry;                                 /* shard index -> file index (NULL = identity) */
    int nq; const int8_t *const *queries; const int64_t *qlens; const scoring_t *sc;
    int32_t **scores_file;                                /* [nq] shared outputs, file order (NULL in top-k mode) */
    int topk; swgpu_hit_t *hits; int32_t *n_hits;         /* [nq * topk], [nq] */
    int want_cigar; char **cigars;                        /* [nq * topk] alignment paths of the hits (SWGPU_CIGAR=1) */
    double ms; int rc; char err[512];
} shard_job_t;

/* "12M2D30M" from run-length operations (length << 2 | code; 0 M, 1 I, 2 D) */
static char *cigar_string(const uint32_t *ops, int64_t n)
{
    char *out = (char *)malloc((size_t)n * 12 + 1);
    size_t w = 0;
    if (!out) return NULL;
    for (int64_t i = 0; i < n; i++) w += (size_t)sprintf(out + w, "%u%c", ops[i] >> 2, "MID?"[ops[i] & 3]);
    out[w] = 0;
    return out;
}

/* one GPU: pack this shard, then all queries through swgpu_search_many() (up to three in flight) */
static void *shard_run(void *arg)
{
    shard_job_t *j = (shard_job_t *)arg;
    swgpu_db_t *db = NULL;
    j->rc = swgpu_db_create(&db, j->device, j->codes, j->off, j->n, 0, NULL);
    if (j->rc == SWGPU_OK) {
        const size_t n1 = (size_t)(j->n ? j->n : 1);
        int32_t **tmp = NULL;
        if (j->scores_file) {
            tmp = (int32_t **)calloc((size_t)j->nq, sizeof(*tmp));
            for (int q = 0; q < j->nq; q++) tmp[q] = j->entry ? (int32_t *)malloc(sizeof(int32_t) * n1) : j->scores_file[q];
        }
        swgpu_batch_stats_t bs;
        j->rc = swgpu_search_many(db, j->nq, j->queries, j->qlens, &j->sc->swap_scores[0][0], j->sc->gap_open,
                                  j->sc->gap_extend, tmp, j->topk, j->hits, j->n_hits, NULL, &bs);
        if (j->rc == SWGPU_OK) {
            j->ms = bs.ms_total;
            for (int q = 0; q < j->nq && tmp && j->entry; q++)
                for (int64_t i = 0; i < j->n; i++) j->scores_file[q][j->entry[i]] = tmp[q][i];
            for (int q = 0; q < j->nq && j->topk > 0 && j->want_cigar && j->rc == SWGPU_OK; q++) {
                /* paths of this shard's hits, while the hits still carry shard-local entry indices */
                const swgpu_hit_t *hq = &j->hits[(size_t)q * (size_t)j->topk];
                int64_t cap = 1;
                for (int h = 0; h < j->n_hits[q]; h++)
                    if (hq[h].q_start >= 0) cap += (hq[h].q_end - hq[h].q_start + 1) + (hq[h].t_end - hq[h].t_start + 1);
                uint32_t *ops = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)cap);
                int64_t *off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(j->n_hits[q] + 1));
                j->rc = swgpu_traceback(db, j->queries[q], j->qlens[q], &j->sc->swap_scores[0][0], j->sc->gap_open,
                                        j->sc->gap_extend, hq, j->n_hits[q], ops, cap, off);
                for (int h = 0; h < j->n_hits[q] && j->rc == SWGPU_OK; h++)
                    j->cigars[(size_t)q * (size_t)j->topk + (size_t)h] = cigar_string(ops + off[h], off[h + 1] - off[h]);
                free(ops); free(off);
            }
            for (int q = 0; q < j->nq && j->topk > 0 && j->entry; q++)
                for (int h = 0; h < j->n_hits[q]; h++) {
                    swgpu_hit_t *hit = &j->hits[(size_t)q * (size_t)j->topk + (size_t)h];
                    hit->entry = j->entry[hit->entry];
                }
        }
        if (tmp) { for (int q = 0; q < j->nq; q++) if (j->entry) free(tmp[q]); free(tmp); }
    }
    if (j->rc != SWGPU_OK) snprintf(j->err, sizeof(j->err), "%s", swgpu_last_error());
    swgpu_db_destroy(db);
    return NULL;
}

typedef struct { swgpu_hit_t h; const char *cigar; } hitx_t;     /* h first: hit_cmp sorts these too */

static int hit_cmp(const void *a, const void *b)
{
    const swgpu_hit_t *x = (const swgpu_hit_t *)a, *y = (const swgpu_hit_t *)b;
    if (x->score != y->score) return x->score > y->score ? -1 : 1;
    return x->entry < y->entry ? -1 : (x->entry > y->entry);
}

static int i64_cmp(const void *a, const void *b)
{
    const int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
    return x < y ? -1 : (x > y);
}

/* length-balanced sharding: deal the targets, longest first, round-robin over the GPUs, so every
 * shard gets the same length mix and (to within one target per round) the same residue count.
 * Inside a shard the targets stay in ascending file order: the engine's tie-break (score desc,
 * shard-local entry asc) is then the global rule and the merged top-k does not depend on the GPU count. */
static void make_shards(const swdb_t *d, int g, int8_t **codes, int64_t **off, int64_t **entry, int64_t *cnt)
{
    const size_t n = (size_t)d->n;
    int64_t maxlen = 0;
    for (size_t t = 0; t < n; t++) { const int64_t l = d->off[t + 1] - d->off[t]; if (l > maxlen) maxlen = l; }
    int64_t *bucket = (int64_t *)calloc((size_t)maxlen + 2, sizeof(int64_t));
    int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (n ? n : 1));
    for (size_t t = 0; t < n; t++) bucket[maxlen - (d->off[t + 1] - d->off[t]) + 1]++;
    for (int64_t l = 0; l <= maxlen; l++) bucket[l + 1] += bucket[l];
    for (size_t t = 0; t < n; t++) order[bucket[maxlen - (d->off[t + 1] - d->off[t])]++] = (int64_t)t;
    free(bucket);
    for (int s = 0; s < g; s++) {
        int64_t m = 0, res = 0;
        for (size_t r = (size_t)s; r < n; r += (size_t)g) { m++; res += d->off[order[r] + 1] - d->off[order[r]]; }
        cnt[s] = m;
        codes[s] = (int8_t *)malloc((size_t)(res ? res : 1));
        off[s] = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m + 1));
        entry[s] = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m ? m : 1));
        int64_t k = 0; off[s][0] = 0;
        for (size_t r = (size_t)s; r < n; r += (size_t)g, k++) entry[s][k] = order[r];
        qsort(entry[s], (size_t)m, sizeof(int64_t), i64_cmp);
        for (k = 0; k < m; k++) {
            const int64_t t = entry[s][k], l = d->off[t + 1] - d->off[t];
            memcpy(codes[s] + off[s][k], d->codes + d->off[t], (size_t)l);
            off[s][k + 1] = off[s][k] + l;
        }
    }
    free(order);
}

void align_from_query_and_db(const char *query_path, const char *db_path, scoring_t *scoring,
                             void (print_alignment)(aligner_t *aligner, size_t total_cnt), bool use_zlib)
{
    swfa_reader_t *qf = swfa_open(query_path, use_zlib);
    if (!qf) { fprintf(stderr, "Error: couldn't open query file %s\n", query_path); fflush(stderr); return; }
    /* The reference aligns the FIRST record of the query file only (src/alignment_cmdline.c:369-380).  With
     * SWGPU_ALL_QUERIES=1 every record is a query (many-to-many): the listing below is repeated per query. */
    const char *eq = getenv("SWGPU_ALL_QUERIES");
    const int all_queries = eq && atoi(eq) > 0;
    swfa_record_t *qrec = NULL; int nq = 0, qcap = 0;
    for (;;) {
        swfa_record_t r = {0};
        const int got = swfa_read(qf, &r);
        if (got <= 0 || r.seq_len == 0) { swfa_record_free(&r); break; }
        if (nq == qcap) { qcap = qcap ? qcap * 2 : 4; qrec = (swfa_record_t *)realloc(qrec, (size_t)qcap * sizeof(*qrec)); }
        qrec[nq++] = r;
        if (!all_queries) break;
    }
    if (nq == 0) {
        fprintf(stderr, "Error: Query file %s is empty or invalid\n", query_path); fflush(stderr);
        swfa_close(qf); free(qrec); return;
    }
    /* query encode; letters whose diagonal the matrix never defined become 'X' (src/alignment_cmdline.c:388-396) */
    int8_t **qi = (int8_t **)calloc((size_t)nq, sizeof(*qi));
    int64_t *qlen = (int64_t *)calloc((size_t)nq, sizeof(int64_t));
    for (int k = 0; k < nq; k++) {
        qlen[k] = (int64_t)qrec[k].seq_len;
        qi[k] = (int8_t *)malloc(qrec[k].seq_len);
        for (size_t i = 0; i < qrec[k].seq_len; i++) {
            qi[k][i] = letters_to_index(qrec[k].seq[i]);
            if (!get_swap_bit(scoring, qi[k][i], qi[k][i])) qi[k][i] = letters_to_index('X');
        }
    }

    /* ---- database: packed cache if asked for and valid, else parse (parallel for plain FASTA files) ---- */
    swdb_t d; memset(&d, 0, sizeof(d));
    const char *ec = getenv("SWGPU_DB_CACHE"), *ev = getenv("SWGPU_VERBOSE"), *et = getenv("SWGPU_PARSE_THREADS");
    char cache_path[4096] = "";
    if (ec && *ec && strcmp(ec, "0") != 0 && strcmp(db_path, "-") != 0) {
        if (!strcmp(ec, "1")) snprintf(cache_path, sizeof(cache_path), "%s.swdb", db_path);
        else snprintf(cache_path, sizeof(cache_path), "%s", ec);
    }
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    int loaded = 0;
    if (cache_path[0] && swdb_cache_open(cache_path, db_path, &d) == 0) loaded = 1;
    if (!loaded) {
        if (swdb_load(db_path, use_zlib, et ? atoi(et) : 0, &d) != 0) {
            fprintf(stderr, "Error: couldn't open database file %s\n", db_path); fflush(stderr);
            for (int k = 0; k < nq; k++) { free(qi[k]); swfa_record_free(&qrec[k]); }
            free(qi); free(qlen); free(qrec); swfa_close(qf); return;
        }
        if (cache_path[0] && swdb_cache_write(cache_path, &d, db_path) != 0)
            fprintf(stderr, "Warning: could not write database cache %s\n", cache_path);
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (ev && atoi(ev) > 0)
        fprintf(stderr, "swgpu: database %s: %lld entries, %lld residues, loaded in %.3f s (%s, %d thread%s)\n", db_path,
                (long long)d.n, (long long)d.n_residues,
                (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec),
                d.from_cache ? "packed cache, mmap" : d.threads_used > 0 ? "mmap FASTA parser" : "streaming reader",
                d.threads_used > 0 ? d.threads_used : 1, d.threads_used > 1 ? "s" : "");

    /* ---- GPU search (the reference's timed `omp parallel for`, src/alignment_cmdline.c:503-509) ---- */
    int ndev = swgpu_device_count();
    if (ndev < 1) { fprintf(stderr, "Error: no CUDA device available (this build has no CPU path)\n"); exit(EXIT_FAILURE); }
    int g = 1;
    const char *eg = getenv("SWGPU_GPUS");
    if (eg && atoi(eg) > 0) g = atoi(eg);
    /* SWGPU_GPUS_SHARE_DEVICE=1: more shards than devices are allowed and share the devices round-robin (the shard /
     * merge path on a box with fewer GPUs; the shards' searches then run concurrently on one device) */
    const char *esh = getenv("SWGPU_GPUS_SHARE_DEVICE");
    const int share = esh && atoi(esh) > 0;
    if (g > ndev && !share) g = ndev;
    if (g > 64) g = 64;
    if ((int64_t)g > d.n) g = d.n ? (int)d.n : 1;
    int topk = 0;
    const char *ek = getenv("SWGPU_TOPK");
    if (ek && atoi(ek) > 0) topk = atoi(ek);
    const char *eci = getenv("SWGPU_CIGAR");
    const int want_cigar = topk > 0 && eci && atoi(eci) > 0;

    int32_t **scores = NULL;
    if (!topk) {
        scores = (int32_t **)calloc((size_t)nq, sizeof(*scores));
        for (int k = 0; k < nq; k++) scores[k] = (int32_t *)calloc(d.n ? (size_t)d.n : 1, sizeof(int32_t));
    }
    shard_job_t *jobs = (shard_job_t *)calloc((size_t)g, sizeof(shard_job_t));
    int8_t **sc_codes = (int8_t **)calloc((size_t)g, sizeof(*sc_codes));
    int64_t **sc_off = (int64_t **)calloc((size_t)g, sizeof(*sc_off)), **sc_entry = (int64_t **)calloc((size_t)g, sizeof(*sc_entry));
    int64_t *sc_cnt = (int64_t *)calloc((size_t)g, sizeof(int64_t));
    if (g > 1) make_shards(&d, g, sc_codes, sc_off, sc_entry, sc_cnt);
    pthread_t *th = (pthread_t *)calloc((size_t)g, sizeof(pthread_t));
    for (int s = 0; s < g; s++) {
        shard_job_t *j = &jobs[s];
        j->device = s % ndev; j->nq = nq; j->queries = (const int8_t *const *)qi; j->qlens = qlen; j->sc = scoring;
        j->scores_file = scores;
        j->topk = topk;
        j->hits = topk ? (swgpu_hit_t *)calloc((size_t)topk * (size_t)nq, sizeof(swgpu_hit_t)) : NULL;
        j->n_hits = topk ? (int32_t *)calloc((size_t)nq, sizeof(int32_t)) : NULL;
        j->want_cigar = want_cigar;
        j->cigars = want_cigar ? (char **)calloc((size_t)topk * (size_t)nq, sizeof(char *)) : NULL;
        if (g > 1) { j->codes = sc_codes[s]; j->off = sc_off[s]; j->n = sc_cnt[s]; j->entry = sc_entry[s]; }
        else { j->codes = d.codes; j->off = d.off; j->n = (int64_t)d.n; j->entry = NULL; }
        if (g > 1) pthread_create(&th[s], NULL, shard_run, j); else shard_run(j);
    }
    double total_ms = 0;
    for (int s = 0; s < g; s++) {
        if (g > 1) pthread_join(th[s], NULL);
        if (jobs[s].rc != SWGPU_OK) { fprintf(stderr, "Error: GPU search failed on device %d: %s\n", s, jobs[s].err); exit(EXIT_FAILURE); }
        if (jobs[s].ms > total_ms) total_ms = jobs[s].ms;          /* GPUs run concurrently: max, not sum */
    }

    /* ---- hand-off: the reference's callback over groups of 16 entries in file order (:511-512) ---- */
    for (int k = 0; k < nq; k++) {
        if (topk > 0) {
            hitx_t *all = (hitx_t *)calloc((size_t)topk * (size_t)g, sizeof(hitx_t));
            int m = 0;
            for (int s = 0; s < g; s++)
                for (int h = 0; h < jobs[s].n_hits[k]; h++) {
                    const size_t at = (size_t)k * (size_t)topk + (size_t)h;
                    all[m].h = jobs[s].hits[at];
                    all[m].cigar = jobs[s].cigars ? jobs[s].cigars[at] : NULL;
                    m++;
                }
            qsort(all, (size_t)m, sizeof(hitx_t), hit_cmp);        /* score desc, entry asc: independent of the GPU count */
            if (m > topk) m = topk;
            if (all_queries) printf("Query #%d: %s\n", k, qrec[k].name);
            for (int h = 0; h < m; h++) {
                const swgpu_hit_t *x = &all[h].h;
                printf("Hit #%d: entry %lld score %d query [%d..%d] target [%d..%d] %s\n", h, (long long)x->entry, x->score,
                       x->q_start, x->q_end, x->t_start, x->t_end, d.names + d.name_off[x->entry]);
                if (want_cigar) printf("  cigar: %s\n", all[h].cigar ? all[h].cigar : "");
            }
            free(all);
        } else {
            aligner_t view; aligner_init(&view);
            score_t s16[16]; int32_t s32[16];
            view.scoring = scoring; view.seq_a_indexes = qi[k]; view.seq_a_str = qrec[k].seq; view.seq_a_fasta = qrec[k].name;
            view.score_width = qrec[k].seq_len + 1; view.max_scores = s16; view.max_scores32 = s32;
            char *b_str[16], *b_name[16];
            char *strbuf = NULL; size_t strcap = 0;
            view.seq_b_str_batch = b_str; view.seq_b_fasta_batch = b_name;
            for (size_t first = 0; first < (size_t)d.n; first += 16) {
                const size_t lanes = (size_t)d.n - first < 16 ? (size_t)d.n - first : 16;
                view.vector_size = lanes;
                /* the 16 sequences as NUL-terminated strings: the file's own text, or (database from a packed
                 * cache) the residue codes decoded to upper-case letters */
                const size_t need = (size_t)(d.off[first + lanes] - d.off[first]) + lanes;
                if (need > strcap) { strcap = need * 2; strbuf = (char *)realloc(strbuf, strcap); if (!strbuf) { fprintf(stderr, "Error: out of memory\n"); exit(EXIT_FAILURE); } }
                size_t maxlen = 0, w = 0;
                for (size_t l = 0; l < lanes; l++) {
                    s32[l] = scores[k][first + l];
                    s16[l] = (score_t)(s32[l] > 32767 ? 32767 : s32[l]);
                    const size_t len = (size_t)(d.off[first + l + 1] - d.off[first + l]);
                    if (len > maxlen) maxlen = len;
                    b_name[l] = d.names + d.name_off[first + l];
                    b_str[l] = strbuf + w;
                    if (d.text) memcpy(strbuf + w, d.text + d.off[first + l], len);
                    else for (size_t i = 0; i < len; i++) strbuf[w + i] = index_to_letters(d.codes[d.off[first + l] + (int64_t)i]);
                    strbuf[w + len] = 0; w += len + 1;
                }
                view.score_height = maxlen + 1;
                print_alignment(&view, first);
            }
            free(strbuf);
        }
    }
    printf("Total Time: %f\n", total_ms * 1e-3);
    printf("Total Entries: %lu\n", (unsigned long)d.n);

    for (int s = 0; s < g; s++) {
        for (size_t c = 0; jobs[s].cigars && c < (size_t)topk * (size_t)nq; c++) free(jobs[s].cigars[c]);
        free(jobs[s].cigars);
    }
    for (int s = 0; s < g; s++) { free(jobs[s].hits); free(jobs[s].n_hits); free(sc_codes[s]); free(sc_off[s]); free(sc_entry[s]); }
    free(jobs); free(sc_codes); free(sc_off); free(sc_entry); free(sc_cnt); free(th);
    for (int k = 0; k < nq && scores; k++) free(scores[k]);
    free(scores);
    swdb_free(&d);
    for (int k = 0; k < nq; k++) { free(qi[k]); swfa_record_free(&qrec[k]); }
    free(qi); free(qlen); free(qrec);
    swfa_close(qf);
}
