/*
 * sw_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE ONLY, NOT PRODUCT CODE)
 *
 * A plain-C restatement of the reference's one-to-many Smith-Waterman scoring
 * path, used only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
 * leg as the checker.  The product path (libswgpu.so + CUDA kernels) never
 * includes, links or calls anything in this file.
 *
 * What is restated, with the reference lines each function follows
 * (paths relative to the reference tree):
 *   swo_score32       3-state affine recurrence, src/alignment.c:117-176,
 *                     boundary rows/cols zero (:77-85, :91-97), running max of the
 *                     match state (:133, :186) -- in int32, one pair, no padding.
 *   swo_score16_wrap  same recurrence in WRAPPING int16 (the reference's
 *                     _mm256_add_epi16 / _mm256_max_epi16 semantics, :124-161) with
 *                     the reference's '*'(=31) padding rows up to the batch's padded
 *                     length (src/alignment_cmdline.c:448-450).
 *   swo_letters_to_index   src/alignment_scoring.c:70-81 (returns -1 instead of exit()).
 *   swo_encode_query  src/alignment_cmdline.c:388-396 (unknown letter -> 'X').
 *   swo_end_start     end / start coordinate rule of SURVEY.md A.7.  The reference
 *                     has NO counterpart (it keeps only a max, alignment.c:133):
 *                     PARITY UNPINNED for coordinates; this file defines the rule.
 *   swo_search32      OpenMP loop over targets, the analogue of
 *                     src/alignment_cmdline.c:504-507 (parallel for, dynamic,1).
 *
 * Pinning: oracle/_ref (the reference's own sources compiled here, see
 * oracle/Makefile) is the authority; tests/test_oracle.py checks this file
 * against it and against tests/golden/ vectors generated from it.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SWO_MAX2(a, b) ((a) > (b) ? (a) : (b))

/* src/alignment_scoring.c:70-81 */
int swo_letters_to_index(int c)
{
    if (c >= 'a' && c <= 'z') return c - 'a' + 1;
    if (c >= 'A' && c <= 'Z') return c - 'A' + 1;
    if (c == '*') return 31;
    return -1; /* the reference prints an error on stdout and exit(1)s */
}

/* src/alignment_cmdline.c:388-396: letters whose diagonal cell was never set by
 * the matrix file are replaced by 'X' (24).  defined_diag[c] != 0 <=> swap bit (c,c) set. */
int swo_encode_query(const char *seq, long len, const uint8_t defined_diag[32], int8_t *out)
{
    for (long i = 0; i < len; i++) {
        int c = swo_letters_to_index((unsigned char)seq[i]);
        if (c < 0) return -1;
        if (!defined_diag[c]) c = 24;
        out[i] = (int8_t)c;
    }
    return 0;
}

/*
 * 3-state recurrence of src/alignment.c:117-176, one (query,target) pair, int32.
 *   M[j][i] = max(0, M[j-1][i-1]+s, E[j-1][i-1]+s, F[j-1][i-1]+s)      :124-129
 *   E[j][i] = max(0, M[j-1][i]+oe,  E[j-1][i]+ge,  F[j-1][i]+oe)       :142-147
 *   F[j][i] = max(0, M[j][i-1]+oe,  E[j][i-1]+oe,  F[j][i-1]+ge)       :156-161
 *   score   = max M                                                    :133
 * j = target position (outer loop, :88), i = query position (inner loop, :103),
 * s = mat[query code][target code] (:33,:41), oe = gap_open + gap_extend (:58).
 */
int32_t swo_score32(const int8_t *q, long lq, const int8_t *t, long lt,
                    const int8_t *mat /* [32][32] row-major */, int gap_open, int gap_extend)
{
    if (lq <= 0 || lt <= 0) return 0;
    const int32_t oe = gap_open + gap_extend, ge = gap_extend;
    int32_t *M = (int32_t *)calloc((size_t)(lq + 1) * 3, sizeof(int32_t));
    int32_t *E = M + (lq + 1), *F = E + (lq + 1);
    int32_t best = 0;
    for (long j = 0; j < lt; j++) {
        int32_t m_left = 0, e_left = 0, f_left = 0;      /* column 0 of this row   */
        int32_t m_ul = 0, e_ul = 0, f_ul = 0;            /* column 0 of row above  */
        const int tc = t[j] & 31;
        for (long i = 0; i < lq; i++) {
            const int32_t s = mat[(q[i] & 31) * 32 + tc];
            const int32_t m_up = M[i + 1], e_up = E[i + 1], f_up = F[i + 1];
            int32_t m = SWO_MAX2(SWO_MAX2(m_ul + s, e_ul + s), SWO_MAX2(f_ul + s, 0));
            if (m > best) best = m;
            int32_t e = SWO_MAX2(SWO_MAX2(m_up + oe, e_up + ge), SWO_MAX2(f_up + oe, 0));
            int32_t f = SWO_MAX2(SWO_MAX2(m_left + oe, e_left + oe), SWO_MAX2(f_left + ge, 0));
            M[i + 1] = m; E[i + 1] = e; F[i + 1] = f;
            m_ul = m_up; e_ul = e_up; f_ul = f_up;
            m_left = m; e_left = e; f_left = f;
        }
    }
    free(M);
    return best;
}

/* wrapping int16 add == _mm256_add_epi16 on one lane */
static inline int16_t add16(int16_t a, int16_t b) { return (int16_t)(uint16_t)((uint16_t)a + (uint16_t)b); }
static inline int16_t max16(int16_t a, int16_t b) { return a > b ? a : b; }

/*
 * The reference's actual arithmetic for one lane: int16 with wrapping adds
 * (src/alignment.c:124-161) and `padded_len - lt` extra rows of code 31 ('*',
 * src/alignment_cmdline.c:448-450).  Lanes of a batch never interact, so one lane
 * at a time reproduces libalign exactly.
 */
int16_t swo_score16_wrap(const int8_t *q, long lq, const int8_t *t, long lt, long padded_len,
                         const int8_t *mat, int gap_open, int gap_extend)
{
    if (lq <= 0) return 0;
    const int16_t oe = (int16_t)(gap_open + gap_extend), ge = (int16_t)gap_extend;
    int16_t *M = (int16_t *)calloc((size_t)(lq + 1) * 3, sizeof(int16_t));
    int16_t *E = M + (lq + 1), *F = E + (lq + 1);
    int16_t best = 0;
    if (padded_len < lt) padded_len = lt;
    for (long j = 0; j < padded_len; j++) {
        int16_t m_left = 0, e_left = 0, f_left = 0, m_ul = 0, e_ul = 0, f_ul = 0;
        const int tc = (j < lt) ? (t[j] & 31) : 31;
        for (long i = 0; i < lq; i++) {
            const int16_t s = mat[(q[i] & 31) * 32 + tc];
            const int16_t m_up = M[i + 1], e_up = E[i + 1], f_up = F[i + 1];
            int16_t m = max16(max16(add16(m_ul, s), add16(e_ul, s)), max16(add16(f_ul, s), 0));
            best = max16(m, best);
            int16_t e = max16(max16(add16(m_up, oe), add16(e_up, ge)), max16(add16(f_up, oe), 0));
            int16_t f = max16(max16(add16(m_left, oe), add16(e_left, oe)), max16(add16(f_left, ge), 0));
            M[i + 1] = m; E[i + 1] = e; F[i + 1] = f;
            m_ul = m_up; e_ul = e_up; f_ul = f_up;
            m_left = m; e_left = e; f_left = f;
        }
    }
    free(M);
    return best;
}

/*
 * Coordinates (no reference counterpart -> PARITY UNPINNED, this
 * restatement is the definition).  End = first cell reaching the maximum in the
 * reference's scan order (target position j outer, query position i inner,
 * strict '>' update, alignment.c:88,103,133).  Start = the same rule applied to
 * the reversed prefixes q[i_end..0], t[j_end..0]: the first cell (reversed scan
 * order) whose match state equals the best score, mapped back.
 * All coordinates 0-based inclusive.  Returns the score; coords = -1 if score 0.
 */
static int32_t swo_scan_first_max(const int8_t *q, long lq, long q_step, const int8_t *t, long lt, long t_step,
                                  const int8_t *mat, int gap_open, int gap_extend,
                                  int32_t stop_at, long *bi, long *bj)
{
    const int32_t oe = gap_open + gap_extend, ge = gap_extend;
    int32_t *M = (int32_t *)calloc((size_t)(lq + 1) * 3, sizeof(int32_t));
    int32_t *E = M + (lq + 1), *F = E + (lq + 1);
    int32_t best = 0; *bi = -1; *bj = -1;
    for (long j = 0; j < lt; j++) {
        int32_t m_left = 0, e_left = 0, f_left = 0, m_ul = 0, e_ul = 0, f_ul = 0;
        const int tc = t[j * t_step] & 31;
        for (long i = 0; i < lq; i++) {
            const int32_t s = mat[(q[i * q_step] & 31) * 32 + tc];
            const int32_t m_up = M[i + 1], e_up = E[i + 1], f_up = F[i + 1];
            int32_t m = SWO_MAX2(SWO_MAX2(m_ul + s, e_ul + s), SWO_MAX2(f_ul + s, 0));
            if (m > best) { best = m; *bi = i; *bj = j;
                if (stop_at > 0 && best >= stop_at) { free(M); return best; } }
            int32_t e = SWO_MAX2(SWO_MAX2(m_up + oe, e_up + ge), SWO_MAX2(f_up + oe, 0));
            int32_t f = SWO_MAX2(SWO_MAX2(m_left + oe, e_left + oe), SWO_MAX2(f_left + ge, 0));
            M[i + 1] = m; E[i + 1] = e; F[i + 1] = f;
            m_ul = m_up; e_ul = e_up; f_ul = f_up;
            m_left = m; e_left = e; f_left = f;
        }
    }
    free(M);
    return best;
}

/* Number of cells whose match state equals the best score (1 = the end cell of the optimum is unique): lets a test
 * decide where an end coordinate printed by ANOTHER implementation (any tie-break) must equal this file's. */
long swo_count_best(const int8_t *q, long lq, const int8_t *t, long lt, const int8_t *mat, int gap_open, int gap_extend)
{
    if (lq <= 0 || lt <= 0) return 0;
    const int32_t oe = gap_open + gap_extend, ge = gap_extend;
    int32_t *M = (int32_t *)calloc((size_t)(lq + 1) * 3, sizeof(int32_t));
    int32_t *E = M + (lq + 1), *F = E + (lq + 1);
    int32_t best = 0; long count = 0;
    for (long j = 0; j < lt; j++) {
        int32_t m_left = 0, e_left = 0, f_left = 0, m_ul = 0, e_ul = 0, f_ul = 0;
        const int tc = t[j] & 31;
        for (long i = 0; i < lq; i++) {
            const int32_t s = mat[(q[i] & 31) * 32 + tc];
            const int32_t m_up = M[i + 1], e_up = E[i + 1], f_up = F[i + 1];
            int32_t m = SWO_MAX2(SWO_MAX2(m_ul + s, e_ul + s), SWO_MAX2(f_ul + s, 0));
            if (m > best) { best = m; count = 1; } else if (m == best && m > 0) count++;
            int32_t e = SWO_MAX2(SWO_MAX2(m_up + oe, e_up + ge), SWO_MAX2(f_up + oe, 0));
            int32_t f = SWO_MAX2(SWO_MAX2(m_left + oe, e_left + oe), SWO_MAX2(f_left + ge, 0));
            M[i + 1] = m; E[i + 1] = e; F[i + 1] = f;
            m_ul = m_up; e_ul = e_up; f_ul = f_up;
            m_left = m; e_left = e; f_left = f;
        }
    }
    free(M);
    return count;
}

int32_t swo_end_start(const int8_t *q, long lq, const int8_t *t, long lt,
                      const int8_t *mat, int gap_open, int gap_extend,
                      int32_t coords[4] /* q_start, q_end, t_start, t_end */)
{
    long ie, je, is, js;
    coords[0] = coords[1] = coords[2] = coords[3] = -1;
    if (lq <= 0 || lt <= 0) return 0;
    int32_t best = swo_scan_first_max(q, lq, 1, t, lt, 1, mat, gap_open, gap_extend, 0, &ie, &je);
    if (best <= 0) return best;
    /* reversed prefixes: q[ie], q[ie-1], ... and t[je], t[je-1], ... */
    int32_t rb = swo_scan_first_max(q + ie, ie + 1, -1, t + je, je + 1, -1, mat, gap_open, gap_extend, best, &is, &js);
    (void)rb;
    coords[0] = (int32_t)(ie - is); coords[1] = (int32_t)ie;
    coords[2] = (int32_t)(je - js); coords[3] = (int32_t)je;
    return best;
}

/*
 * One-to-many search: scores[n] for n targets stored back to back in `residues`
 * with offsets[n+1].  OpenMP over targets, dynamic schedule, like
 * src/alignment_cmdline.c:504-507.  Used as bench.py's cpu_baseline "port" leg
 * and as the bulk checker.  threads <= 0 -> omp default.
 */
void swo_search32(const int8_t *q, long lq, const int8_t *residues, const int64_t *offsets, long n,
                  const int8_t *mat, int gap_open, int gap_extend, int threads, int32_t *scores)
{
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#else
    (void)threads;
#endif
#pragma omp parallel for schedule(dynamic, 16)
    for (long k = 0; k < n; k++)
        scores[k] = swo_score32(q, lq, residues + offsets[k], (long)(offsets[k + 1] - offsets[k]), mat, gap_open, gap_extend);
}

/* Alignment path of a hit (no reference counterpart: PARITY UNPINNED; the rule is the one documented at
 * include/swgpu.h::swgpu_traceback and this function is its CPU restatement).  Affine GLOBAL alignment of
 * q[0..m) and t[0..n) forced to start and end in a match state; a cell prefers the diagonal, then a gap in the
 * query (consumes a target residue, 'D' = code 2), then a gap in the target ('I' = code 1); a gap prefers to
 * open rather than extend.  Writes run-length encoded operations (length << 2 | code, code 0 = 'M') in path
 * order into ops[cap]; returns the number of operations (or -1 if cap is too small / out of memory) and the
 * score of the path in *score. */
long swo_traceback(const int8_t *q, long m, const int8_t *t, long n, const int8_t *table, int gap_open, int gap_extend,
                   uint32_t *ops, long cap, int32_t *score)
{
    const int32_t NEG = -(1 << 29);
    const int oe = gap_open + gap_extend, ge = gap_extend;
    if (m <= 0 || n <= 0) { *score = 0; return 0; }
    uint8_t *dir = (uint8_t *)malloc((size_t)m * (size_t)n);
    int32_t *Hrow = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));    /* H(i-1, .) then H(i, .) */
    int32_t *Frow = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));    /* F(i-1, .) then F(i, .) */
    if (!dir || !Hrow || !Frow) { free(dir); free(Hrow); free(Frow); return -1; }
    for (long j = 0; j < n; j++) { Hrow[j] = NEG; Frow[j] = NEG; }
    for (long i = 0; i < m; i++) {
        int32_t Hl = NEG, El = NEG;                          /* H(i, j-1), E(i, j-1) */
        int32_t diag = (i == 0) ? 0 : NEG;                   /* H(i-1, j-1): the forced corner */
        const int8_t *row = table + (q[i] & 31) * 32;
        for (long j = 0; j < n; j++) {
            const int32_t up_h = Hrow[j], up_f = Frow[j];
            const int32_t D = diag > NEG / 2 ? diag + row[t[j] & 31] : NEG;
            const int32_t e_open = Hl + oe, e_ext = El + ge;
            const int32_t E = e_open > e_ext ? e_open : e_ext;
            const int32_t f_open = up_h + oe, f_ext = up_f + ge;
            const int32_t F = f_open > f_ext ? f_open : f_ext;
            int32_t H = D; uint8_t code = 0;
            if (E > H) { H = E; code = 1; }
            if (F > H) { H = F; code = 2; }
            if (e_ext > e_open) code |= 4;
            if (f_ext > f_open) code |= 8;
            if (H < NEG) H = NEG;
            dir[(size_t)i * (size_t)n + (size_t)j] = code;
            diag = up_h;
            Hl = H; El = E < NEG ? NEG : E;
            Hrow[j] = H; Frow[j] = F < NEG ? NEG : F;
        }
    }
    *score = Hrow[n - 1];
    /* walk back, collecting reversed run-length operations */
    long nops = 0, i = m - 1, j = n - 1; int state = 0;
    uint32_t cur_op = 3, cur_len = 0;
    uint32_t *rev = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(m + n));
    if (!rev) { free(dir); free(Hrow); free(Frow); return -1; }
    while (i >= 0 && j >= 0) {
        const uint8_t c = dir[(size_t)i * (size_t)n + (size_t)j];
        uint32_t op;
        if (state == 0) {
            const int src = c & 3;
            if (src == 1) { state = 1; continue; }
            if (src == 2) { state = 2; continue; }
            op = 0; i--; j--;
        } else if (state == 1) { op = 2; state = (c & 4) ? 1 : 0; j--; }
        else { op = 1; state = (c & 8) ? 2 : 0; i--; }
        if (op == cur_op) cur_len++;
        else { if (cur_len) rev[nops++] = cur_len << 2 | cur_op; cur_op = op; cur_len = 1; }
    }
    if (cur_len) rev[nops++] = cur_len << 2 | cur_op;
    long ret = nops;
    if (i != -1 || j != -1 || state != 0 || nops > cap) ret = -1;
    else for (long x = 0; x < nops; x++) ops[x] = rev[nops - 1 - x];
    free(rev); free(dir); free(Hrow); free(Frow);
    return ret;
}

int swo_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
