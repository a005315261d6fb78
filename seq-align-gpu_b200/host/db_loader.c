/* db_loader.c -- see db_loader.h */
#include "db_loader.h"
#include "fasta_reader.h"
#include "alignment_scoring.h"

#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define SWDB_MAGIC   0x42445753u   /* "SWDB" */
#define SWDB_VERSION 1u

static void oom(void) { fprintf(stderr, "Error: out of memory\n"); exit(EXIT_FAILURE); }
static void *xmalloc(size_t n) { void *p = malloc(n ? n : 1); if (!p) oom(); return p; }

/* residue LUT: letters_to_index() without its side effects; -1 = illegal */
static int8_t g_lut[256];
static pthread_once_t g_lut_once = PTHREAD_ONCE_INIT;
static void lut_init(void)
{
    for (int c = 0; c < 256; c++) g_lut[c] = -1;
    for (int c = 'a'; c <= 'z'; c++) g_lut[c] = (int8_t)(c - 'a' + 1);
    for (int c = 'A'; c <= 'Z'; c++) g_lut[c] = (int8_t)(c - 'A' + 1);
    g_lut['*'] = 31;
}

void swdb_free(swdb_t *db)
{
    if (!db) return;
    if (db->map) munmap(db->map, db->map_len);
    else { free(db->codes); free(db->off); free(db->names); free(db->name_off); free(db->text); }
    memset(db, 0, sizeof(*db));
}

/* ------------------------------------------------------------------ streaming path (gz, stdin, FASTQ, plain) */

static int load_streaming(const char *path, bool use_zlib, swdb_t *out)
{
    swfa_reader_t *r = swfa_open(path, use_zlib);
    if (!r) return -1;
    size_t cap = 0, res_cap = 0, name_cap = 0;
    int64_t n = 0, res = 0, nb = 0;
    swfa_record_t rec = {0};
    out->off = NULL; out->name_off = NULL;
    while (swfa_read(r, &rec) > 0) {
        if ((size_t)n + 1 >= cap) {
            cap = cap ? cap * 2 : 4096;
            out->off = (int64_t *)realloc(out->off, (cap + 1) * sizeof(int64_t));
            out->name_off = (int64_t *)realloc(out->name_off, (cap + 1) * sizeof(int64_t));
            if (!out->off || !out->name_off) oom();
        }
        if ((size_t)res + rec.seq_len + 1 > res_cap) {
            res_cap = res_cap ? res_cap : (1u << 20);
            while ((size_t)res + rec.seq_len + 1 > res_cap) res_cap *= 2;
            out->codes = (int8_t *)realloc(out->codes, res_cap);
            out->text = (char *)realloc(out->text, res_cap);
            if (!out->codes || !out->text) oom();
        }
        if ((size_t)nb + rec.name_len + 1 > name_cap) {
            name_cap = name_cap ? name_cap : (1u << 16);
            while ((size_t)nb + rec.name_len + 1 > name_cap) name_cap *= 2;
            out->names = (char *)realloc(out->names, name_cap);
            if (!out->names) oom();
        }
        /* same encoding, same failure mode (stdout message + exit(1)) as src/alignment_cmdline.c:444-446 */
        for (size_t i = 0; i < rec.seq_len; i++) out->codes[res + (int64_t)i] = letters_to_index(rec.seq[i]);
        memcpy(out->text + res, rec.seq, rec.seq_len);
        memcpy(out->names + nb, rec.name, rec.name_len + 1);
        out->off[n] = res; out->name_off[n] = nb;
        res += (int64_t)rec.seq_len; nb += (int64_t)rec.name_len + 1;
        n++;
    }
    if (!out->off) { out->off = (int64_t *)xmalloc(sizeof(int64_t)); out->name_off = (int64_t *)xmalloc(sizeof(int64_t)); }
    out->off[n] = res; out->name_off[n] = nb;
    out->n = n; out->n_residues = res; out->threads_used = 0;
    swfa_record_free(&rec);
    swfa_close(r);
    return 0;
}

/* ------------------------------------------------------------------ parallel path (plain FASTA, mmap) */

typedef struct {
    const unsigned char *base; size_t begin, end;   /* byte range, begins at a '>' */
    int64_t n, res, nb;                             /* pass 1: counts */
    int64_t n0, res0, nb0;                          /* pass 2: where this range's output starts */
    swdb_t *out;
    size_t bad_pos; int pass;
} range_t;

static inline size_t chomp_cr(const unsigned char *s, size_t len) { while (len && s[len - 1] == '\r') len--; return len; }

static void *range_run(void *arg)
{
    range_t *g = (range_t *)arg;
    const unsigned char *b = g->base;
    size_t p = g->begin;
    int64_t n = 0, res = 0, nb = 0;
    swdb_t *o = g->out;
    g->bad_pos = (size_t)-1;
    while (p < g->end) {
        const unsigned char *nl = (const unsigned char *)memchr(b + p, '\n', g->end - p);
        const size_t le = nl ? (size_t)(nl - b) : g->end;
        if (b[p] == '>') {
            const size_t len = chomp_cr(b + p + 1, le - p - 1);
            if (g->pass == 2) {
                o->off[g->n0 + n] = g->res0 + res;
                o->name_off[g->n0 + n] = g->nb0 + nb;
                memcpy(o->names + g->nb0 + nb, b + p + 1, len);
                o->names[g->nb0 + nb + (int64_t)len] = 0;
            }
            n++; nb += (int64_t)len + 1;
        } else {
            const size_t len = chomp_cr(b + p, le - p);
            if (g->pass == 2) {
                int8_t *dst = o->codes + g->res0 + res;
                for (size_t i = 0; i < len; i++) {
                    const int8_t c = g_lut[b[p + i]];
                    if (c < 0 && g->bad_pos == (size_t)-1) g->bad_pos = p + i;
                    dst[i] = c;
                }
                memcpy(o->text + g->res0 + res, b + p, len);
            }
            res += (int64_t)len;
        }
        p = le + 1;
    }
    g->n = n; g->res = res; g->nb = nb;
    return NULL;
}

static int load_parallel(const unsigned char *base, size_t size, size_t first, int threads, swdb_t *out)
{
    if (threads < 1) threads = 1;
    if (size - first < ((size_t)1 << 20)) threads = 1;
    range_t *rg = (range_t *)calloc((size_t)threads, sizeof(range_t));
    pthread_t *th = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    if (!rg || !th) oom();
    /* range t starts at the first record start ('>' at the beginning of a line) at or after t/T of the file */
    int T = 1;
    rg[0].begin = first;
    for (int t = 1; t < threads; t++) {
        const size_t want = first + (size_t)((double)(size - first) * t / threads);
        if (want <= rg[T - 1].begin) continue;
        size_t b = size;
        for (size_t p = want - 1; p < size; ) {
            const unsigned char *nl = (const unsigned char *)memchr(base + p, '\n', size - p);
            if (!nl) break;
            const size_t q = (size_t)(nl - base) + 1;
            if (q < size && base[q] == '>') { b = q; break; }
            p = q;
        }
        if (b >= size) break;
        if (b > rg[T - 1].begin) rg[T++].begin = b;
    }
    for (int t = 0; t < T; t++) { rg[t].base = base; rg[t].out = out; rg[t].end = t + 1 < T ? rg[t + 1].begin : size; }
    for (int pass = 1; pass <= 2; pass++) {
        for (int t = 0; t < T; t++) { rg[t].pass = pass; if (T > 1) pthread_create(&th[t], NULL, range_run, &rg[t]); else range_run(&rg[t]); }
        for (int t = 0; t < T && T > 1; t++) pthread_join(th[t], NULL);
        if (pass == 1) {
            int64_t n = 0, res = 0, nb = 0;
            for (int t = 0; t < T; t++) { rg[t].n0 = n; rg[t].res0 = res; rg[t].nb0 = nb; n += rg[t].n; res += rg[t].res; nb += rg[t].nb; }
            out->n = n; out->n_residues = res;
            out->off = (int64_t *)xmalloc(sizeof(int64_t) * (size_t)(n + 1));
            out->name_off = (int64_t *)xmalloc(sizeof(int64_t) * (size_t)(n + 1));
            out->codes = (int8_t *)xmalloc((size_t)res);
            out->text = (char *)xmalloc((size_t)res + 1);
            out->names = (char *)xmalloc((size_t)nb + 1);
            out->off[n] = res; out->name_off[n] = nb;
        }
    }
    size_t bad = (size_t)-1;
    for (int t = 0; t < T; t++) if (rg[t].bad_pos < bad) bad = rg[t].bad_pos;
    out->threads_used = T;
    free(rg); free(th);
    if (bad != (size_t)-1) (void)letters_to_index((char)base[bad]);   /* prints the reference's message and exits */
    return 0;
}

int swdb_load(const char *path, bool use_zlib, int threads, swdb_t *out)
{
    memset(out, 0, sizeof(*out));
    pthread_once(&g_lut_once, lut_init);
    if (threads <= 0) {
        long nc = sysconf(_SC_NPROCESSORS_ONLN);
        threads = (int)(nc < 1 ? 1 : nc > 32 ? 32 : nc);
    }
    if (strcmp(path, "-") != 0) {
        const int fd = open(path, O_RDONLY);
        if (fd < 0) return -1;
        struct stat sb;
        if (fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) {
            const size_t size = (size_t)sb.st_size;
            unsigned char *base = (unsigned char *)mmap(NULL, size, PROT_READ, MAP_PRIVATE, fd, 0);
            if (base != MAP_FAILED) {
                size_t first = 0;
                while (first < size && (base[first] == '\n' || base[first] == '\r')) first++;
                const int gz = size >= 2 && base[0] == 0x1f && base[1] == 0x8b;
                if (!gz && first < size && base[first] == '>') {
                    (void)posix_madvise(base, size, POSIX_MADV_SEQUENTIAL);
                    const int rc = load_parallel(base, size, first, threads, out);
                    munmap(base, size); close(fd);
                    return rc;
                }
                munmap(base, size);
            }
        }
        close(fd);
    }
    return load_streaming(path, use_zlib, out);
}

/* ------------------------------------------------------------------ cache file */

typedef struct {
    uint32_t magic, version;
    int64_t n, n_residues, names_bytes, src_size, src_mtime_ns;
    int64_t reserved[2];
} swdb_header_t;                                   /* 64 bytes */

static size_t align64(size_t x) { return (x + 63) & ~(size_t)63; }

static int src_stamp(const char *src_path, int64_t *size, int64_t *mtime_ns)
{
    struct stat sb;
    *size = -1; *mtime_ns = -1;
    if (!src_path) return 0;
    if (stat(src_path, &sb) != 0) return -1;
    *size = (int64_t)sb.st_size;
    *mtime_ns = (int64_t)sb.st_mtim.tv_sec * 1000000000ll + (int64_t)sb.st_mtim.tv_nsec;
    return 0;
}

static int write_section(FILE *f, const void *p, size_t bytes)
{
    static const char zero[64] = {0};
    if (bytes && fwrite(p, 1, bytes, f) != bytes) return -1;
    const size_t pad = align64(bytes) - bytes;
    if (pad && fwrite(zero, 1, pad, f) != pad) return -1;
    return 0;
}

int swdb_cache_write(const char *cache_path, const swdb_t *db, const char *src_path)
{
    swdb_header_t h; memset(&h, 0, sizeof(h));
    h.magic = SWDB_MAGIC; h.version = SWDB_VERSION; h.n = db->n; h.n_residues = db->n_residues;
    h.names_bytes = db->name_off[db->n];
    if (src_stamp(src_path, &h.src_size, &h.src_mtime_ns) != 0) return -1;
    char tmp[4096];
    if (snprintf(tmp, sizeof(tmp), "%s.tmp.%ld", cache_path, (long)getpid()) >= (int)sizeof(tmp)) return -1;
    FILE *f = fopen(tmp, "wb");
    if (!f) return -1;
    int rc = 0;
    rc |= write_section(f, &h, sizeof(h));
    rc |= write_section(f, db->off, sizeof(int64_t) * (size_t)(db->n + 1));
    rc |= write_section(f, db->name_off, sizeof(int64_t) * (size_t)(db->n + 1));
    rc |= write_section(f, db->codes, (size_t)db->n_residues);
    rc |= write_section(f, db->names, (size_t)h.names_bytes);
    if (fclose(f) != 0) rc = -1;
    if (rc == 0 && rename(tmp, cache_path) != 0) rc = -1;    /* atomic publish: readers never see a partial file */
    if (rc != 0) unlink(tmp);
    return rc;
}

int swdb_cache_open(const char *cache_path, const char *src_path, swdb_t *out)
{
    memset(out, 0, sizeof(*out));
    const int fd = open(cache_path, O_RDONLY);
    if (fd < 0) return -1;
    struct stat sb;
    if (fstat(fd, &sb) != 0 || (size_t)sb.st_size < sizeof(swdb_header_t)) { close(fd); return -1; }
    const size_t size = (size_t)sb.st_size;
    unsigned char *base = (unsigned char *)mmap(NULL, size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (base == MAP_FAILED) return -1;
    const swdb_header_t *h = (const swdb_header_t *)base;
    int64_t ssize, smtime;
    int ok = h->magic == SWDB_MAGIC && h->version == SWDB_VERSION && h->n >= 0 && h->n_residues >= 0 && h->names_bytes >= 0;
    if (ok && src_path) ok = src_stamp(src_path, &ssize, &smtime) == 0 && ssize == h->src_size && smtime == h->src_mtime_ns;
    size_t o_off = 0, o_noff = 0, o_codes = 0, o_names = 0, total = 0;
    if (ok) {
        o_off = align64(sizeof(swdb_header_t));
        o_noff = o_off + align64(sizeof(int64_t) * (size_t)(h->n + 1));
        o_codes = o_noff + align64(sizeof(int64_t) * (size_t)(h->n + 1));
        o_names = o_codes + align64((size_t)h->n_residues);
        total = o_names + align64((size_t)h->names_bytes);
        ok = total == size;
    }
    if (ok) {
        const int64_t *off = (const int64_t *)(base + o_off), *noff = (const int64_t *)(base + o_noff);
        ok = off[0] == 0 && off[h->n] == h->n_residues && noff[0] == 0 && noff[h->n] == h->names_bytes;
    }
    if (!ok) { munmap(base, size); return -1; }
    out->n = h->n; out->n_residues = h->n_residues;
    out->off = (int64_t *)(base + o_off); out->name_off = (int64_t *)(base + o_noff);
    out->codes = (int8_t *)(base + o_codes); out->names = (char *)(base + o_names);
    out->text = NULL; out->from_cache = 1; out->threads_used = 0;
    out->map = base; out->map_len = size;
    return 0;
}
