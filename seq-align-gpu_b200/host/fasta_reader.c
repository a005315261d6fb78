/* fasta_reader.c -- see fasta_reader.h */
#include "fasta_reader.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#define SWFA_BUF (1u << 20)     /* 1 MiB read buffer, like the reference's stream buffer */

struct swfa_reader {
    gzFile gz; FILE *fp;
    unsigned char *buf; size_t pos, end;
    int fmt;                    /* 0 unknown, 1 FASTA, 2 FASTQ, 3 plain */
    int eof;
};

static int fill(swfa_reader_t *r)
{
    if (r->eof) return 0;
    int n = r->gz ? gzread(r->gz, r->buf, SWFA_BUF) : (int)fread(r->buf, 1, SWFA_BUF, r->fp);
    r->pos = 0; r->end = n > 0 ? (size_t)n : 0;
    if (n <= 0) r->eof = 1;
    return n > 0;
}

static inline int peekc(swfa_reader_t *r) { if (r->pos >= r->end && !fill(r)) return -1; return r->buf[r->pos]; }
static inline int getc_(swfa_reader_t *r) { const int c = peekc(r); if (c >= 0) r->pos++; return c; }

static void push(char **b, size_t *len, size_t *cap, const unsigned char *src, size_t n)
{
    if (*len + n + 1 > *cap) {
        size_t nc = *cap ? *cap : 256;
        while (nc < *len + n + 1) nc *= 2;
        *b = (char *)realloc(*b, nc); *cap = nc;
        if (!*b) { fprintf(stderr, "Error: out of memory reading sequence file\n"); exit(EXIT_FAILURE); }
    }
    if (n) memcpy(*b + *len, src, n);
    *len += n; (*b)[*len] = 0;
}

/* appends the rest of the current line (without CR/LF) to the buffer; consumes the newline */
static void read_line_into(swfa_reader_t *r, char **b, size_t *len, size_t *cap)
{
    for (;;) {
        if (r->pos >= r->end && !fill(r)) break;
        unsigned char *s = r->buf + r->pos;
        unsigned char *nl = (unsigned char *)memchr(s, '\n', r->end - r->pos);
        const size_t n = nl ? (size_t)(nl - s) : r->end - r->pos;
        push(b, len, cap, s, n);
        r->pos += n;
        if (nl) { r->pos++; break; }
    }
    while (*len && ((*b)[*len - 1] == '\r' || (*b)[*len - 1] == '\n')) (*b)[--*len] = 0;
}

swfa_reader_t *swfa_open(const char *path, bool use_zlib)
{
    swfa_reader_t *r = (swfa_reader_t *)calloc(1, sizeof(*r));
    if (!r) return NULL;
    const int is_stdin = strcmp(path, "-") == 0;
    if (use_zlib) r->gz = is_stdin ? gzdopen(fileno(stdin), "r") : gzopen(path, "r");
    else r->fp = is_stdin ? stdin : fopen(path, "r");
    if (!r->gz && !r->fp) { free(r); return NULL; }
    r->buf = (unsigned char *)malloc(SWFA_BUF);
    if (!r->buf) { swfa_close(r); return NULL; }
    return r;
}

void swfa_close(swfa_reader_t *r)
{
    if (!r) return;
    if (r->gz) gzclose(r->gz);
    if (r->fp && r->fp != stdin) fclose(r->fp);
    free(r->buf); free(r);
}

void swfa_record_free(swfa_record_t *rec) { free(rec->name); free(rec->seq); memset(rec, 0, sizeof(*rec)); }

int swfa_read(swfa_reader_t *r, swfa_record_t *rec)
{
    rec->name_len = rec->seq_len = 0;
    push(&rec->name, &rec->name_len, &rec->name_cap, NULL, 0);
    push(&rec->seq, &rec->seq_len, &rec->seq_cap, NULL, 0);
    int c;
    /* skip blank lines between records */
    while ((c = peekc(r)) == '\n' || c == '\r') r->pos++;
    if (c < 0) return 0;
    if (r->fmt == 0) r->fmt = (c == '>') ? 1 : (c == '@') ? 2 : 3;
    if (r->fmt == 1) {
        if (c != '>') return -1;
        r->pos++;
        read_line_into(r, &rec->name, &rec->name_len, &rec->name_cap);
        while ((c = peekc(r)) >= 0 && c != '>') read_line_into(r, &rec->seq, &rec->seq_len, &rec->seq_cap);
        return 1;
    }
    if (r->fmt == 2) {
        if (c != '@') return -1;
        r->pos++;
        read_line_into(r, &rec->name, &rec->name_len, &rec->name_cap);
        while ((c = peekc(r)) >= 0 && c != '+') read_line_into(r, &rec->seq, &rec->seq_len, &rec->seq_cap);
        if (c < 0) return -1;
        size_t qn = 0, qlen = 0, qcap = 0; char *q = NULL;
        read_line_into(r, &q, &qlen, &qcap);                    /* the '+' line */
        qlen = 0;
        while (qn < rec->seq_len && peekc(r) >= 0) { size_t before = qlen; read_line_into(r, &q, &qlen, &qcap); qn += qlen - before; }
        free(q);
        return 1;
    }
    read_line_into(r, &rec->seq, &rec->seq_len, &rec->seq_cap);
    (void)getc_;
    return 1;
}
