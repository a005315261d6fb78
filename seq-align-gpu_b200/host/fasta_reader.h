/*
 * fasta_reader.h -- minimal sequence-file reader for the host driver (gzip-transparent).
 * Implements the record semantics the reference gets from its vendored seq_file library
 * (libs/seq_file/seq_file.h:274-295,311-323 of the reference tree):
 *   FASTA  name = the whole header line after '>' (CR/LF chomped), sequence = all following
 *          lines up to the next line starting with '>', each chomped and concatenated;
 *   FASTQ  '@name', sequence lines, '+', quality (skipped);
 *   plain  one sequence per non-empty line, empty name.
 * The format is sniffed from the first non-blank character.
 */
#ifndef SWGPU_FASTA_READER_H
#define SWGPU_FASTA_READER_H
#include <stddef.h>
#include <stdbool.h>

typedef struct swfa_reader swfa_reader_t;

typedef struct {
    char *name; size_t name_len, name_cap;
    char *seq;  size_t seq_len,  seq_cap;
} swfa_record_t;

/* path "-" = stdin.  use_zlib=false reads raw bytes (interactive stdin).  NULL if it cannot be opened. */
swfa_reader_t *swfa_open(const char *path, bool use_zlib);
/* 1 = record read, 0 = end of file, -1 = malformed */
int  swfa_read(swfa_reader_t *r, swfa_record_t *rec);
void swfa_close(swfa_reader_t *r);
void swfa_record_free(swfa_record_t *rec);

#endif
