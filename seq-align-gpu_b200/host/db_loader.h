/*
 * db_loader.h -- database side of the host driver: FASTA -> encoded residues, and the packed on-disk cache.
 *
 * Replaces the serial, per-run parse + encode of the reference (src/alignment_cmdline.c:422-457 of the
 * reference tree: one record at a time through seq_file, strdup of name and sequence, letters_to_index per
 * residue), which is untimed there but dominates end-to-end once the DP takes milliseconds:
 *   * plain (uncompressed) FASTA files are mmap'ed and parsed by several threads, each owning a byte range
 *     that starts at a record boundary; two passes (count, fill) with a prefix sum between them;
 *   * gzip / stdin / FASTQ / plain-text inputs go through the streaming reader (fasta_reader.h);
 *   * a parsed database can be written to `<db>.swdb` and mmap'ed back on later runs.
 * Record semantics are those of fasta_reader.h; residue encoding and the failure mode of an
 * illegal character are letters_to_index()'s (message on stdout, exit status 1), reported for the FIRST
 * offending character in file order whatever the thread count.
 */
#ifndef SWGPU_DB_LOADER_H
#define SWGPU_DB_LOADER_H
#include <stdint.h>
#include <stddef.h>
#include <stdbool.h>

typedef struct {
    int64_t  n, n_residues;
    int8_t  *codes;        /* n_residues residue codes, entries back to back, file order */
    int64_t *off;          /* n + 1 prefix offsets into codes (and into text) */
    char    *names;        /* header lines back to back, each NUL-terminated */
    int64_t *name_off;     /* n + 1 offsets into names (name i = names + name_off[i]) */
    char    *text;         /* the residues as written in the file (case preserved), same offsets as codes;
                              NULL when the database came from a cache file (decode codes instead) */
    int      from_cache;   /* 1 = mmap'ed from a .swdb file */
    int      threads_used; /* threads of the mmap parser; 0 = streaming reader or cache */
    void    *map; size_t map_len;   /* private: backing mapping of a cache file */
} swdb_t;

/* Parse a sequence file.  threads <= 0: one per online core (capped at 32).  Returns 0, or -1 if the file
 * cannot be opened.  Exits like the reference on an illegal residue character. */
int  swdb_load(const char *path, bool use_zlib, int threads, swdb_t *out);

/* Cache file: little-endian, 64-byte header {"SWDB", version, n, n_residues, names_bytes, source size,
 * source mtime (ns)}, then off[n+1], name_off[n+1], codes, names -- each section 64-byte aligned.
 * swdb_cache_open() returns 0 only if the file is a valid cache OF `src_path` in its current state
 * (size and mtime match); otherwise -1 and the caller re-parses.  src_path may be NULL (no staleness check). */
int  swdb_cache_write(const char *cache_path, const swdb_t *db, const char *src_path);
int  swdb_cache_open(const char *cache_path, const char *src_path, swdb_t *out);

void swdb_free(swdb_t *db);

#endif
