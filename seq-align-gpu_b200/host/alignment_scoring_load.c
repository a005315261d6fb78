/*
 * alignment_scoring_load.c -- substitution-matrix text parser.
 * Format accepted = the reference's (src/alignment_scoring_load.c:57-214):
 * '#' lines and blank lines skipped; the first real line lists the column letters; if it starts
 * with whitespace the file is whitespace-separated, otherwise its first character is the
 * separator; every following line is a row letter followed by one integer per column.
 * Errors: message on stderr, exit(EXIT_FAILURE).
 */
#include "alignment_scoring_load.h"
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static void load_fail(const char *msg, const char *path, long line)
{
    fprintf(stderr, "Error: substitution matrix : %s\n", msg);
    if (path) fprintf(stderr, "File: %s\n", path);
    if (line >= 0) fprintf(stderr, "Line: %ld\n", line);
    exit(EXIT_FAILURE);
}

/* reads one line of any length (without the line terminator); returns 0 at EOF */
static int read_line(gzFile f, char **buf, size_t *cap)
{
    size_t len = 0;
    for (;;) {
        if (*cap - len < 256) { *cap = *cap ? *cap * 2 : 1024; *buf = (char *)realloc(*buf, *cap); if (!*buf) exit(EXIT_FAILURE); }
        if (gzgets(f, *buf + len, (int)(*cap - len)) == NULL) { if (len == 0) return 0; break; }
        len += strlen(*buf + len);
        if (len && (*buf)[len - 1] == '\n') break;
    }
    while (len && ((*buf)[len - 1] == '\n' || (*buf)[len - 1] == '\r')) (*buf)[--len] = 0;
    return 1;
}

static int skippable(const char *s)
{
    if (*s == '#') return 1;
    while (*s) { if (!isspace((unsigned char)*s)) return 0; s++; }
    return 1;
}

void align_scoring_load_matrix(gzFile file, const char *path, scoring_t *sc, char case_sensitive)
{
    (void)case_sensitive;   /* letters_to_index folds case, as in the reference */
    char *line = NULL; size_t cap = 0; long lineno = 0;
    char cols[256]; int ncols = 0;
    int ws_separated = 0; char sep = 0;

    /* header */
    for (;;) {
        if (!read_line(file, &line, &cap)) load_fail("Empty file", path, -1);
        lineno++;
        if (!skippable(line)) break;
    }
    if (isspace((unsigned char)line[0])) {
        ws_separated = 1;
        for (const char *p = line; *p; p++) if (!isspace((unsigned char)*p)) { if (ncols == 255) load_fail("Too many columns", path, lineno); cols[ncols++] = *p; }
    } else {
        sep = line[0];
        for (const char *p = line; *p; ) {
            if (*p != sep) load_fail("Separator missing from line", path, lineno);
            p++;
            if (!*p) load_fail("Column letter missing", path, lineno);
            if (ncols == 255) load_fail("Too many columns", path, lineno);
            cols[ncols++] = *p++;
        }
    }
    if (ncols == 0) load_fail("No columns in substitution matrix", path, lineno);

    /* rows */
    while (read_line(file, &line, &cap)) {
        lineno++;
        if (skippable(line)) continue;
        const char *p = line;
        while (isspace((unsigned char)*p)) p++;
        const char row = *p++;
        for (int c = 0; c < ncols; c++) {
            if (ws_separated) { while (isspace((unsigned char)*p)) p++; }
            else { if (*p != sep) load_fail("Separator missing from line", path, lineno); p++; }
            char *end = NULL;
            const long v = strtol(p, &end, 10);
            if (end == p) load_fail("Missing number value on line", path, lineno);
            if (v <= -128 || v >= 128) load_fail("Score out of the int8 range", path, lineno);
            scoring_add_mutation(sc, row, cols[c], (int)v);
            p = end;
        }
        while (isspace((unsigned char)*p)) p++;
        if (*p) load_fail("Too many columns on row", path, lineno);
    }
    free(line);
}
