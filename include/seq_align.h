/*
 * seq_align.h -- version macros of the drop-in host library.
 * Replaces src/seq_align.h:12-13 of the reference (same macro names and values, so code that
 * tests SEQ_ALIGN_VERSION keeps compiling); SEQ_ALIGN_BACKEND is new.
 */
#ifndef SEQ_ALIGN_HEADER_SEEN
#define SEQ_ALIGN_HEADER_SEEN

#define SEQ_ALIGN_VERSION_STR  "1.0.0"
#define SEQ_ALIGN_VERSION      0x100
#define SEQ_ALIGN_BACKEND      "swgpu/sm_100a"

#endif
