// swgpu_internal.h -- engine-private state shared by swgpu.cu and swgpu_topk.cu (not part of the C-ABI).
#pragma once
#include "../../include/swgpu.h"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdarg>
#include <cstdint>
#include <vector>

namespace swi {

inline char* err_buf() { static thread_local char buf[512] = ""; return buf; }

inline int fail(int code, const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(err_buf(), 512, fmt, ap); va_end(ap);
  return code;
}

}  // namespace swi

// Guard zones (debug mode, SWGPU_GUARD=1 in the environment): every device allocation of the engine gets a 512-byte tail filled
// with 0xA5, and swgpu_guard_check() (include/swgpu.h) reads all live tails back and counts the damaged ones -- a write past the
// end of any engine buffer shows up there.  This pool refuses compute-sanitizer (profiles/r2_sanitize_gpu.txt); this is what
// stands in for its memcheck on the write side.  Off (the default), the two wrappers are cudaMalloc / cudaFree.
namespace swi {
cudaError_t guard_malloc(void** p, size_t bytes);
cudaError_t guard_free(void* p);
int guard_check();                     // damaged tails among this process's live engine allocations; -1 when the mode is off
}
#define cudaMalloc(p, n) swi::guard_malloc((void**)(p), (n))
#define cudaFree(p) swi::guard_free((void*)(p))

#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return swi::fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? SWGPU_ENODEV : SWGPU_ECUDA, \
              "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

constexpr int SWGPU_SLOTS = 3;        // query slots: up to 3 searches of one batch are in flight (swgpu_search_many)

// Everything one search writes: scratch in HBM, streams, events, pinned read-back staging.  Slot 0 serves the
// single-query entry points; swgpu_search_many() rotates queries over all slots so that the tail of one query
// (long streams, finalize, rescoring, top-k) overlaps the bulk kernel of the next.
struct swgpu_slot {
  bool ready = false;                  // buffers allocated
  int32_t*  d_scores = nullptr;        // file order
  unsigned* d_raw_s = nullptr;         // n_pairs * 32 packed running maxima
  int*      d_rescore_list = nullptr;  // up to n
  int*      d_row_bound = nullptr;     // [n] by sorted position: last query row that can hold the target's score (k1s_finalize)
  unsigned* d_counters = nullptr;      // [0] k1s queue, [1] k1s long queue, [2] rescore count, [3] k3 queue, [4] k3x queue,
                                       // [5] k3 all-targets queue, [6] k3x error flag, [7] top-k error flag,
                                       // [8] top-k coordinate queue phase 0, [9] phase 1, [12..13] residues of the rescored targets (u64)
  uint2*    d_bnd_s = nullptr;         // k1s multi-stripe carry rows
  // k1s X items (long streams, one stripe per item): ping-pong rows, progress words, per-stripe maxima
  uint2*    d_xs_bnd = nullptr;  int64_t xs_bnd_cap = 0;       // two rows per X stream (2 * its columns)
  unsigned* d_xs_progress = nullptr;  unsigned* d_xs_raw = nullptr;  int64_t xs_prog_cap = 0, xs_raw_cap = 0;
  int2*     d_bnd3 = nullptr;          // k3_intra carry rows
  int2*     d_bnd3s = nullptr;  int64_t bnd3s_cap = 0;   // k3_three (positive gap scores) carry rows: full target length, allocated on first use
  // k3x (32-bit wavefront over many warps): target slots
  int4*     d_x_rows = nullptr;        // [K3X_SLOTS][2][x_row_stride] (H, tag, V, tag)
  int4*     d_x_partial = nullptr;  int x_stripes_cap = 0;   // [K3X_SLOTS][x_stripes_cap]
  unsigned  x_epoch = 0;               // k3x launches on this slot so far (tags are unique per launch)
  unsigned* d_x_done = nullptr;        // [K3X_SLOTS]
  int8_t*   d_query = nullptr;  int64_t query_cap = 0;
  int8_t*   d_table = nullptr;
  unsigned* d_sprof = nullptr;  size_t sprof_cap = 0;     // k1s profile (bulk plan)
  unsigned* d_sprof2 = nullptr; size_t sprof2_cap = 0;    // k1s profile (long-stream plan)
  cudaStream_t s_main = nullptr;       // bulk K1S (low priority)
  cudaStream_t s_aux = nullptr;        // long-stream K1S, then rescoring / top-k / read-back (high priority)
  cudaEvent_t ev[12] = {};             // 0 start, 1 16-bit done, 2 scores final, 3 fork, 4 long done, 5 bulk done,
                                       // 6/7 h2d, 8/9 top-k, 10 d2h start, 11 everything done (slot reusable)
  // top-k scratch
  unsigned long long* d_keys = nullptr;  int* d_hit_sidx = nullptr;  void* d_sel = nullptr;  int topk_cap = 0;
  int4* d_coords = nullptr;
  // pinned host staging for asynchronous read-back
  unsigned* h_ctr = nullptr;                                // 16 counters
  unsigned long long* h_keys = nullptr;  int4* h_coords = nullptr;  int* h_best = nullptr;  int* h_sidx = nullptr;  int h_cap = 0;
  int32_t* h_scores = nullptr;  int64_t h_scores_cap = 0;   // staging when the caller's score buffer is pageable
  // host-side state of the search in flight / last finished
  bool in_flight = false, have_scores = false, use16 = false, split_long = false;
  bool three_state = false;                    // last search ran the reference's three-state recurrence (a gap score was positive)
  int64_t n_split_pairs = 0;                   // pairs scored by the separate long-stream launch of this search
  int64_t lq = 0;  int oe = 0, ge = 0;
  int topk_k = 0, topk_launches = 0;
  int32_t* user_scores = nullptr;  bool scores_staged = false;
  swgpu_stats_t st = {};
};

struct swgpu_db {
  int device = 0;
  int sm_count = 0;
  int64_t n = 0, n_residues = 0, entry_base = 0;
  int32_t max_len = 0;
  int32_t long_threshold = 0;          // > 0: targets longer than this always take the 32-bit kernels
  // database, resident in HBM
  int8_t*   d_plain = nullptr;         // residues, sorted order, back to back
  int64_t*  d_plain_off = nullptr;     // n + 1
  int64_t*  d_order = nullptr;         // sorted position -> file index
  uint16_t* d_scols = nullptr;         // streams of paired targets (db_pack.h)
  int64_t*  d_stream_off = nullptr;
  int32_t*  d_stream_ncols = nullptr;
  int64_t*  d_stream_pair0 = nullptr;
  int64_t   n_pairs = 0, n_streams = 0;
  int32_t   max_stream_cols = 0;
  int64_t   n_long_streams = 0, n_long_pairs = 0;   // leading long streams get their own low-latency launch
  int64_t   n_vlong_streams = 0, n_vlong_pairs = 0; // the few streams long enough to outlast the whole bulk launch (swgpu.cu)
  int64_t   long_cols = 0, scols_len = 0;           // stream columns of the long streams / of all streams
  long long bnd_s_stride = 0;  int k1s_grid = 0;    // k1s multi-stripe carry rows: per warp slot
  long long bnd3_stride = 0;   int k3_grid = 0;     // k3_intra carry rows
  long long x_row_stride = 0;                       // k3x rows: longest target, rounded up
  int64_t   device_bytes = 0;
  std::vector<int64_t> h_len_sorted_prefix;         // plain_off on the host (statistics, traceback)
  std::vector<int32_t> h_inv_order;                 // file index -> sorted position (traceback)
  swgpu_slot slots[SWGPU_SLOTS];
  cudaEvent_t ev_batch[1] = {};                     // start of the current swgpu_search_many() batch
  swgpu_stats_t last_stats = {};
};

namespace swi {
int slot_prepare(swgpu_db* db, swgpu_slot& S);                        // allocate a slot's buffers on first use
void slot_release(swgpu_slot& S);
int k3x_prepare(swgpu_db* db, swgpu_slot& S, int lq, cudaStream_t stream);   // (re)allocate + zero the k3x slot state
int enqueue_topk(swgpu_db* db, swgpu_slot& S, int k);                  // no host synchronisation
int collect_topk(swgpu_db* db, swgpu_slot& S, swgpu_hit_t* hits, int* n_hits);   // after S.ev[11] has completed
int fill_batch16(int device, const int8_t* query, int64_t query_len, const int8_t* batch, int64_t padded_len, int lanes,
                 const int8_t* table, int gap_open, int gap_extend, int32_t* scores16);   // swgpu.cu: per-thread cached context
// swgpu_synth.cu: generate the residues of a layout-only database into d_plain, then pair them into d_scols
cudaError_t synth_fill(swgpu_db* db, const swgpu_synth_t& gen, const int64_t* d_entry_ids, int64_t entry_base,
                       const int64_t* d_pair_col);
}  // namespace swi
