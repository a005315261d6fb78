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

#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return swi::fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? SWGPU_ENODEV : SWGPU_ECUDA, \
              "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

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
  // per-search scratch
  int32_t*  d_scores = nullptr;        // file order
  unsigned* d_raw_s = nullptr;         // n_pairs * 32 packed running maxima
  int*      d_rescore_list = nullptr;  // up to n
  unsigned* d_counters = nullptr;      // [0] k1s queue, [1] k1s long queue, [2] rescore count, [3] k3 queue, [4] k3w queue,
                                       // [5] k3 all-targets queue, [6] k3w error flag, [7] top-k error flag
  uint2*    d_bnd_s = nullptr;  long long bnd_s_stride = 0;  int k1s_grid = 0;   // k1s multi-stripe carry rows
  int2*     d_bnd3 = nullptr;   long long bnd3_stride = 0;   int k3_grid = 0;    // k3 / k3w carry rows
  int8_t*   d_query = nullptr;  int64_t query_cap = 0;
  int8_t*   d_table = nullptr;
  unsigned* d_sprof = nullptr;  size_t sprof_cap = 0;     // k1s profile (bulk plan)
  unsigned* d_sprof2 = nullptr; size_t sprof2_cap = 0;    // k1s profile (long-stream plan)
  int64_t   device_bytes = 0;
  std::vector<int64_t> h_len_sorted_prefix;               // plain_off on the host (statistics, top-k buffers)
  cudaStream_t s_main = nullptr, s_aux = nullptr;
  cudaEvent_t ev[8] = {};
  int64_t last_lq = 0;
  bool have_scores = false;
  int last_oe = 0, last_ge = 0;        // gaps of the last search (top-k coordinates are recomputed against it)
  // top-k scratch
  unsigned long long* d_keys = nullptr;  int* d_hit_sidx = nullptr;  void* d_sel = nullptr;  int topk_cap = 0;
  int4* d_coords = nullptr;
  int2* d_bnd_loc = nullptr;  long long bnd_loc_stride = 0;  int loc_slots = 0;
  swgpu_stats_t last_stats = {};
};
