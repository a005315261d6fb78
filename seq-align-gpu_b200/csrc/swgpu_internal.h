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
  int64_t n = 0, n_residues = 0, n_long = 0, n_chunks = 0, entry_base = 0;
  int32_t max_len = 0, max_chunk_cols = 0;
  // device arrays (database, resident)
  int8_t*   d_plain = nullptr;
  int64_t*  d_plain_off = nullptr;
  int64_t*  d_order = nullptr;
  uint2*    d_cols = nullptr;
  int64_t*  d_chunk_off = nullptr;
  int32_t*  d_chunk_cols = nullptr;
  // streams of paired targets (systolic kernel K1S)
  uint16_t* d_scols = nullptr;
  int64_t*  d_stream_off = nullptr;
  int32_t*  d_stream_ncols = nullptr;
  int64_t*  d_stream_pair0 = nullptr;
  int64_t   n_pairs = 0, n_streams = 0;
  int32_t   max_stream_cols = 0;
  unsigned* d_raw_s = nullptr;         // n_pairs * 32
  int64_t   n_long_streams = 0, n_long_pairs = 0;   // leading long streams get their own low-latency launch
  unsigned* d_sprof2 = nullptr;  size_t sprof2_cap = 0;
  uint2*    d_bnd_s = nullptr;  long long bnd_s_stride = 0;  int k1s_grid = 0;
  unsigned* d_sprof = nullptr;  size_t sprof_cap = 0;
  int       use_old_k1 = 0;
  // device scratch (per search)
  int32_t*  d_scores = nullptr;        // file order
  unsigned* d_raw = nullptr;           // n_chunks*32
  int*      d_rescore_list = nullptr;  // up to n
  unsigned* d_counters = nullptr;      // [0]=k1 queue, [1]=k3 long queue, [2]=rescore count, [3]=k3 rescore queue
  uint2*    d_bnd1 = nullptr;  long long bnd1_stride = 0;  int k1_grid = 0;
  int2*     d_bnd3 = nullptr;  long long bnd3_stride = 0;  int k3_grid = 0;
  int8_t*   d_query = nullptr; int64_t query_cap = 0;
  int8_t*   d_table = nullptr;
  unsigned* d_prof = nullptr;  size_t prof_cap = 0;
  int64_t   device_bytes = 0;
  // host copies needed after packing
  std::vector<int64_t> h_order;
  std::vector<int64_t> h_len_sorted_prefix;   // plain_off (sorted)
  cudaStream_t s_main = nullptr, s_aux = nullptr;
  cudaEvent_t ev[8] = {};

  int64_t last_lq = 0;
  bool have_scores = false;
  // gaps of the last search (top-k coordinates are recomputed against the last query / table)
  int last_oe = 0, last_ge = 0;
  // top-k scratch
  unsigned long long* d_keys = nullptr;  int* d_hit_sidx = nullptr;  void* d_sel = nullptr;  int topk_cap = 0;
  int4* d_coords = nullptr;
  int2* d_bnd_loc = nullptr;  long long bnd_loc_stride = 0;  int loc_slots = 0;
  swgpu_stats_t last_stats = {};
};
