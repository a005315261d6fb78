// swgpu_topk.cu -- device-side top-hit selection and linear-space coordinate recovery.
//
// The reference has no counterpart: it keeps one maximum per target (src/alignment.c:133,186 of
// the reference tree) and prints every entry (src/tools/sw_cmdline.c:55-72).  The rules are the
// ones SURVEY.md A.7 defines:
//   order   score descending, then file-order entry index ascending;
//   end     first cell attaining the maximum in the reference's scan order (target position j
//           outer loop, query position i inner loop, alignment.c:88,103) = smallest j, then smallest i;
//   start   the same rule on the reversed prefixes q[i_end..0], t[j_end..0].
//
// Selection: an 8-pass MSB-first radix select (8 bits per pass) over 64-bit keys
// (score << 32 | ~entry) finds the k-th largest key exactly; a compaction pass collects the k keys
// >= it together with their position in the length-sorted database; a rank sort orders them.
// Coordinates: one warp per hit runs the 32-bit systolic recurrence of K3 (sw_kernels.cuh) twice,
// tracking the first maximum.
#include "swgpu_internal.h"
#include "sw_kernels.cuh"
#include "sw_k3w.cuh"

#include <algorithm>
#include <cstring>

namespace swk {

struct SelState {
  unsigned long long prefix;     // decided high bits of the k-th largest key
  unsigned           k_rem;      // how many keys still to take among those matching the prefix
  unsigned           n_out;      // compaction cursor
  unsigned           hist[8][256];
};

__device__ __forceinline__ unsigned long long make_key(const int32_t* scores_file, const int64_t* order, long s) {
  const long e = (long)order[s];
  return ((unsigned long long)(unsigned)scores_file[e] << 32) | (unsigned long long)(0xffffffffu - (unsigned)e);
}

// pass p looks at byte (7 - p).  Keys whose higher bytes equal the decided prefix are histogrammed.
__global__ void topk_hist(const int32_t* __restrict__ scores_file, const int64_t* __restrict__ order, long n,
                          int pass, SelState* st) {
  __shared__ unsigned h[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const int shift = 56 - 8 * pass;
  const unsigned long long prefix = st->prefix;
  const unsigned long long himask = pass == 0 ? 0ull : (~0ull << (shift + 8));
  for (long s = blockIdx.x * (long)blockDim.x + threadIdx.x; s < n; s += (long)gridDim.x * blockDim.x) {
    const unsigned long long key = make_key(scores_file, order, s);
    if ((key & himask) == (prefix & himask)) atomicAdd(&h[(unsigned)(key >> shift) & 255u], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x) if (h[i]) atomicAdd(&st->hist[pass][i], h[i]);
}

// one thread: walk the histogram from the top, fix the next byte of the k-th largest key
__global__ void topk_pick(int pass, SelState* st) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  unsigned rem = st->k_rem;
  int d = 255;
  for (; d > 0; --d) {
    const unsigned c = st->hist[pass][d];
    if (c >= rem) break;
    rem -= c;
  }
  st->k_rem = rem;
  st->prefix |= (unsigned long long)(unsigned)d << (56 - 8 * pass);
}

__global__ void topk_compact(const int32_t* __restrict__ scores_file, const int64_t* __restrict__ order, long n,
                             SelState* st, unsigned long long* __restrict__ keys, int* __restrict__ sidx, int k) {
  const unsigned long long kth = st->prefix;
  for (long s = blockIdx.x * (long)blockDim.x + threadIdx.x; s < n; s += (long)gridDim.x * blockDim.x) {
    const unsigned long long key = make_key(scores_file, order, s);
    if (key >= kth) {
      const unsigned slot = atomicAdd(&st->n_out, 1u);
      if (slot < (unsigned)k) { keys[slot] = key; sidx[slot] = (int)s; }
    }
  }
}

// rank sort of k distinct keys, descending: out position = number of larger keys
__global__ void topk_ranksort(const unsigned long long* __restrict__ keys, const int* __restrict__ sidx, int k,
                              unsigned long long* __restrict__ keys_out, int* __restrict__ sidx_out) {
  __shared__ unsigned long long tile[256];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long mine = i < k ? keys[i] : 0ull;
  int rank = 0;
  for (int base = 0; base < k; base += 256) {
    __syncthreads();
    tile[threadIdx.x] = (base + threadIdx.x < k) ? keys[base + threadIdx.x] : 0ull;
    __syncthreads();
    const int lim = min(256, k - base);
    for (int j = 0; j < lim; ++j) rank += (tile[j] > mine);
  }
  if (i < k) { keys_out[rank] = mine; sidx_out[rank] = sidx[i]; }
}

// Coordinates: the wavefront kernel k3w<TRACK> (sw_k3w.cuh), one CTA per hit; phase 0 = forward (end cell),
// phase 1 = the reversed prefixes (start cell).

}  // namespace swk

using swi::fail;

extern "C" int swgpu_topk(swgpu_db_t* db, int k, swgpu_hit_t* hits, int* n_hits) {
  if (n_hits) *n_hits = 0;
  if (!db || k < 0 || (k > 0 && !hits)) return fail(SWGPU_EINVAL, "swgpu_topk: bad argument");
  if (!db->have_scores) return fail(SWGPU_EINVAL, "swgpu_topk: no search result resident");
  if (k > 65536) return fail(SWGPU_EINVAL, "swgpu_topk: k > 65536 not supported");
  k = (int)std::min<int64_t>(k, db->n);
  db->last_stats.ms_topk = 0; db->last_stats.topk_launches = 0;
  if (k == 0) return SWGPU_OK;
  CU(cudaSetDevice(db->device));
  cudaStream_t st = db->s_main;
  if (k > db->topk_cap) {
    cudaFree(db->d_keys); cudaFree(db->d_hit_sidx); cudaFree(db->d_coords); cudaFree(db->d_bnd_loc);
    db->d_keys = nullptr; db->d_hit_sidx = nullptr; db->d_coords = nullptr; db->d_bnd_loc = nullptr; db->topk_cap = 0;
    const int cap = std::max(k, 128);
    CU(cudaMalloc((void**)&db->d_keys, sizeof(unsigned long long) * 2 * (size_t)cap));
    CU(cudaMalloc((void**)&db->d_hit_sidx, sizeof(int) * 3 * (size_t)cap));       // unsorted, sorted, best_out
    CU(cudaMalloc((void**)&db->d_coords, sizeof(int4) * (size_t)cap));
    db->topk_cap = cap;
  }
  if (!db->d_sel) CU(cudaMalloc(&db->d_sel, sizeof(swk::SelState)));
  swk::SelState* sel = (swk::SelState*)db->d_sel;
  const int cap = db->topk_cap;
  unsigned long long* keys_u = db->d_keys; unsigned long long* keys_s = db->d_keys + cap;
  int* sidx_u = db->d_hit_sidx; int* sidx_s = db->d_hit_sidx + cap; int* best_out = db->d_hit_sidx + 2 * cap;

  int launches = 0;
  CU(cudaEventRecord(db->ev[4], st));
  swk::SelState init; memset(&init, 0, sizeof(init)); init.k_rem = (unsigned)k;
  CU(cudaMemcpyAsync(sel, &init, sizeof(init), cudaMemcpyHostToDevice, st));   // pageable source: copied before return
  const int grid = (int)std::min<int64_t>((db->n + 255) / 256, (int64_t)db->sm_count * 8);
  for (int pass = 0; pass < 8; ++pass) {
    swk::topk_hist<<<grid, 256, 0, st>>>(db->d_scores, db->d_order, (long)db->n, pass, sel);
    swk::topk_pick<<<1, 32, 0, st>>>(pass, sel);
    launches += 2;
  }
  swk::topk_compact<<<grid, 256, 0, st>>>(db->d_scores, db->d_order, (long)db->n, sel, keys_u, sidx_u, k);
  swk::topk_ranksort<<<(k + 255) / 256, 256, 0, st>>>(keys_u, sidx_u, k, keys_s, sidx_s);
  launches += 2;
  CU(cudaGetLastError());

  // coordinates of the k hits against the query of the last search
  std::vector<unsigned long long> hk(k);
  std::vector<int> hs(k);
  CU(cudaMemcpyAsync(hk.data(), keys_s, sizeof(unsigned long long) * k, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(hs.data(), sidx_s, sizeof(int) * k, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  std::vector<int4> hc(k, make_int4(-1, -1, -1, -1));
  std::vector<int> hb(k, 0);
  unsigned loc_err = 0;
  const int lq = (int)db->last_lq;
  int n_pos = 0;
  while (n_pos < k && (hk[n_pos] >> 32) != 0) ++n_pos;          // hits with score 0 have no coordinates
  if (lq > 0 && n_pos > 0) {
    // linear-space carry buffer: one row of (H, V) per hit, as long as the longest hit target
    long long need = 32;
    for (int i = 0; i < n_pos; ++i)
      need = std::max<long long>(need, db->h_len_sorted_prefix[hs[i] + 1] - db->h_len_sorted_prefix[hs[i]]);
    if (need > db->bnd_loc_stride || n_pos > db->loc_slots) {
      cudaFree(db->d_bnd_loc); db->d_bnd_loc = nullptr;
      db->bnd_loc_stride = std::max(need, db->bnd_loc_stride); db->loc_slots = std::max(n_pos, db->loc_slots);
      CU(cudaMalloc((void**)&db->d_bnd_loc, sizeof(int2) * (size_t)db->bnd_loc_stride * (size_t)db->loc_slots));
    }
    swk::K3WParams P; memset(&P, 0, sizeof(P));
    P.sidx_list = sidx_s; P.n_work = n_pos; P.plain = db->d_plain; P.plain_off = db->d_plain_off; P.query = db->d_query;
    P.lq = lq; P.table = db->d_table; P.oe = db->last_oe; P.ge = db->last_ge; P.bnd = db->d_bnd_loc;
    P.bnd_stride = db->bnd_loc_stride; P.coords = db->d_coords; P.best_out = best_out;
    P.error_flag = db->d_counters + 7;
    CU(cudaMemsetAsync(db->d_counters + 7, 0, sizeof(unsigned), st));
    P.phase = 0; swk::k3w<true><<<n_pos, swk::K3W_WARPS * 32, 0, st>>>(P);
    P.phase = 1; swk::k3w<true><<<n_pos, swk::K3W_WARPS * 32, 0, st>>>(P);
    launches += 2;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(&loc_err, db->d_counters + 7, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hc.data(), db->d_coords, sizeof(int4) * n_pos, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hb.data(), best_out, sizeof(int) * n_pos, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaEventRecord(db->ev[5], st));
  CU(cudaStreamSynchronize(st));
  float ms = 0; CU(cudaEventElapsedTime(&ms, db->ev[4], db->ev[5]));
  db->last_stats.ms_topk = ms; db->last_stats.topk_launches = launches;
  if (loc_err) return fail(SWGPU_ECUDA, "swgpu_topk: coordinate pipeline stalled (internal error)");
  for (int i = 0; i < k; ++i) {
    swgpu_hit_t& h = hits[i];
    h.score = (int32_t)(hk[i] >> 32);
    h.entry = db->entry_base + (int64_t)(0xffffffffu - (unsigned)(hk[i] & 0xffffffffu));
    h.q_start = hc[i].x; h.q_end = hc[i].y; h.t_start = hc[i].z; h.t_end = hc[i].w;
    if (i < n_pos && lq > 0 && hb[i] != h.score)
      return fail(SWGPU_ECUDA, "swgpu_topk: coordinate pass found score %d for a hit scored %d (internal error)", hb[i], h.score);
  }
  if (n_hits) *n_hits = k;
  return SWGPU_OK;
}

extern "C" int swgpu_db_last_stats(const swgpu_db_t* db, swgpu_stats_t* out) {
  if (!db || !out) return fail(SWGPU_EINVAL, "swgpu_db_last_stats: bad argument");
  *out = db->last_stats;
  return SWGPU_OK;
}
