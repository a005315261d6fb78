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
// Coordinates: the wavefront kernel k3x<TRACK> (sw_k3x.cuh) runs the 32-bit recurrence twice per hit,
// tracking the first maximum; (hit, stripe) items are pulled from a device-side queue, so nothing returns to the host
// between selection and coordinates (swgpu_search_many keeps several queries in flight).
#include "swgpu_internal.h"
#include "sw_kernels.cuh"
#include "sw_k3x.cuh"

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

// Coordinates: the wavefront kernel k3x<TRACK> (sw_k3x.cuh), one warp per 256-row stripe of a hit; phase 0 =
// forward (end cell), phase 1 = the reversed prefixes (start cell).

}  // namespace swk

using swi::fail;

namespace swk {
__global__ void topk_init(SelState* st, unsigned k) {
  if (threadIdx.x == 0 && blockIdx.x == 0) { st->prefix = 0ull; st->k_rem = k; st->n_out = 0u; }
}
}  // namespace swk

namespace swi {

// Queue selection + coordinate passes + read-back of the k hits on the slot's high-priority stream, behind the
// search queued on it.  No host synchronisation (buffers grow on first use only); 1 <= k <= min(65536, n).
int enqueue_topk(swgpu_db* db, swgpu_slot& S, int k) {
  if (S.three_state)
    return fail(SWGPU_EINVAL, "top-k with coordinates is not available with positive gap scores (the coordinate passes use the "
                "two-gap-state form); the per-entry scores are");
  cudaStream_t st = S.s_aux;
  if (k > S.topk_cap) {
    cudaFree(S.d_keys); cudaFree(S.d_hit_sidx); cudaFree(S.d_coords);
    S.d_keys = nullptr; S.d_hit_sidx = nullptr; S.d_coords = nullptr; S.topk_cap = 0;
    const int cap = std::max(k, 128);
    CU(cudaMalloc((void**)&S.d_keys, sizeof(unsigned long long) * 2 * (size_t)cap));
    CU(cudaMalloc((void**)&S.d_hit_sidx, sizeof(int) * 5 * (size_t)cap));       // unsorted, sorted, best_out, stop columns of the two coordinate passes
    CU(cudaMalloc((void**)&S.d_coords, sizeof(int4) * (size_t)cap));
    S.topk_cap = cap;
  }
  if (k > S.h_cap) {
    cudaFreeHost(S.h_keys); cudaFreeHost(S.h_coords); cudaFreeHost(S.h_best); cudaFreeHost(S.h_sidx);
    S.h_keys = nullptr; S.h_coords = nullptr; S.h_best = nullptr; S.h_sidx = nullptr; S.h_cap = 0;
    const int cap = std::max(k, 128);
    CU(cudaMallocHost((void**)&S.h_keys, sizeof(unsigned long long) * (size_t)cap));
    CU(cudaMallocHost((void**)&S.h_coords, sizeof(int4) * (size_t)cap));
    CU(cudaMallocHost((void**)&S.h_best, sizeof(int) * (size_t)cap));
    CU(cudaMallocHost((void**)&S.h_sidx, sizeof(int) * (size_t)cap));
    S.h_cap = cap;
  }
  if (!S.d_sel) CU(cudaMalloc(&S.d_sel, sizeof(swk::SelState)));
  swk::SelState* sel = (swk::SelState*)S.d_sel;
  const int cap = S.topk_cap;
  unsigned long long* keys_u = S.d_keys; unsigned long long* keys_s = S.d_keys + cap;
  int* sidx_u = S.d_hit_sidx; int* sidx_s = S.d_hit_sidx + cap; int* best_out = S.d_hit_sidx + 2 * cap;

  int launches = 0;
  CU(cudaEventRecord(S.ev[8], st));
  CU(cudaMemsetAsync(sel, 0, sizeof(swk::SelState), st));
  CU(cudaMemsetAsync(S.d_counters + 7, 0, 3 * sizeof(unsigned), st));
  swk::topk_init<<<1, 32, 0, st>>>(sel, (unsigned)k);
  const int grid = (int)std::min<int64_t>((db->n + 255) / 256, (int64_t)db->sm_count * 8);
  for (int pass = 0; pass < 8; ++pass) {
    swk::topk_hist<<<grid, 256, 0, st>>>(S.d_scores, db->d_order, (long)db->n, pass, sel);
    swk::topk_pick<<<1, 32, 0, st>>>(pass, sel);
    launches += 2;
  }
  swk::topk_compact<<<grid, 256, 0, st>>>(S.d_scores, db->d_order, (long)db->n, sel, keys_u, sidx_u, k);
  swk::topk_ranksort<<<(k + 255) / 256, 256, 0, st>>>(keys_u, sidx_u, k, keys_s, sidx_s);
  launches += 3;
  CU(cudaGetLastError());
  if (S.lq > 0) {
    // coordinates of the k hits against the query of this slot's search (hits with score 0 come out as -1)
    swk::K3XParams P; memset(&P, 0, sizeof(P));
    P.sidx_list = sidx_s; P.n_targets = k; P.plain = db->d_plain; P.plain_off = db->d_plain_off; P.query = S.d_query;
    P.lq = (int)S.lq; P.nstripes_max = ((int)S.lq + swk::K3X_ROWS - 1) / swk::K3X_ROWS; P.table = S.d_table; P.oe = S.oe; P.ge = S.ge;
    P.rows = S.d_x_rows; P.row_stride = db->x_row_stride; P.coords = S.d_coords; P.best_out = best_out;
    P.error_flag = S.d_counters + 7;
    P.row_bound = (S.use16 && !getenv("SWGPU_TOPK_NO_ROW_BOUND")) ? S.d_row_bound : nullptr;      // A/B knob
    P.hit_keys = getenv("SWGPU_TOPK_NO_EARLY_STOP") ? nullptr : keys_s;     // A/B knob: the coordinate passes scan their whole rectangles
    CU(cudaMemsetAsync(S.d_hit_sidx + 3 * cap, 0x7f, sizeof(int) * 2 * (size_t)cap, st));        // stop columns: 0x7f7f7f7f = none yet
    const int grid = db->sm_count * swk::k3x_ctas_per_sm<true>(db->device);    // the whole machine: CTAs beyond the levels the items need leave at once (sw_k3x.cuh)
    for (int phase = 0; phase < 2; ++phase) {
      int r = swi::k3x_prepare(db, S, (int)S.lq, st);
      if (r) return r;
      P.epoch = S.x_epoch; P.done = S.d_x_done; P.partial = S.d_x_partial; P.rows = S.d_x_rows;
      P.sm_slots = S.d_x_done + swk::K3X_SLOTS; P.level_ctas = db->sm_count;
      P.phase = phase; P.counter = S.d_counters + 8 + phase; P.stop_col = S.d_hit_sidx + (3 + phase) * cap;
      int g2 = grid;                                       // the reversed pass sizes its levels from the hits' scores (sw_k3x.cuh)
      if (phase == 1) { if (const char* ev = getenv("SWGPU_K3X_REV_CTAS")) g2 = std::min(grid, db->sm_count * std::max(1, std::min(4, atoi(ev)))); }   // A/B knob
      swk::k3x<true><<<g2, swk::K3X_WARPS * 32, 0, st>>>(P);
    }
    launches += 2;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(S.h_coords, S.d_coords, sizeof(int4) * k, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(S.h_best, best_out, sizeof(int) * k, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaMemcpyAsync(S.h_keys, keys_s, sizeof(unsigned long long) * k, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(S.h_sidx, sidx_s, sizeof(int) * k, cudaMemcpyDeviceToHost, st));
  CU(cudaEventRecord(S.ev[9], st));
  S.topk_k = k; S.topk_launches = launches;
  return SWGPU_OK;
}

// Decode the hits of a finished slot (its completion event has been waited for; S.h_ctr holds the counters).
int collect_topk(swgpu_db* db, swgpu_slot& S, swgpu_hit_t* hits, int* n_hits) {
  const int k = S.topk_k;
  if (n_hits) *n_hits = 0;
  if (S.h_ctr[7]) return fail(SWGPU_ECUDA, "swgpu_topk: coordinate pipeline stalled (internal error)");
  for (int i = 0; i < k; ++i) {
    swgpu_hit_t& h = hits[i];
    const unsigned long long key = S.h_keys[i];
    h.score = (int32_t)(key >> 32);
    h.entry = db->entry_base + (int64_t)(0xffffffffu - (unsigned)(key & 0xffffffffu));
    if (S.lq > 0 && h.score > 0) {
      const int4 c = S.h_coords[i];
      h.q_start = c.x; h.q_end = c.y; h.t_start = c.z; h.t_end = c.w;
      if (S.h_best[i] != h.score)
        return fail(SWGPU_ECUDA, "swgpu_topk: coordinate pass found score %d for a hit scored %d (internal error)",
                    S.h_best[i], h.score);
    } else {
      h.q_start = h.q_end = h.t_start = h.t_end = -1;         // hits with score 0 have no coordinates
    }
  }
  if (n_hits) *n_hits = k;
  return SWGPU_OK;
}

}  // namespace swi

extern "C" int swgpu_topk(swgpu_db_t* db, int k, swgpu_hit_t* hits, int* n_hits) {
  if (n_hits) *n_hits = 0;
  if (!db || k < 0 || (k > 0 && !hits)) return fail(SWGPU_EINVAL, "swgpu_topk: bad argument");
  swgpu_slot& S = db->slots[0];
  if (!S.have_scores || S.in_flight) return fail(SWGPU_EINVAL, "swgpu_topk: no search result resident");
  if (k > 65536) return fail(SWGPU_EINVAL, "swgpu_topk: k > 65536 not supported");
  k = (int)std::min<int64_t>(k, db->n);
  db->last_stats.ms_topk = 0; db->last_stats.topk_launches = 0;
  if (k == 0) return SWGPU_OK;
  CU(cudaSetDevice(db->device));
  int r = swi::enqueue_topk(db, S, k);
  if (r) { cudaDeviceSynchronize(); return r; }
  CU(cudaMemcpyAsync(S.h_ctr, S.d_counters, 16 * sizeof(unsigned), cudaMemcpyDeviceToHost, S.s_aux));
  CU(cudaStreamSynchronize(S.s_aux));
  float ms = 0; CU(cudaEventElapsedTime(&ms, S.ev[8], S.ev[9]));
  S.st.ms_topk = ms; S.st.topk_launches = S.topk_launches;
  db->last_stats.ms_topk = ms; db->last_stats.topk_launches = S.topk_launches;
  return swi::collect_topk(db, S, hits, n_hits);
}

extern "C" int swgpu_db_last_stats(const swgpu_db_t* db, swgpu_stats_t* out) {
  if (!db || !out) return fail(SWGPU_EINVAL, "swgpu_db_last_stats: bad argument");
  *out = db->last_stats;
  return SWGPU_OK;
}
