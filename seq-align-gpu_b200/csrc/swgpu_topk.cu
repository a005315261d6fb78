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

// ------------------------------------------------------------------------------------------
// Coordinates.  Same systolic 32-bit recurrence as k3_intra (one warp per hit, lane l owns R3 rows
// of the current 32*R3-row stripe, the target streams through the lanes, the stripe's bottom row is
// carried in a linear-space buffer), plus first-maximum tracking.  phase 0 = forward (finds the
// end cell), phase 1 = the reversed prefixes (finds the start cell).
// ------------------------------------------------------------------------------------------
struct LocParams {
  const int*     sidx;         // [k] sorted positions of the hits
  int            k;
  const int8_t*  plain;
  const int64_t* plain_off;
  const int8_t*  query;
  int            lq;
  const int8_t*  table;
  int            oe, ge;
  int2*          bnd;
  long long      bnd_stride;
  int4*          coords;       // [k] (q_start, q_end, t_start, t_end); phase 0 writes y,w; phase 1 writes x,z
  int*           best_out;     // [k] score found by phase 0 (cross-check against the search result)
  int            phase;
  unsigned       one;
};

constexpr int LOC_WARPS = 4;
constexpr int LOC_R = 8;

__device__ __forceinline__ bool loc_better(int s, int j, int i, int bs, int bj, int bi) {
  return s > bs || (s == bs && (j < bj || (j == bj && i < bi)));
}

__global__ void __launch_bounds__(LOC_WARPS * 32) k_locate(const LocParams P) {
  constexpr int R3 = LOC_R, W = R3 / 4;
  __shared__ unsigned pl_all[LOC_WARPS][32 * W * 32];
  __shared__ int8_t tab[32 * 32];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hit = blockIdx.x * LOC_WARPS + warp;
  if (hit >= P.k) return;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  int2* bnd = P.bnd + (size_t)hit * (size_t)P.bnd_stride;
  const int sidx = P.sidx[hit];
  const int8_t* tp = P.plain + P.plain_off[sidx];
  int lt = (int)(P.plain_off[sidx + 1] - P.plain_off[sidx]);
  int lq = P.lq;
  const int8_t* qp = P.query;
  int qstep = 1, tstep = 1;
  if (P.phase == 1) {
    const int4 c = P.coords[hit];
    if (c.y < 0) { if (lane == 0) P.coords[hit] = make_int4(-1, -1, -1, -1); return; }
    lq = c.y + 1; lt = c.w + 1; qp += c.y; tp += c.w; qstep = -1; tstep = -1;
  }
  const int oe = P.oe, ge = P.ge;
  const int npass = (lq + 32 * R3 - 1) / (32 * R3);
  int best = 0, bj = 0x7fffffff, bi = 0x7fffffff;

  for (int pass = 0; pass < npass; ++pass) {
    const int row0 = pass * 32 * R3 + lane * R3;
    __syncwarp();
    for (int c = 0; c < 32; ++c) {
#pragma unroll
      for (int ww = 0; ww < W; ++ww) {
        unsigned word = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int i = row0 + ww * 4 + b;
          const int s = (c == 0 || i >= lq) ? -128 : (int)tab[(qp[(long)i * qstep] & 31) * 32 + c];
          word |= (unsigned)(s & 0xff) << (8 * b);
        }
        pl[(c * W + ww) * 32 + lane] = word;
      }
    }
    __syncwarp();
    int H[R3], G[R3];
#pragma unroll
    for (int k = 0; k < R3; ++k) { H[k] = 0; G[k] = 0; }
    int diag_in = 0, out_h = 0, out_v = 0, my_code = 0;
    int pre_code = 0, pre_h = 0, pre_v = 0;
    const bool first = (pass == 0), last = (pass == npass - 1);
    const int nsteps = lt + 31;
    for (int s = 0; s < nsteps; ++s) {
      if ((s & 31) == 0) {
        const int j = s + lane;
        pre_code = (j < lt) ? (int)tp[(long)j * tstep] : 0;
        int2 b = make_int2(0, 0);
        if (!first && j < lt) b = bnd[j];
        pre_h = b.x; pre_v = b.y;
      }
      int code = __shfl_up_sync(0xffffffffu, my_code, 1);
      int h_up = __shfl_up_sync(0xffffffffu, out_h, 1);
      int v_in = __shfl_up_sync(0xffffffffu, out_v, 1);
      const int c0 = __shfl_sync(0xffffffffu, pre_code, s & 31);
      const int h0 = __shfl_sync(0xffffffffu, pre_h, s & 31);
      const int v0 = __shfl_sync(0xffffffffu, pre_v, s & 31);
      if (lane == 0) { code = c0; h_up = h0; v_in = v0; }
      my_code = code;
      const int j = s - lane;
      if (j >= 0 && j < lt) {
        int d = diag_in;
        diag_in = h_up;
        int V = v_in;
        unsigned wreg[W];
#pragma unroll
        for (int ww = 0; ww < W; ++ww) wreg[ww] = lds32(pl_addr + ((code & 31) * W + ww) * 128);
        int colbest = 0, colk = 0;
#pragma unroll
        for (int k = 0; k < R3; ++k) {
          const int t = (int)add_byte(wreg[k >> 2], 1u << (8 * (k & 3)), (unsigned)d);
          const int h = __vimax3_s32_relu(t, G[k], V);
          d = H[k];
          H[k] = h;
          const int u = h + oe;
          G[k] = __viaddmax_s32(G[k], ge, u);
          V = __viaddmax_s32(V, ge, u);
          if (h > colbest) { colbest = h; colk = k; }          // strict: first (smallest i) maximum of the column
        }
        if (loc_better(colbest, j, row0 + colk, best, bj, bi)) { best = colbest; bj = j; bi = row0 + colk; }
        out_h = H[R3 - 1];
        out_v = V;
        if (lane == 31 && !last) bnd[j] = make_int2(out_h, out_v);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int s2 = __shfl_xor_sync(0xffffffffu, best, o);
    const int j2 = __shfl_xor_sync(0xffffffffu, bj, o);
    const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
    if (loc_better(s2, j2, i2, best, bj, bi)) { best = s2; bj = j2; bi = i2; }
  }
  if (lane == 0) {
    int4 c = P.coords[hit];
    if (P.phase == 0) {
      if (best > 0) { c.y = bi; c.w = bj; } else { c = make_int4(-1, -1, -1, -1); }
      P.best_out[hit] = best;
    } else {
      c.x = c.y - bi; c.z = c.w - bj;
    }
    P.coords[hit] = c;
  }
}

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
    swk::LocParams P; memset(&P, 0, sizeof(P));
    P.sidx = sidx_s; P.k = n_pos; P.plain = db->d_plain; P.plain_off = db->d_plain_off; P.query = db->d_query; P.lq = lq;
    P.table = db->d_table; P.oe = db->last_oe; P.ge = db->last_ge; P.bnd = db->d_bnd_loc; P.bnd_stride = db->bnd_loc_stride;
    P.coords = db->d_coords; P.best_out = best_out; P.one = 1u;
    const int lgrid = (n_pos + swk::LOC_WARPS - 1) / swk::LOC_WARPS;
    P.phase = 0; swk::k_locate<<<lgrid, swk::LOC_WARPS * 32, 0, st>>>(P);
    P.phase = 1; swk::k_locate<<<lgrid, swk::LOC_WARPS * 32, 0, st>>>(P);
    launches += 2;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(hc.data(), db->d_coords, sizeof(int4) * n_pos, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hb.data(), best_out, sizeof(int) * n_pos, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaEventRecord(db->ev[5], st));
  CU(cudaStreamSynchronize(st));
  float ms = 0; CU(cudaEventElapsedTime(&ms, db->ev[4], db->ev[5]));
  db->last_stats.ms_topk = ms; db->last_stats.topk_launches = launches;
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
