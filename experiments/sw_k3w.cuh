// sw_k3w.cuh -- K3W: 32-bit intra-sequence WAVEFRONT kernel, one CTA (8 warps) per target.
//
// Same recurrence as k3_intra (sw_kernels.cuh; reference src/alignment.c:117-176), but built for
// LATENCY: a single (query, target) pair is a chain of Lt dependent column steps per stripe of
// query rows, and k3_intra / k_locate ran the stripes one after the other on one warp.  Here the
// 8 warps of a CTA work on 8 consecutive stripes of the SAME target at the same time, as a
// pipeline: warp w consumes the bottom row (H, V per column) that warp w-1 produced a few columns
// earlier.
//   * stripe = 32 lanes x K3W_R rows; warp w owns stripes w, w+8, w+16, ...
//   * warp w-1 -> warp w: a 128-column ring in shared memory per warp, with producer/consumer
//     column counters (volatile shared ints, __threadfence_block before publishing);
//   * warp 7 -> warp 0 (next group of 8 stripes): a full-length row in global memory (linear
//     space), same counters;
//   * inside a warp the scheme is k3_intra's: lane l owns K3W_R rows, the target streams through
//     the lanes by __shfl_up_sync.
// Used for (a) the coordinate passes of swgpu_topk (TRACK: first maximum in the reference's scan
// order) and (b) 32-bit rescoring of long targets.
#pragma once
#include "sw_kernels.cuh"

namespace swk {

constexpr int K3W_WARPS = 8;
constexpr int K3W_R = 4;                      // rows per lane: 128 rows per stripe, 1024 rows per group of stripes
constexpr int K3W_RING = 128;                 // columns of slack between neighbouring warps
constexpr unsigned K3W_SPIN_LIMIT = 1u << 26; // a stuck pipeline raises the error flag instead of hanging the GPU

struct K3WParams {
  const int*      sidx_list;    // work items (sorted positions); nullptr -> item index itself
  int             n_work;
  const unsigned* work_count;   // if non-null the item count is read from the device
  unsigned*       counter;      // non-null: CTAs pull items from a queue; null: CTA b handles item b
  const int8_t*   plain;
  const int64_t*  plain_off;
  const int8_t*   query;
  int             lq;
  const int8_t*   table;
  int             oe, ge;
  int2*           bnd;          // per CTA: bnd_stride columns (H, V) for the warp 7 -> warp 0 hand-over
  long long       bnd_stride;
  const int64_t*  order;        // !TRACK: scores_file[order[sidx]] = score
  int32_t*        scores_file;
  int4*           coords;       // TRACK: per item (q_start, q_end, t_start, t_end)
  int*            best_out;     // TRACK phase 0: score found
  int             phase;        // TRACK: 0 = forward (end cell), 1 = reversed prefixes (start cell)
  unsigned*       error_flag;
  int             min_len;      // queue mode: targets shorter than this are left to k3_intra
};

__device__ __forceinline__ bool k3w_better(int s, int j, int i, int bs, int bj, int bi) {
  return s > bs || (s == bs && (j < bj || (j == bj && i < bi)));
}

template <bool TRACK>
__global__ void __launch_bounds__(K3W_WARPS * 32) k3w(const K3WParams P) {
  constexpr int R3 = K3W_R, W = K3W_WARPS;
  __shared__ int8_t tab[32 * 32];
  __shared__ unsigned pl_all[W][32 * 32];                 // [code][lane], one int8x4 word (R3 == 4)
  __shared__ int2 ring[W][K3W_RING];
  __shared__ volatile int prod[W], cons[W];
  __shared__ int s_item;
  __shared__ int red[W][3];
  static_assert(R3 == 4, "profile layout assumes 4 rows per lane");
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  int2* bnd = P.bnd + (size_t)blockIdx.x * (size_t)P.bnd_stride;
  const int n_work = P.work_count ? (int)*P.work_count : P.n_work;
  const int oe = P.oe, ge = P.ge;
  const int pw = (warp + W - 1) % W;                      // the warp that produces my input row

  for (int round = 0;; ++round) {
    __syncthreads();
    if (threadIdx.x == 0) s_item = P.counter ? (int)atomicAdd(P.counter, 1u) : (round == 0 ? (int)blockIdx.x : n_work);
    if (threadIdx.x < W) { prod[threadIdx.x] = 0; cons[threadIdx.x] = 0; }
    __syncthreads();
    const int item = s_item;
    if (item >= n_work) break;
    const int sidx = P.sidx_list ? P.sidx_list[item] : item;
    const int8_t* tp = P.plain + P.plain_off[sidx];
    int lt = (int)(P.plain_off[sidx + 1] - P.plain_off[sidx]);
    int lq = P.lq;
    const int8_t* qp = P.query;
    int qstep = 1, tstep = 1;
    bool skip = false;
    if (TRACK && P.phase == 1) {
      const int4 c = P.coords[item];
      if (c.y < 0) skip = true;
      else { lq = c.y + 1; lt = c.w + 1; qp += c.y; tp += c.w; qstep = -1; tstep = -1; }
    }
    if (!TRACK && lt < P.min_len) continue;                // CTA-uniform: this target belongs to k3_intra
    const int nstripes = skip ? 0 : (lq + 32 * R3 - 1) / (32 * R3);
    int best = 0, bj = 0x7fffffff, bi = 0x7fffffff;

    for (int stripe = warp; stripe < nstripes; stripe += W) {
      const int row0 = stripe * 32 * R3 + lane * R3;
      __syncwarp();
      for (int c = 0; c < 32; ++c) {
        unsigned word = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int i = row0 + b;
          const int s = (c == 0 || i >= lq) ? -128 : (int)tab[(qp[(long)i * qstep] & 31) * 32 + c];
          word |= (unsigned)(s & 0xff) << (8 * b);
        }
        pl[c * 32 + lane] = word;
      }
      __syncwarp();
      const bool has_in = stripe > 0, in_global = has_in && warp == 0;
      const bool has_out = stripe + 1 < nstripes, out_global = has_out && warp == W - 1;
      const int base_in = ((stripe - 1) / W) * lt;        // producer's column counter at the start of its stripe
      const int base_me = (stripe / W) * lt;
      int H[R3], G[R3];
#pragma unroll
      for (int k = 0; k < R3; ++k) { H[k] = 0; G[k] = 0; }
      int diag_in = 0, out_h = 0, out_v = 0, my_code = 0;
      int pre_code = 0, pre_h = 0, pre_v = 0;
      const int nsteps = lt + 31;
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {
          const int j = s + lane;
          if (has_in && s < lt) {                         // wait until the row above has produced columns [s, s+32)
            const int need = base_in + min(s + 32, lt);
            if (lane == 0) {
              unsigned spins = 0;
              while (prod[pw] < need) { if (++spins > K3W_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; } }
            }
            __syncwarp();
          }
          int2 b = make_int2(0, 0);
          if (has_in && j < lt) b = in_global ? __ldcg(&bnd[j]) : ring[pw][(base_in + j) & (K3W_RING - 1)];
          pre_h = b.x; pre_v = b.y;
          pre_code = (j < lt) ? (int)tp[(long)j * tstep] : 0;
          __syncwarp();
          if (has_in && !in_global && lane == 0 && s < lt) cons[pw] = base_in + min(s + 32, lt);
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
          const unsigned w0 = lds32(pl_addr + (code & 31) * 128);
          int colbest = 0, colk = 0;
#pragma unroll
          for (int k = 0; k < R3; ++k) {
            const int t = (int)add_byte(w0, 1u << (8 * k), (unsigned)d);
            const int h = __vimax3_s32_relu(t, G[k], V);
            d = H[k];
            H[k] = h;
            const int u = h + oe;
            G[k] = __viaddmax_s32(G[k], ge, u);
            V = __viaddmax_s32(V, ge, u);
            if (TRACK) { if (h > colbest) { colbest = h; colk = k; } } else best = max(best, h);
          }
          if (TRACK && k3w_better(colbest, j, row0 + colk, best, bj, bi)) { best = colbest; bj = j; bi = row0 + colk; }
          out_h = H[R3 - 1];
          out_v = V;
          if (lane == 31 && has_out) {
            if (out_global) {
              bnd[j] = make_int2(out_h, out_v);
            } else {
              if ((j & 31) == 0) {                        // never run more than a ring ahead of the consumer
                const int need = base_me + j + 32 - K3W_RING;
                unsigned spins = 0;
                while (cons[warp] < need) { if (++spins > K3W_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; } }
              }
              ring[warp][(base_me + j) & (K3W_RING - 1)] = make_int2(out_h, out_v);
            }
            if ((j & 31) == 31 || j == lt - 1) {
              if (out_global) __threadfence(); else __threadfence_block();
              prod[warp] = base_me + j + 1;
            }
          }
        }
      }
    }

    // result of this item: reduce over lanes, then over the warps of the team
    if (TRACK) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const int s2 = __shfl_xor_sync(0xffffffffu, best, o);
        const int j2 = __shfl_xor_sync(0xffffffffu, bj, o);
        const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
        if (k3w_better(s2, j2, i2, best, bj, bi)) { best = s2; bj = j2; bi = i2; }
      }
    } else {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    }
    if (lane == 0) { red[warp][0] = best; red[warp][1] = bj; red[warp][2] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < W; ++w)
        if (TRACK ? k3w_better(red[w][0], red[w][1], red[w][2], best, bj, bi) : red[w][0] > best) {
          best = red[w][0]; bj = red[w][1]; bi = red[w][2];
        }
      if (TRACK) {
        int4 c = P.coords[item];
        if (P.phase == 0) {
          if (best > 0) { c.y = bi; c.w = bj; } else { c = make_int4(-1, -1, -1, -1); }
          P.best_out[item] = best;
          P.coords[item] = c;
        } else if (!skip) {
          c.x = c.y - bi; c.z = c.w - bj;
          P.coords[item] = c;
        } else {
          P.coords[item] = make_int4(-1, -1, -1, -1);
        }
      } else {
        P.scores_file[P.order[sidx]] = best;
      }
    }
  }
}

}  // namespace swk
