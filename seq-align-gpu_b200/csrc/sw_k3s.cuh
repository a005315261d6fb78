// sw_k3s.cuh -- K3S: the reference's THREE-state recurrence, 32-bit, one warp per target (sm_100a).
//
// Used when a gap score is positive.  The reference accepts any --gapopen / --gapextend (src/alignment_cmdline.c:255-267)
// and runs its three-state recurrence on them (src/alignment.c:124-161), all three states clamped at 0:
//     M[j][i] = max(0, max(M, E, F)[j-1][i-1] + s)                                   :124-129
//     E[j][i] = max(0, M[j-1][i] + oe, E[j-1][i] + ge, F[j-1][i] + oe)               :142-147  (previous target residue, same query row)
//     F[j][i] = max(0, M[j][i-1] + oe, E[j][i-1] + oe, F[j][i-1] + ge)               :156-161  (previous query row, same target column)
//     score   = max M                                                                :133, :186
// with boundary row / column zero (:77-85, :91-97) -- so E of the first target column and F of the first query row are
// max(0, oe, ge), computed from those zeros.  For oe, ge <= 0 this equals the two-gap-state form every other kernel of the
// engine uses; for positive values it does not (E can follow F and the clamps matter), hence this
// kernel.  A rare path: plain 32-bit code, no packing.
//
// Same decomposition as k3_intra (sw_kernels.cuh): lane l owns R3 consecutive query rows of the current stripe, the
// target streams through the lanes systolically (lane l works on column s - l at step s), the stripe's bottom row is
// carried through a per-warp row of (D, F) in global memory.  Per row the state kept from one column to the next is
//     D[k] = max(M, E, F) of the previous column (the diagonal input of row k + 1)   and   EN[k] = E of the next column.
// Unlike k3_intra a lane may NOT run on padding before its first column (with positive gap scores the state would grow
// out of nothing), so out-of-range columns reset the lane's state, and rows past the query end do not feed the maximum.
#pragma once
#include "sw_kernels.cuh"

namespace swk {

template <int R3>
__global__ void __launch_bounds__(K3_WARPS * 32) k3_three(const K3Params P) {
  static_assert(R3 % 4 == 0, "R3 must be a multiple of 4");
  constexpr int W = R3 / 4;
  __shared__ unsigned pl_all[K3_WARPS][32 * W * 32];
  __shared__ int8_t tab[32 * 32];
  __shared__ unsigned ringc_all[K3_WARPS][64];
  __shared__ int2 ringb_all[K3_WARPS][32];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  unsigned* rc = ringc_all[warp];
  int2* rb = ringb_all[warp];
  int2* bnd = P.bnd + (size_t)(blockIdx.x * K3_WARPS + warp) * (size_t)P.bnd_stride;
  const int n_work = P.work_count ? (int)*P.work_count : P.n_work;
  const int oe = P.oe, ge = P.ge, lq = P.lq;
  const int e0 = max(0, max(oe, ge));                      // E / F next to the zero boundary
  const int npass = (lq + 32 * R3 - 1) / (32 * R3);

  for (;;) {
    int w = 0;
    if (lane == 0) w = (int)atomicAdd(P.counter, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= n_work) break;
    const int sidx = P.worklist ? P.worklist[w] : w;
    const int8_t* tp = P.plain + P.plain_off[sidx];
    const int lt = (int)(P.plain_off[sidx + 1] - P.plain_off[sidx]);
    int best = 0;
    for (int pass = 0; pass < npass && lt > 0; ++pass) {
      const int row0 = pass * 32 * R3 + lane * R3;
      const int rows_valid = min(max(lq - row0, 0), R3);   // rows of this lane that exist in the query
      __syncwarp();
      for (int c = 0; c < 32; ++c) {
#pragma unroll
        for (int ww = 0; ww < W; ++ww) {
          unsigned word = 0;
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const int i = row0 + ww * 4 + b;
            const int s = (c == 0 || i >= lq) ? -128 : (int)tab[(P.query[i] & 31) * 32 + c];
            word |= (unsigned)(s & 0xff) << (8 * b);
          }
          pl[(c * W + ww) * 32 + lane] = word;
        }
      }
      __syncwarp();
      int D[R3], EN[R3];
#pragma unroll
      for (int k = 0; k < R3; ++k) { D[k] = 0; EN[k] = e0; }
      int diag_in = 0, out_d = 0, out_f = e0;
      const bool first = (pass == 0), last = (pass == npass - 1);
      const int nsteps = lt + 31;
      rc[lane] = 0; rc[lane + 32] = 0;
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {
          const int j = s + lane;
          int2 b = make_int2(0, e0);                       // stripe 0: the zero boundary row above the query
          if (!first && j < lt) b = bnd[j];
          __syncwarp();
          rb[lane] = b;
          rc[j & 63] = (j < lt) ? ((unsigned)tp[j] & 31u) * (unsigned)(W * 128) : 0u;
          __syncwarp();
        }
        const int j = s - lane;
        int d_up = __shfl_up_sync(0xffffffffu, out_d, 1);  // max(M, E, F) of the row above, this column
        int f_in = __shfl_up_sync(0xffffffffu, out_f, 1);  // F of this lane's first row, this column
        const int2 b0 = rb[s & 31];
        if (lane == 0) { d_up = b0.x; f_in = b0.y; }
        if ((unsigned)j < (unsigned)lt) {
          const unsigned poff = pl_addr + rc[j & 63];
          int d = diag_in;
          diag_in = d_up;
          int f = f_in;
          unsigned wreg[W];
#pragma unroll
          for (int ww = 0; ww < W; ++ww) wreg[ww] = lds32_ordered(poff + ww * 128);
#pragma unroll
          for (int k = 0; k < R3; ++k) {
            const int sc = (int)(signed char)((wreg[k >> 2] >> (8 * (k & 3))) & 0xffu);
            const int m = max(d + sc, 0);
            const int e = EN[k];
            if (k < rows_valid) best = max(best, m);
            d = D[k];
            D[k] = max(m, max(e, f));
            EN[k] = max(max(max(m, f) + oe, e + ge), 0);
            f = max(max(max(m, e) + oe, f + ge), 0);
          }
          out_d = D[R3 - 1];
          out_f = f;
          if (lane == 31 && !last) bnd[j] = make_int2(out_d, out_f);
        } else {                                           // before the first / past the last column: the zero boundary column
          diag_in = 0;
#pragma unroll
          for (int k = 0; k < R3; ++k) { D[k] = 0; EN[k] = e0; }
          out_d = 0; out_f = e0;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) P.scores_file[P.order[sidx]] = best;
  }
}

}  // namespace swk
