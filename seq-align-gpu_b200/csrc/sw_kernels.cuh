// sw_kernels.cuh -- shared device helpers and K3, the 32-bit intra-sequence kernel (sm_100a).
//
// The recurrence is the reference's 3-state affine form (src/alignment.c:117-176 of the
// reference tree), evaluated in the score-equivalent 2-gap-state (Gotoh) form that
// SURVEY.md A.1/B.7 proves identical for gap_open, gap_extend <= 0:
//     H(i,j) = max(0, H(i-1,j-1) + s(q_i,t_j), G(i,j), V(i,j))
//     G(i,j) = max(G(i,j-1) + ge, H(i,j-1) + oe)      gap along the target  (reference "gap_a", :142-147)
//     V(i,j) = max(V(i-1,j) + ge, H(i-1,j) + oe)      gap along the query   (reference "gap_b", :156-161)
//     score  = max H                                   (:133, :186)
// i = query position, j = target position, oe = gap_open + gap_extend (:58), ge = gap_extend (:59).
//
// K1S (sw_k1s.cuh)  inter-sequence, packed 2 x u16, the dominant kernel
// K3  k3_intra<R3>  32-bit, one warp per target: rescoring of targets whose 16-bit result
//                   saturated -- the reference's 16-bit / 32-bit split, README.md:11
// K3X (sw_k3x.cuh)  32-bit wavefront, one warp per 256-row stripe: long targets, top-k coordinates
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace swk {

// ------------------------------------------------------------------------------------------
// Small device helpers.  The point of the odd-looking ones: the integer ALU pipe (VIMNMX*,
// VIADDMNMX*, IADD3, LOP3, PRMT) issues 64 lanes/clk/SM on B200, and so does the FMA-heavy pipe
// (IMAD, IDP) -- measured by bench_micro/ubench_int.cu -- and they issue concurrently.  Every
// add that can be expressed as IMAD/IDP is therefore moved off the ALU pipe.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned lds32(unsigned saddr) {
  unsigned v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr)); return v;
}
// same, ordered against ordinary shared-memory stores (data another lane wrote before a __syncwarp)
__device__ __forceinline__ unsigned lds32_ordered(unsigned saddr) {
  unsigned v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory"); return v;
}
__device__ __forceinline__ void lds64(unsigned saddr, unsigned& lo, unsigned& hi) {
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(saddr));
}
__device__ __forceinline__ void lds64_ordered(unsigned saddr, unsigned& lo, unsigned& hi) {
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(saddr) : "memory");
}
// c + (signed byte `sel` of a): IDP.4A on the FMA-heavy pipe.  sel_mask = 1 << (8*byte).
__device__ __forceinline__ unsigned add_byte(unsigned a, unsigned sel_mask, unsigned c) {
  return (unsigned)__dp4a((int)a, (int)sel_mask, (int)c);
}
// c + 65536 * (signed byte 0 of b)  [bytes 0 and 1 of b hold the same value; 32768*b0 + 32768*b1]
__device__ __forceinline__ unsigned add_hi_lo(unsigned b, unsigned c) {
  unsigned d; asm("dp2a.lo.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(0x80008000u), "r"(b), "r"(c)); return d;
}
// c + 65536 * (signed byte 2 of b)  [bytes 2 and 3 equal]
__device__ __forceinline__ unsigned add_hi_hi(unsigned b, unsigned c) {
  unsigned d; asm("dp2a.hi.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(0x80008000u), "r"(b), "r"(c)); return d;
}
// value of the lane below within a `width`-lane segment; the segment's first lane (no lane below) gets `dflt`.
// The shuffle's own range predicate drives the select: no separate compare.
template <int WIDTH>
__device__ __forceinline__ unsigned shfl_up1_or(unsigned v, unsigned dflt) {
  unsigned d;
  asm volatile("{ .reg .pred p; shfl.sync.up.b32 %0|p, %1, 1, %2, 0xffffffff; @!p mov.b32 %0, %3; }"
               : "=r"(d) : "r"(v), "n"((32 - WIDTH) << 8), "r"(dflt));
  return d;
}
// c + lo16(m) * byte0(b) + hi16(m) * byte1(b): IDP.2A on the FMA-heavy pipe (bytes of b below 128)
__device__ __forceinline__ unsigned mul_bytes01(unsigned m, unsigned b, unsigned c) {
  unsigned d; asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(m), "r"(b), "r"(c)); return d;
}
// a + b as IMAD (FMA-heavy pipe); `one` is an opaque register holding 1.
__device__ __forceinline__ unsigned add_fma(unsigned a, unsigned one, unsigned b) {
  unsigned d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b)); return d;
}

// ------------------------------------------------------------------------------------------
// K3: intra-sequence 32-bit kernel, one warp per target (long targets, and rescoring of
// targets whose 16-bit result saturated).
//
// Lane l owns R3 consecutive query rows of the current stripe (32*R3 rows per pass); the
// target streams through the lanes systolically: at step s lane l works on column s-l and
// takes H(row above, col) / V(first own row, col) and the residue code from lane l-1's
// previous step by __shfl_up_sync.  The bottom row of a stripe is carried to the next
// stripe through a per-warp linear-space buffer (H, V per column) in global memory.
// Substitution scores come from a per-lane private profile in shared memory,
// PL[code][word][lane] = int8 x 4 of s(own row k, code): bank == lane, conflict-free although
// every lane is at a different column; IDP.4A extracts and adds the byte.
// Unbiased signed 32-bit arithmetic: VIMNMX3.RELU / VIADDMNMX on the ALU pipe.
// ------------------------------------------------------------------------------------------
struct K3Params {
  const int*     worklist;     // sorted indices to score; nullptr -> 0..n_work-1
  int            n_work;
  const unsigned* work_count;  // if non-null, n_work is read from here (device-produced list)
  unsigned*      counter;
  const int8_t*  plain;        // residues, sorted order, back to back
  const int64_t* plain_off;    // n_sorted + 1
  const int8_t*  query;
  int            lq;
  const int8_t*  table;        // [32*32]
  int            oe, ge;
  int2*          bnd;          // per warp slot: bnd_stride int2 (H, V)
  long long      bnd_stride;
  const int64_t* order;
  int32_t*       scores_file;
  unsigned       one;
  int            max_len;      // targets longer than this are left to the wavefront kernel k3x (0 = no limit)
  unsigned*      sm_slots;     // [256], zero at launch, or nullptr: CTAs started per SM (placement by level, as in k3x)
  int            level_ctas;   // CTAs per level = number of SMs
};

constexpr int K3_WARPS = 4;

template <int R3>
__global__ void __launch_bounds__(K3_WARPS * 32) k3_intra(const K3Params P) {
  static_assert(R3 % 4 == 0, "R3 must be a multiple of 4");
  constexpr int W = R3 / 4;                        // profile words per (lane, code)
  __shared__ unsigned pl_all[K3_WARPS][32 * W * 32];
  __shared__ int8_t tab[32 * 32];
  // Placement by level (see k3x): a rescoring list is often far shorter than the grid has warps, and the queue hands the
  // items to whichever warps ask first -- a few SMs with four busy warps per scheduler while the rest of the machine idles.
  // CTAs number themselves per SM in arrival order; level L takes part only when the levels below cannot give every item
  // a warp of its own, and lets them ask first.
  __shared__ int cta_level;
  if (P.sm_slots && threadIdx.x == 0) {
    unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    cta_level = (int)atomicAdd(P.sm_slots + (smid & 255u), 1u);
  }
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  if (P.sm_slots) {
    const long long items = P.work_count ? (long long)*P.work_count : (long long)P.n_work;
    const long long per_level = (long long)P.level_ctas * K3_WARPS;
    if (items <= (long long)cta_level * per_level) return;
    if (cta_level > 0 && items < (long long)gridDim.x * K3_WARPS) __nanosleep(3000u * (unsigned)min(cta_level, 8));
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  // per-warp rings (same scheme as k3x, sw_k3x.cuh): profile byte offset of the last 64 target columns, and the (H, V)
  // of the stripe above for the current block of 32 columns -- one shared load each per step instead of a lane-to-lane
  // code shuffle and three broadcast shuffles
  __shared__ unsigned ringc_all[K3_WARPS][64];
  __shared__ int2 ringb_all[K3_WARPS][32];
  unsigned* rc = ringc_all[warp];
  int2* rb = ringb_all[warp];
  int2* bnd = P.bnd + (size_t)(blockIdx.x * K3_WARPS + warp) * (size_t)P.bnd_stride;
  const int n_work = P.work_count ? (int)*P.work_count : P.n_work;
  const int oe = P.oe, ge = P.ge, lq = P.lq;
  const int npass = (lq + 32 * R3 - 1) / (32 * R3);

  for (;;) {
    int w = 0;
    if (lane == 0) w = (int)atomicAdd(P.counter, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= n_work) break;
    const int sidx = P.worklist ? P.worklist[w] : w;
    const int8_t* tp = P.plain + P.plain_off[sidx];
    const int lt = (int)(P.plain_off[sidx + 1] - P.plain_off[sidx]);
    if (P.max_len > 0 && lt > P.max_len) continue;
    int best = 0;

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
            const int s = (c == 0 || i >= lq) ? -128 : (int)tab[(P.query[i] & 31) * 32 + c];
            word |= (unsigned)(s & 0xff) << (8 * b);
          }
          pl[(c * W + ww) * 32 + lane] = word;
        }
      }
      __syncwarp();
      int H[R3], G[R3];
#pragma unroll
      for (int k = 0; k < R3; ++k) { H[k] = 0; G[k] = 0; }
      int diag_in = 0, out_h = 0, out_v = 0;
      const bool first = (pass == 0), last = (pass == npass - 1);
      const int nsteps = lt + 31;
      rc[lane] = 0; rc[lane + 32] = 0;             // columns before the first one: padding code
      int2* pcol = bnd - lane;                     // &bnd[s - lane], advanced every step
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {                       // coalesced fetch of the next 32 columns
          const int j = s + lane;
          int2 b = make_int2(0, 0);
          if (!first && j < lt) b = bnd[j];
          __syncwarp();
          rb[lane] = b;
          rc[j & 63] = (j < lt) ? ((unsigned)tp[j] & 31u) * (unsigned)(W * 128) : 0u;
          __syncwarp();
        }
        // Column j = s - lane; no range test around the cell update (see k3x): before its first column a lane sees
        // zeros and the padding code, past its last column it computes on the padding code, which cannot raise the
        // maximum.  Only the hand-over store is guarded.
        const int j = s - lane;
        int h_up = __shfl_up_sync(0xffffffffu, out_h, 1);
        int v_in = __shfl_up_sync(0xffffffffu, out_v, 1);
        const int2 b0 = rb[s & 31];
        if (lane == 0) { h_up = b0.x; v_in = b0.y; }
        const unsigned poff = pl_addr + rc[j & 63];
        {
          int d = diag_in;
          diag_in = h_up;
          int V = v_in;
          unsigned wreg[W];
#pragma unroll
          for (int ww = 0; ww < W; ++ww) wreg[ww] = lds32_ordered(poff + ww * 128);
#pragma unroll
          for (int k = 0; k < R3; ++k) {
            // gap states shifted by -oe (G holds G - oe, V holds V - oe): one add per operation and a row-to-row chain of ONE
            // operation (V' from Hx, which does not depend on V; valid for oe <= ge) -- see sw_k3x.cuh
            const int t = (int)add_byte(wreg[k >> 2], 1u << (8 * (k & 3)), (unsigned)d);
            const int hx = __viaddmax_s32_relu(G[k], oe, t);
            const int h = __viaddmax_s32(V, oe, hx);
            V = __viaddmax_s32(V, ge, hx);
            G[k] = __viaddmax_s32(G[k], ge, h);
            d = H[k];
            H[k] = h;
            if (k & 1) best = __vimax3_s32(best, H[k - 1], h);
          }
          out_h = H[R3 - 1];
          out_v = V;
          if (lane == 31 && !last && (unsigned)j < (unsigned)lt) *pcol = make_int2(out_h, out_v);
        }
        ++pcol;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) P.scores_file[P.order[sidx]] = best;
  }
}

}  // namespace swk
