// sw_kernels.cuh -- sm_100a kernels of the Smith-Waterman database-search engine.
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
// K1  k1_inter<R>   inter-sequence, packed 2 x u16 per register, one lane = one PAIR of targets.
// K3  k3_intra<R3>  intra-sequence, 32-bit, one warp = one target (long targets and
//                   rescoring of targets whose 16-bit result saturated -- the reference's
//                   16-bit / 32-bit split, README.md:11).
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
// a + b as IMAD (FMA-heavy pipe); `one` is an opaque register holding 1.
__device__ __forceinline__ unsigned add_fma(unsigned a, unsigned one, unsigned b) {
  unsigned d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b)); return d;
}

// ------------------------------------------------------------------------------------------
// Query-profile tiles for K1.
// One tile = R consecutive query rows.  Per tile and per residue code c (0..31):
//   PA[c][k/4]  int8 x 4 : s(q_{i0+k}, c)                      (target in the LOW  u16 half)
//   PB[c][k/2]  int8 x 4 : s(k), s(k), s(k+1), s(k+1)            (target in the HIGH u16 half)
// Row strides SA = (R/4)|1 and SB = (R/2)|1 words are odd, so the 32 lanes of a warp, which
// read the SAME k with DIFFERENT codes, hit 32 different banks (code -> code*odd mod 32 is a
// bijection) or the same word (broadcast): the lookups are bank-conflict-free by construction.
// Code 0 (padding) and rows past the query end score -128 (below every legal matrix entry,
// src/alignment_scoring.c:61), so padded cells can never raise the running maximum.
// ------------------------------------------------------------------------------------------
template <int R> struct Tile {
  static constexpr int SA = (R / 4) | 1;
  static constexpr int SB = (R / 2) | 1;
  static constexpr int WORDS = 32 * (SA + SB);
};

static __global__ void build_profile(const int8_t* __restrict__ query, int lq, const int8_t* __restrict__ table,
                              int R, int SA, int SB, int ntiles, unsigned* __restrict__ prof) {
  const int words = 32 * (SA + SB);
  const long total = (long)ntiles * words;
  for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total; idx += (long)gridDim.x * blockDim.x) {
    const int tile = (int)(idx / words), w = (int)(idx % words);
    unsigned out = 0;
    auto score = [&](int c, int k) -> unsigned {
      const long i = (long)tile * R + k;
      int s = (c == 0 || i >= lq || k >= R) ? -128 : (int)table[(query[i] & 31) * 32 + c];
      return (unsigned)(s & 0xff);
    };
    if (w < 32 * SA) {
      const int c = w / SA, k4 = w % SA;
      if (k4 < R / 4) for (int b = 0; b < 4; ++b) out |= score(c, 4 * k4 + b) << (8 * b);
    } else {
      const int v = w - 32 * SA, c = v / SB, k2 = v % SB;
      if (k2 < R / 2) {
        const unsigned s0 = score(c, 2 * k2), s1 = score(c, 2 * k2 + 1);
        out = s0 | (s0 << 8) | (s1 << 16) | (s1 << 24);
      }
    }
    prof[idx] = out;
  }
}

// ------------------------------------------------------------------------------------------
// K1: inter-sequence kernel.
//
// A 64-target chunk of the length-sorted database is owned by one warp: lane l holds targets
// 2l (low u16 half) and 2l+1 (high half) of the chunk.  The warp walks the chunk's columns
// (target positions); per column every lane updates R query rows kept in registers
// (H and G per row).  All lanes work on the SAME query rows at the same time, which is what
// makes the profile lookups conflict-free.  The query is covered in ceil(Lq/R) passes; the
// bottom row (H, V) of a pass is carried to the next pass through a per-warp column buffer
// in global memory, [column][lane] of uint2, coalesced.
//
// Arithmetic is BIASED unsigned 16-bit: every stored value is true value + B with
// B >= 128 + |oe| + |ge|.  With the bias no partial result is ever negative, so a 32-bit
// integer add of (sA + 65536*sB) to the packed register cannot borrow across the halves;
// that is what lets the substitution add (IDP.4A + IDP.2A) and the H+oe add (IMAD) run on
// the FMA-heavy pipe while the ALU pipe does only the four max-type DPX ops per cell pair:
//     ef = VIMNMX3.U16x2(G, V, BB)        floor at the bias == the reference's clamp at 0
//     h  = VIMNMX.U16x2(t, ef)
//     G' = VIADDMNMX.U16x2(G, ge, u)      u = h + oe
//     V' = VIADDMNMX.U16x2(V, ge, u)
// plus half a VIMNMX3 for the running maximum.  A half that reaches 65535-128-1 may have
// carried into its neighbour: both targets of such a lane are rescored in 32 bits (K3).
// ------------------------------------------------------------------------------------------
struct K1Params {
  const uint2*   cols;        // [chunk][col/4][lane] : 4 columns x (codeA | codeB << 8)
  const int64_t* chunk_off;   // per chunk, offset into cols (uint2 units)
  const int32_t* chunk_cols;  // per chunk, padded column count (multiple of 4)
  int            n_chunks;
  const unsigned* prof;       // ntiles x Tile<R>::WORDS
  int            ntiles;
  uint2*         bnd;         // per warp slot: bnd_stride uint2
  long long      bnd_stride;
  unsigned*      counter;     // work-queue head (chunks are sorted longest first)
  unsigned*      raw_out;     // [chunk][lane] packed biased best
  unsigned       BB, OE32, GE2, one;
};

constexpr int K1_WARPS = 8;

template <int R>
__global__ void __launch_bounds__(K1_WARPS * 32, 2) k1_inter(const K1Params P) {
  using T = Tile<R>;
  extern __shared__ unsigned smem_u32[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pw = smem_u32 + warp * T::WORDS;
  const unsigned pa_addr = (unsigned)__cvta_generic_to_shared(pw);
  const unsigned pb_addr = pa_addr + 32 * T::SA * 4;
  uint2* bnd = P.bnd + (size_t)(blockIdx.x * K1_WARPS + warp) * (size_t)P.bnd_stride + lane;
  const unsigned BB = P.BB, OE32 = P.OE32, GE2 = P.GE2, one = P.one;

  for (;;) {
    int chunk = 0;
    if (lane == 0) chunk = (int)atomicAdd(P.counter, 1u);
    chunk = __shfl_sync(0xffffffffu, chunk, 0);
    if (chunk >= P.n_chunks) break;
    const int ngroups = P.chunk_cols[chunk] >> 2;
    const uint2* cp = P.cols + P.chunk_off[chunk] + lane;
    unsigned best = BB;

    for (int tile = 0; tile < P.ntiles; ++tile) {
      __syncwarp();
      {
        const unsigned* g = P.prof + (size_t)tile * T::WORDS;
        for (int i = lane; i < T::WORDS; i += 32) pw[i] = g[i];
      }
      __syncwarp();
      const bool first = (tile == 0), last = (tile == P.ntiles - 1);
      unsigned H[R], G[R];
#pragma unroll
      for (int k = 0; k < R; ++k) { H[k] = BB; G[k] = BB; }
      unsigned hin_prev = BB;                       // H(i0-1, j-1)
      uint2 letters = cp[0];
      uint2 bin = first ? make_uint2(BB, BB) : bnd[0];

      for (int g4 = 0; g4 < ngroups; ++g4) {
        const uint2 letters_next = (g4 + 1 < ngroups) ? cp[(size_t)(g4 + 1) * 32] : make_uint2(0u, 0u);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int col = g4 * 4 + c;
          uint2 bin_next = make_uint2(BB, BB);
          if (!first && (c < 3 || g4 + 1 < ngroups)) bin_next = bnd[(size_t)(col + 1) * 32];
          const unsigned lw = ((c < 2) ? letters.x : letters.y) >> (16 * (c & 1));
          const unsigned aoff = pa_addr + (lw & 0xffu) * (T::SA * 4);
          const unsigned boff = pb_addr + ((lw >> 8) & 0xffu) * (T::SB * 4);
          unsigned d = hin_prev;
          hin_prev = bin.x;
          unsigned V = bin.y;
          unsigned wA = 0, wB = 0, hprev = BB;
#pragma unroll
          for (int k = 0; k < R; ++k) {
            if ((k & 3) == 0) wA = lds32(aoff + k);          // word k/4
            if ((k & 1) == 0) wB = lds32(boff + 2 * k);      // word k/2
            unsigned t = add_byte(wA, 1u << (8 * (k & 3)), d);
            t = (k & 1) ? add_hi_hi(wB, t) : add_hi_lo(wB, t);
            const unsigned ef = __vimax3_u16x2(G[k], V, BB);
            const unsigned h = __vmaxu2(t, ef);
            d = H[k];
            H[k] = h;
            const unsigned u = add_fma(h, one, OE32);
            G[k] = __viaddmax_u16x2(G[k], GE2, u);
            V = __viaddmax_u16x2(V, GE2, u);
            if (k & 1) best = __vimax3_u16x2(best, hprev, h); else hprev = h;
          }
          if (R & 1) best = __vmaxu2(best, hprev);
          if (!last) bnd[(size_t)col * 32] = make_uint2(H[R - 1], V);
          bin = bin_next;
        }
        letters = letters_next;
      }
    }
    P.raw_out[(size_t)chunk * 32 + lane] = best;
  }
}

// ------------------------------------------------------------------------------------------
// K1 epilogue: un-bias, detect saturation, scatter to file order.
// sorted index s = first + chunk*64 + lane*2 + half.  A lane whose either half reached
// `limit` sends BOTH its targets to the 32-bit rescoring list (a carry out of the low half
// would have corrupted the high half).
// ------------------------------------------------------------------------------------------
static __global__ void k1_finalize(const unsigned* __restrict__ raw, long n_lanes, long first, long n_sorted,
                            unsigned B, unsigned limit, const int64_t* __restrict__ order,
                            int32_t* __restrict__ scores_file, int* __restrict__ rescore_list,
                            unsigned* __restrict__ rescore_count) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n_lanes; i += (long)gridDim.x * blockDim.x) {
    const unsigned r = raw[i], lo = r & 0xffffu, hi = r >> 16;
    const bool sat = (lo >= limit) || (hi >= limit);
    const long s0 = first + 2 * i;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const long s = s0 + h;
      if (s >= n_sorted) continue;
      if (sat) {
        const unsigned slot = atomicAdd(rescore_count, 1u);
        rescore_list[slot] = (int)s;
      } else {
        scores_file[order[s]] = (int32_t)((h ? hi : lo) - B);
      }
    }
  }
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
  int            max_len;      // targets longer than this are left to the wavefront kernel k3w (0 = no limit)
};

constexpr int K3_WARPS = 4;

template <int R3>
__global__ void __launch_bounds__(K3_WARPS * 32) k3_intra(const K3Params P) {
  static_assert(R3 % 4 == 0, "R3 must be a multiple of 4");
  constexpr int W = R3 / 4;                        // profile words per (lane, code)
  __shared__ unsigned pl_all[K3_WARPS][32 * W * 32];
  __shared__ int8_t tab[32 * 32];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
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
      int diag_in = 0, out_h = 0, out_v = 0, my_code = 0;
      int pre_code = 0, pre_h = 0, pre_v = 0;
      const bool first = (pass == 0), last = (pass == npass - 1);
      const int nsteps = lt + 31;
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {                       // coalesced prefetch of the next 32 columns for lane 0
          const int j = s + lane;
          pre_code = (j < lt) ? (int)tp[j] : 0;
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
#pragma unroll
          for (int k = 0; k < R3; ++k) {
            const int t = (int)add_byte(wreg[k >> 2], 1u << (8 * (k & 3)), (unsigned)d);
            const int h = __vimax3_s32_relu(t, G[k], V);
            d = H[k];
            H[k] = h;
            const int u = (int)add_fma((unsigned)h, P.one, (unsigned)oe);
            G[k] = __viaddmax_s32(G[k], ge, u);
            V = __viaddmax_s32(V, ge, u);
            best = max(best, h);
          }
          out_h = H[R3 - 1];
          out_v = V;
          if (lane == 31 && !last) bnd[j] = make_int2(out_h, out_v);
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) P.scores_file[P.order[sidx]] = best;
  }
}

}  // namespace swk
