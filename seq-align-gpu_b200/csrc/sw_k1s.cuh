// sw_k1s.cuh -- K1S: systolic inter-sequence kernel, packed 2 x u16 (sm_100a).
//
// Same recurrence and the same biased-u16 cell update as k1_inter (sw_kernels.cuh; reference
// src/alignment.c:117-176), but a different decomposition, chosen after profiling k1_inter
// (profiles/r1a_*: 31 GB of DRAM traffic per launch for the per-column boundary rows, 35 % of
// warp stalls on those loads):
//
//   * one PAIR of targets (the two u16 halves of every register) is owned by a group of g lanes
//     (g = 8, 16 or 32); lane l of the group keeps R consecutive query rows of the current
//     g*R-row stripe in registers, so a query of up to 32*32 = 1024 residues lives entirely in
//     the register file of one warp and NO boundary row ever goes to memory;
//   * the target streams through the lanes systolically: at step s lane l works on column
//     s - l and takes H(row above) / V from lane l-1 by __shfl_up_sync;
//   * pairs are concatenated into streams (db_pack.h), a flag bit in the column word marks the
//     first column of a pair: the lane flushes its running maximum for the finished pair and
//     resets its rows.  The pipeline therefore fills and drains once per STREAM, not once per
//     target, which is what keeps short targets efficient;
//   * the query profile is private per lane, PL[code][word][lane] in shared memory: every lane
//     is at a different column (different residue), bank == lane, conflict-free by construction;
//   * queries longer than 32*R rows take several passes (stripes); only then the stripe's bottom
//     row is carried through a per-warp linear-space buffer (8 bytes per column per pass).
//
// Round 2 (profiles/r2a_*):
//   * GAPSEL: the gap penalties as template constants.  VIMNMX3 / VIADDMNMX take a 32-bit immediate in their
//     second source, so with compile-time penalties three of the four DPX instructions of a cell have TWO register
//     sources instead of three.  That matters because a DPX instruction whose three source registers sit in the
//     same register-file bank (even / odd register number) takes 3 issue cycles on the ALU pipe instead of 2
//     (bench_micro/ubench_bank.cu: 1.37 vs 1.99 warp-instructions / clk / SM), ptxas cannot avoid all of them at
//     120+ live registers, and the three constants (and their per-step LDC reloads) leave the register file.
//     GAPSEL 0 is the generic kernel (penalties in registers), 1 = --gapopen -10 --gapextend -1, 2 = -11 / -1.
//   * the column loop is unrolled by two where that pays (K1S_UNROLL_*): the new H of a row no longer has to be moved
//     back into the register the next column reads it from (one IMAD.MOV per row per column before).
//   * multi-stripe: the first lane takes the boundary values of the stripe above from a broadcast shared load that
//     feeds the shuffle's default operand (same code for every stripe, no per-step branch), and lane 31 hands its
//     bottom row over through a 32-column shared ring that the whole warp writes out coalesced every 32 steps
//     (before: a bounds-checked 8-byte global store by one lane per step).
#pragma once
#include "sw_kernels.cuh"

namespace swk {

// Profile layout.  From K1S_PAIRED_MIN_R rows per lane on, the words of a lane are stored in PAIRS (8 bytes per lane,
// [code][word pair][lane][2]) and fetched with one 64-bit shared load per pair: half the LDS instructions for the
// same bytes.  Below that (the 3-CTA/SM plans, at their register limit) it is [code][word][lane], 32-bit loads.
#ifndef K1S_PAIRED_MIN_R
#define K1S_PAIRED_MIN_R 20
#endif
// Measured per instantiation (scripts/ab_lib.py): pairs gain 4-5 % at R = 28 and 32, 0-2 % at R = 24, and lose 1 % at
// R = 22 and in the multi-stripe R = 30 kernel, which therefore keep the 32-bit layout.
__host__ __device__ constexpr bool sprof_paired(int R, bool multi) {
  return R >= K1S_PAIRED_MIN_R && R != 21 && R != 22 && !(multi && R == 30);
}
__host__ __device__ constexpr int sprof_wa(int R, bool multi) { return sprof_paired(R, multi) ? 2 * (((R + 3) / 4 + 1) / 2) : (R + 3) / 4; }
__host__ __device__ constexpr int sprof_wb(int R, bool multi) { return sprof_paired(R, multi) ? 2 * (((R + 1) / 2 + 1) / 2) : (R + 1) / 2; }
__host__ __device__ constexpr int sprof_words(int R, bool multi) { return 32 * (sprof_wa(R, multi) + sprof_wb(R, multi)) * 32; }

template <int R, bool MULTI> struct SProf {
  static_assert(R >= 4 && R <= 32, "R must be 4..32");
  static constexpr bool PAIRED = sprof_paired(R, MULTI);
  static constexpr int WA = sprof_wa(R, MULTI);            // int8x4 words: s(k..k+3) for the LOW-half target (padded to even if PAIRED)
  static constexpr int WB = sprof_wb(R, MULTI);          // int8x4 words: s(k),s(k),s(k+1),s(k+1) for the HIGH-half target
  static constexpr int WORDS = sprof_words(R, MULTI);    // A section (all codes), then B section
};

// prof[pass][A section | B section]; lane l of a g-lane group owns rows pass*g*R + (l % g)*R + k.
// Code 0 (padding) and rows past the query end (or past R) score -128.
static __global__ void build_sprofile(const int8_t* __restrict__ query, int lq, const int8_t* __restrict__ table,
                                      int R, int g, int npass, unsigned* __restrict__ prof) {
  const bool multi = npass > 1;
  const int WA = sprof_wa(R, multi), WB = sprof_wb(R, multi);
  const bool paired = sprof_paired(R, multi);
  const int words = 32 * (WA + WB) * 32;
  const long total = (long)npass * words;
  for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total; idx += (long)gridDim.x * blockDim.x) {
    const int pass = (int)(idx / words), w = (int)(idx % words);
    const bool isA = w < 32 * WA * 32;
    const int v = isA ? w : w - 32 * WA * 32, W = isA ? WA : WB;
    const int c = v / (W * 32), r = v % (W * 32);
    // word index within the code and lane: [word][lane], or [word pair][lane][2]
    const int word = paired ? 2 * (r / 64) + (r & 1) : r / 32;
    const int lane = paired ? (r % 64) / 2 : r % 32;
    const long row0 = (long)pass * g * R + (long)(lane & (g - 1)) * R;
    auto score = [&](int k) -> unsigned {
      const long i = row0 + k;
      const int s = (c == 0 || k >= R || i >= lq) ? -128 : (int)table[(query[i] & 31) * 32 + c];
      return (unsigned)(s & 0xff);
    };
    unsigned out;
    if (isA) {
      out = score(4 * word) | (score(4 * word + 1) << 8) | (score(4 * word + 2) << 16) | (score(4 * word + 3) << 24);
    } else {
      const unsigned s0 = score(2 * word), s1 = score(2 * word + 1);
      out = s0 | (s0 << 8) | (s1 << 16) | (s1 << 24);
    }
    prof[idx] = out;
  }
}

struct K1SParams {
  const uint16_t* scols;        // stream columns (db_pack.h)
  const int64_t*  stream_off;
  const int32_t*  stream_ncols;
  const int64_t*  stream_pair0;
  int             stream_begin; // first stream of this launch
  int             n_streams;    // streams in this launch
  int             n_items;      // warp-items: ceil(n_streams / (32/g))
  const unsigned* prof;         // npass x SProf<R>::WORDS
  int             npass;
  int             g, g_log2;    // lanes per stream
  int             active_warps; // single-pass only: warps of a CTA that pull work (1 for the long-stream launch)
  uint2*          bnd;          // multi-pass only: per warp slot, bnd_stride columns of (H, V)
  long long       bnd_stride;
  unsigned*       counter;      // work queue head
  unsigned*       raw;          // [pair][g] packed biased running maxima
  unsigned        BB, BBM, OE2, GE2;   // bias, bias - oe, packed oe, packed ge (u16 x 2); a GAPSEL kernel has them as immediates
  // multi-pass only -- X items: the first n_x_streams streams (the long ones) are cut into (8 streams, ONE stripe)
  // items, so the stripes of a titin-length pair run at the same time on different CTAs instead of one after the
  // other on one warp.  Stripe p-1 -> stripe p through x_bnd (two full-length rows per stream, ping-pong by stripe
  // parity) and one progress word per (stream, stripe): st.release.gpu by the producer's lane 31 every 32 columns
  // (after the warp's coalesced write of those columns and a __syncwarp), ld.acquire.gpu by the consumer's lane 0.
  // Items are ordered (group, stripe) ascending, so a producer always has the smaller item index and is already
  // running when its consumer starts.
  int             n_x_streams;
  uint2*          x_bnd;        // per X stream two rows of its own length, at 2 * stream_off[stream]
  unsigned*       x_progress;   // [n_x_streams][npass], zeroed per search
  unsigned*       x_raw;        // [pair][npass][32] running maxima of the X pairs (one slot set per stripe)
  unsigned*       error_flag;
};

// Gap penalties known at compile time (GAPSEL > 0): oe = gap_open + gap_extend, ge = gap_extend.
template <int GAPSEL> struct K1SGap { static constexpr int OE = 0, GE = 0; };
template <> struct K1SGap<1> { static constexpr int OE = -11, GE = -1; };    // --gapopen -10 --gapextend -1 (the BASELINE configs)
template <> struct K1SGap<2> { static constexpr int OE = -12, GE = -1; };    // --gapopen -11 --gapextend -1 (BLAST's BLOSUM62 default)
constexpr int K1S_GAPSELS = 3;
__host__ __device__ constexpr int k1s_gapsel(int oe, int ge) { return (oe == -11 && ge == -1) ? 1 : (oe == -12 && ge == -1) ? 2 : 0; }
// bias of the unsigned 16-bit arithmetic: stored = true + B, no partial result is negative
__host__ __device__ constexpr unsigned k1s_bias(int oe, int ge) { return 128u + (unsigned)(-oe) + (unsigned)(-ge); }

__device__ __forceinline__ unsigned k1s_ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void k1s_st_release(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
constexpr unsigned K1S_SPIN_LIMIT = 1u << 24;

constexpr int K1S_WARPS = 8;
constexpr int K1S_RING_WORDS = 160;           // column ring per warp: (32 / g) group segments of 2 * (2g) + g words

#ifndef K1S_R3CTA
#define K1S_R3CTA 16                         // R up to this: three CTAs per SM (<= 85 registers per thread); measured: R = 18 runs
                                             // 2-8 % faster with two CTAs of 86 registers, R = 13 and 16 1-2 % faster with three
#endif
// columns per loop iteration (A/B knobs, scripts/build_variant.sh)
#ifndef K1S_UNROLL_SINGLE
#define K1S_UNROLL_SINGLE 4                  // four columns: 0.23 register moves per row left (two columns: 0.47, one: 1.03); 1-4 % faster at
#endif                                       // every single-stripe length of the C2 sweep (profiles/r2d_ab_unroll4.txt)
#ifndef K1S_UNROLL_MULTI
#define K1S_UNROLL_MULTI 1
#endif
// 1: single-stripe kernels load the ring word of step s + 1 at the end of step s, so that the shared-memory latency is not at
// the head of every step's dependency chain (1-2 % at most single-stripe lengths; the multi-stripe kernel, which would also
// prefetch the boundary pair of the stripe above, measured 0.2-1.8 % SLOWER with it and keeps the direct loads)
#ifndef K1S_PREFETCH
#define K1S_PREFETCH 1
#endif

template <int R, bool MULTI>
constexpr size_t k1s_smem_bytes() {
  return (size_t)SProf<R, MULTI>::WORDS * 4 + (size_t)K1S_WARPS * (K1S_RING_WORDS * 4 + (MULTI ? 2 : 1) * 32 * 8) + 16;
}

// LANES = lanes per target pair (g), a compile-time constant (shuffle widths, ring masks and the block period fold into
// immediates: measured 5-7 % over a run-time g).
// MULTI = false: the whole query fits one stripe; warps pull items independently.
// MULTI = true : g == 32, several stripes; the CTA pulls 8 items at a time and reloads the
//                shared profile between stripes.
// GAPSEL: see the header of this file.
template <int R, int LANES, bool MULTI, int GAPSEL>
__global__ void __launch_bounds__(K1S_WARPS * 32, (R <= K1S_R3CTA ? 3 : 2)) k1s(const K1SParams P) {
  static_assert(LANES == 8 || LANES == 16 || LANES == 32, "lanes per pair");
  static_assert(!MULTI || LANES == 32, "multi-stripe mode uses all 32 lanes");
  static_assert(GAPSEL >= 0 && GAPSEL < K1S_GAPSELS, "gap specialisation");
  using S = SProf<R, MULTI>;
  extern __shared__ __align__(16) unsigned smem_u32[];
  unsigned* prof_s = smem_u32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Column ring: K1S_RING_WORDS words per warp.  A group's segment holds RN = 2g columns, each stored TWICE (at p and p + RN), so
  // that the read index (s mod RN) - lane_g + RN never wraps: the address is a per-lane constant plus a warp-uniform
  // term, no per-lane arithmetic per step.  A ring word is (code of the low target | code of the high target << 8 |
  // first-column flag << 31): the two profile addresses come out of it with one IDP.2A each (FMA pipe).
  unsigned* ring = smem_u32 + S::WORDS + warp * K1S_RING_WORDS;
  uint2* ring2 = reinterpret_cast<uint2*>(smem_u32 + S::WORDS + K1S_WARPS * K1S_RING_WORDS) + warp * 32;   // (H, V) of the stripe above, this block of 32 columns
  uint2* ring3 = ring2 + K1S_WARPS * 32;                                                                     // MULTI: (H, V) of this stripe's bottom row, last 32 steps
  __shared__ unsigned cta_item;

  constexpr int g = LANES, g_log2 = (LANES == 8 ? 3 : LANES == 16 ? 4 : 5);
  constexpr int UN = MULTI ? K1S_UNROLL_MULTI : K1S_UNROLL_SINGLE;
  constexpr bool PF = K1S_PREFETCH != 0 && (!MULTI || K1S_PREFETCH == 2);       // single-stripe only (measured: the multi-stripe kernel is 0.2-1.8 % slower with it; -DK1S_PREFETCH=2 turns it on there)
  static_assert(UN == 1 || UN == 2 || UN == 4, "columns per loop iteration");
  const int lane_g = lane & (g - 1), grp = lane >> g_log2, per_warp = 32 >> g_log2;
  constexpr int RN = 2 * g;
  // segments are g words apart modulo the 32 banks: the groups of a warp read their rings in the same instruction, each
  // through a window of g consecutive words -- without the skew lanes of different groups would hit the same bank
  // (ncu at 8 lanes per pair: 18 % of the shared-memory wavefronts were conflicts)
  unsigned* gring = ring + grp * (2 * RN + g);
  const unsigned rd_lane = (unsigned)__cvta_generic_to_shared(gring + (RN - lane_g));
  constexpr bool IMM = GAPSEL != 0;
  constexpr unsigned cB = k1s_bias(K1SGap<GAPSEL>::OE, K1SGap<GAPSEL>::GE);
  const unsigned BB = IMM ? cB * 0x10001u : P.BB;
  const unsigned BBM = IMM ? (cB + (unsigned)(-K1SGap<GAPSEL>::OE)) * 0x10001u : P.BBM;
  const unsigned OE2 = IMM ? ((unsigned)K1SGap<GAPSEL>::OE & 0xffffu) * 0x10001u : P.OE2;
  const unsigned GE2 = IMM ? ((unsigned)K1SGap<GAPSEL>::GE & 0xffffu) * 0x10001u : P.GE2;
  const unsigned pa_lane = (unsigned)__cvta_generic_to_shared(prof_s) + lane * (S::PAIRED ? 8 : 4);
  constexpr unsigned PB_OFF = 32 * S::WA * 128;         // B section: a compile-time offset in every load address
  constexpr unsigned MUL_A = (unsigned)(S::WA * 128), MUL_B = (unsigned)(S::WB * 128) << 16;
  const unsigned r2_addr = (unsigned)__cvta_generic_to_shared(ring2);
  const bool is31 = lane == 31;
  uint2* bnd = MULTI ? P.bnd + (size_t)(blockIdx.x * K1S_WARPS + warp) * (size_t)P.bnd_stride : nullptr;

  auto load_profile = [&](int pass) {
    const uint4* src = reinterpret_cast<const uint4*>(P.prof + (size_t)pass * S::WORDS);
    uint4* dst = reinterpret_cast<uint4*>(prof_s);
    for (int i = threadIdx.x; i < S::WORDS / 4; i += blockDim.x) dst[i] = src[i];
  };
  if (!MULTI) { load_profile(0); __syncthreads(); if (warp >= P.active_warps) return; }

  for (;;) {
    int item;
    if (MULTI) {
      __syncthreads();
      if (threadIdx.x == 0) cta_item = atomicAdd(P.counter, 1u);
      __syncthreads();
      if (cta_item >= (unsigned)P.n_items) break;
      item = (int)cta_item;
    } else {
      item = 0;
      if (lane == 0) item = (int)atomicAdd(P.counter, 1u);
      item = __shfl_sync(0xffffffffu, item, 0);
      if (item >= P.n_items) break;
    }
    // MULTI: P.n_items counts CTA items: first the X items (8 long streams x one stripe), then 8 streams x all stripes
    const int n_x_items = MULTI ? ((P.n_x_streams + K1S_WARPS - 1) / K1S_WARPS) * P.npass : 0;
    const bool x_item = MULTI && item < n_x_items;
    const int x_pass = x_item ? item % P.npass : 0;
    long st_local;
    if (!MULTI) st_local = (long)item * per_warp + grp;
    else if (x_item) st_local = (long)(item / P.npass) * K1S_WARPS + warp;
    else st_local = (long)P.n_x_streams + (long)(item - n_x_items) * K1S_WARPS + warp;
    const bool valid = x_item ? st_local < P.n_x_streams : st_local < P.n_streams;
    const long st = P.stream_begin + st_local;
    const int ncols = valid ? P.stream_ncols[st] : 0;
    const unsigned short* cp = P.scols + (valid ? P.stream_off[st] : 0);
    int maxcols = ncols;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) maxcols = max(maxcols, __shfl_xor_sync(0xffffffffu, maxcols, o));
    // column steps of this item, rounded up to the unroll factor: a step past the end works on the padding code
    // (score -128), which cannot raise a maximum; its hand-over column is past the end and is not written
    const int nsteps = maxcols > 0 ? (maxcols + g - 1 + UN - 1) / UN * UN : 0;

    const int pass_begin = x_item ? x_pass : 0, pass_end = x_item ? x_pass + 1 : (MULTI ? P.npass : 1);
    for (int pass = pass_begin; pass < pass_end; ++pass) {
      if (MULTI) { __syncthreads(); load_profile(pass); __syncthreads(); }
      const bool first = (pass == 0), last = !MULTI || (pass == P.npass - 1);
      // hand-over rows: own per-warp row (all stripes on this warp), or the stream's ping-pong rows (X item: two rows of
      // ncols columns, read from row (pass + 1) & 1, written to row pass & 1).  One base pointer is kept; the row offsets
      // and the progress words of the X item (stripe p publishes into word p, reads word p - 1) are derived where they
      // are used -- once per 32 steps -- so that they do not occupy registers in the cell loop.
      const bool xv = MULTI && x_item && valid;
      uint2* rows = xv ? P.x_bnd + 2 * (size_t)P.stream_off[st] : bnd;
      unsigned* prog = xv ? P.x_progress + (size_t)st_local * P.npass + pass : nullptr;
      const bool has_in = xv && !first, has_out = xv && !last;
      auto bnd_in = [&]() -> const uint2* { return rows + (xv ? (size_t)((pass + 1) & 1) * (size_t)ncols : 0); };
      auto bnd_out = [&]() -> uint2* { return rows + (xv ? (size_t)(pass & 1) * (size_t)ncols : 0); };
      auto wait_cols = [&](int upto) {                 // X item: the stripe above has published columns [0, upto)
        if (has_in) {
          if (lane == 0) {
            const unsigned need = (unsigned)min(upto, ncols);
            unsigned spins = 0;
            while (k1s_ld_acquire(prog - 1) < need) {
              __nanosleep(100);
              if (++spins > K1S_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; }
            }
          }
          __syncwarp();
        }
      };
      // MULTI: the warp writes the hand-over columns of steps [sb, se) (lane 31 left them in ring3) in one go
      auto flush_bnd = [&](int sb, int se) {
        const int j = sb + lane - 31;                  // lane 31 worked on column step - 31
        if (sb + lane < se && j >= 0 && j < ncols) bnd_out()[j] = ring3[lane];
      };
      unsigned H[R], G[R];
#pragma unroll
      for (int k = 0; k < R; ++k) { H[k] = BB; G[k] = BBM; }
      unsigned out_h = BB, out_v = BBM, diag_in = BB, best = BB;
      int pair_rel = -1;                           // pair of the stream being scored (relative to the stream's first pair)
      __syncwarp();
      for (int i = lane; i < K1S_RING_WORDS; i += 32) ring[i] = 0;
      auto ring_word = [](unsigned c) { return (c & 0x1f1fu) | ((c & 0x80u) << 24); };
      unsigned pre = (lane_g < ncols) ? (unsigned)cp[lane_g] : 0u;
      uint2 pre_b = make_uint2(BB, BBM);
      if (MULTI && !first) {
        wait_cols(32);
        if (lane < ncols) pre_b = __ldcg(bnd_in() + lane);
      }
      __syncwarp();

      // where this lane's maxima go: slot of the stream's first pair and the distance to the next pair's, in registers -- the
      // flush (once per pair) then touches no kernel parameter (its constant-bank loads sat in the column loop's body).
      // A/B per plan on the C2 database (profiles/r2y_ab_slot_ptr.txt, r2w2_ab_nested_x_slot.txt): 1-3 % faster for the multi-stripe
      // kernels except R = 26 and 28 (1-2.5 % slower), within +-0.7 % for the single-stripe ones except g x R = 32 x 32 (1.8 % slower).
      // Also tried there: the four IDP byte selectors in registers instead of re-materialised uniform registers (4 UMOV per
      // column step): 2-4 % SLOWER in the multi-stripe kernels, not kept.
#ifndef K1S_SLOT_PTR
#define K1S_SLOT_PTR (-1)
#endif
      constexpr bool SLOTP = K1S_SLOT_PTR >= 0 ? K1S_SLOT_PTR != 0 : (MULTI ? (R != 26 && R != 28) : !(R == 32 && g == 32));
      unsigned* slot0 = nullptr; int slot_stride = 0;
      if (SLOTP && valid) {
        const size_t p0 = (size_t)P.stream_pair0[st];
        if (MULTI && x_item) { slot0 = P.x_raw + (((p0 * P.npass + pass) << 5) + lane); slot_stride = P.npass << 5; }
        else { slot0 = P.raw + ((p0 << g_log2) + lane_g); slot_stride = 1 << g_log2; }
      }
      auto flush = [&]() {
        if (pair_rel >= 0) {
          if (SLOTP) {
            unsigned* slot = slot0 + (size_t)pair_rel * (size_t)slot_stride;
            if (MULTI && x_item) *slot = best;
            else *slot = (MULTI && !first) ? __vmaxu2(*slot, best) : best;
          } else {
            const size_t pair_idx = (size_t)P.stream_pair0[st] + (size_t)pair_rel;
            if (MULTI && x_item) {
              P.x_raw[((pair_idx * P.npass + pass) << 5) + lane] = best;
            } else {
              unsigned* slot = P.raw + (pair_idx << g_log2) + lane_g;
              *slot = (MULTI && !first) ? __vmaxu2(*slot, best) : best;
            }
          }
        }
      };

      unsigned cw_next = 0;
      // every g steps: publish the prefetched block, fetch the next
      auto boundary = [&](const int s0) {
          __syncwarp();
          { const int p = (s0 & (RN - 1)) + lane_g; const unsigned w = ring_word(pre); gring[p] = w; gring[p + RN] = w; }
          const int c = s0 + g + lane_g;
          pre = (c < ncols) ? (unsigned)cp[c] : 0u;
          if (MULTI) {
            if (!last && s0 > 0) flush_bnd(s0 - 32, s0);
            ring2[lane] = pre_b;
            pre_b = make_uint2(BB, BBM);
            if (!first) {
              wait_cols(s0 + 2 * g);
              if (c < ncols) pre_b = __ldcg(bnd_in() + c);
            }
          }
          __syncwarp();
          // columns [0, s0 - 31) of this stripe's bottom row are in memory (written by the whole warp above, ordered by the __syncwarp)
          if (MULTI && has_out && is31 && s0 > 31) k1s_st_release(prog, (unsigned)min(s0 - 31, ncols));
          if (PF) cw_next = lds32_ordered(rd_lane + ((unsigned)(s0 & (RN - 1)) << 2));   // the word prefetched for this step predates the publish
      };
      auto step = [&](const int s) {
          const unsigned cw = PF ? cw_next : lds32_ordered(rd_lane + ((unsigned)(s & (RN - 1)) << 2));
          unsigned h_up, v_in;
          if (MULTI) {                              // first lane: the stripe above (constants in the first stripe), via the shuffle's default
            unsigned bx, by;
            lds64_ordered(r2_addr + ((unsigned)(s & 31) << 3), bx, by);
            h_up = shfl_up1_or<g>(out_h, bx);
            v_in = shfl_up1_or<g>(out_v, by);
          } else {
            h_up = shfl_up1_or<g>(out_h, BB);       // first lane of the group: row above the stripe = (0, no gap)
            v_in = shfl_up1_or<g>(out_v, BBM);
          }
          if ((int)cw < 0) {                        // first column of a new pair: flush, reset the rows
            flush();
            ++pair_rel; best = BB; diag_in = BB;
#pragma unroll
            for (int k = 0; k < R; ++k) { H[k] = BB; G[k] = BBM; }
          }
          // pa_lane + code(low target) * WA * 128, pa_lane + code(high target) * WB * 128 (+ PB_OFF in the load offsets)
          const unsigned aoff = mul_bytes01(MUL_A, cw, pa_lane);
          const unsigned boff = mul_bytes01(MUL_B, cw, pa_lane);
          // Cell update.  G and V are kept SHIFTED by -oe (Gs = G - oe, Vs = V - oe), which turns
          //   G' = max(G + ge, H + oe)  into  Gs' = max(Gs + ge, H)          (no separate H + oe add)
          //   H  = max(0, t, G, V)      into  H   = max(max3(Gs, Vs, 0 - oe) + oe, t)
          // i.e. per cell pair: 2 IDP (FMA pipe) + VIMNMX3 + 3 VIADDMNMX + half a VIMNMX3 (ALU pipe).
          // t of row k+1 is formed from H[k] of the previous column BEFORE H[k] is overwritten.
          unsigned V = v_in;
          unsigned wA, wB, wA2 = 0, wB2 = 0, hprev = BB;
          if (S::PAIRED) { lds64(aoff, wA, wA2); lds64(boff + PB_OFF, wB, wB2); }
          else { wA = lds32(aoff); wB = lds32(boff + PB_OFF); }
          unsigned t = add_hi_lo(wB, add_byte(wA, 1u, diag_in));
          diag_in = h_up;
#pragma unroll
          for (int k = 0; k < R; ++k) {
            unsigned tn = 0;
            if (k + 1 < R) {
              const int k1 = k + 1;
              if (S::PAIRED) {
                if ((k1 & 7) == 0) lds64(aoff + (k1 >> 3) * 256, wA, wA2); else if ((k1 & 3) == 0) wA = wA2;
                if ((k1 & 3) == 0) lds64(boff + PB_OFF + (k1 >> 2) * 256, wB, wB2); else if ((k1 & 1) == 0) wB = wB2;
              } else {
                if ((k1 & 3) == 0) wA = lds32(aoff + (k1 >> 2) * 128);
                if ((k1 & 1) == 0) wB = lds32(boff + PB_OFF + (k1 >> 1) * 128);
              }
              tn = add_byte(wA, 1u << (8 * (k1 & 3)), H[k]);
              tn = (k1 & 1) ? add_hi_hi(wB, tn) : add_hi_lo(wB, tn);
            }
            const unsigned ef = __vimax3_u16x2(G[k], BBM, V);
            const unsigned h = __viaddmax_u16x2(ef, OE2, t);
            H[k] = h;
            G[k] = __viaddmax_u16x2(G[k], GE2, h);
            V = __viaddmax_u16x2(V, GE2, h);
            if (k & 1) best = __vimax3_u16x2(best, hprev, h); else hprev = h;
            t = tn;
          }
          if (R & 1) best = __vmaxu2(best, hprev);           // odd R: the last row has no partner
          out_h = H[R - 1];
          out_v = V;
          if (MULTI && !last && is31) ring3[s & 31] = make_uint2(out_h, out_v);
          if (PF) cw_next = lds32_ordered(rd_lane + ((unsigned)((s + 1) & (RN - 1)) << 2));   // next step's ring word (re-loaded after a publish, above)
      };
      // Multi-stripe kernels: blocks of 32 steps without the per-step block test, for the rows-per-lane values where that is
      // faster: every multi-stripe plan timed on the C2 database under the four combinations of this and the slot pointer below
      // (profiles/r2w_ab_nested_per_R.txt, r2w2_ab_nested_x_slot.txt): nested for every R but 26, 27, 31.  -DK1S_NESTED=0 / 1 forces one form.
#ifndef K1S_NESTED
#define K1S_NESTED (-1)
#endif
      constexpr bool NEST = MULTI && (K1S_NESTED >= 0 ? K1S_NESTED != 0 : (R != 26 && R != 27 && R != 31));
      if (NEST) {
        for (int s0 = 0; s0 < nsteps; s0 += 32) {
          boundary(s0);
          const int se = min(s0 + 32, nsteps);
#pragma unroll 1
          for (int s = s0; s < se; ++s) step(s);
        }
      } else {
        for (int s0 = 0; s0 < nsteps; s0 += UN) {
          if ((s0 & (g - 1)) == 0) boundary(s0);
#pragma unroll
          for (int u = 0; u < UN; ++u) step(s0 + u);
        }
      }
      if (MULTI && !last && nsteps > 0) {          // the columns of the last (partial) block of steps
        __syncwarp();
        flush_bnd((nsteps - 1) & ~31, nsteps);
        __syncwarp();
        if (has_out && is31) k1s_st_release(prog, (unsigned)ncols);
      }
      flush();
    }
  }
}

// reduce the `slots` per-lane (and, for X pairs, per-stripe) maxima of every pair, un-bias, detect saturation, scatter to file order.
// A pair whose either half reached `limit` sends BOTH members to the 32-bit list (a carry out of the low
// half would have corrupted the high half); long_thr > 0 also sends every target longer than that.
static __global__ void k1s_finalize(const unsigned* __restrict__ raw, long pair_begin, long n_pairs, int slots, long n_sorted,
                                    unsigned B, unsigned limit, const int64_t* __restrict__ order,
                                    const int64_t* __restrict__ plain_off, int long_thr,
                                    int32_t* __restrict__ scores_file, int* __restrict__ rescore_list,
                                    unsigned* __restrict__ rescore_count, unsigned long long* __restrict__ rescore_residues,
                                    int slot_rows, int lq, int* __restrict__ row_bound) {
  for (long p = pair_begin + blockIdx.x * (long)blockDim.x + threadIdx.x; p < n_pairs; p += (long)gridDim.x * blockDim.x) {
    const unsigned* r = raw + (size_t)p * (size_t)slots;
    // maximum of each half and the LAST slot that holds it.  Slot i of a launch whose slots are not merged over the stripes
    // covers the query rows [i * slot_rows, (i + 1) * slot_rows): no row below that slot's last one holds the target's score,
    // which is all the forward coordinate pass of a reported hit has to look at (swgpu_topk.cu).
    unsigned lo = 0, hi = 0; int ilo = 0, ihi = 0;
    for (int i = 0; i < slots; ++i) {
      const unsigned v = r[i], l = v & 0xffffu, h = v >> 16;
      if (l >= lo) { lo = l; ilo = i; }
      if (h >= hi) { hi = h; ihi = i; }
    }
    const bool sat = (lo >= limit) || (hi >= limit);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const long s = 2 * p + h;
      if (s >= n_sorted) continue;
      const bool force = long_thr > 0 && (plain_off[s + 1] - plain_off[s]) > long_thr;
      int bound = lq - 1;
      if (sat || force) {
        const unsigned slot = atomicAdd(rescore_count, 1u);
        rescore_list[slot] = (int)s;
        atomicAdd(rescore_residues, (unsigned long long)(plain_off[s + 1] - plain_off[s]));      // statistics: cells of the 32-bit pass
      } else {
        scores_file[order[s]] = (int32_t)((h ? hi : lo) - B);
        if (slot_rows > 0) bound = min(bound, ((h ? ihi : ilo) + 1) * slot_rows - 1);
      }
      if (row_bound) row_bound[s] = bound;
    }
  }
}

// Launchers of one gap specialisation (sw_k1s_inst.cu is compiled once per GAPSEL and per ptxas option set; the
// instantiations live in those objects, nowhere else).  cudaErrorInvalidValue for a plan that has no kernel.
template <int GAPSEL>
cudaError_t launch_k1s_o3(int g, int R, bool multi, const K1SParams& P, cudaStream_t stream, int grid, int device);
template <int GAPSEL>
cudaError_t launch_k1s_o1(int R, const K1SParams& P, cudaStream_t stream, int grid, int device);     // multi-stripe, R = 24..32

}  // namespace swk
