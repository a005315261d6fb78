// sw_k3x.cuh -- K3X: 32-bit intra-sequence WAVEFRONT kernel, one WARP per stripe of query rows, any number of
// warps (from any CTA) on the same target at the same time.
//
// Same recurrence as k3_intra (sw_kernels.cuh; reference src/alignment.c:117-176).  A single (query, target)
// pair is a chain of Lt dependent column steps per stripe of query rows; a 35 213 x 35 213 pair has 138 stripes
// of 256 rows.  Here every stripe is a work item of its own:
//   * items = (target, stripe), enumerated stripe-ascending per target (stripe-major inside a generation of K3X_SLOTS
//     targets) and pulled from one atomic queue, so the producer of an item's input row always has a smaller item
//     index: it is already running (or done) when the consumer starts -- no deadlock whatever the number of resident warps;
//   * stripe s-1 -> stripe s: the bottom row goes through a full-length row in global memory (two rows per target
//     slot, ping-pong by stripe parity).  Every column is ONE aligned 16-byte store (H, tag, V, tag) by the
//     producer's lane 31 and one 16-byte volatile load by the consumer, which simply re-reads until both tags are
//     the producer's (two item-specific tag words, see k3x_tag_a / k3x_tag_b: neither an older stripe of the same parity
//     nor an earlier launch nor a torn record can be mistaken for it).  The flag travels with the data: no fence, no separate progress word
//     (the first version published a counter with st.release.gpu every 32 columns and polled it with
//     ld.acquire.gpu: the MEMBAR and the L1 invalidate were the two hottest instructions in ncu);
//   * target slots (rows, per-stripe results) are reused round-robin: a warp starts an item of
//     target t only after target t - K3X_SLOTS has completed (done counter per slot);
//   * the last warp to finish a target reduces the per-stripe results and writes the target's output.
// Inside a warp the scheme is k3_intra's: lane l owns 8 rows, the target streams through the lanes by
// __shfl_up_sync.  The cell update is the latency-optimised form: with Hx = max(0, t, G) (independent of V)
//   V' = max(V + ge, Hx + oe)     -- the only row-to-row dependency, one VIADDMNMX
//   H  = max(Hx, V)                 G' = max(G + ge, H + oe)
// evaluated on gap states shifted by -oe so that every operation has one add (4 VIADDMNMX + 1 IDP per cell, see the kernel).
// Placement of the items on the machine (CTA levels per SM), the block-ahead fetches and the warp-uniform waits are described
// where they are implemented; DESIGN.md section 5 has the measurements that led to each of them.
// Used for (a) the coordinate passes of top-k (TRACK: first maximum in the reference's scan order)
// and (b) 32-bit rescoring of long targets.
#pragma once
#include "sw_kernels.cuh"

namespace swk {

constexpr int K3X_WARPS = 4;                   // 36 KB of static shared memory per CTA (a lane-private profile of 8 KB per warp)
#ifndef K3X_CTAS
#define K3X_CTAS 4
#endif
constexpr int K3X_CTAS_PER_SM = K3X_CTAS;      // 16 warps per SM (at most: the launch asks the occupancy calculator)
constexpr int K3X_R = 8;                       // rows per lane
constexpr int K3X_ROWS = 32 * K3X_R;           // 256 rows per stripe
constexpr int K3X_PREFETCH_AT = 24;            // step of a 32-step block at which the next block's input columns are requested
constexpr int K3X_SLOTS = 128;                 // targets in flight
constexpr unsigned K3X_SPIN_LIMIT = 1u << 24;  // polls (with back-off) before the error flag is raised instead of hanging

struct K3XParams {
  const int*      sidx_list;    // targets (sorted positions); nullptr -> target index itself
  int             n_targets;
  const unsigned* work_count;   // if non-null the target count is read from the device
  unsigned*       counter;      // item queue
  const int8_t*   plain;
  const int64_t*  plain_off;
  const int8_t*   query;
  int             lq;
  int             nstripes_max; // ceil(lq / K3X_ROWS)
  const int8_t*   table;
  int             oe, ge;
  int4*           rows;         // [K3X_SLOTS][2][row_stride] (H, tag, V, tag)
  long long       row_stride;
  unsigned        epoch;        // launch counter of this slot set: makes tags unique across launches
  unsigned*       done;         // [K3X_SLOTS]: items finished on this slot since the launch
  int4*           partial;      // [K3X_SLOTS][nstripes_max]: (best, j, i, -) per stripe
  const int64_t*  order;        // !TRACK: scores_file[order[sidx]] = score
  int32_t*        scores_file;
  int4*           coords;       // TRACK: per target (q_start, q_end, t_start, t_end)
  int*            best_out;     // TRACK phase 0: score found
  int             phase;        // TRACK: 0 = forward (end cell), 1 = reversed prefixes (start cell)
  // TRACK, early stop: the hit's score S is known (hit_keys[t] >> 32), so a pass is over as soon as every column up to the
  // first column that holds a cell with H == S has been evaluated -- later columns cannot win the tie-break (SURVEY.md A.7:
  // smallest target position first).  stop_col[t] (initialised to a huge value) is lowered with atomicMin by whichever
  // stripe sees S; a warp leaves when all its remaining columns lie beyond it, a lane stops waiting for an input column
  // beyond it (it computes on zeros instead: the recurrence is monotone in its boundary values, so such cells can only come out
  // LOWER than the truth, never above S), and candidates beyond it are ignored.  The reversed pass
  // typically ends after the ~100 columns of the alignment itself instead of the whole prefix.
  const unsigned long long* hit_keys;   // TRACK: sorted keys of the hits (score << 32 | ~entry); nullptr = no early stop
  int*            stop_col;     // TRACK: [n_targets] for this phase
  const int*      row_bound;    // TRACK phase 0: by sorted position, the last query row that can hold the hit's score (nullptr: none)
  unsigned*       error_flag;
  unsigned*       sm_slots;     // [256], zero at launch: CTAs that have started on each SM (a CTA's level = its arrival number there)
  int             level_ctas;   // CTAs per level = number of SMs
  int             min_len;      // !TRACK: targets shorter than this are left to k3_intra
#ifdef K3X_TRACE
  int4*           trace;        // profiling build only: per item (begin, end, time spent waiting for input; 0.1 us units, SM id)
#endif
};

__device__ __forceinline__ int4 ld_volatile_v4(const int4* p) {
  int4 v;
  asm volatile("ld.volatile.global.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_v4(int4* p, int a, int b, int c, int d) {
  asm volatile("st.volatile.global.v4.s32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// Hand-over tags of a column: TWO words, one in each 8-byte half of the 16-byte (H, tagA, V, tagB) record, both specific
// to the item, so a half that is stale (an older stripe of the same parity, an earlier launch) or torn cannot pass:
//   tagA = item + 1                     (items are numbered below 2^32 - 1, checked by the host: never 0 = a zeroed row)
//   tagB = tagA ^ (epoch * odd + 1)     (injective in the launch epoch for a given item)
// Equality on both is exact -- no hash, no collision (round 1 used one 31-bit hash in both words).
__device__ __forceinline__ int k3x_tag_a(long long item) { return (int)(unsigned)(item + 1); }
__device__ __forceinline__ int k3x_tag_b(long long item, unsigned epoch) { return (int)((unsigned)(item + 1) ^ (epoch * 0x9E3779B9u + 1u)); }
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ int ld_volatile_s32(const int* p) {
  int v;
  asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ bool k3x_better(int s, int j, int i, int bs, int bj, int bi) {
  return s > bs || (s == bs && (j < bj || (j == bj && i < bi)));
}

template <bool TRACK>
__global__ void __launch_bounds__(K3X_WARPS * 32) k3x(const K3XParams P) {
  constexpr int R3 = K3X_R;
  static_assert(R3 == 8, "profile layout: two int8x4 words per (code, lane)");
  __shared__ int8_t tab[32 * 32];
  __shared__ uint2 pl_all[K3X_WARPS][32 * 32];             // [code][lane]: int8 x 8 = s(own row k, code), k = 0..7
  // Placement.  The grid is the whole machine (K3X_CTAS_PER_SM CTAs on every SM) but a job may have far fewer items than
  // warps (ONE 35 213 x 35 213 pair is 138 stripes), and the queue hands consecutive items to whichever warps ask first:
  // left alone, 138 items landed on 31 SMs, up to four per scheduler, and the whole chain ran at the pace of the most
  // crowded scheduler.  So CTAs are numbered per SM in arrival order (their level); level L only takes part when the
  // levels below cannot give every item a warp of its own, and waits a moment so that the lower levels ask first.
  // Items still come from the one queue, in order, whoever asks: the no-deadlock argument above is untouched.
  __shared__ int cta_level;
  if (threadIdx.x == 0) {
    unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    cta_level = (int)atomicAdd(P.sm_slots + (smid & 255u), 1u);
  }
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  {
    long long items = (long long)(P.work_count ? (int)*P.work_count : P.n_targets) * P.nstripes_max;
    if (TRACK && P.phase == 1 && P.hit_keys) {
      // The reversed pass stops at the alignment's start, so most of a hit's stripes leave after one block; what stays busy
      // is about one stripe per 256 rows of the alignment, i.e. per ~1 024 score units (4 per aligned row, a low estimate
      // of the score density), plus two.  The levels are sized for THAT: with 100 unrelated hits and one full-length
      // alignment (C4) the pass is a single long chain that wants a scheduler per stripe, with 100 long alignments (C3) it is
      // throughput work that wants every warp.
      __shared__ unsigned long long busy_est;
      if (threadIdx.x == 0) busy_est = 0ull;
      __syncthreads();
      unsigned long long mine = 0;
      const int nt = P.n_targets;
      for (int h = threadIdx.x; h < nt; h += blockDim.x) mine += (unsigned long long)((unsigned)(P.hit_keys[h] >> 32) >> 10) + 2ull;
      atomicAdd(&busy_est, mine);
      __syncthreads();
      items = min(items, (long long)busy_est);
    }
    const long long per_level = (long long)P.level_ctas * K3X_WARPS;
    if (items <= (long long)cta_level * per_level) return;
    if (cta_level > 0 && items < (long long)gridDim.x * K3X_WARPS) __nanosleep(3000u * (unsigned)min(cta_level, 8));
  }
  // lane id through a volatile shared-memory round trip: ptxas otherwise re-derives it from the special register with an
  // S2R (tens of cycles) on the critical path of the step's first-lane select; a volatile load cannot be re-materialised
  __shared__ unsigned lane_ids[K3X_WARPS * 32];
  lane_ids[threadIdx.x] = threadIdx.x & 31;
  unsigned lane_v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(lane_v) : "r"((unsigned)__cvta_generic_to_shared(&lane_ids[threadIdx.x])) : "memory");
  const int warp = threadIdx.x >> 5, lane = (int)lane_v;
  uint2* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 8;
  // per-warp rings: the profile byte offset (code * 256) of 128 consecutive target columns -- written ONE BLOCK of 32
  // columns ahead, so a lane can fetch its profile words a step early and the residue load from global memory is a
  // whole block off the critical path -- and the (H, Vs) of the stripe above for the current block of 32 columns.
  __shared__ unsigned ringc_all[K3X_WARPS][128];
  __shared__ int2 ringb_all[K3X_WARPS][32];
  unsigned* rc = ringc_all[warp];
  int2* rb = ringb_all[warp];
  const unsigned rc_addr = (unsigned)__cvta_generic_to_shared(rc);
  const unsigned rb_addr = (unsigned)__cvta_generic_to_shared(rb);
  const int n_targets = P.work_count ? (int)*P.work_count : P.n_targets;
  const long long n_items = (long long)n_targets * P.nstripes_max;
  // opaque copies: the penalties stay in registers (as uniform-register operands they were re-loaded from the
  // constant bank inside every step)
  int oe, ge;
  asm volatile("mov.s32 %0, %1;" : "=r"(oe) : "r"(P.oe));
  asm volatile("mov.s32 %0, %1;" : "=r"(ge) : "r"(P.ge));

#ifdef K3X_TRACE
  unsigned long long tr_entry; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_entry));
#endif
  for (;;) {
    long long item = 0;
    if (lane == 0) item = (long long)atomicAdd(P.counter, 1u);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= n_items) break;
    // item -> (target, stripe): targets in generations of K3X_SLOTS, and inside a generation STRIPE-major (stripe 0 of all its
    // targets, then stripe 1, ...): the warps of the machine then work on the leading stripes of many targets, not on all
    // 138 stripes of a few, most of which would only wait for their turn.  The producer of (t, stripe) is (t, stripe - 1),
    // n_in_gen items earlier: still a smaller item number.
    const long long per_gen = (long long)K3X_SLOTS * P.nstripes_max;
    const int gen = (int)(item / per_gen);
    const int in_gen = (int)(item - (long long)gen * per_gen);
    const int n_in_gen = min(K3X_SLOTS, n_targets - gen * K3X_SLOTS);
    const int stripe = in_gen / n_in_gen;
    const int slot = in_gen - stripe * n_in_gen;
    const int t = gen * K3X_SLOTS + slot;
    // the slot's previous user (target t - K3X_SLOTS) must be completely done before its rows are touched
    if (gen > 0) {
      if (lane == 0) {
        const unsigned need = (unsigned)gen * (unsigned)P.nstripes_max;
        unsigned spins = 0;
        while (ld_acquire_u32(P.done + slot) < need) {
          __nanosleep(200);
          if (++spins > K3X_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; }
        }
      }
      __syncwarp();
    }
    const int sidx = P.sidx_list ? P.sidx_list[t] : t;
    const int8_t* tp = P.plain + P.plain_off[sidx];
    int lt = (int)(P.plain_off[sidx + 1] - P.plain_off[sidx]);
    int lq = P.lq;
    const int8_t* qp = P.query;
    int qstep = 1, tstep = 1;
    bool skip = false;
    if (TRACK && P.phase == 0 && P.row_bound) lq = min(lq, P.row_bound[sidx] + 1);   // rows below hold lower scores: not needed
    if (TRACK && P.phase == 1) {
      const int4 c = P.coords[t];
      if (c.y < 0) skip = true;
      else { lq = c.y + 1; lt = c.w + 1; qp += c.y; tp += c.w; qstep = -1; tstep = -1; }
    }
    if (!TRACK && lt < P.min_len) skip = true;             // this target belongs to k3_intra
    const int nstripes = skip ? 0 : (lq + K3X_ROWS - 1) / K3X_ROWS;
    int best = 0, bj = 0x7fffffff, bi = 0x7fffffff;
    const bool early = TRACK && P.hit_keys != nullptr;
    const int known = early ? (int)(P.hit_keys[t] >> 32) : 0x7fffffff;      // the hit's score
    int* stop_p = early ? P.stop_col + t : nullptr;
    int stop = 0x7fffffff;                                  // this lane's view of stop_col[t] (only ever decreases)

    if (stripe < nstripes) {
      const int row0 = stripe * K3X_ROWS + lane * R3;
      {
        int qc[R3];
#pragma unroll
        for (int b = 0; b < R3; ++b) { const int i = row0 + b; qc[b] = (i < lq) ? ((int)qp[(long)i * qstep] & 31) : -1; }
        __syncwarp();                                        // the previous item of this warp is done with pl and the rings
        for (int c = 0; c < 32; ++c) {
          unsigned w[2] = {0u, 0u};
#pragma unroll
          for (int b = 0; b < R3; ++b) {
            const int sc = (c == 0 || qc[b] < 0) ? -128 : (int)tab[qc[b] * 32 + c];
            w[b >> 2] |= (unsigned)(sc & 0xff) << (8 * (b & 3));
          }
          pl[c * 32 + lane] = make_uint2(w[0], w[1]);
        }
      }
      const bool has_in = stripe > 0, has_out = stripe + 1 < nstripes;
      const int4* row_in = P.rows + ((size_t)slot * 2 + ((stripe + 1) & 1)) * (size_t)P.row_stride;
      int4* row_out = P.rows + ((size_t)slot * 2 + (stripe & 1)) * (size_t)P.row_stride;
      const int tag_in_a = k3x_tag_a(item - n_in_gen), tag_in_b = k3x_tag_b(item - n_in_gen, P.epoch);
      const int tag_out_a = k3x_tag_a(item), tag_out_b = k3x_tag_b(item, P.epoch);
      // State of the lane's 8 rows.  Gs = G - oe and Vs = V - oe are kept SHIFTED, which leaves one add per operation:
      //   Hx = max(0, t, G)          = max(Gs + oe, t, 0)     VIADDMNMX.RELU   (independent of V: off the row-to-row chain)
      //   H  = max(Hx, V)            = max(Vs + oe, Hx)       VIADDMNMX
      //   V' = max(V + ge, Hx + oe)  : Vs' = max(Vs + ge, Hx) VIADDMNMX        (the chain: one operation per row; oe <= ge)
      //   G' = max(G + ge, H + oe)   : Gs' = max(Gs + ge, H)  VIADDMNMX
      // i.e. 4 ALU-pipe operations + 1 IDP per cell.  A boundary value of 0 for Gs / Vs stands for G = V = oe <= 0: as good as
      // "no gap" (H >= 0 everywhere).  The hand-over rows carry (H, Vs).
      int H[R3], Gs[R3];
#pragma unroll
      for (int k = 0; k < R3; ++k) { H[k] = 0; Gs[k] = 0; }
      int diag_in = 0, out_h = 0, out_v = 0;
      // code ring: columns < 0 and >= lt are the padding code 0 (profile -128); columns [0, 32) now, [32, 64) in `creg`
      auto code_of = [&](int j) -> unsigned { return (j < lt) ? ((unsigned)tp[(long)j * tstep] & 31u) * 256u : 0u; };
      rc[lane] = code_of(lane); rc[lane + 32] = 0; rc[lane + 64] = 0; rc[lane + 96] = 0;
      unsigned creg = code_of(32 + lane);
      __syncwarp();
      const int nsteps = lt + 31;
      int4* pcol = row_out - lane;                          // &row_out[s - lane], advanced every step
#ifdef K3X_TRACE
      unsigned long long tr_begin, tr_wait = 0, tr_m0 = 0, tr_m1 = 0, tr_m2 = 0; unsigned tr_miss = 0; int tr_b[16]; for (int q = 0; q < 16; ++q) tr_b[q] = 0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_begin));
#endif
      // software pipeline: the profile words of step s are fetched during step s - 1, their ring entry during step s - 2
      unsigned w0, w1, cn;
      int4 bpre = make_int4(0, 0, 0, 0);                    // tag 0 is nobody's: the first block polls
      lds64(pl_addr + lds32_ordered(rc_addr + (((unsigned)(0 - lane) & 127u) << 2)), w0, w1);
      cn = lds32_ordered(rc_addr + (((unsigned)(1 - lane) & 127u) << 2));
      for (int s0 = 0; s0 < nsteps; s0 += 32) {
        {
          const int j = s0 + lane;
          int4 b = make_int4(0, 0, 0, 0);
          if (early) {
            stop = min(stop, ld_volatile_s32(stop_p));
            stop = __reduce_min_sync(0xffffffffu, stop);   // one view per warp: the exit below must be uniform
            if (s0 - 31 > stop) break;                     // every column this warp still has lies beyond the stop column
          }
#ifdef K3X_TRACE
          unsigned long long tr_w0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_w0));
#endif
          const bool want = has_in && j < lt;                // my column of the stripe above
          if (want) b = bpre;                                // requested during the previous block (nothing before the first one)
          bool ok = !want || (b.y == tag_in_a && b.w == tag_in_b);
          if (!__all_sync(0xffffffffu, ok)) {
            // Not there yet.  Both waits below are loops the WHOLE warp leaves together: with per-lane exits the lanes came out
            // of the wait on different paths and ran the block that followed un-converged, every shuffle on its slow path
            // (measured: 12.5 us for a stripe's first 32 steps instead of 2.5, i.e. 15 us per stripe for a long pair to get going).
            // First the block's LAST column, one 16-byte record read by all lanes, with a growing pause: while a long pair starts
            // up every stripe below the front waits for its first block, on the same few lines of the two rows.  Then every lane
            // re-reads its own column until it carries its tags (normally the first read; the order in which the producer's
            // stores become visible is not relied upon).
            {
              const int jl = min(s0 + 31, lt - 1);
              unsigned spins = 0;
              for (;;) {
                const int4 l = ld_volatile_v4(&row_in[jl]);
                bool seen = (l.y == tag_in_a && l.w == tag_in_b);
                if (early) { stop = min(stop, ld_volatile_s32(stop_p)); seen |= jl > stop; }   // may never come: decided per lane below
                if (++spins > K3X_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); seen = true; }
                if (__any_sync(0xffffffffu, seen)) break;
                __nanosleep(64u << min(spins, 4u));
              }
            }
#ifdef K3X_TRACE
            if (s0 > 0) ++tr_miss;
#endif
            unsigned spins = 0;
            for (;;) {
              if (!ok) {
                b = ld_volatile_v4(&row_in[j]);
                ok = (b.y == tag_in_a && b.w == tag_in_b);
                if (!ok && early) {                        // the producer may have left already: columns beyond the stop column never come
                  stop = min(stop, ld_volatile_s32(stop_p));
                  if (j > stop) { b = make_int4(0, 0, 0, 0); ok = true; }
                }
              }
              if (__all_sync(0xffffffffu, ok)) break;
              __nanosleep(40);
              if (++spins > K3X_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; }
            }
          }
          __syncwarp();                                    // the previous block's rb has been read by the first lane
#ifdef K3X_TRACE
          { unsigned long long tr_w1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_w1)); tr_wait += tr_w1 - tr_w0;
            if (s0 == 0) tr_m0 = tr_w1; if (s0 == 64) tr_m1 = tr_w1; if (s0 == 1024) tr_m2 = tr_w1;
#pragma unroll
            for (int q = 0; q < 8; ++q) if (s0 == 32 * q) { tr_b[2 * q] = (int)((tr_w0 - tr_begin) / 10); tr_b[2 * q + 1] = (int)((tr_w1 - tr_begin) / 10); } }
#endif
          rb[lane] = make_int2(b.x, b.z);
          rc[(s0 + 32 + lane) & 127] = creg;               // the NEXT block's codes (this block's went in a block ago)
          creg = code_of(s0 + 64 + lane);                  // in flight during the 32 steps below
          __syncwarp();
        }
        unsigned bx, by;
        lds64_ordered(rb_addr, bx, by);                     // the first lane's boundary pair of this block's first column
        const int send = min(32, nsteps - s0);
        auto step = [&](const int u) {
          const int s = s0 + u;
          const int j = s - lane;                           // this lane's column
          int h_up = __shfl_up_sync(0xffffffffu, out_h, 1);
          int v_in = __shfl_up_sync(0xffffffffu, out_v, 1);
          if (lane == 0) { h_up = (int)bx; v_in = (int)by; }
          // fetches for the next steps (nothing below depends on them)
          unsigned wn0, wn1;
          lds64(pl_addr + cn, wn0, wn1);
          const unsigned cnn = lds32_ordered(rc_addr + (((unsigned)(j + 2) & 127u) << 2));
          lds64_ordered(rb_addr + (((unsigned)(s + 1) & 31u) << 3), bx, by);   // (stale in a block's last step: re-read above)
          // No range test around the cell update: before its first column a lane sees zeros from the lane below and the
          // padding code (H stays 0); past its last column it computes on the padding code, which can only produce values
          // below those already seen.  Only the hand-over store and the tracked maximum are guarded.
          int d = diag_in;
          diag_in = h_up;
          int Vs = v_in;
#pragma unroll
          for (int k = 0; k < R3; ++k) {
            const int tt = (int)add_byte(k < 4 ? w0 : w1, 1u << (8 * (k & 3)), (unsigned)d);
            const int hx = __viaddmax_s32_relu(Gs[k], oe, tt);
            const int h = __viaddmax_s32(Vs, oe, hx);
            Vs = __viaddmax_s32(Vs, ge, hx);                 // the row-to-row chain: one operation
            Gs[k] = __viaddmax_s32(Gs[k], ge, h);
            d = H[k];
            H[k] = h;
          }
          out_h = H[R3 - 1];
          out_v = Vs;
          const bool in_range = (unsigned)j < (unsigned)lt;
          if (TRACK) {
            int cb = __vimax3_s32(H[0], H[1], H[2]);
            cb = __vimax3_s32(cb, H[3], H[4]);
            cb = __vimax3_s32(cb, H[5], H[6]);
            cb = max(cb, H[7]);
            // a lane's columns come in increasing order: a later column wins only with a strictly higher score; within the
            // column the smallest row wins
            if (cb > best && in_range && j <= stop) {
              int kk = R3 - 1;
#pragma unroll
              for (int k = R3 - 2; k >= 0; --k) if (H[k] == cb) kk = k;
              best = cb; bj = j; bi = row0 + kk;
              if (early && cb == known) { atomicMin(stop_p, j); stop = j; }   // nothing right of this column can win any more
            }
          } else {
            best = __vimax3_s32(best, H[0], H[1]);
            best = __vimax3_s32(best, H[2], H[3]);
            best = __vimax3_s32(best, H[4], H[5]);
            best = __vimax3_s32(best, H[6], H[7]);
          }
          if (lane == 31 && has_out && in_range) st_volatile_v4(pcol, out_h, tag_out_a, out_v, tag_out_b);
          ++pcol;
          w0 = wn0; w1 = wn1; cn = cnn;
        };
        const int upre = min(send, K3X_PREFETCH_AT);
        for (int u = 0; u < upre; ++u) step(u);
        // The next block's input columns are requested a few steps before they are needed: a consumer that is not right on
        // its producer's heels finds them complete and pays no load latency at the block boundary (a stripe that failed the
        // check once has fallen behind by the re-poll and succeeds from then on); an incomplete record is re-polled above.
        if (has_in) {
          const int jn = s0 + 32 + lane;
          bpre = make_int4(0, 0, 0, 0);
          if (jn < lt) bpre = ld_volatile_v4(&row_in[jn]);
        }
        for (int u = upre; u < send; ++u) step(u);
      }
      // this stripe's candidate
      if (TRACK) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const int s2 = __shfl_xor_sync(0xffffffffu, best, o);
          const int j2 = __shfl_xor_sync(0xffffffffu, bj, o);
          const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
          if (k3x_better(s2, j2, i2, best, bj, bi)) { best = s2; bj = j2; bi = i2; }
        }
      } else {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
      }
      if (lane == 0) P.partial[(size_t)slot * P.nstripes_max + stripe] = make_int4(best, bj, bi, 0);
#ifdef K3X_TRACE
      if (lane == 0 && P.trace) {
        unsigned long long tr_end; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_end));
        unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        P.trace[6 * item] = make_int4((int)((tr_begin - tr_entry) / 100), (int)((tr_end - tr_entry) / 100), (int)(tr_wait / 100), (int)smid);
        P.trace[6 * item + 1] = make_int4((int)((tr_m0 - tr_entry) / 100), (int)((tr_m1 - tr_entry) / 100), (int)((tr_m2 - tr_entry) / 100), (int)tr_miss);
#pragma unroll
        for (int q = 0; q < 4; ++q) P.trace[6 * item + 2 + q] = make_int4(tr_b[4 * q], tr_b[4 * q + 1], tr_b[4 * q + 2], tr_b[4 * q + 3]);
      }
#endif
    }

    // completion: the last item of this target reduces the per-stripe candidates and writes the result
    unsigned fin = 0;
    if (lane == 0) {
      __threadfence();
      fin = atomicAdd(P.done + slot, 1u) + 1u;
      __threadfence();
    }
    fin = __shfl_sync(0xffffffffu, fin, 0);
    if (fin == (unsigned)(gen + 1) * (unsigned)P.nstripes_max) {
      int rb = 0, rj = 0x7fffffff, ri = 0x7fffffff;
      for (int s = lane; s < nstripes; s += 32) {
        const int4 c = __ldcg(&P.partial[(size_t)slot * P.nstripes_max + s]);
        if (TRACK ? k3x_better(c.x, c.y, c.z, rb, rj, ri) : c.x > rb) { rb = c.x; rj = c.y; ri = c.z; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const int s2 = __shfl_xor_sync(0xffffffffu, rb, o);
        const int j2 = __shfl_xor_sync(0xffffffffu, rj, o);
        const int i2 = __shfl_xor_sync(0xffffffffu, ri, o);
        if (TRACK ? k3x_better(s2, j2, i2, rb, rj, ri) : s2 > rb) { rb = s2; rj = j2; ri = i2; }
      }
      if (lane == 0) {
        if (TRACK) {
          int4 c = P.coords[t];
          if (P.phase == 0) {
            if (rb > 0) { c.y = ri; c.w = rj; } else { c = make_int4(-1, -1, -1, -1); }
            P.best_out[t] = rb;
            P.coords[t] = c;
          } else if (!skip) {
            c.x = c.y - ri; c.z = c.w - rj;
            P.coords[t] = c;
          } else {
            P.coords[t] = make_int4(-1, -1, -1, -1);
          }
        } else if (!skip) {
          P.scores_file[P.order[sidx]] = rb;
        }
      }
    }
  }
}

// CTAs of k3x<TRACK> that fit an SM of `device` (the current device), at most K3X_CTAS_PER_SM: the launch grid is
// sm_count times this, so that every CTA of the grid is resident at once when the GPU is otherwise idle.
template <bool TRACK>
inline int k3x_ctas_per_sm(int device) {
  static int cached[64] = {0};
  if (device >= 0 && device < 64 && cached[device]) return cached[device];
  cudaFuncSetAttribute(k3x<TRACK>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k3x<TRACK>, K3X_WARPS * 32, 0) != cudaSuccess || n < 1) n = 1;
  if (n > K3X_CTAS_PER_SM) n = K3X_CTAS_PER_SM;
  if (device >= 0 && device < 64) cached[device] = n;
  return n;
}

}  // namespace swk
