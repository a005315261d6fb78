// sw_k3x.cuh -- K3X: 32-bit intra-sequence WAVEFRONT kernel, one WARP per stripe of query rows, any number of
// warps (from any CTA) on the same target at the same time.
//
// Same recurrence as k3_intra (sw_kernels.cuh; reference src/alignment.c:117-176).  A single (query, target)
// pair is a chain of Lt dependent column steps per stripe of query rows; a 35 213 x 35 213 pair has 276 stripes
// of 128 rows.  Here every stripe is a work item of its own:
//   * items = (target, stripe), enumerated target-major / stripe-ascending and pulled from one atomic queue, so
//     the producer of an item's input row always has a smaller item index: it is already running (or done) when
//     the consumer starts -- no deadlock whatever the number of resident warps;
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
// Inside a warp the scheme is k3_intra's: lane l owns 4 rows, the target streams through the lanes by
// __shfl_up_sync.  The cell update is the latency-optimised form: with Hx = max(0, t, G) (independent of V)
//   V' = max(V + ge, Hx + oe)     -- the only row-to-row dependency, one VIADDMNMX
//   H  = max(Hx, V)                 G' = max(G + ge, H + oe)
// Used for (a) the coordinate passes of top-k (TRACK: first maximum in the reference's scan order)
// and (b) 32-bit rescoring of long targets.
#pragma once
#include "sw_kernels.cuh"

namespace swk {

constexpr int K3X_WARPS = 8;
constexpr int K3X_R = 4;                       // rows per lane: 128 rows per stripe
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
  int             nstripes_max; // ceil(lq / 128)
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
  unsigned*       error_flag;
  int             min_len;      // !TRACK: targets shorter than this are left to k3_intra
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
  __shared__ int8_t tab[32 * 32];
  __shared__ unsigned pl_all[K3X_WARPS][32 * 32];          // [code][lane], one int8x4 word (R3 == 4)
  static_assert(R3 == 4, "profile layout assumes 4 rows per lane");
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  // lane id through a volatile shared-memory round trip: ptxas otherwise re-derives it from the special register with an
  // S2R (tens of cycles) twice per column step, on the critical path of the step's range test and first-lane select;
  // a volatile load cannot be re-materialised, so the value stays in a register
  __shared__ unsigned lane_ids[K3X_WARPS * 32];
  lane_ids[threadIdx.x] = threadIdx.x & 31;
  unsigned lane_v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(lane_v) : "r"((unsigned)__cvta_generic_to_shared(&lane_ids[threadIdx.x])) : "memory");
  const int warp = threadIdx.x >> 5, lane = (int)lane_v;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  // per-warp rings: the profile byte offset (code * 128) of the last 64 target columns, and the (H, V) of the stripe
  // above for the current block of 32 columns.  A lane reads its column's code and the first lane its boundary inputs
  // with one shared load each, instead of passing the code from lane to lane and broadcasting three values by shuffle.
  __shared__ unsigned ringc_all[K3X_WARPS][64];
  __shared__ int2 ringb_all[K3X_WARPS][32];
  unsigned* rc = ringc_all[warp];
  int2* rb = ringb_all[warp];
  const int n_targets = P.work_count ? (int)*P.work_count : P.n_targets;
  const long long n_items = (long long)n_targets * P.nstripes_max;
  const int oe = P.oe, ge = P.ge;

  for (;;) {
    long long item = 0;
    if (lane == 0) item = (long long)atomicAdd(P.counter, 1u);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= n_items) break;
    const int t = (int)(item / P.nstripes_max), stripe = (int)(item - (long long)t * P.nstripes_max);
    const int slot = t % K3X_SLOTS, gen = t / K3X_SLOTS;
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
    if (TRACK && P.phase == 1) {
      const int4 c = P.coords[t];
      if (c.y < 0) skip = true;
      else { lq = c.y + 1; lt = c.w + 1; qp += c.y; tp += c.w; qstep = -1; tstep = -1; }
    }
    if (!TRACK && lt < P.min_len) skip = true;             // this target belongs to k3_intra
    const int nstripes = skip ? 0 : (lq + 32 * R3 - 1) / (32 * R3);
    int best = 0, bj = 0x7fffffff, bi = 0x7fffffff;
    const bool early = TRACK && P.hit_keys != nullptr;
    const int known = early ? (int)(P.hit_keys[t] >> 32) : 0x7fffffff;      // the hit's score
    int* stop_p = early ? P.stop_col + t : nullptr;
    int stop = 0x7fffffff;                                  // this lane's view of stop_col[t] (only ever decreases)

    if (stripe < nstripes) {
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
      const bool has_in = stripe > 0, has_out = stripe + 1 < nstripes;
      const int4* row_in = P.rows + ((size_t)slot * 2 + ((stripe + 1) & 1)) * (size_t)P.row_stride;
      int4* row_out = P.rows + ((size_t)slot * 2 + (stripe & 1)) * (size_t)P.row_stride;
      const int tag_in_a = k3x_tag_a(item - 1), tag_in_b = k3x_tag_b(item - 1, P.epoch);
      const int tag_out_a = k3x_tag_a(item), tag_out_b = k3x_tag_b(item, P.epoch);
      int H[R3], G[R3];
#pragma unroll
      for (int k = 0; k < R3; ++k) { H[k] = 0; G[k] = 0; }
      int diag_in = 0, out_h = 0, out_v = 0;
      rc[lane] = 0; rc[lane + 32] = 0;                      // columns before the first one: padding code
      const int nsteps = lt + 31;
      int4* pcol = row_out - lane;                          // &row_out[s - lane], advanced every step (a loop-carried pointer
                                                            // is not re-derived from the slot/stripe arithmetic each step)
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {
          const int j = s + lane;
          int4 b = make_int4(0, 0, 0, 0);
          if (early) {
            stop = min(stop, ld_volatile_s32(stop_p));
            stop = __reduce_min_sync(0xffffffffu, stop);   // one view per warp: the exit below must be uniform
            if (s - 31 > stop) break;                      // every column this warp still has lies beyond the stop column
          }
          if (has_in && j < lt) {                          // my column of the stripe above: re-read until it carries its tag
            unsigned spins = 0;
            for (;;) {
              b = ld_volatile_v4(&row_in[j]);
              if (b.y == tag_in_a && b.w == tag_in_b) break;
              if (early) {                                 // the producer may have left already: columns beyond the stop column never come
                stop = min(stop, ld_volatile_s32(stop_p));
                if (j > stop) { b = make_int4(0, 0, 0, 0); break; }
              }
              __nanosleep(40);
              if (++spins > K3X_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; }
            }
          }
          __syncwarp();                                    // the previous block's rb has been read by the first lane
          rb[lane] = make_int2(b.x, b.z);
          rc[j & 63] = (j < lt) ? ((unsigned)tp[(long)j * tstep] & 31u) * 128u : 0u;
          __syncwarp();
        }
        // Column j = s - lane.  No range test around the cell update: before its first column a lane sees zeros from the
        // lane below and the padding code (H stays 0); past its last column it computes on the padding code, which can
        // only produce values below those already seen (the profile of padding is -128, gaps cost).  Only the hand-over
        // store and the tracked maximum are guarded.
        const int j = s - lane;
        int h_up = __shfl_up_sync(0xffffffffu, out_h, 1);
        int v_in = __shfl_up_sync(0xffffffffu, out_v, 1);
        const int2 b0 = rb[s & 31];
        if (lane == 0) { h_up = b0.x; v_in = b0.y; }
        const unsigned w0 = lds32_ordered(pl_addr + rc[j & 63]);
        const bool in_range = (unsigned)j < (unsigned)lt;
        {
          int d = diag_in;
          diag_in = h_up;
          int V = v_in;
          int colbest = 0, colk = 0;
#pragma unroll
          for (int k = 0; k < R3; ++k) {
            const int tt = (int)add_byte(w0, 1u << (8 * k), (unsigned)d);
            const int hx = __vimax3_s32_relu(tt, G[k], 0);   // max(0, t, G): independent of V
            const int h = max(hx, V);
            V = __viaddmax_s32(V, ge, hx + oe);              // the row-to-row chain: one op
            d = H[k];
            H[k] = h;
            G[k] = __viaddmax_s32(G[k], ge, h + oe);
            if (TRACK) { if (h > colbest) { colbest = h; colk = k; } } else best = max(best, h);
          }
          if (TRACK && in_range && j <= stop && k3x_better(colbest, j, row0 + colk, best, bj, bi)) {
            best = colbest; bj = j; bi = row0 + colk;
            if (early && colbest == known) { atomicMin(stop_p, j); stop = j; }      // nothing right of this column can win any more
          }
          out_h = H[R3 - 1];
          out_v = V;
          if (lane == 31 && has_out && in_range) st_volatile_v4(pcol, out_h, tag_out_a, out_v, tag_out_b);
        }
        ++pcol;
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

}  // namespace swk
