// sw_k1sw.cuh -- K1SW: packed 2 x u16 WAVEFRONT kernel for long streams, one CTA (8 warps) per stream.
//
// A titin-length target pair (tens of thousands of columns) is a serial chain of column steps; inside
// k1s one warp would walk it alone while the rest of the GPU has finished (measured: the 35 213-column
// stream of the Swiss-Prot-shaped database bounded every query shorter than ~600 residues).  Here the
// query rows are spread over the 8 warps of a CTA as a pipeline of stripes, exactly like k3w
// (sw_k3w.cuh), with the packed 16-bit arithmetic of k1s (sw_k1s.cuh):
//   * stripe = 32 lanes x 4 rows; warp w owns stripes w, w+8, ...; 1024 query rows are in flight at once;
//   * warp w-1 -> warp w through a 128-column ring in shared memory, warp 7 -> warp 0 through a
//     full-length row in global memory; producer/consumer column counters in shared memory;
//   * the cell update is the LATENCY-optimised form: with Hx = max(0, t, G) (independent of V) and Vs = V - oe,
//       Vs' = max(Vs + ge, Hx)   -- the only dependency from row to row, one VIADDMNMX
//       H   = max(Vs + oe, Hx)      G' = max(G + ge, H + oe)
//     so a column step costs about one shuffle plus four dependent DPX ops.
// Pair flags, biased-u16 representation, saturation limit and the raw[pair][slot] result layout are
// those of k1s; a pair gets 256 slots (8 warps x 32 lanes), reduced by k1s_finalize.
#pragma once
#include "sw_k1s.cuh"
#include "sw_k3w.cuh"

namespace swk {

struct K1SWParams {
  const uint16_t* scols;
  const int64_t*  stream_off;
  const int32_t*  stream_ncols;
  const int64_t*  stream_pair0;
  int             n_streams;      // the leading long streams
  unsigned*       counter;        // queue of streams
  const int8_t*   query;
  int             lq;
  const int8_t*   table;
  uint2*          bnd;            // per CTA: bnd_stride columns (H, Vs) for the warp 7 -> warp 0 hand-over
  long long       bnd_stride;
  unsigned*       raw;            // [pair][256], zero-initialised (biased values are > 0)
  unsigned*       error_flag;
  unsigned*       sm_flags;       // optional: this CTA marks its SM as reserved (the bulk kernel then leaves it)
  unsigned        BB, BBM, OE2, GE2, OE32, one;
};

constexpr size_t K1SW_SMEM = (size_t)K3W_WARPS * 3 * 32 * 32 * sizeof(unsigned);

__global__ void __launch_bounds__(K3W_WARPS * 32) k1sw(const K1SWParams P) {
  constexpr int R = 4, W = K3W_WARPS;
  __shared__ int8_t tab[32 * 32];
  extern __shared__ __align__(16) unsigned pl_dyn[];      // per warp: A[code][lane], B[code][2][lane] (K1SW_SMEM bytes)
  __shared__ uint2 ring[W][K3W_RING];
  __shared__ volatile int prod[W], cons[W];
  __shared__ int s_item;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  if (P.sm_flags && threadIdx.x == 0) {
    unsigned smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    atomicExch(P.sm_flags + smid, 1u);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pl = pl_dyn + warp * (3 * 32 * 32);
  const unsigned pa_lane = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  const unsigned pb_lane = pa_lane + 32 * 128;
  uint2* bnd = P.bnd + (size_t)blockIdx.x * (size_t)P.bnd_stride;
  const unsigned BB = P.BB, BBM = P.BBM, OE2 = P.OE2, GE2 = P.GE2, OE32 = P.OE32, one = P.one;
  const int lq = P.lq;
  const int pw = (warp + W - 1) % W;
  const int nstripes = (lq + 32 * R - 1) / (32 * R);

  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_item = (int)atomicAdd(P.counter, 1u);
    if (threadIdx.x < W) { prod[threadIdx.x] = 0; cons[threadIdx.x] = 0; }
    __syncthreads();
    const int st = s_item;
    if (st >= P.n_streams) break;
    const int lt = P.stream_ncols[st];
    const long pair0 = (long)P.stream_pair0[st];
    const unsigned short* cp = P.scols + P.stream_off[st];

    for (int stripe = warp; stripe < nstripes; stripe += W) {
      const int row0 = stripe * 32 * R + lane * R;
      __syncwarp();
      for (int c = 0; c < 32; ++c) {
        unsigned sc[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int i = row0 + b;
          const int s = (c == 0 || i >= lq) ? -128 : (int)tab[(P.query[i] & 31) * 32 + c];
          sc[b] = (unsigned)(s & 0xff);
        }
        pl[c * 32 + lane] = sc[0] | (sc[1] << 8) | (sc[2] << 16) | (sc[3] << 24);
        pl[32 * 32 + (c * 2 + 0) * 32 + lane] = sc[0] | (sc[0] << 8) | (sc[1] << 16) | (sc[1] << 24);
        pl[32 * 32 + (c * 2 + 1) * 32 + lane] = sc[2] | (sc[2] << 8) | (sc[3] << 16) | (sc[3] << 24);
      }
      __syncwarp();
      const bool has_in = stripe > 0, in_global = has_in && warp == 0;
      const bool has_out = stripe + 1 < nstripes, out_global = has_out && warp == W - 1;
      const int base_in = ((stripe - 1) / W) * lt;
      const int base_me = (stripe / W) * lt;
      unsigned H[R], G[R];
#pragma unroll
      for (int k = 0; k < R; ++k) { H[k] = BB; G[k] = BB; }
      unsigned diag_in = BB, out_h = BB, out_v = BBM, my_code = 0, best = BB;
      unsigned pre_code = 0, pre_h = BB, pre_v = BBM;
      long pair_idx = pair0 - 1;
      auto flush = [&]() {
        if (pair_idx >= pair0) {
          unsigned* slot = P.raw + (size_t)pair_idx * 256 + warp * 32 + lane;
          *slot = __vmaxu2(*slot, best);
        }
      };
      const int nsteps = lt + 31;
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {
          const int j = s + lane;
          if (has_in && s < lt) {
            const int need = base_in + min(s + 32, lt);
            if (lane == 0) {
              unsigned spins = 0;
              while (prod[pw] < need) { if (++spins > K3W_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; } }
            }
            __syncwarp();
          }
          uint2 b = make_uint2(BB, BBM);
          if (has_in && j < lt) b = in_global ? __ldcg(&bnd[j]) : ring[pw][(base_in + j) & (K3W_RING - 1)];
          pre_h = b.x; pre_v = b.y;
          pre_code = (j < lt) ? (unsigned)cp[j] : 0u;
          __syncwarp();
          if (has_in && !in_global && lane == 0 && s < lt) cons[pw] = base_in + min(s + 32, lt);
        }
        unsigned code = __shfl_up_sync(0xffffffffu, my_code, 1);
        unsigned h_up = __shfl_up_sync(0xffffffffu, out_h, 1);
        unsigned v_in = __shfl_up_sync(0xffffffffu, out_v, 1);
        const unsigned c0 = __shfl_sync(0xffffffffu, pre_code, s & 31);
        const unsigned h0 = __shfl_sync(0xffffffffu, pre_h, s & 31);
        const unsigned v0 = __shfl_sync(0xffffffffu, pre_v, s & 31);
        if (lane == 0) { code = c0; h_up = h0; v_in = v0; }
        my_code = code;
        const int j = s - lane;
        if (j >= 0 && j < lt) {
          if (code & 0x80u) {                             // first column of a new pair
            flush();
            ++pair_idx; best = BB; diag_in = BB;
#pragma unroll
            for (int k = 0; k < R; ++k) { H[k] = BB; G[k] = BB; }
          }
          const unsigned wA = lds32(pa_lane + (code & 0x1fu) * 128);
          const unsigned boff = pb_lane + ((code >> 8) & 0x1fu) * 256;
          const unsigned wB0 = lds32(boff), wB1 = lds32(boff + 128);
          unsigned d = diag_in;
          diag_in = h_up;
          unsigned V = v_in;
#pragma unroll
          for (int k = 0; k < R; ++k) {
            unsigned t = add_byte(wA, 1u << (8 * k), d);
            t = (k & 1) ? add_hi_hi(k < 2 ? wB0 : wB1, t) : add_hi_lo(k < 2 ? wB0 : wB1, t);
            const unsigned hx = __vimax3_u16x2(G[k], t, BB);
            const unsigned h = __viaddmax_u16x2(V, OE2, hx);
            V = __viaddmax_u16x2(V, GE2, hx);
            d = H[k];
            H[k] = h;
            G[k] = __viaddmax_u16x2(G[k], GE2, add_fma(h, one, OE32));
            best = __vmaxu2(best, h);
          }
          out_h = H[R - 1];
          out_v = V;
          if (lane == 31 && has_out) {
            if (out_global) {
              bnd[j] = make_uint2(out_h, out_v);
            } else {
              if ((j & 31) == 0) {
                const int need = base_me + j + 32 - K3W_RING;
                unsigned spins = 0;
                while (cons[warp] < need) { if (++spins > K3W_SPIN_LIMIT) { atomicExch(P.error_flag, 1u); break; } }
              }
              ring[warp][(base_me + j) & (K3W_RING - 1)] = make_uint2(out_h, out_v);
            }
            if ((j & 31) == 31 || j == lt - 1) {
              if (out_global) __threadfence(); else __threadfence_block();
              prod[warp] = base_me + j + 1;
            }
          }
        }
      }
      flush();
    }
  }
}

}  // namespace swk
