// swgpu_noalloc.cu -- swgpu_score_targets_noalloc(): an entry point that allocates nothing (include/swgpu.h).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid allocation).  The kernel is a variant of k3_intra
// (sw_kernels.cuh: 32-bit, one warp per target, same recurrence as src/alignment.c:117-176 of the reference in its
// score-equivalent two-gap-state form) whose scratch memory -- the per-warp carry rows between query stripes and the
// work-queue counter -- is a file-scope static __device__ array instead of cudaMalloc'ed buffers of a database handle.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the host gets the address once per device,
// with that device current, from cudaGetSymbolAddress() and passes that pointer to the kernel.
// The array is shared by every call on a GPU and keeps the last call's contents: the queue counter is cleared on the
// caller's stream before every launch; the carry rows need no clearing (a stripe only reads columns the stripe above
// wrote in the same call).  Calls that overlap on the GPU race on it (documented in the header).
#include "swgpu_internal.h"
#include "sw_kernels.cuh"
#include <algorithm>
#include <mutex>

namespace {

constexpr int NA_WARPS = 4;                                  // warps per CTA (as k3_intra)
constexpr int NA_CTAS = 64;                                  // persistent CTAs: 256 warp slots
constexpr int NA_SLOTS = NA_WARPS * NA_CTAS;
constexpr int NA_MAXLEN = SWGPU_NOALLOC_MAX_TARGET_LEN;
constexpr int NA_HDR = 32;                                   // int2 words of header (256 bytes): [0].x = work-queue counter
static_assert(sizeof(int2) * (size_t)(NA_HDR + NA_SLOTS * NA_MAXLEN) == (size_t)SWGPU_NOALLOC_SCRATCH_BYTES, "header and array disagree");

// ~8 MiB per GPU that loads the library, held until the process exits
static __device__ __align__(256) int2 g_noalloc_scratch[NA_HDR + NA_SLOTS * NA_MAXLEN];

struct NAParams {
  const int8_t*  query;   int lq;
  const int8_t*  targets; const int64_t* offsets; int n_targets;
  const int8_t*  table;   int oe, ge;
  int32_t*       scores;
  int2*          scratch;                                    // device address of g_noalloc_scratch (cudaGetSymbolAddress)
  unsigned       one;
};

// Lane l owns R3 consecutive query rows of the current stripe (32 * R3 rows per pass); the target streams through the
// lanes systolically (lane l works on column s - l at step s); the bottom row of a stripe goes to the warp's carry row
// in the scratch array and comes back as the top boundary of the next stripe.
template <int R3>
__global__ void __launch_bounds__(NA_WARPS * 32) k3_noalloc(const NAParams P) {
  static_assert(R3 % 4 == 0, "R3 must be a multiple of 4");
  constexpr int W = R3 / 4;
  __shared__ unsigned pl_all[NA_WARPS][32 * W * 32];        // per-lane private profile: [code][word][lane]
  __shared__ int8_t tab[32 * 32];
  __shared__ unsigned ringc_all[NA_WARPS][64];               // profile byte offset of the last 64 target columns
  __shared__ int2 ringb_all[NA_WARPS][32];                   // (H, V) of the stripe above, current block of 32 columns
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* pl = pl_all[warp];
  const unsigned pl_addr = (unsigned)__cvta_generic_to_shared(pl) + lane * 4;
  unsigned* rc = ringc_all[warp];
  int2* rb = ringb_all[warp];
  unsigned* counter = reinterpret_cast<unsigned*>(P.scratch);
  int2* bnd = P.scratch + NA_HDR + (size_t)(blockIdx.x * NA_WARPS + warp) * (size_t)NA_MAXLEN;
  const int oe = P.oe, ge = P.ge, lq = P.lq;
  const int npass = (lq + 32 * R3 - 1) / (32 * R3);

  for (;;) {
    int w = 0;
    if (lane == 0) w = (int)atomicAdd(counter, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= P.n_targets) break;
    const int8_t* tp = P.targets + P.offsets[w];
    const long long lt64 = (long long)(P.offsets[w + 1] - P.offsets[w]);
    if (lt64 > NA_MAXLEN) { if (lane == 0) P.scores[w] = -1; continue; }      // beyond the scratch rows: not scored
    const int lt = lt64 > 0 ? (int)lt64 : 0;
    int best = 0;
    for (int pass = 0; pass < npass && lt > 0; ++pass) {
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
      rc[lane] = 0; rc[lane + 32] = 0;
      for (int s = 0; s < nsteps; ++s) {
        if ((s & 31) == 0) {                                   // coalesced fetch of the next 32 columns
          const int j = s + lane;
          int2 b = make_int2(0, 0);
          if (!first && j < lt) b = bnd[j];
          __syncwarp();
          rb[lane] = b;
          rc[j & 63] = (j < lt) ? ((unsigned)tp[j] & 31u) * (unsigned)(W * 128) : 0u;
          __syncwarp();
        }
        const int j = s - lane;                                // this lane's column; out of range = padding code, cannot raise the maximum
        int h_up = __shfl_up_sync(0xffffffffu, out_h, 1);
        int v_in = __shfl_up_sync(0xffffffffu, out_v, 1);
        const int2 b0 = rb[s & 31];
        if (lane == 0) { h_up = b0.x; v_in = b0.y; }
        const unsigned poff = pl_addr + rc[j & 63];
        int d = diag_in;
        diag_in = h_up;
        int V = v_in;
        unsigned wreg[W];
#pragma unroll
        for (int ww = 0; ww < W; ++ww) wreg[ww] = swk::lds32_ordered(poff + ww * 128);
#pragma unroll
        for (int k = 0; k < R3; ++k) {
          const int t = (int)swk::add_byte(wreg[k >> 2], 1u << (8 * (k & 3)), (unsigned)d);
          const int h = __vimax3_s32_relu(t, G[k], V);
          d = H[k];
          H[k] = h;
          const int u = (int)swk::add_fma((unsigned)h, P.one, (unsigned)oe);
          G[k] = __viaddmax_s32(G[k], ge, u);
          V = __viaddmax_s32(V, ge, u);
          best = max(best, h);
        }
        out_h = H[R3 - 1];
        out_v = V;
        if (lane == 31 && !last && (unsigned)j < (unsigned)lt) bnd[j] = make_int2(out_h, out_v);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) P.scores[w] = best;
  }
}

}  // namespace

namespace {

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
int2* g_na_ptr[NA_MAX_DEVICES] = {};                         // device address of g_noalloc_scratch, per device

// Once per device, with that device current: the device address of the scratch array.  (In host code the name
// g_noalloc_scratch is a stand-in for the symbol, not an address a kernel could use.)
int noalloc_scratch(int device, int2** out) {
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[device]) {
    int2* p = nullptr;
    CU(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (!p) return swi::fail(SWGPU_ECUDA, "cudaGetSymbolAddress returned a null address for the scratch array");
    g_na_ptr[device] = p;
  }
  *out = g_na_ptr[device];
  return SWGPU_OK;
}

}  // namespace

extern "C" int swgpu_score_targets_noalloc(int device, const int8_t* d_query, int32_t query_len, const int8_t* d_targets,
                                           const int64_t* d_offsets, int32_t n_targets, int32_t max_target_len,
                                           const int8_t* d_table, int gap_open, int gap_extend, int32_t* d_scores, void* stream) {
  using swi::fail;
  if (n_targets < 0 || query_len < 0 || max_target_len < 0) return fail(SWGPU_EINVAL, "swgpu_score_targets_noalloc: negative size");
  if (max_target_len > SWGPU_NOALLOC_MAX_TARGET_LEN)
    return fail(SWGPU_EINVAL, "swgpu_score_targets_noalloc: max_target_len %d exceeds SWGPU_NOALLOC_MAX_TARGET_LEN = %d (the static "
                "scratch array holds carry rows of that many columns; use swgpu_db_create + swgpu_search for longer targets)",
                (int)max_target_len, (int)SWGPU_NOALLOC_MAX_TARGET_LEN);
  if (n_targets > 0 && (!d_offsets || !d_scores || !d_table || (query_len > 0 && !d_query) || (max_target_len > 0 && !d_targets)))
    return fail(SWGPU_EINVAL, "swgpu_score_targets_noalloc: null pointer");
  const int oe = gap_open + gap_extend, ge = gap_extend;
  if (oe > 0 || ge > 0 || gap_open > 0)
    return fail(SWGPU_EINVAL, "positive gap scores are outside the supported (and the reference-equivalent) range");
  if (oe < -100000 || ge < -100000) return fail(SWGPU_EINVAL, "gap penalty out of range");
  int ndev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) return fail(SWGPU_ENODEV, "no CUDA device");
  if (device < 0 || device >= ndev || device >= NA_MAX_DEVICES) return fail(SWGPU_EINVAL, "device %d out of range (%d devices)", device, ndev);
  CU(cudaSetDevice(device));
  if (n_targets == 0) return SWGPU_OK;
  int2* scratch = nullptr;
  if (int r = noalloc_scratch(device, &scratch)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's contents: the work-queue counter must start at zero
  CU(cudaMemsetAsync(scratch, 0, sizeof(int2) * NA_HDR, st));
  if (query_len == 0 || max_target_len == 0) {               // nothing to align: every score is 0
    CU(cudaMemsetAsync(d_scores, 0, sizeof(int32_t) * (size_t)n_targets, st));
    return SWGPU_OK;
  }
  NAParams P;
  P.query = d_query; P.lq = (int)query_len; P.targets = d_targets; P.offsets = d_offsets; P.n_targets = (int)n_targets;
  P.table = d_table; P.oe = oe; P.ge = ge; P.scores = d_scores; P.scratch = scratch; P.one = 1u;
  const int grid = (int)std::min<long long>(NA_CTAS, ((long long)n_targets + NA_WARPS - 1) / NA_WARPS);
  if (query_len <= 128) k3_noalloc<4><<<grid, NA_WARPS * 32, 0, st>>>(P);
  else                  k3_noalloc<8><<<grid, NA_WARPS * 32, 0, st>>>(P);
  CU(cudaGetLastError());
  return SWGPU_OK;
}
