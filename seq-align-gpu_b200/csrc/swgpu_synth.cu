// swgpu_synth.cu -- device-side generator of synthetic databases (swgpu_db_create_synthetic, include/swgpu.h).
//
// a UniRef50-shaped database of 10 G residues is generated on the device, shard by shard,
// never as a FASTA file.  The host builds the layout from the lengths alone (db_pack.c::swpack_layout); here the
// residues are written straight into the two packed representations:
//   synth_plain  -- one warp per sorted target: residue j of global entry e from the counter hash of (seed, e, j)
//   synth_scols  -- one warp per target pair: the pair's 16-bit stream columns from `plain` (same rule as
//                   db_pack.c::fill_streams: low byte = first member, bit 7 on the first column, high byte = second)
// Both are pure streaming writes (1 B + 2 B per residue); the 10 G-residue database takes well under a second.
#include "swgpu_internal.h"

namespace {

struct SynthTab { unsigned cdf[32]; signed char codes[32]; int n; };

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {   // splitmix64 finaliser
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}

__global__ void __launch_bounds__(256) synth_plain(int8_t* __restrict__ plain, const int64_t* __restrict__ plain_off,
                                                   const int64_t* __restrict__ order, const int64_t* __restrict__ entry_ids,
                                                   long long entry_base, long long n, unsigned long long seed, SynthTab T) {
  __shared__ unsigned cdf_s[32];
  __shared__ signed char code_s[32];
  if (threadIdx.x < 32) { cdf_s[threadIdx.x] = threadIdx.x < T.n ? T.cdf[threadIdx.x] : 0xffffffffu; code_s[threadIdx.x] = T.codes[threadIdx.x]; }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long s = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < n; s += warps) {
    const long long t = order[s];
    const unsigned long long e = (unsigned long long)(entry_ids ? entry_ids[t] : entry_base + t);
    const unsigned long long key = seed ^ (e * 0x9E3779B97F4A7C15ull);
    const long long b = plain_off[s], len = plain_off[s + 1] - b;
    for (long long j = lane; j < len; j += 32) {
      const unsigned u = (unsigned)(mix64(key + (unsigned long long)(j + 1) * 0xBF58476D1CE4E5B9ull) >> 48);
      int i = 0;
      while (i < T.n - 1 && u >= cdf_s[i]) ++i;
      plain[b + j] = code_s[i];
    }
  }
}

__global__ void __launch_bounds__(256) synth_scols(uint16_t* __restrict__ scols, const int8_t* __restrict__ plain,
                                                   const int64_t* __restrict__ plain_off, const int64_t* __restrict__ pair_col,
                                                   long long n, long long n_pairs) {
  const int lane = threadIdx.x & 31;
  const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long q = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); q < n_pairs; q += warps) {
    const long long a = plain_off[2 * q], la = plain_off[2 * q + 1] - a;
    const long long lb = (2 * q + 1 < n) ? plain_off[2 * q + 2] - plain_off[2 * q + 1] : 0;
    const long long len = la < 1 ? 1 : la;        // an empty pair still takes one flagged, all-padding column
    uint16_t* dst = scols + pair_col[q];
    for (long long j = lane; j < len; j += 32) {
      const unsigned ca = j < la ? (unsigned char)plain[a + j] : 0u, cb = j < lb ? (unsigned char)plain[a + la + j] : 0u;
      dst[j] = (uint16_t)(ca | (cb << 8) | (j == 0 ? 0x80u : 0u));
    }
  }
}

}  // namespace

namespace swi {

cudaError_t synth_fill(swgpu_db* db, const swgpu_synth_t& gen, const int64_t* d_entry_ids, int64_t entry_base,
                       const int64_t* d_pair_col) {
  if (db->n == 0) return cudaSuccess;
  SynthTab T;
  T.n = gen.n_letters;
  for (int i = 0; i < 32; ++i) { T.cdf[i] = gen.cdf[i]; T.codes[i] = gen.codes[i]; }
  const int grid = db->sm_count * 8;
  synth_plain<<<grid, 256>>>(db->d_plain, db->d_plain_off, db->d_order, d_entry_ids, (long long)entry_base, (long long)db->n,
                             (unsigned long long)gen.seed, T);
  synth_scols<<<grid, 256>>>(db->d_scols, db->d_plain, db->d_plain_off, d_pair_col, (long long)db->n, (long long)db->n_pairs);
  return cudaGetLastError();
}

}  // namespace swi
