// sw_k1s_inst.cu -- the K1S instantiations of ONE gap specialisation (-DK1S_GAPSEL=0|1|2) and ONE ptxas option set
// (-DK1S_PART=0: everything built at -O3; -DK1S_PART=1: the multi-stripe kernels with 24..32 rows per lane, built at
// ptxas -O1).  The Makefile compiles this file six times, in parallel.
//
// Why a part of its own at -O1: at -O3 the multi-stripe kernels with the most rows per lane sit on the 128-register
// ceiling of two CTAs per SM and spill 24-56 bytes in the cell loop; at -O1 they take 120-126 registers and do not
// spill (round 1: 4.5-5.5 % faster at R = 30, 32; re-checked with the round-2 kernel by `nvcc -Xptxas -v`).  Every
// other instantiation is as fast or faster at -O3 (scripts/ab_lib.py, two builds interleaved on one box).  Round 2, every
// multi-stripe plan timed at both levels (profiles/r2y_ab_ptxas_levels.txt): R = 26..32 are 1.4-5.8 % faster at -O1, R = 24, 25
// 1.0-1.8 % (moved here), R = 16, 20 and the single-stripe kernels are faster at -O3.
#include "sw_k1s.cuh"

#ifndef K1S_GAPSEL
#error "compile with -DK1S_GAPSEL=0|1|2"
#endif
#ifndef K1S_PART
#error "compile with -DK1S_PART=0|1"
#endif

namespace swk {
namespace {

template <int R, int G, bool MULTI>
cudaError_t launch_t(const K1SParams& P, cudaStream_t stream, int grid, int device) {
  constexpr size_t smem = k1s_smem_bytes<R, MULTI>();
  static thread_local bool attr_done[16] = {};
  if (!attr_done[device & 15]) {
    const cudaError_t e = cudaFuncSetAttribute(k1s<R, G, MULTI, K1S_GAPSEL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_done[device & 15] = true;
  }
  k1s<R, G, MULTI, K1S_GAPSEL><<<grid, K1S_WARPS * 32, smem, stream>>>(P);
  return cudaGetLastError();
}

#if K1S_PART == 0
template <int G>
cudaError_t launch_single(int R, const K1SParams& P, cudaStream_t stream, int grid, int device) {
#define K1S_CASE(r) case r: return launch_t<r, G, false>(P, stream, grid, device);
  switch (R) {
    K1S_CASE(8) K1S_CASE(9) K1S_CASE(10) K1S_CASE(11) K1S_CASE(12) K1S_CASE(13) K1S_CASE(14) K1S_CASE(15) K1S_CASE(16)
    K1S_CASE(17) K1S_CASE(18) K1S_CASE(19) K1S_CASE(20) K1S_CASE(21) K1S_CASE(22) K1S_CASE(23) K1S_CASE(24) K1S_CASE(25)
    K1S_CASE(26) K1S_CASE(27) K1S_CASE(28) K1S_CASE(29) K1S_CASE(30) K1S_CASE(31) K1S_CASE(32)
    default: return cudaErrorInvalidValue;
  }
#undef K1S_CASE
}
#endif

}  // namespace

#if K1S_PART == 0
// Instantiated: every R in 8..32 for g = 8, 16, 32; multi-stripe (g = 32) with R in {16, 20} (24..32: part 1).
template <>
cudaError_t launch_k1s_o3<K1S_GAPSEL>(int g, int R, bool multi, const K1SParams& P, cudaStream_t stream, int grid, int device) {
  if (multi) {
    switch (R) {
      case 16: return launch_t<16, 32, true>(P, stream, grid, device);
      case 20: return launch_t<20, 32, true>(P, stream, grid, device);
      default: return cudaErrorInvalidValue;
    }
  }
  if (g == 8) return launch_single<8>(R, P, stream, grid, device);
  if (g == 16) return launch_single<16>(R, P, stream, grid, device);
  if (g == 32) return launch_single<32>(R, P, stream, grid, device);
  return cudaErrorInvalidValue;
}
#else
template <>
cudaError_t launch_k1s_o1<K1S_GAPSEL>(int R, const K1SParams& P, cudaStream_t stream, int grid, int device) {
  switch (R) {
    case 24: return launch_t<24, 32, true>(P, stream, grid, device);
    case 25: return launch_t<25, 32, true>(P, stream, grid, device);
    case 26: return launch_t<26, 32, true>(P, stream, grid, device);
    case 27: return launch_t<27, 32, true>(P, stream, grid, device);
    case 28: return launch_t<28, 32, true>(P, stream, grid, device);
    case 29: return launch_t<29, 32, true>(P, stream, grid, device);
    case 30: return launch_t<30, 32, true>(P, stream, grid, device);
    case 31: return launch_t<31, 32, true>(P, stream, grid, device);
    case 32: return launch_t<32, 32, true>(P, stream, grid, device);
    default: return cudaErrorInvalidValue;
  }
}
#endif

}  // namespace swk
