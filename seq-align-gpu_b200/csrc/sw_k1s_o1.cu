// sw_k1s_o1.cu -- the multi-stripe K1S instantiations with the most rows per lane (R = 26..32), built in a
// translation unit of their own so that the Makefile can give ptxas a different optimisation level for them.
//
// Why: at ptxas -O3 these three kernels hit the 128-register ceiling of two CTAs per SM and spill (24/32/40 bytes
// of stack in the cell loop) and carry 1.2-1.5 register moves per row; at -O1 they need 116-122 registers, do not
// spill and carry 0.4-0.8 moves per row.  Measured on the C2 database (scripts/ab_lib.py, same box, interleaved):
// R = 32 and R = 30 are 4.5-5.5 % faster, R = 28 is 1 % faster.  Every other instantiation is as fast or faster at
// -O3 (single-stripe R = 24: -O1 issues 8 % fewer instructions but at 0.62 instead of 0.73 per cycle), so those
// stay in swgpu.cu.
#include "sw_k1s.cuh"

namespace swk {

template <int R>
static cudaError_t launch_multi(const K1SParams& P, cudaStream_t stream, int grid, int device) {
  const size_t smem = k1s_smem_bytes<R, true>();
  static thread_local bool attr_done[16] = {};
  if (!attr_done[device & 15]) {
    const cudaError_t e = cudaFuncSetAttribute(k1s<R, 32, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_done[device & 15] = true;
  }
  k1s<R, 32, true><<<grid, K1S_WARPS * 32, smem, stream>>>(P);
  return cudaGetLastError();
}

// R must be 26..32 (cudaErrorInvalidValue otherwise).
cudaError_t launch_k1s_multi_o1(int R, const K1SParams& P, cudaStream_t stream, int grid, int device) {
  switch (R) {
    case 26: return launch_multi<26>(P, stream, grid, device);
    case 27: return launch_multi<27>(P, stream, grid, device);
    case 28: return launch_multi<28>(P, stream, grid, device);
    case 29: return launch_multi<29>(P, stream, grid, device);
    case 30: return launch_multi<30>(P, stream, grid, device);
    case 31: return launch_multi<31>(P, stream, grid, device);
    case 32: return launch_multi<32>(P, stream, grid, device);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace swk
