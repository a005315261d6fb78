// ubench_u8x4.cu -- SURVEY.md 8(f)4: is there a tier below 16 bits on sm_100a?  The cell update needs packed MAX and packed
// (saturating) ADD.  For 2 x 16 bit both are single DPX instructions (VIMNMX3.U16x2, VIADDMNMX.U16x2).  For 4 x 8 bit the CUDA
// intrinsics exist (__vmaxu4, __vaddus4, __vmaxs4, __vaddss4) -- this file measures what they cost: the SASS they compile to
// (scripts: cuobjdump -sass bench_micro/ubench_u8x4, committed as profiles/r2_u8x4_sass.txt) and their issue rate per SM,
// next to the 16x2 DPX forms on the same chains.  Cells per instruction decide: 16x2 does 2 cells in 1 ALU-pipe slot;
// 8x4 would have to do 4 cells in fewer than 2 slots to pay.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o ubench_u8x4 ubench_u8x4.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int CH = 8, ITERS = 4096;
enum { OP_MAXU4 = 0, OP_ADDUS4, OP_MAXS4, OP_ADDSS4, OP_MAX3_U16X2, OP_ADDMAX_U16X2, OP_CELL_U8X4, OP_CELL_U16X2 };

template <int OP>
__device__ __forceinline__ unsigned op(unsigned x, unsigned a, unsigned b) {
  if (OP == OP_MAXU4) return __vmaxu4(x, a);
  if (OP == OP_ADDUS4) return __vaddus4(x, a);
  if (OP == OP_MAXS4) return __vmaxs4(x, a);
  if (OP == OP_ADDSS4) return __vaddss4(x, a);
  if (OP == OP_MAX3_U16X2) return __vimax3_u16x2(x, a, b);
  if (OP == OP_ADDMAX_U16X2) return __viaddmax_u16x2(x, a, b);
  if (OP == OP_CELL_U8X4) {        // one affine-gap cell in 4 x u8: h = max(max(E, F), t); E' = max(E - ge, h - oe) (saturating)
    const unsigned h = __vmaxu4(__vmaxu4(x, a), b);
    return __vmaxu4(__vsubus4(x, 0x01010101u), __vsubus4(h, 0x0b0b0b0bu));
  }
  if (OP == OP_CELL_U16X2) {       // the same two lines in 2 x u16 DPX: 2 instructions
    const unsigned h = __vimax3_u16x2(x, a, b);
    return __viaddmax_u16x2(x, 0xffffffffu, h);
  }
  return x;
}

template <int OP>
__global__ void __launch_bounds__(256) k_u8(const unsigned* __restrict__ in, unsigned* __restrict__ out) {
  unsigned x[CH];
  const unsigned a = in[threadIdx.x & 255], b = in[(threadIdx.x + 7) & 255];
#pragma unroll
  for (int c = 0; c < CH; ++c) x[c] = in[(threadIdx.x + 3 * c) & 255];
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int c = 0; c < CH; ++c) x[c] = op<OP>(x[c], a, b);
    }
  }
  unsigned acc = 0;
#pragma unroll
  for (int c = 0; c < CH; ++c) acc ^= x[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP>
static void run(const unsigned* in, unsigned* out, int sms, double max_hz, const char* name, int cells_per_op) {
  const int grid = sms * 4, block = 256;
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int i = 0; i < 4; ++i) k_u8<OP><<<grid, block>>>(in, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 16; ++i) k_u8<OP><<<grid, block>>>(in, out);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 16;
  const double calls = (double)ITERS * 4 * CH * (block / 32.0) * grid;          // warp-level calls of op<>
  const double rate = calls / (ms * 1e-3) / (sms * max_hz);
  printf("{\"op\":\"%s\",\"ms\":%.4f,\"warp_calls_per_maxclk_sm\":%.3f,\"cells_per_call\":%d,\"warp_cells_per_maxclk_sm\":%.3f}\n",
         name, ms, rate, cells_per_op, rate * cells_per_op);
  fflush(stdout);
}

int main() {
  CK(cudaSetDevice(0));
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  unsigned h[256]; srand(5);
  for (int i = 0; i < 256; ++i) h[i] = (unsigned)rand();
  unsigned *in, *out;
  CK(cudaMalloc(&in, sizeof(h))); CK(cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&out, sizeof(unsigned) * p.multiProcessorCount * 4 * 256));
  const int sms = p.multiProcessorCount; const double hz = p.clockRate * 1e3;
  run<OP_MAXU4>(in, out, sms, hz, "__vmaxu4", 4);
  run<OP_ADDUS4>(in, out, sms, hz, "__vaddus4", 4);
  run<OP_MAXS4>(in, out, sms, hz, "__vmaxs4", 4);
  run<OP_ADDSS4>(in, out, sms, hz, "__vaddss4", 4);
  run<OP_MAX3_U16X2>(in, out, sms, hz, "__vimax3_u16x2", 2);
  run<OP_ADDMAX_U16X2>(in, out, sms, hz, "__viaddmax_u16x2", 2);
  run<OP_CELL_U8X4>(in, out, sms, hz, "cell fragment, 4 x u8 (h = max3, E' = max(E - ge, h - oe))", 4);
  run<OP_CELL_U16X2>(in, out, sms, hz, "cell fragment, 2 x u16 DPX (VIMNMX3 + VIADDMNMX)", 2);
  return 0;
}
