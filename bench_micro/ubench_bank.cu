// ubench_bank.cu -- does a DPX instruction whose three source registers sit in the SAME register-file bank
// (even / odd register number) issue slower than one with mixed banks?  (sm_100a)
//
// Sixteen accumulators v[0..15]; every instruction is v[i] = op(v[i], v[(i+D1)&15], v[(i+D2)&15]).  ptxas gives the
// accumulators sixteen registers; which of the instructions end up with three same-parity sources is read back from
// the SASS (scripts/sass_banks.py) -- the (D1, D2) patterns below are chosen so that, for the usual consecutive
// allocation, (2,4) is all-same-bank and (1,2) / (1,3) are mixed.  Also: the immediate forms (two register sources).
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o ubench_bank ubench_bank.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int NV = 16, ITERS = 4096;

enum { K_MAX3 = 0, K_ADDMAX = 1, K_MAX3_IMM = 2, K_ADDMAX_IMM = 3, K_MIX = 4, K_MIX_IMM = 5 };

template <int KIND>
__device__ __forceinline__ unsigned op3(unsigned x, unsigned a, unsigned b) {
  if (KIND == K_MAX3) return __vimax3_u16x2(x, a, b);
  if (KIND == K_ADDMAX) return __viaddmax_u16x2(x, a, b);
  if (KIND == K_MAX3_IMM) return __vimax3_u16x2(x, 0x00970097u, b);
  if (KIND == K_ADDMAX_IMM) return __viaddmax_u16x2(x, 0xfff5fff5u, b);
  return x;
}

template <int KIND, int D1, int D2>
__global__ void __launch_bounds__(256) k_bank(const unsigned* __restrict__ in, unsigned* __restrict__ out) {
  unsigned v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = in[(threadIdx.x + 17 * i) & 255];
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        if (KIND == K_MIX) {             // the K1S cell: max3 + 3 addmax, register operands
          v[i] = (i & 3) == 0 ? __vimax3_u16x2(v[i], v[(i + D1) & 15], v[(i + D2) & 15])
                              : __viaddmax_u16x2(v[i], v[(i + D1) & 15], v[(i + D2) & 15]);
        } else if (KIND == K_MIX_IMM) {
          v[i] = (i & 3) == 0 ? __vimax3_u16x2(v[i], 0x00970097u, v[(i + D2) & 15])
                              : __viaddmax_u16x2(v[i], 0xfffffff5u, v[(i + D2) & 15]);
        } else {
          v[i] = op3<KIND>(v[i], v[(i + D1) & 15], v[(i + D2) & 15]);
        }
      }
    }
  }
  unsigned acc = 0;
#pragma unroll
  for (int i = 0; i < NV; ++i) acc ^= v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// Forced banks: eight 64-bit accumulators (aligned even/odd register pairs).  Every instruction reads one half of three
// of them and writes the same half of the first back, so the three source registers have a FIXED parity:
// SEL 0 = all low halves (three even registers), 7 = all high halves (three odd), 2 = low / high / low (mixed).
template <int SEL>
__global__ void __launch_bounds__(256) k_forced(const unsigned long long* __restrict__ in, unsigned long long* __restrict__ out) {
  unsigned long long x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = in[(threadIdx.x + 17 * i) & 127];
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        unsigned long long &d = x[c];
        const unsigned long long b = x[(c + 1) & 7], e = x[(c + 2) & 7];
        if (SEL == 0)
          asm volatile("{.reg .b32 l0,h0,l1,h1,l2,h2,t1;\n\tmov.b64 {l0,h0}, %0;\n\tmov.b64 {l1,h1}, %1;\n\tmov.b64 {l2,h2}, %2;\n\t"
                       "max.u16x2 t1, l0, l1;\n\tmax.u16x2 l0, t1, l2;\n\tmov.b64 %0, {l0,h0};}" : "+l"(d) : "l"(b), "l"(e));
        else if (SEL == 2)
          asm volatile("{.reg .b32 l0,h0,l1,h1,l2,h2,t1;\n\tmov.b64 {l0,h0}, %0;\n\tmov.b64 {l1,h1}, %1;\n\tmov.b64 {l2,h2}, %2;\n\t"
                       "max.u16x2 t1, l0, h1;\n\tmax.u16x2 l0, t1, l2;\n\tmov.b64 %0, {l0,h0};}" : "+l"(d) : "l"(b), "l"(e));
        else
          asm volatile("{.reg .b32 l0,h0,l1,h1,l2,h2,t1;\n\tmov.b64 {l0,h0}, %0;\n\tmov.b64 {l1,h1}, %1;\n\tmov.b64 {l2,h2}, %2;\n\t"
                       "max.u16x2 t1, h0, h1;\n\tmax.u16x2 h0, t1, h2;\n\tmov.b64 %0, {l0,h0};}" : "+l"(d) : "l"(b), "l"(e));
      }
    }
  }
  unsigned long long acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc += x[i] * (unsigned long long)(i + 3);     // the accumulators stay 64-bit values
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

struct Ctx { unsigned* in; unsigned* out; int sms; double max_hz; };

template <int KIND, int D1, int D2>
static void run(const Ctx& c, const char* name) {
  const int grid = c.sms * 4, block = 256;               // 32 warps per SM, 8 per scheduler
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  constexpr int REPS = 16;
  for (int i = 0; i < 4; ++i) k_bank<KIND, D1, D2><<<grid, block>>>(c.in, c.out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < REPS; ++i) k_bank<KIND, D1, D2><<<grid, block>>>(c.in, c.out);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= REPS;
  const double winst = (double)ITERS * 4 * NV * (block / 32.0) * grid;
  const double rate = winst / (ms * 1e-3) / (c.sms * c.max_hz);
  printf("{\"kernel\":\"k_bank<%d,%d,%d>\",\"what\":\"%s\",\"ms\":%.4f,\"winst_per_maxclk_sm\":%.3f}\n", KIND, D1, D2, name, ms, rate);
  fflush(stdout);
}

template <int SEL>
static void run_forced(const Ctx& c, const char* name) {
  const int grid = c.sms * 4, block = 256;
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  constexpr int REPS = 16;
  const unsigned long long* in64 = reinterpret_cast<const unsigned long long*>(c.in);
  for (int i = 0; i < 4; ++i) k_forced<SEL><<<grid, block>>>(in64, reinterpret_cast<unsigned long long*>(c.out));
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < REPS; ++i) k_forced<SEL><<<grid, block>>>(in64, reinterpret_cast<unsigned long long*>(c.out));
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= REPS;
  const double winst = (double)ITERS * 8 * 8 * (block / 32.0) * grid;
  printf("{\"kernel\":\"k_forced<%d>\",\"what\":\"%s\",\"ms\":%.4f,\"winst_per_maxclk_sm\":%.3f}\n", SEL, name, ms,
         winst / (ms * 1e-3) / (c.sms * c.max_hz));
  fflush(stdout);
}

int main() {
  CK(cudaSetDevice(0));
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  Ctx c; c.sms = p.multiProcessorCount; c.max_hz = p.clockRate * 1e3;
  unsigned h[256];
  srand(7);
  for (int i = 0; i < 256; ++i) h[i] = (unsigned)(rand() & 0x3ff) | ((unsigned)(rand() & 0x3ff) << 16);
  CK(cudaMalloc(&c.in, sizeof(h))); CK(cudaMemcpy(c.in, h, sizeof(h), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&c.out, sizeof(unsigned long long) * c.sms * 4 * 256));
  run<K_MAX3, 2, 4>(c, "VIMNMX3.U16x2 sources i,i+2,i+4");
  run<K_MAX3, 1, 2>(c, "VIMNMX3.U16x2 sources i,i+1,i+2");
  run<K_MAX3, 1, 3>(c, "VIMNMX3.U16x2 sources i,i+1,i+3");
  run<K_ADDMAX, 2, 4>(c, "VIADDMNMX.U16x2 sources i,i+2,i+4");
  run<K_ADDMAX, 1, 2>(c, "VIADDMNMX.U16x2 sources i,i+1,i+2");
  run<K_ADDMAX, 1, 3>(c, "VIADDMNMX.U16x2 sources i,i+1,i+3");
  run<K_MAX3_IMM, 0, 2>(c, "VIMNMX3.U16x2 imm, sources i,i+2");
  run<K_MAX3_IMM, 0, 1>(c, "VIMNMX3.U16x2 imm, sources i,i+1");
  run<K_ADDMAX_IMM, 0, 2>(c, "VIADDMNMX.U16x2 imm, sources i,i+2");
  run<K_ADDMAX_IMM, 0, 1>(c, "VIADDMNMX.U16x2 imm, sources i,i+1");
  run<K_MIX, 2, 4>(c, "max3 + 3 addmax, sources i,i+2,i+4");
  run<K_MIX, 1, 2>(c, "max3 + 3 addmax, sources i,i+1,i+2");
  run<K_MIX_IMM, 0, 2>(c, "max3 + 3 addmax imm, sources i,i+2");
  run<K_MIX_IMM, 0, 1>(c, "max3 + 3 addmax imm, sources i,i+1");
  run_forced<0>(c, "VIMNMX3.U16x2, three EVEN source registers (low halves of 64-bit pairs)");
  run_forced<7>(c, "VIMNMX3.U16x2, three ODD source registers (high halves)");
  run_forced<2>(c, "VIMNMX3.U16x2, even/odd/even source registers");
  return 0;
}
