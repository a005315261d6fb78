// ubench_int.cu -- integer / DPX / data-movement issue-rate microbenchmark for sm_100a.
// Measures warp-instructions per clock per SM for the instruction classes the
// Smith-Waterman cell update is built from, alone and in pairs (co-issue test).
// The measured DPX rate is the denominator of the cell-update roofline
// (MEASURED_PEAKS.json has no integer figure).
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o ubench_int ubench_int.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

enum Op { OP_NONE = 0, OP_VIADD2, OP_VIMNMX2, OP_VIMNMX3R, OP_VIADDMNMX, OP_VIADDMNMXR,
          OP_IADD3, OP_IMAD, OP_LOP3, OP_PRMT, OP_IDP4A, OP_IDP2A, OP_LDS, OP_SHFL,
          OP_VIMNMX3_32, OP_VIADDMNMX_32, OP_FFMA, OP_HADD2, OP_HMNMX2, OP_LDS_RAND, OP_COUNT };

static const char* op_name[] = { "none", "VIADD.16x2", "VIMNMX.S16x2", "VIMNMX3.S16x2.RELU", "VIADDMNMX.S16x2",
  "VIADDMNMX.S16x2.RELU", "IADD3", "IMAD", "LOP3", "PRMT", "IDP.4A", "IDP.2A", "LDS.32(cf)", "SHFL.IDX",
  "VIMNMX3.S32.RELU", "VIADDMNMX.S32", "FFMA", "HADD2", "HMNMX2", "LDS.32(rand)" };

template <int OP>
__device__ __forceinline__ unsigned do_op(unsigned x, unsigned a, unsigned b, unsigned m,
                                          const unsigned* sm, unsigned lane) {
  if (OP == OP_VIADD2) return __vadd2(x, a);
  if (OP == OP_VIMNMX2) return __vmaxs2(x, a);
  if (OP == OP_VIMNMX3R) return __vimax3_s16x2_relu(x, a, b);
  if (OP == OP_VIADDMNMX) return __viaddmax_s16x2(x, a, b);
  if (OP == OP_VIADDMNMXR) return __viaddmax_s16x2_relu(x, a, b);
  if (OP == OP_IADD3) { unsigned r; asm volatile("add.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(a)); return r; }
  if (OP == OP_IMAD) { unsigned r; asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(m), "r"(a)); return r; }
  if (OP == OP_LOP3) { unsigned r; asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
  if (OP == OP_PRMT) return __byte_perm(x, a, b & 0x7777);
  if (OP == OP_IDP4A) return (unsigned)__dp4a((int)a, (int)m, (int)x);
  if (OP == OP_IDP2A) return (unsigned)__dp2a_lo((int)a, (int)m, (int)x);
  // pointer-chasing loads: the loaded word IS the next byte offset, so no address arithmetic is issued
  if (OP == OP_LDS) { return *(const unsigned*)((const char*)sm + x); }        // [row][lane] -> lane-distinct banks
  if (OP == OP_LDS_RAND) { return *(const unsigned*)((const char*)sm + x); }   // random word -> random bank
  if (OP == OP_SHFL) return __shfl_sync(0xffffffffu, x, (int)(a & 31));
  if (OP == OP_VIMNMX3_32) return (unsigned)__vimax3_s32_relu((int)x, (int)a, (int)b);
  if (OP == OP_VIADDMNMX_32) return (unsigned)__viaddmax_s32((int)x, (int)a, (int)b);
  if (OP == OP_FFMA) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(__uint_as_float(x)), "f"(__uint_as_float(m)), "f"(__uint_as_float(a))); return __float_as_uint(r); }
  if (OP == OP_HADD2) { unsigned r; asm volatile("add.f16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(a)); return r; }
  if (OP == OP_HMNMX2) { unsigned r; asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(a)); return r; }
  return x;
}

constexpr int CH = 8;      // independent chains per op class
constexpr int ITERS = 8192;

template <int OP1, int N1, int OP2, int N2>
__global__ void __launch_bounds__(256) k_mix(const unsigned* __restrict__ in, unsigned* __restrict__ out,
                                             long long* __restrict__ cyc) {
  __shared__ unsigned sm[1024];
  constexpr bool kRand = (OP1 == OP_LDS_RAND || OP2 == OP_LDS_RAND);
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
    const unsigned r = in[(i * 7 + 3) & 255] & 0x3ff;                   // pseudo-random word index 0..1023
    sm[i] = kRand ? (r << 2) : ((((r >> 5) << 5) + (i & 31)) << 2);     // byte offset of the next hop
  }
  __syncthreads();
  const unsigned lane = threadIdx.x & 31;
  unsigned x1[CH], x2[CH];
  unsigned a = in[threadIdx.x & 255], b = in[(threadIdx.x + 7) & 255], m = in[256];
#pragma unroll
  for (int c = 0; c < CH; ++c) { x1[c] = in[(threadIdx.x + c) & 255]; x2[c] = in[(threadIdx.x + 3 * c + 1) & 255]; }
  if (OP1 == OP_LDS || OP1 == OP_LDS_RAND) {
#pragma unroll
    for (int c = 0; c < CH; ++c) x1[c] = (((x1[c] & 31) << 5) + lane) << 2;
  }
  if (OP2 == OP_LDS || OP2 == OP_LDS_RAND) {
#pragma unroll
    for (int c = 0; c < CH; ++c) x2[c] = (((x2[c] & 31) << 5) + lane) << 2;
  }
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int j = 0; j < (N1 > N2 ? N1 : N2); ++j) {
        if (j < N1) {
#pragma unroll
          for (int c = 0; c < CH; ++c) x1[c] = do_op<OP1>(x1[c], a, b, m, sm, lane);
        }
        if (j < N2) {
#pragma unroll
          for (int c = 0; c < CH; ++c) x2[c] = do_op<OP2>(x2[c], a, b, m, sm, lane);
        }
      }
    }
  }
  long long t1 = clock64();
  unsigned acc = 0;
#pragma unroll
  for (int c = 0; c < CH; ++c) acc ^= x1[c] ^ x2[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

struct Ctx { unsigned* in; unsigned* out; long long* cyc; int sms; int ctas_per_sm; double max_hz; };

template <int OP1, int N1, int OP2, int N2>
static void run(const Ctx& c) {
  const int grid = c.sms * c.ctas_per_sm, block = 256;
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  constexpr int REPS = 24;                                          // long enough for the clocks to settle
  for (int i = 0; i < 8; ++i) k_mix<OP1, N1, OP2, N2><<<grid, block>>>(c.in, c.out, c.cyc);     // warm-up
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < REPS; ++i) k_mix<OP1, N1, OP2, N2><<<grid, block>>>(c.in, c.out, c.cyc);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= REPS;
  long long* h = (long long*)malloc(sizeof(long long) * grid);
  CK(cudaMemcpy(h, c.cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost));
  double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid; free(h);
  // warp-instructions per CTA for each class
  const double warps = block / 32.0;
  const double w1 = (double)ITERS * 4 * N1 * CH * warps, w2 = (double)ITERS * 4 * N2 * CH * warps;
  // all resident CTAs of an SM run concurrently (ctas_per_sm*8 warps <= 64), so per-SM rate = ctas_per_sm * w / cycles
  const double r1 = c.ctas_per_sm * w1 / avg, r2 = c.ctas_per_sm * w2 / avg;
  const double ghz = (avg / (ms * 1e-3)) * 1e-9;
  // time-based absolute rates (what a roofline needs): thread-instructions per second over the whole chip,
  // and the same expressed per SM per clock of the max SM clock (clock64 need not tick at the SM clock)
  const double t1 = w1 * grid * 32.0 / (ms * 1e-3), t2 = w2 * grid * 32.0 / (ms * 1e-3);
  const double per = 1.0 / (32.0 * c.sms * c.max_hz);
  printf("{\"op1\":\"%s\",\"n1\":%d,\"op2\":\"%s\",\"n2\":%d,\"ms\":%.4f,\"cycles\":%.0f,\"clock64_ghz\":%.3f,"
         "\"winst_per_clk64_sm_total\":%.3f,\"tinst_per_s_op1\":%.4e,\"tinst_per_s_op2\":%.4e,"
         "\"winst_per_maxclk_sm_op1\":%.3f,\"winst_per_maxclk_sm_op2\":%.3f,\"winst_per_maxclk_sm_total\":%.3f}\n",
         op_name[OP1], N1, op_name[OP2], N2, ms, avg, ghz, r1 + r2, t1, t2, t1 * per, t2 * per, (t1 + t2) * per);
  fflush(stdout);
  CK(cudaEventDestroy(e0)); CK(cudaEventDestroy(e1));
}

int main(int argc, char** argv) {
  int dev = 0; CK(cudaSetDevice(dev));
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
  Ctx c; c.sms = p.multiProcessorCount; c.max_hz = p.clockRate * 1e3; c.ctas_per_sm = (argc > 1) ? atoi(argv[1]) : 4;   // 4 x 256 thr = 32 warps/SM
  fprintf(stderr, "device %s, %d SMs, clock %d kHz, ctas/SM %d\n", p.name, c.sms, p.clockRate, c.ctas_per_sm);
  unsigned h[1024 + 64];
  srand(12345);
  for (int i = 0; i < 1024 + 64; ++i) h[i] = ((unsigned)(rand() & 0x3ff)) | ((unsigned)(rand() & 0x3ff) << 16);
  h[256] = 1u;   // multiplier for IMAD / selector for IDP
  CK(cudaMalloc(&c.in, sizeof(h))); CK(cudaMemcpy(c.in, h, sizeof(h), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&c.out, sizeof(unsigned) * c.sms * 8 * 256));
  CK(cudaMalloc(&c.cyc, sizeof(long long) * c.sms * 8));
  // singles
  run<OP_VIADD2, 1, OP_NONE, 0>(c);
  run<OP_VIMNMX3R, 1, OP_NONE, 0>(c);
  run<OP_VIADDMNMX, 1, OP_NONE, 0>(c);
  run<OP_VIADDMNMXR, 1, OP_NONE, 0>(c);
  run<OP_VIMNMX3_32, 1, OP_NONE, 0>(c);
  run<OP_VIADDMNMX_32, 1, OP_NONE, 0>(c);
  run<OP_IADD3, 1, OP_NONE, 0>(c);
  run<OP_IMAD, 1, OP_NONE, 0>(c);
  run<OP_LOP3, 1, OP_NONE, 0>(c);
  run<OP_PRMT, 1, OP_NONE, 0>(c);
  run<OP_IDP4A, 1, OP_NONE, 0>(c);
  run<OP_IDP2A, 1, OP_NONE, 0>(c);
  run<OP_FFMA, 1, OP_NONE, 0>(c);
  run<OP_HADD2, 1, OP_NONE, 0>(c);
  run<OP_HMNMX2, 1, OP_NONE, 0>(c);
  run<OP_LDS, 1, OP_NONE, 0>(c);
  run<OP_LDS_RAND, 1, OP_NONE, 0>(c);
  run<OP_SHFL, 1, OP_NONE, 0>(c);
  // pairs: does class 2 issue in the shadow of the DPX pipe?
  run<OP_VIADDMNMX, 1, OP_VIMNMX3R, 1>(c);
  run<OP_VIADDMNMX, 1, OP_IMAD, 1>(c);
  run<OP_VIADDMNMX, 2, OP_IMAD, 1>(c);
  run<OP_VIADDMNMX, 1, OP_IADD3, 1>(c);
  run<OP_VIADDMNMX, 1, OP_PRMT, 1>(c);
  run<OP_VIADDMNMX, 1, OP_IDP4A, 1>(c);
  run<OP_VIADDMNMX, 2, OP_IDP4A, 1>(c);
  run<OP_VIADDMNMX, 1, OP_FFMA, 1>(c);
  run<OP_VIADDMNMX, 1, OP_HADD2, 1>(c);
  run<OP_VIADDMNMX, 1, OP_HMNMX2, 1>(c);
  run<OP_VIADDMNMX, 2, OP_LDS, 1>(c);
  run<OP_VIADDMNMX, 4, OP_LDS, 1>(c);
  run<OP_VIADDMNMX, 2, OP_SHFL, 1>(c);
  run<OP_VIADDMNMX, 4, OP_SHFL, 1>(c);
  run<OP_LDS, 1, OP_SHFL, 1>(c);
  run<OP_IMAD, 1, OP_FFMA, 1>(c);
  run<OP_IMAD, 1, OP_IDP4A, 1>(c);
  run<OP_HADD2, 1, OP_HMNMX2, 1>(c);
  return 0;
}
