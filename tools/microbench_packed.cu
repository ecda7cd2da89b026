// tools/microbench_packed.cu -- latency and throughput of scalar and packed FP32 operations on the device (sm_100a):
//   FFMA / FMUL / FADD and their packed forms FFMA2 / FMUL2 (two IEEE operations per lane and issue slot).
// latency: one warp, one dependent chain, cycles per operation (clock64 around 4096 operations)
// throughput: 148 x 8 CTAs x 256 threads, 8 independent chains per thread, lane-operations per cycle and SM
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/microbench_packed.cu -o scratch/microbench_packed
#include <cstdio>
#include <cuda_runtime.h>

template<int OP> __device__ __forceinline__ void step(float2 &a, float2 m, float2 c) {
  if(OP == 0) { a.x = __fmaf_rn(a.x, m.x, c.x); }
  if(OP == 1) { a.x = __fmul_rn(a.x, m.x); }
  if(OP == 2) { a.x = __fadd_rn(a.x, c.x); }
  if(OP == 3) { a = __ffma2_rn(a, m, c); }
  if(OP == 4) { a = __fmul2_rn(a, m); }
}

template<int OP> __global__ void latency(float2 *out, float2 m, float2 c, long long *cycles) {
  float2 a = make_float2(threadIdx.x * 1e-3f, 1.0f);
  const long long t0 = clock64();
#pragma unroll 64
  for(int i = 0; i < 4096; ++i) { step<OP>(a, m, c); }
  const long long t1 = clock64();
  if(threadIdx.x == 0) { cycles[0] = t1 - t0; }
  out[threadIdx.x] = a;
}

template<int OP> __global__ void throughput(float2 *out, float2 m, float2 c, int iters) {
  float2 a[8];
#pragma unroll
  for(int k = 0; k < 8; ++k) { a[k] = make_float2(threadIdx.x * 1e-3f + k, 1.0f + k); }
  for(int i = 0; i < iters; ++i) {
#pragma unroll
    for(int u = 0; u < 8; ++u) {
#pragma unroll
      for(int k = 0; k < 8; ++k) { step<OP>(a[k], m, c); }
    }
  }
  float2 r = make_float2(0.0f, 0.0f);
#pragma unroll
  for(int k = 0; k < 8; ++k) { r.x += a[k].x; r.y += a[k].y; }
  if(r.x == 123.456f) { out[0] = r; }
}

template<int OP> void run(const char *name, int ops_per_instruction) {
  float2 *out; long long *cycles; long long host = 0;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cycles, 8);
  const float2 m = make_float2(0.999f, 1.001f), c = make_float2(1e-4f, -1e-4f);
  latency<OP><<<1, 32>>>(out, m, c, cycles);
  cudaMemcpy(&host, cycles, 8, cudaMemcpyDeviceToHost);
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int iters = 4096, grid = prop.multiProcessorCount * 8;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  throughput<OP><<<grid, 256>>>(out, m, c, iters);
  cudaEventRecord(a);
  throughput<OP><<<grid, 256>>>(out, m, c, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms = 0.0f; cudaEventElapsedTime(&ms, a, b);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double instr = (double)grid * 256.0 * iters * 64.0;                // thread instructions
  const double per_cycle_sm = instr / (ms * 1e-3 * khz * 1e3) / prop.multiProcessorCount;
  printf("%-6s latency %.2f cycles per dependent instruction | %.1f thread-instructions per cycle and SM = %.1f lane-operations per cycle and SM (%.1f ms)\n",
         name, host / 4096.0, per_cycle_sm, per_cycle_sm * ops_per_instruction, ms);
}

int main() {
  run<0>("FFMA", 1); run<1>("FMUL", 1); run<2>("FADD", 1); run<3>("FFMA2", 2); run<4>("FMUL2", 2);
  return 0;
}
