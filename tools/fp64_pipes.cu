// Are the FP64 vector pipe (DFMA) and the FP64 tensor pipe (DMMA, mma.sync m8n8k4 f64) of a B200 SM separate
// resources or one?  Times DFMA-only, DMMA-only and two mixes (same warp / alternate warps) of the SAME per-kind
// instruction counts: mixed ~ max(a, b) -> the pipes overlap; mixed ~ a + b -> one shared resource.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipes tools/fp64_pipes.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// MODE 0: DFMA only; 1: DMMA only; 2: both in every warp; 3: even warps DFMA, odd warps DMMA (each doing 2x)
template <int MODE>
__global__ void __launch_bounds__(256) k(double* out, int iters, double a, double b) {
  double x[16], c[8][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3 + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
  const bool odd = (threadIdx.x >> 5) & 1;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 2 || (MODE == 3 && !odd)) {
      // 16 warp-level DFMA = 16 x 32 FMA = 512 FMA
#pragma unroll
      for (int r = 0; r < (MODE == 3 ? 2 : 1); ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = fma(x[i], a, b);
    }
    if (MODE == 1 || MODE == 2 || (MODE == 3 && odd)) {
      // 2 warp-level DMMA = 2 x 256 FMA = 512 FMA  (same FMA count as the DFMA block)
#pragma unroll
      for (int r = 0; r < (MODE == 3 ? 2 : 1); ++r)
#pragma unroll
        for (int i = 0; i < 2; ++i) dmma(c[i + 2 * r][0], c[i + 2 * r][1], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
float run(double* out, int blocks, int iters) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount, iters = 40000;
  double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  for (int per_sm = 2; per_sm <= 8; per_sm *= 2) {
    const int blocks = sms * per_sm;
    const float t0 = run<0>(out, blocks, iters), t1 = run<1>(out, blocks, iters), t2 = run<2>(out, blocks, iters),
                t3 = run<3>(out, blocks, iters);
    const double fma_per = 512.0 * iters * blocks * 8;   // FMAs of ONE kind in modes 0, 1, 2, 3
    printf("{\"ctas_per_sm\": %d, \"dfma_ms\": %.3f, \"dmma_ms\": %.3f, \"mixed_same_warp_ms\": %.3f, \"mixed_alt_warps_ms\": %.3f, "
           "\"dfma_tflops\": %.2f, \"dmma_tflops\": %.2f, \"mixed_same_warp_tflops\": %.2f, \"mixed_alt_warps_tflops\": %.2f}\n",
           per_sm, t0, t1, t2, t3, 2 * fma_per / t0 / 1e9, 2 * fma_per / t1 / 1e9, 4 * fma_per / t2 / 1e9, 4 * fma_per / t3 / 1e9);
  }
  printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
