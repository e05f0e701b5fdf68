// FP64 roofline denominators measured on the device the bench runs on (MEASURED_PEAKS.json carries no fp64 figure):
// dependent-free DFMA chains (vector pipe) and mma.sync m8n8k4 f64 (DMMA, the Blackwell FP64 tensor path: tcgen05
// has no f64 kind).  On B200 both run on ONE pipe (tools/fp64_pipes.cu: a mix of the two takes the SUM of their
// times), so the higher of the two is the FP64 peak a kernel that uses both can reach.
#include "../../include/b2ba.h"

#include <cuda_runtime.h>

namespace {

__global__ void __launch_bounds__(512) peak_dfma_kernel(double* out, int iters, double a, double b) {
  double x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = fma(x[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(512) peak_dmma_kernel(double* out, int iters, double a, double b) {
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

extern "C" int b2ba_peak_flops(int device, double* dfma_tflops, double* dmma_tflops) {
  if (!dfma_tflops || !dmma_tflops) return B2BA_E_INVALID;
  int prev = -1, sms = 0;
  if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) return B2BA_E_CUDA;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const int thr = 512, blocks = sms * 4, iters = 4000;
  double* out = nullptr;
  cudaEvent_t e0, e1;
  int rc = B2BA_OK;
  if (cudaMalloc(&out, sizeof(double) * blocks * thr) != cudaSuccess) rc = B2BA_E_CUDA;
  if (rc == B2BA_OK && (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess)) rc = B2BA_E_CUDA;
  for (int which = 0; which < 2 && rc == B2BA_OK; ++which) {
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {      // first repetition warms the clocks up
      cudaEventRecord(e0);
      if (which == 0) peak_dfma_kernel<<<blocks, thr>>>(out, iters, 1.0000001, 1e-9);
      else peak_dmma_kernel<<<blocks, thr>>>(out, iters, 1.0000001, 1e-9);
      cudaEventRecord(e1);
      if (cudaEventSynchronize(e1) != cudaSuccess) { rc = B2BA_E_CUDA; break; }
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep > 0 && ms < best) best = ms;
    }
    const double flops = which == 0 ? 2.0 * 16 * iters * (double)blocks * thr
                                    : 2.0 * 8 * 8 * 4 * 8 * iters * (double)blocks * (thr / 32);
    (which == 0 ? *dfma_tflops : *dmma_tflops) = flops / best / 1e9;
  }
  if (out) cudaFree(out);
  if (prev >= 0) cudaSetDevice(prev);
  return rc;
}
