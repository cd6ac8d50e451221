// FP32 pipe micro-benchmark: the roofline denominator for the fp32 LSTM kernels is not in
// MEASURED_PEAKS.json (which has HBM and bf16 tensor peaks), so bench.py measures it here.
// mode 0: scalar FFMA chains; mode 1: packed fma.rn.f32x2 (FFMA2) chains.
#include <cuda_runtime.h>

#include "../../include/gnnqc.h"

namespace {

constexpr int kChains = 16;

template <int MODE>
__global__ void __launch_bounds__(256) fma_probe_kernel(float* out, int iters, float b, float c) {
  float a[kChains];
#pragma unroll
  for (int i = 0; i < kChains; ++i) a[i] = 0.001f * static_cast<float>(threadIdx.x + i);
  if (MODE == 0) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < kChains; ++i) a[i] = fmaf(a[i], b, c);
    }
  } else {
    unsigned long long p[kChains / 2], bb, cc;
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
    asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
#pragma unroll
    for (int i = 0; i < kChains / 2; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < kChains / 2; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(bb), "l"(cc));
    }
#pragma unroll
    for (int i = 0; i < kChains / 2; ++i) asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * i]), "=f"(a[2 * i + 1]) : "l"(p[i]));
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kChains; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

extern "C" int gnnqc_probe_fp32_peak(int device, int mode, double* tflops_out) {
  if (!tflops_out || mode < 0 || mode > 1) return GNNQC_ERR_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return GNNQC_ERR_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return GNNQC_ERR_CUDA;
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 20000;
  float* out = nullptr;
  if (cudaMalloc(&out, sizeof(float) * blocks * threads) != cudaSuccess) return GNNQC_ERR_CUDA;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    if (mode == 0) fma_probe_kernel<0><<<blocks, threads>>>(out, iters, 0.9999f, 0.0001f);
    else fma_probe_kernel<1><<<blocks, threads>>>(out, iters, 0.9999f, 0.0001f);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return GNNQC_ERR_CUDA; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * kChains * double(iters) * double(blocks) * threads;
    if (rep > 0 && ms > 0.f) best = best > flops / (ms * 1e-3) ? best : flops / (ms * 1e-3);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  *tflops_out = best * 1e-12;
  return GNNQC_OK;
}
