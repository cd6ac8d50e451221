// Micro-benchmark (scratch tool, not part of the library): how many LSTM cell updates per clock one SM sustains with
// the cell arithmetic of lstm_tc.cuh alone - no tensor core, no barriers.  16 warps of 128-register threads per SM,
// like the scoring kernels.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DGNNQC_POLY_EXPS=k
#include <cstdio>
#include <cuda_runtime.h>
#include "../../gnn-xai-timeseries-qualitycontrol_b200/csrc/lstm_tc.cuh"
using namespace gnnqc;

template <int UNITS, bool SPLIT>
__global__ void __launch_bounds__(512, 1) cell_kernel(float* out, int steps, float s0) {
  __shared__ uint4 sink[2][512];
  f32x2 c[UNITS / 2];
  float h[UNITS];
#pragma unroll
  for (int j = 0; j < UNITS / 2; ++j) c[j] = 0ull;
#pragma unroll
  for (int j = 0; j < UNITS; ++j) h[j] = 0.001f * (threadIdx.x + j);
  for (int t = 0; t < steps; ++t) {
#pragma unroll
    for (int b8 = 0; b8 < UNITS / 8; ++b8) {
      float hv[8];
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        // pre-activations that depend on the previous outputs (as in the recurrence), cheap to form
        const float a0 = h[b8 * 8 + i] * s0, a1 = h[b8 * 8 + i + 1] * s0;
        f32x2 sv[5];
        up2(lstm_cell2<false>(pk2(a0 + 0.3f, a1 - 0.2f), pk2(a0 - 1.1f, a1 + 0.7f), pk2(a1 + 0.1f, a0 - 0.4f), pk2(a1 - 0.6f, a0 + 0.9f), c[b8 * 4 + i / 2], sv),
            hv[i], hv[i + 1]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) h[b8 * 8 + i] = hv[i];
      if (SPLIT) {
        uint4 hi, lo;
        split8(hv, hi, lo);
        sink[0][threadIdx.x] = hi;
        sink[1][threadIdx.x] = lo;
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < UNITS; ++j) s += h[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + (SPLIT ? __uint_as_float(sink[0][(threadIdx.x + 1) & 511].x) : 0.f);
}

int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  float* out;
  cudaMalloc(&out, sizeof(float) * prop.multiProcessorCount * 512);
  const int steps = 4000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](auto kernel, const char* name, int units) {
    kernel<<<prop.multiProcessorCount, 512>>>(out, 10, 0.5f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    kernel<<<prop.multiProcessorCount, 512>>>(out, steps, 0.5f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    int khz;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double cells_per_smsp = double(steps) * units * 4;          // 4 warps per scheduler
    const double cycles = ms * 1e-3 * khz * 1e3;
    printf("poly=%d %-22s %.3f ms  %.1f cycles per warp-cell per scheduler (nominal clock %d MHz)  err=%s\n", GNNQC_POLY_EXPS, name, ms, cycles / cells_per_smsp,
           khz / 1000, cudaGetErrorString(cudaGetLastError()));
  };
  run(cell_kernel<16, true>, "16 units + split", 16);
  run(cell_kernel<16, false>, "16 units", 16);
  run(cell_kernel<32, true>, "32 units + split", 32);
  return 0;
}
