// Self-test of the tcgen05 building blocks (csrc/umma.cuh): one 128 x 64 x 48 bf16 GEMM through
// TMEM, checked on the host against a plain float loop.  Exercised by tests/test_gpu_umma.py before
// anything else relies on the descriptor encodings.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <vector>

#include "../../include/gnnqc.h"
#include "umma.cuh"

namespace {

constexpr int M = 128, N = 64, K = 48;

__global__ void __launch_bounds__(128) umma_selftest_kernel(const __nv_bfloat16* __restrict__ A,   // [M][K] row-major
                                                            const __nv_bfloat16* __restrict__ B,   // [N][K] row-major
                                                            float* __restrict__ D) {               // [M][N]
  using namespace gnnqc;
  __shared__ __align__(128) __nv_bfloat16 As[(K / 8) * M * 8];
  __shared__ __align__(128) __nv_bfloat16 Bs[(K / 8) * N * 8];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) umma::tmem_alloc(&tmem_base, 64);
  if (tid == 0) {
    ptx::mbar_init(&bar, 1);
    ptx::mbar_init_fence();
  }
  for (int i = tid; i < M * K; i += 128) {
    const int r = i / K, k = i % K;
    As[(k / 8) * M * 8 + r * 8 + (k % 8)] = A[i];
  }
  for (int i = tid; i < N * K; i += 128) {
    const int r = i / K, k = i % K;
    Bs[(k / 8) * N * 8 + r * 8 + (k % 8)] = B[i];
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = tmem_base;
  if (tid == 0) {
    constexpr uint32_t idesc = umma::instr_desc_bf16(M, N);
    for (int kc = 0; kc < K / 16; ++kc) {
      const uint64_t da = umma::smem_desc(As + kc * 2 * M * 8, M * 16, 128);
      const uint64_t db = umma::smem_desc(Bs + kc * 2 * N * 8, N * 16, 128);
      umma::mma_bf16(tbase, da, db, idesc, kc > 0);
    }
    umma::commit(&bar);
  }
  ptx::mbar_wait(&bar, 0);
  umma::fence_after_sync();
  const int row = warp * 32 + (tid & 31);
#pragma unroll
  for (int c = 0; c < N; c += 16) {
    float v[16];
    umma::tmem_ld16(tbase + (static_cast<uint32_t>(warp * 32) << 16) + c, v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D[row * N + c + i] = v[i];
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, 64);
}

}  // namespace

extern "C" int gnnqc_selftest_umma(int device, float* max_abs_err) {
  if (!max_abs_err) return GNNQC_ERR_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return GNNQC_ERR_CUDA;
  std::vector<__nv_bfloat16> hA(M * K), hB(N * K);
  std::vector<float> fA(M * K), fB(N * K), want(M * N, 0.f), got(M * N, -1.f);
  unsigned s = 12345u;
  auto rnd = [&s]() { s = s * 1664525u + 1013904223u; return (static_cast<float>((s >> 8) & 0xFFFF) / 65536.f) * 2.f - 1.f; };
  for (int i = 0; i < M * K; ++i) { hA[i] = __float2bfloat16(rnd()); fA[i] = __bfloat162float(hA[i]); }
  for (int i = 0; i < N * K; ++i) { hB[i] = __float2bfloat16(rnd()); fB[i] = __bfloat162float(hB[i]); }
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      float acc = 0.f;
      for (int k = 0; k < K; ++k) acc += fA[m * K + k] * fB[n * K + k];
      want[m * N + n] = acc;
    }
  __nv_bfloat16 *dA = nullptr, *dB = nullptr;
  float* dD = nullptr;
  cudaMalloc(&dA, hA.size() * 2);
  cudaMalloc(&dB, hB.size() * 2);
  cudaMalloc(&dD, got.size() * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xFF, got.size() * 4);
  umma_selftest_kernel<<<1, 128>>>(dA, dB, dD);
  cudaError_t err = cudaDeviceSynchronize();
  if (err == cudaSuccess) err = cudaMemcpy(got.data(), dD, got.size() * 4, cudaMemcpyDeviceToHost);
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  if (err != cudaSuccess) return GNNQC_ERR_CUDA;
  float worst = 0.f;
  for (int i = 0; i < M * N; ++i) {
    const float e = std::fabs(got[i] - want[i]);
    worst = (e > worst || std::isnan(e)) ? (std::isnan(e) ? 1e30f : e) : worst;
  }
  *max_abs_err = worst;
  return GNNQC_OK;
}
