// Shared arithmetic of the SoilNet graph front-end (GeneralConv with mean aggregation over a sensor's
// neighbours, reference libs/create_model.py:173-189,224-231), used by the materialised-window kernel
// (soilnet.cu) and by the resident-series one (soil_series.cu) so that both give the same bits.
// prelu(z) = k1*z + k2*|z| (k1 = (1+alpha)/2, k2 = (1-alpha)/2) and z is affine in the three features, so the
// neighbour sum needs |z| per (edge, channel) only: z two channels at a time (fma.rn.f32x2), one add with |z| as
// operand modifier; the linear part closes from the three feature sums.
#pragma once
#include "frontend.cuh"

namespace gnnqc {

struct SoilAcc {
  float sabs[kGcnChannelsDev];
  float s0 = 0.f, s1 = 0.f, s2 = 0.f;
  __device__ __forceinline__ SoilAcc() {
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) sabs[c] = 0.f;
  }
  __device__ __forceinline__ void add(const GcnFolded& p, float v0, float v1, float v2) {
    s0 += v0; s1 += v1; s2 += v2;
    unsigned long long x0, x1, x2;
    asm("mov.b64 %0, {%1, %1};" : "=l"(x0) : "f"(v0));
    asm("mov.b64 %0, {%1, %1};" : "=l"(x1) : "f"(v1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(x2) : "f"(v2));
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; c += 2) {
      unsigned long long z, a0, a1, a2;
      asm("mov.b64 %0, {%1, %2};" : "=l"(z) : "f"(p.c[c]), "f"(p.c[c + 1]));
      asm("mov.b64 %0, {%1, %2};" : "=l"(a0) : "f"(p.a[0][c]), "f"(p.a[0][c + 1]));
      asm("mov.b64 %0, {%1, %2};" : "=l"(a1) : "f"(p.a[1][c]), "f"(p.a[1][c + 1]));
      asm("mov.b64 %0, {%1, %2};" : "=l"(a2) : "f"(p.a[2][c]), "f"(p.a[2][c + 1]));
      asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(z) : "l"(x2), "l"(a2));
      asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(z) : "l"(x1), "l"(a1));
      asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(z) : "l"(x0), "l"(a0));
      float z0, z1;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(z0), "=f"(z1) : "l"(z));
      sabs[c] += fabsf(z0);
      sabs[c + 1] += fabsf(z1);
    }
  }
  // mean over the `d` neighbours added so far (inv = 1/d, 0 for an isolated sensor)
  __device__ __forceinline__ float mean(const GcnFolded& p, int c, float d, float inv) const {
    const float lin = fmaf(p.a[0][c], s0, fmaf(p.a[1][c], s1, fmaf(p.a[2][c], s2, p.c[c] * d)));
    return fmaf(0.5f * (1.f + p.alpha[c]), lin, 0.5f * (1.f - p.alpha[c]) * sabs[c]) * inv;
  }
};

}  // namespace gnnqc
