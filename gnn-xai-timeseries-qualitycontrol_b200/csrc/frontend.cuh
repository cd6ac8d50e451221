// Kernels either side of the LSTM stack: layout conversion into sequence tiles, the CML graph
// front-end (GeneralConv + per-window node pooling + concat), and the dense head.
#pragma once
#include <cstdint>

#include "lstm_fwd.cuh"

namespace gnnqc {

constexpr int kGcnChannelsDev = 16;  // GeneralConv channels (model_config.yml graph_convolution.units)

// ---- [n_seq, T, C] row-major  ->  sequence tiles [tile][t][c][64], t < t_run ---------------
// Input contract of the baseline models: libs/preprocessing_functions.py:756-760 (CML,
// [B,181,2]) and graph_reshape's output (libs/create_model.py:242-258, SoilNet [S,337,3]).
template <int C>
__global__ void __launch_bounds__(256) seq_to_tiles_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                           long long n_seq, int T, int t_run) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long seq = tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    float v[C];
#pragma unroll
    for (int c = 0; c < C; ++c) v[c] = (seq < n_seq) ? __ldg(x + (seq * T + t) * C + c) : 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) out[((tile * t_run + t) * C + c) * kTile + b] = v[c];
  }
}

// ---- CML node coefficients -----------------------------------------------------------------
// GeneralConv aggregates with a mean over each receiving node's edges (spektral 1.3.0
// MessagePassing: gather by a.indices[:,1], unsorted_segment_mean by a.indices[:,0]; call
// site libs/create_model.py:185-189,219) and timeseries_pooling then averages over the
// window's nodes (libs/create_model.py:23-27).  Both are linear, so
//   pooled[t] = sum_j coef_j * z_j[t],   coef_j = (1/N) * sum_{(i,j) in E} 1/deg_i,
// with deg_i the number of edges received by node i.  One warp per window; `deg` is integer
// scratch that must be zero on entry.
static __global__ void __launch_bounds__(256) cml_node_coef_kernel(const long long* __restrict__ edges,
                                                            const int* __restrict__ node_off,
                                                            const long long* __restrict__ edge_off,
                                                            int* __restrict__ deg, float* __restrict__ coef,
                                                            long long n_windows) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_windows) return;
  const int n0 = node_off[w], nn = node_off[w + 1] - n0;
  const long long e0 = edge_off[w], ne = edge_off[w + 1] - e0;
  for (long long e = lane; e < ne; e += 32) {
    const long long row = edges[2 * (e0 + e)];
    if (row >= 0 && row < nn) atomicAdd(&deg[n0 + row], 1);
  }
  __syncwarp();
  __threadfence_block();
  const float inv_n = nn > 0 ? 1.f / static_cast<float>(nn) : 0.f;
  for (int j = lane; j < nn; j += 32) {
    float s = 0.f;
    for (long long e = 0; e < ne; ++e) {
      const long long row = edges[2 * (e0 + e)], col = edges[2 * (e0 + e) + 1];
      if (col == j && row >= 0 && row < nn) s += 1.f / static_cast<float>(__ldcg(deg + n0 + row));
    }
    coef[n0 + j] = s * inv_n;
  }
}

// BatchNorm (inference) folded into the GeneralConv projection: z = prelu(x0*a0 + x1*a1 (+ x2*a2) + c).
struct GcnFolded {
  float a[3][kGcnChannelsDev];
  float c[kGcnChannelsDev];
  float alpha[kGcnChannelsDev];
};

// ---- CML front-end: features (b,t,n) + coef + target series -> x1 tiles [tile][t][18][64] --
// Restates GCNClassifier.call for 'cml' up to the Concatenate (libs/create_model.py:217-223):
// channels 0..1 = anomalous_sensor_features, 2..17 = pooled GeneralConv output.
static __global__ void __launch_bounds__(256) cml_frontend_kernel(const float2* __restrict__ feat, const float2* __restrict__ anom,
                                                           const float* __restrict__ coef, const int* __restrict__ node_off,
                                                           const GcnFolded p, float* __restrict__ out,
                                                           long long n_windows, int T, int t_run) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long win = tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
  const bool live = win < n_windows;
  const int n0 = live ? node_off[win] : 0;
  const int nn = live ? node_off[win + 1] - n0 : 0;
  constexpr int CH = 2 + kGcnChannelsDev;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    float acc[kGcnChannelsDev];
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) acc[c] = 0.f;
    const float2* row = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn);
    for (int n = 0; n < nn; ++n) {
      const float2 v = __ldg(row + n);
      const float cf = __ldg(coef + n0 + n);
#pragma unroll
      for (int c = 0; c < kGcnChannelsDev; ++c) {
        float z = fmaf(v.x, p.a[0][c], fmaf(v.y, p.a[1][c], p.c[c]));
        z = fmaxf(z, 0.f) + p.alpha[c] * fminf(z, 0.f);
        acc[c] = fmaf(cf, z, acc[c]);
      }
    }
    const float2 av = live ? __ldg(anom + win * T + t) : make_float2(0.f, 0.f);
    float* o = out + ((tile * t_run + t) * CH) * kTile + b;
    o[0] = av.x;
    o[kTile] = av.y;
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) o[(2 + c) * kTile] = acc[c];
  }
}

// ---- dense head: [tile][128][64] -> sigmoid(Dense(LeakyReLU(Dense(LeakyReLU(Dense))))) ------
// libs/create_model.py:235-239 (GCN) and :372-376 (baseline).  One CTA per tile.
template <int HL, int D>
__global__ void __launch_bounds__(256) head_kernel(const float* __restrict__ y7, const float* __restrict__ w1,
                                                   const float* __restrict__ b1, const float* __restrict__ w2,
                                                   const float* __restrict__ b2, const float* __restrict__ wo,
                                                   const float* __restrict__ bo, float leaky, float* __restrict__ scores,
                                                   long long n_seq) {
  __shared__ float d1_s[D * kTile];
  __shared__ float d2_s[D * kTile];
  const long long tile = blockIdx.x;
  const int b = threadIdx.x & 63, q = threadIdx.x >> 6;  // q: which quarter of the D outputs
  constexpr int DQ = D / 4;
  float acc[DQ];
#pragma unroll
  for (int o = 0; o < DQ; ++o) acc[o] = __ldg(b1 + q * DQ + o);
  for (int k = 0; k < HL; ++k) {
    const float h = __ldg(y7 + (tile * HL + k) * kTile + b);
#pragma unroll
    for (int o = 0; o < DQ; ++o) acc[o] = fmaf(h, __ldg(w1 + k * D + q * DQ + o), acc[o]);
  }
#pragma unroll
  for (int o = 0; o < DQ; ++o) d1_s[(q * DQ + o) * kTile + b] = acc[o] > 0.f ? acc[o] : leaky * acc[o];
  __syncthreads();
#pragma unroll
  for (int o = 0; o < DQ; ++o) acc[o] = __ldg(b2 + q * DQ + o);
  for (int k = 0; k < D; ++k) {
    const float h = d1_s[k * kTile + b];
#pragma unroll
    for (int o = 0; o < DQ; ++o) acc[o] = fmaf(h, __ldg(w2 + k * D + q * DQ + o), acc[o]);
  }
#pragma unroll
  for (int o = 0; o < DQ; ++o) d2_s[(q * DQ + o) * kTile + b] = acc[o] > 0.f ? acc[o] : leaky * acc[o];
  __syncthreads();
  if (q == 0) {
    float z = __ldg(bo);
    for (int k = 0; k < D; ++k) z = fmaf(d2_s[k * kTile + b], __ldg(wo + k), z);
    const long long seq = tile * kTile + b;
    if (seq < n_seq) scores[seq] = 1.f / (1.f + expf(-z));
  }
}

}  // namespace gnnqc
