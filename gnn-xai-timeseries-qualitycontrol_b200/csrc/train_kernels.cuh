// Kernels used only by the training step (csrc/train.cu).
#pragma once
#include <cstdint>

#include "frontend.cuh"
#include "lstm_fwd.cuh"

namespace gnnqc {

// ---- weight gradients ------------------------------------------------------------------------
// G[a][c] = sum over items i, windows b < 64 of A_i[a][b] * B_i[c][b], where item i = (tile, step)
// and A_i / B_i are [rows][64] blocks of sequence tiles:
//   A_i = A + tile*a_tile + step*a_step,   B_i = B + tile*b_tile + step*b_step.
// With BIAS the output has one extra row NA holding sum_i sum_b B_i[c][b].
// This is how dW = X^T dZ, dU = H_prev^T dZ, db = sum dZ of every LSTM layer and the dense-layer
// gradients are formed.  Two stages keep it deterministic: each CTA reduces a slice of the items
// into partial[slice][a][c]; wgrad_reduce_kernel then sums the slices in order.
// CTA = 64 threads computing a 32 x 64 block; thread = 4 rows x 8 columns, packed FFMA2 with the
// two lanes of each pair holding the even / odd windows' partial sums.
constexpr int kWgA = 32, kWgC = 64, kWgStride = kTile + 4;
constexpr size_t kWgradSmem = size_t(2) * (kWgA + kWgC) * kWgStride * sizeof(float);

struct WgradArgs {
  const float* A;
  const float* B;
  long long a_tile, a_step, b_tile, b_step;
  int NA, NC;        // valid rows of A / B blocks
  int n_steps;       // items per tile
  long long n_items; // tiles * n_steps
  int bias;          // 1: append the ones row
  float* partial;    // [slices][NA + bias][NC]
};

using ptx::cp_async16;
using ptx::cp_async_commit;
using ptx::cp_async_wait;

static __global__ void __launch_bounds__(64) wgrad_kernel(const WgradArgs g) {
  extern __shared__ __align__(16) float wg_smem[];
  float* As[2] = {wg_smem, wg_smem + (kWgA + kWgC) * kWgStride};
  float* Bs[2] = {As[0] + kWgA * kWgStride, As[1] + kWgA * kWgStride};
  const int tid = threadIdx.x, ta = tid & 7, tc = tid >> 3;
  const int a0 = blockIdx.x * kWgA, c0 = blockIdx.y * kWgC;
  const int n_slices = gridDim.z, slice = blockIdx.z;
  const int rows_out = g.NA + g.bias;
  // zero both stages once (rows that are never loaded must read as 0), then the ones row
  for (int i = tid; i < 2 * (kWgA + kWgC) * kWgStride; i += 64) wg_smem[i] = 0.f;
  __syncthreads();
  if (g.bias) {
    const int r = g.NA - a0;   // local index of the ones row, if it falls in this block
    if (r >= 0 && r < kWgA)
      for (int i = tid; i < kTile; i += 64) { As[0][r * kWgStride + i] = 1.f; As[1][r * kWgStride + i] = 1.f; }
  }
  const long long per = (g.n_items + n_slices - 1) / n_slices;
  const long long i_begin = slice * per, i_end = (i_begin + per < g.n_items) ? i_begin + per : g.n_items;

  auto stage_load = [&](long long item, int s) {
    const long long tile = item / g.n_steps, step = item % g.n_steps;
    const float* Ai = g.A + tile * g.a_tile + step * g.a_step;
    const float* Bi = g.B + tile * g.b_tile + step * g.b_step;
    for (int i = tid; i < kWgA * (kTile / 4); i += 64) {
      const int r = i / (kTile / 4), q = i % (kTile / 4);
      if (a0 + r < g.NA) cp_async16(As[s] + r * kWgStride + 4 * q, Ai + static_cast<long long>(a0 + r) * kTile + 4 * q);
    }
    for (int i = tid; i < kWgC * (kTile / 4); i += 64) {
      const int r = i / (kTile / 4), q = i % (kTile / 4);
      if (c0 + r < g.NC) cp_async16(Bs[s] + r * kWgStride + 4 * q, Bi + static_cast<long long>(c0 + r) * kTile + 4 * q);
    }
    cp_async_commit();
  };

  unsigned long long acc[4][8];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[r][q] = 0ull;

  if (i_begin < i_end) stage_load(i_begin, 0);
  for (long long it = i_begin; it < i_end; ++it) {
    const int s = static_cast<int>((it - i_begin) & 1);
    if (it + 1 < i_end) { stage_load(it + 1, s ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const ulonglong2* A2 = reinterpret_cast<const ulonglong2*>(As[s]);
    const ulonglong2* B2 = reinterpret_cast<const ulonglong2*>(Bs[s]);
#pragma unroll 2
    for (int b4 = 0; b4 < kTile / 4; ++b4) {
      ulonglong2 av[4], bv[8];
#pragma unroll
      for (int r = 0; r < 4; ++r) av[r] = A2[(ta + 8 * r) * (kWgStride / 4) + b4];
#pragma unroll
      for (int q = 0; q < 8; ++q) bv[q] = B2[(tc + 8 * q) * (kWgStride / 4) + b4];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          ffma2(acc[r][q], av[r].x, bv[q].x);
          ffma2(acc[r][q], av[r].y, bv[q].y);
        }
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int a = a0 + ta + 8 * r;
    if (a >= rows_out) continue;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = c0 + tc + 8 * q;
      if (c >= g.NC) continue;
      float lo, hi;
      unpack2(acc[r][q], lo, hi);
      g.partial[(static_cast<long long>(slice) * rows_out + a) * g.NC + c] = lo + hi;
    }
  }
}

// out[a][c] = sum_s partial[s][a][c].  One warp per output element: lane l adds slices l, l+32, ...
// and the 32 partial sums are combined with a fixed shuffle tree, so the result does not depend on
// scheduling (deterministic) and no thread walks hundreds of slices serially.
static __global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partial, int slices, int rows, int cols,
                                                                  float* __restrict__ out_main, int main_rows, float* __restrict__ out_bias) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= rows * cols) return;
  float s = 0.f;
  for (int k = lane; k < slices; k += 32) s += partial[static_cast<long long>(k) * rows * cols + i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if (lane != 0) return;
  const int a = i / cols, c = i % cols;
  if (a < main_rows) out_main[i] = s;
  else if (out_bias != nullptr) out_bias[c] = s;
}

// ---- batch-norm statistics of the GeneralConv pre-activation -------------------------------------
// z = x W + b is linear in the two input channels, so its batch mean / variance over all node rows
// follow from five moments of x: sum x0, sum x1, sum x0^2, sum x1^2, sum x0 x1 (double precision).
static __global__ void __launch_bounds__(256) feat_moments_kernel(const float2* __restrict__ feat, long long n_rows, double* __restrict__ moments) {
  double s[5] = {0, 0, 0, 0, 0};
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_rows; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float2 v = __ldg(feat + i);
    const double a = v.x, b = v.y;
    s[0] += a; s[1] += b; s[2] += a * a; s[3] += b * b; s[4] += a * b;
  }
  __shared__ double red[5][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    double v = s[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) red[k][w] = v;
  }
  __syncthreads();
  if (threadIdx.x < 5) {
    double v = 0;
    for (int i = 0; i < 8; ++i) v += red[threadIdx.x][i];
    atomicAdd(&moments[threadIdx.x], v);
  }
}

// Batch statistics + folded projection for this step.  moments = {Sx0, Sx1, Sx00, Sx11, Sx01, N}.
struct BnBatch {
  float mu[kGcnChannelsDev], var[kGcnChannelsDev];
};
static __global__ void bn_prepare_kernel(const float* __restrict__ w, const float* __restrict__ bias, const float* __restrict__ prelu,
                                         const float* __restrict__ gamma, const float* __restrict__ beta, const double* __restrict__ m,
                                         float eps, GcnFolded* __restrict__ folded, BnBatch* __restrict__ stats) {
  const int c = threadIdx.x;
  if (c >= kGcnChannelsDev) return;
  const double n = m[5], w0 = w[c], w1 = w[kGcnChannelsDev + c], b = bias[c];
  const double lin = (w0 * m[0] + w1 * m[1]) / n;
  const double mu = lin + b;
  const double ez2 = (w0 * w0 * m[2] + w1 * w1 * m[3] + 2.0 * w0 * w1 * m[4]) / n + 2.0 * b * lin + b * b;
  double var = ez2 - mu * mu;
  if (var < 0) var = 0;
  const double scale = gamma[c] / sqrt(var + eps);
  folded->a[0][c] = static_cast<float>(w0 * scale);
  folded->a[1][c] = static_cast<float>(w1 * scale);
  folded->a[2][c] = 0.f;
  folded->c[c] = static_cast<float>((b - mu) * scale + beta[c]);
  folded->alpha[c] = prelu[c];
  stats->mu[c] = static_cast<float>(mu);
  stats->var[c] = static_cast<float>(var);
}

// ---- head, training mode ---------------------------------------------------------------------
// Forward, class-weighted BCE-with-logits, backward to dy7; leaves the activations and the
// pre-activation gradients of the three dense layers as [rows][64] tiles for wgrad_kernel.
// tiles_out (per tile, floats): d1[64][64] d2[64][64] dd1[64][64] dd2[64][64] dz3[1][64].
constexpr int kHeadTrainFloats = 4 * 64 * kTile + kTile;
template <int HL, int D>
__global__ void __launch_bounds__(256) head_train_kernel(const float* __restrict__ y7, const float* __restrict__ w1,
                                                         const float* __restrict__ b1, const float* __restrict__ w2,
                                                         const float* __restrict__ b2, const float* __restrict__ wo,
                                                         const float* __restrict__ bo, float leaky, const long long* __restrict__ labels,
                                                         float cw0, float cw1, float inv_batch, float* __restrict__ dy7,
                                                         float* __restrict__ tiles_out, float* __restrict__ loss_partial,
                                                         float* __restrict__ scores, long long n_seq) {
  __shared__ float d1_s[D * kTile];
  __shared__ float d2_s[D * kTile];
  __shared__ float m1_s[D * kTile];
  const long long tile = blockIdx.x;
  const int b = threadIdx.x & 63, q = threadIdx.x >> 6;
  constexpr int DQ = D / 4;
  float* out = tiles_out + tile * kHeadTrainFloats;
  float acc[DQ];
#pragma unroll
  for (int o = 0; o < DQ; ++o) acc[o] = __ldg(b1 + q * DQ + o);
  for (int k = 0; k < HL; ++k) {
    const float h = __ldg(y7 + (tile * HL + k) * kTile + b);
#pragma unroll
    for (int o = 0; o < DQ; ++o) acc[o] = fmaf(h, __ldg(w1 + k * D + q * DQ + o), acc[o]);
  }
#pragma unroll
  for (int o = 0; o < DQ; ++o) {
    const float v = acc[o] > 0.f ? acc[o] : leaky * acc[o];
    d1_s[(q * DQ + o) * kTile + b] = v;
    out[(0 * D + q * DQ + o) * kTile + b] = v;
    m1_s[(q * DQ + o) * kTile + b] = acc[o] > 0.f ? 1.f : leaky;
  }
  __syncthreads();
#pragma unroll
  for (int o = 0; o < DQ; ++o) acc[o] = __ldg(b2 + q * DQ + o);
  for (int k = 0; k < D; ++k) {
    const float h = d1_s[k * kTile + b];
#pragma unroll
    for (int o = 0; o < DQ; ++o) acc[o] = fmaf(h, __ldg(w2 + k * D + q * DQ + o), acc[o]);
  }
  float m2[DQ];
#pragma unroll
  for (int o = 0; o < DQ; ++o) {
    const float v = acc[o] > 0.f ? acc[o] : leaky * acc[o];
    d2_s[(q * DQ + o) * kTile + b] = v;
    out[(1 * D + q * DQ + o) * kTile + b] = v;
    m2[o] = acc[o] > 0.f ? 1.f : leaky;
  }
  __syncthreads();
  float z = __ldg(bo);
  for (int k = 0; k < D; ++k) z = fmaf(d2_s[k * kTile + b], __ldg(wo + k), z);
  const float pr = 1.f / (1.f + expf(-z));
  const long long seq = tile * kTile + b;
  const bool live = seq < n_seq;
  const float y = live ? static_cast<float>(labels[seq] > 0) : 0.f;
  const float sw = live ? (y > 0.f ? cw1 : cw0) : 0.f;
  // sigmoid_cross_entropy_with_logits: max(z,0) - z*y + log(1 + exp(-|z|))
  const float bce = fmaxf(z, 0.f) - z * y + log1pf(expf(-fabsf(z)));
  const float dz3 = sw * (pr - y) * inv_batch;
  if (q == 0) {
    out[4 * D * kTile + b] = dz3;
    if (scores != nullptr && live) scores[seq] = pr;
  }
  __syncthreads();   // all reads of d2_s (the logit) are done before it is overwritten
#pragma unroll
  for (int o = 0; o < DQ; ++o) {
    const float v = dz3 * __ldg(wo + q * DQ + o) * m2[o];
    d2_s[(q * DQ + o) * kTile + b] = v;
    out[(3 * D + q * DQ + o) * kTile + b] = v;
  }
  __syncthreads();
#pragma unroll
  for (int kk = 0; kk < DQ; ++kk) {
    const int k = q * DQ + kk;
    float s = 0.f;
    for (int o = 0; o < D; ++o) s = fmaf(d2_s[o * kTile + b], __ldg(w2 + k * D + o), s);
    acc[kk] = s * m1_s[k * kTile + b];
  }
  __syncthreads();
#pragma unroll
  for (int kk = 0; kk < DQ; ++kk) {
    d1_s[(q * DQ + kk) * kTile + b] = acc[kk];
    out[(2 * D + q * DQ + kk) * kTile + b] = acc[kk];
  }
  __syncthreads();
  for (int kk = 0; kk < HL / 4; ++kk) {
    const int k = q * (HL / 4) + kk;
    float s = 0.f;
    for (int o = 0; o < D; ++o) s = fmaf(d1_s[o * kTile + b], __ldg(w1 + k * D + o), s);
    dy7[(tile * HL + k) * kTile + b] = s;
  }
  // deterministic per-tile loss: warp 0 and 1 (q == 0, lanes b = 0..63) reduce by shuffles, lane 0
  // of warp 0 adds the two halves through shared memory.
  if (q == 0) {
    float l = sw * bce * inv_batch;
    for (int o = 16; o > 0; o >>= 1) l += __shfl_down_sync(0xffffffffu, l, o);
    if ((b & 31) == 0) m1_s[b >> 5] = l;
  }
  __syncthreads();
  if (threadIdx.x == 0) loss_partial[tile] = m1_s[0] + m1_s[1];
}

// ---- GCN parameter gradients: raw sums -------------------------------------------------------------
// With dy = d loss / d (BN output) at every node row (dy_c = dpool_c * coef_n * prelu'(y_c)) the
// kernel accumulates, over all rows with t < t_run:
//   sums[0..15]  S_dy[c]        sums[16..31] S_dy_zhat[c]     sums[32..47] S_x0dy[c]
//   sums[48..63] S_x1dy[c]      sums[64..79] d prelu alpha[c]
// One partial row per CTA; summed afterwards in double in a fixed order.
constexpr int kGcnSums = 5 * kGcnChannelsDev;
constexpr int kFeBwdSteps = 8;   // timesteps per CTA: 4 thread rows x 2 steps (small batches need the CTAs)
static __global__ void __launch_bounds__(256) cml_frontend_train_bwd_kernel(const float2* __restrict__ feat, const float* __restrict__ coef,
                                                                            const int* __restrict__ node_off, const float* __restrict__ w,
                                                                            const float* __restrict__ bias, const GcnFolded* __restrict__ pp,
                                                                            const BnBatch* __restrict__ stats, float eps,
                                                                            const float* __restrict__ dx1, float* __restrict__ partial,
                                                                            long long n_windows, int T, int t_run) {
  __shared__ float red[kGcnSums][8];
  // per-channel constants, staged once: y = x0*ya0 + x1*ya1 + yc (BN folded, pp), z_hat the same way with
  // (w - mu) / sigma, and the PReLU slope
  __shared__ float cst_s[7][kGcnChannelsDev];
  if (threadIdx.x < kGcnChannelsDev) {
    const int c = threadIdx.x;
    const float rstd = rsqrtf(stats->var[c] + eps);
    cst_s[0][c] = pp->a[0][c];
    cst_s[1][c] = pp->a[1][c];
    cst_s[2][c] = pp->c[c];
    cst_s[3][c] = w[c] * rstd;
    cst_s[4][c] = w[kGcnChannelsDev + c] * rstd;
    cst_s[5][c] = (bias[c] - stats->mu[c]) * rstd;
    cst_s[6][c] = pp->alpha[c];
  }
  __syncthreads();
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long win = tile * kTile + b;
  const int t0 = blockIdx.y * kFeBwdSteps + tq * (kFeBwdSteps / 4);
  constexpr int CH = 2 + kGcnChannelsDev;
  float s_dy[kGcnChannelsDev], s_dyz[kGcnChannelsDev], s_x0[kGcnChannelsDev], s_x1[kGcnChannelsDev], s_al[kGcnChannelsDev];
#pragma unroll
  for (int c = 0; c < kGcnChannelsDev; ++c) s_dy[c] = s_dyz[c] = s_x0[c] = s_x1[c] = s_al[c] = 0.f;
  if (win < n_windows) {
    const int n0 = node_off[win], nn = node_off[win + 1] - n0;
#pragma unroll 1
    for (int dt = 0; dt < kFeBwdSteps / 4; ++dt) {
      const int t = t0 + dt;
      if (t >= t_run) break;
      const float* g = dx1 + ((tile * t_run + t) * CH) * kTile + b;
      float gv[kGcnChannelsDev];
#pragma unroll
      for (int c = 0; c < kGcnChannelsDev; ++c) gv[c] = __ldg(g + (2 + c) * kTile);
      const float2* row = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn);
      for (int n = 0; n < nn; ++n) {
        const float2 v = __ldg(row + n);
        const float cf = __ldg(coef + n0 + n);
#pragma unroll
        for (int c = 0; c < kGcnChannelsDev; ++c) {
          const float zh = fmaf(v.x, cst_s[3][c], fmaf(v.y, cst_s[4][c], cst_s[5][c]));
          const float y = fmaf(v.x, cst_s[0][c], fmaf(v.y, cst_s[1][c], cst_s[2][c]));
          const float up = gv[c] * cf;
          const float dy = up * (y > 0.f ? 1.f : cst_s[6][c]);
          s_dy[c] += dy;
          s_dyz[c] = fmaf(dy, zh, s_dyz[c]);
          s_x0[c] = fmaf(dy, v.x, s_x0[c]);
          s_x1[c] = fmaf(dy, v.y, s_x1[c]);
          s_al[c] = fmaf(up, fminf(y, 0.f), s_al[c]);
        }
      }
    }
  }
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < kGcnChannelsDev; ++c) {
    float v[5] = {s_dy[c], s_dyz[c], s_x0[c], s_x1[c], s_al[c]};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
      if (lane == 0) red[k * kGcnChannelsDev + c][wp] = v[k];
    }
  }
  __syncthreads();
  if (threadIdx.x < kGcnSums) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += red[threadIdx.x][i];
    partial[(static_cast<long long>(blockIdx.y) * gridDim.x + blockIdx.x) * kGcnSums + threadIdx.x] = s;
  }
}

// ---- SoilNet (three input channels) -------------------------------------------------------------------------------
// The same closure with F = 3: z = x W + b is linear in the three channels, so the batch statistics over all node
// rows follow from nine moments, moments = {Sx0, Sx1, Sx2, Sx00, Sx11, Sx22, Sx01, Sx02, Sx12, N} (double precision).
constexpr int kSoilMoments = 10;
static __global__ void __launch_bounds__(256) feat_moments3_kernel(const float* __restrict__ feat, long long n_rows, double* __restrict__ moments) {
  double s[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_rows; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const double a = __ldg(feat + 3 * i), b = __ldg(feat + 3 * i + 1), c = __ldg(feat + 3 * i + 2);
    s[0] += a; s[1] += b; s[2] += c; s[3] += a * a; s[4] += b * b; s[5] += c * c; s[6] += a * b; s[7] += a * c; s[8] += b * c;
  }
  __shared__ double red[9][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    double v = s[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) red[k][w] = v;
  }
  __syncthreads();
  if (threadIdx.x < 9) {
    double v = 0;
    for (int i = 0; i < 8; ++i) v += red[threadIdx.x][i];
    atomicAdd(&moments[threadIdx.x], v);
  }
}

// S_x[f] and S_xx[f][g] out of the ten moments
struct Moments3 {
  double n, sx[3], sxx[3][3];
  __device__ explicit Moments3(const double* m) {
    n = m[9];
    sx[0] = m[0]; sx[1] = m[1]; sx[2] = m[2];
    sxx[0][0] = m[3]; sxx[1][1] = m[4]; sxx[2][2] = m[5];
    sxx[0][1] = sxx[1][0] = m[6]; sxx[0][2] = sxx[2][0] = m[7]; sxx[1][2] = sxx[2][1] = m[8];
  }
};
static __global__ void bn_prepare3_kernel(const float* __restrict__ w, const float* __restrict__ bias, const float* __restrict__ prelu,
                                          const float* __restrict__ gamma, const float* __restrict__ beta, const double* __restrict__ m,
                                          float eps, GcnFolded* __restrict__ folded, BnBatch* __restrict__ stats) {
  const int c = threadIdx.x;
  if (c >= kGcnChannelsDev) return;
  const Moments3 mo(m);
  const double b = bias[c];
  double wf[3], lin = 0, quad = 0;
  for (int f = 0; f < 3; ++f) wf[f] = w[f * kGcnChannelsDev + c];
  for (int f = 0; f < 3; ++f) {
    lin += wf[f] * mo.sx[f] / mo.n;
    for (int g = 0; g < 3; ++g) quad += wf[f] * wf[g] * mo.sxx[f][g] / mo.n;
  }
  const double mu = lin + b;
  double var = quad + 2.0 * b * lin + b * b - mu * mu;
  if (var < 0) var = 0;
  const double scale = gamma[c] / sqrt(var + eps);
  for (int f = 0; f < 3; ++f) folded->a[f][c] = static_cast<float>(wf[f] * scale);
  folded->c[c] = static_cast<float>((b - mu) * scale + beta[c]);
  folded->alpha[c] = prelu[c];
  stats->mu[c] = static_cast<float>(mu);
  stats->var[c] = static_cast<float>(var);
}

// Raw sums of the SoilNet GeneralConv backward.  The front-end computes h_i = mean_{j in N(i)} prelu(y_j) for every
// sensor i (y = BN output), so the gradient that reaches sensor j's activation is up_j = sum_{i : j in N(i)} dh_i / deg_i:
// a gather over the TRANSPOSED adjacency (rows = sending node).  With dy = up * prelu'(y) the kernel accumulates over
// all (sensor, t < t_run):
//   sums[0..15] S_dy   [16..31] S_dy_zhat   [32..47] S_x0dy   [48..63] S_x1dy   [64..79] S_x2dy   [80..95] d prelu alpha
// One thread per (sequence, step group); one partial row per CTA, summed afterwards in double in a fixed order.
constexpr int kSoilGcnSums = 6 * kGcnChannelsDev;
constexpr int kRvecGcn = kSoilGcnSums;            // reduction-vector slots reserved for the raw GCN sums (CML uses the first 80)
static __global__ void __launch_bounds__(256) soil_frontend_train_bwd_kernel(const float* __restrict__ feat, const int* __restrict__ node_off,
                                                                             const int* __restrict__ seq_win, const int* __restrict__ rowT_ptr,
                                                                             const int* __restrict__ degT, const int* __restrict__ nbrT,
                                                                             const int* __restrict__ deg, const float* __restrict__ w,
                                                                             const float* __restrict__ bias, const GcnFolded* __restrict__ pp,
                                                                             const BnBatch* __restrict__ stats, float eps,
                                                                             const float* __restrict__ dx1, float* __restrict__ partial,
                                                                             long long n_seq, int T, int t_run) {
  __shared__ float red[kSoilGcnSums][8];
  __shared__ float cst_s[9][kGcnChannelsDev];   // y = sum_f x_f ya_f + yc; z_hat = sum_f x_f za_f + zc; prelu slope
  if (threadIdx.x < kGcnChannelsDev) {
    const int c = threadIdx.x;
    const float rstd = rsqrtf(stats->var[c] + eps);
    for (int f = 0; f < 3; ++f) {
      cst_s[f][c] = pp->a[f][c];
      cst_s[4 + f][c] = w[f * kGcnChannelsDev + c] * rstd;
    }
    cst_s[3][c] = pp->c[c];
    cst_s[7][c] = (bias[c] - stats->mu[c]) * rstd;
    cst_s[8][c] = pp->alpha[c];
  }
  __syncthreads();
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long s = tile * kTile + b;
  const int t0 = blockIdx.y * kFeBwdSteps + tq * (kFeBwdSteps / 4);
  constexpr int CH = kGcnChannelsDev + 3;
  float acc[6][kGcnChannelsDev];
#pragma unroll
  for (int k = 0; k < 6; ++k)
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) acc[k][c] = 0.f;
  if (s < n_seq) {
    const int wdw = seq_win[s];
    const int n0 = node_off[wdw], nn = node_off[wdw + 1] - n0;
    const int me = static_cast<int>(s - n0), dT = degT[s], rp = rowT_ptr[s];
#pragma unroll 1
    for (int dt = 0; dt < kFeBwdSteps / 4; ++dt) {
      const int t = t0 + dt;
      if (t >= t_run) break;
      float up[kGcnChannelsDev];
#pragma unroll
      for (int c = 0; c < kGcnChannelsDev; ++c) up[c] = 0.f;
      for (int e = 0; e < dT; ++e) {                       // receivers of this sensor's activation (ascending order: deterministic)
        const long long si = n0 + __ldg(nbrT + rp + e);
        const float inv = 1.f / static_cast<float>(__ldg(deg + si));
        const float* g = dx1 + (((si / kTile) * t_run + t) * CH) * kTile + (si % kTile);
#pragma unroll
        for (int c = 0; c < kGcnChannelsDev; ++c) up[c] = fmaf(__ldg(g + c * kTile), inv, up[c]);
      }
      const float* x = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn + me) * 3;
      const float x0 = __ldg(x), x1 = __ldg(x + 1), x2 = __ldg(x + 2);
#pragma unroll
      for (int c = 0; c < kGcnChannelsDev; ++c) {
        const float y = fmaf(x0, cst_s[0][c], fmaf(x1, cst_s[1][c], fmaf(x2, cst_s[2][c], cst_s[3][c])));
        const float zh = fmaf(x0, cst_s[4][c], fmaf(x1, cst_s[5][c], fmaf(x2, cst_s[6][c], cst_s[7][c])));
        const float dy = up[c] * (y > 0.f ? 1.f : cst_s[8][c]);
        acc[0][c] += dy;
        acc[1][c] = fmaf(dy, zh, acc[1][c]);
        acc[2][c] = fmaf(dy, x0, acc[2][c]);
        acc[3][c] = fmaf(dy, x1, acc[3][c]);
        acc[4][c] = fmaf(dy, x2, acc[4][c]);
        acc[5][c] = fmaf(up[c], fminf(y, 0.f), acc[5][c]);
      }
    }
  }
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 6; ++k)
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) {
      float v = acc[k][c];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
      if (lane == 0) red[k * kGcnChannelsDev + c][wp] = v;
    }
  __syncthreads();
  if (threadIdx.x < kSoilGcnSums) {
    float v = 0.f;
    for (int i = 0; i < 8; ++i) v += red[threadIdx.x][i];
    partial[(static_cast<long long>(blockIdx.y) * gridDim.x + blockIdx.x) * kSoilGcnSums + threadIdx.x] = v;
  }
}

static __global__ void __launch_bounds__(128) sum_partials_kernel(const float* __restrict__ partial, long long n_rows, int width,
                                                                  float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= width) return;
  double s = 0;
  for (long long r = 0; r < n_rows; ++r) s += partial[r * width + c];
  out[c] = static_cast<float>(s);
}

// ---- finalize + optimizer -----------------------------------------------------------------------------
// rvec layout: [0, n_params) gradients in flat-parameter order, [n_params, n_params+80) the raw
// GCN sums above, [n_params+80] the loss.  After the cross-rank sum this kernel turns the raw sums
// into the GeneralConv / BatchNorm / PReLU gradients (BN backward with batch statistics):
//   d beta = S_dy, d gamma = S_dy_zhat, d alpha as accumulated,
//   dW[f][c] = gamma/sigma * (S_xfdy - d beta/N * S_xf - d gamma/N * S_xf_zhat),  d bias = 0.
struct GcnOffsets {
  long long kernel, bias, prelu, gamma, beta, mean, var;
};
static __global__ void gcn_finalize_kernel(float* __restrict__ rvec, long long n_params, const float* __restrict__ params, GcnOffsets o,
                                           const double* __restrict__ m, const BnBatch* __restrict__ stats, float eps) {
  const int c = threadIdx.x;
  if (c >= kGcnChannelsDev) return;
  const float* raw = rvec + n_params;
  const double n = m[5];
  const double w0 = params[o.kernel + c], w1 = params[o.kernel + kGcnChannelsDev + c], b = params[o.bias + c];
  const double mu = stats->mu[c], sigma = sqrt(static_cast<double>(stats->var[c]) + eps);
  const double dbeta = raw[c], dgamma = raw[kGcnChannelsDev + c];
  const double sx[2] = {m[0], m[1]};
  const double sxx[2][2] = {{m[2], m[4]}, {m[4], m[3]}};
  const double gos = params[o.gamma + c] / sigma;
  for (int f = 0; f < 2; ++f) {
    const double sxz = w0 * sxx[f][0] + w1 * sxx[f][1] + b * sx[f];
    const double sxzh = (sxz - mu * sx[f]) / sigma;
    const double sxdy = raw[(2 + f) * kGcnChannelsDev + c];
    rvec[o.kernel + f * kGcnChannelsDev + c] = static_cast<float>(gos * (sxdy - dbeta / n * sx[f] - dgamma / n * sxzh));
  }
  rvec[o.bias + c] = 0.f;
  rvec[o.prelu + c] = raw[4 * kGcnChannelsDev + c];
  rvec[o.gamma + c] = static_cast<float>(dgamma);
  rvec[o.beta + c] = static_cast<float>(dbeta);
}

// The same closure for three input channels (raw sums laid out as in soil_frontend_train_bwd_kernel).
static __global__ void gcn_finalize3_kernel(float* __restrict__ rvec, long long n_params, const float* __restrict__ params, GcnOffsets o,
                                            const double* __restrict__ m, const BnBatch* __restrict__ stats, float eps) {
  const int c = threadIdx.x;
  if (c >= kGcnChannelsDev) return;
  const float* raw = rvec + n_params;
  const Moments3 mo(m);
  double wf[3];
  for (int f = 0; f < 3; ++f) wf[f] = params[o.kernel + f * kGcnChannelsDev + c];
  const double b = params[o.bias + c];
  const double mu = stats->mu[c], sigma = sqrt(static_cast<double>(stats->var[c]) + eps);
  const double dbeta = raw[c], dgamma = raw[kGcnChannelsDev + c];
  const double gos = params[o.gamma + c] / sigma;
  for (int f = 0; f < 3; ++f) {
    double sxz = b * mo.sx[f];
    for (int g = 0; g < 3; ++g) sxz += wf[g] * mo.sxx[f][g];
    const double sxzh = (sxz - mu * mo.sx[f]) / sigma;
    const double sxdy = raw[(2 + f) * kGcnChannelsDev + c];
    rvec[o.kernel + f * kGcnChannelsDev + c] = static_cast<float>(gos * (sxdy - dbeta / mo.n * mo.sx[f] - dgamma / mo.n * sxzh));
  }
  rvec[o.bias + c] = 0.f;
  rvec[o.prelu + c] = raw[5 * kGcnChannelsDev + c];
  rvec[o.gamma + c] = static_cast<float>(dgamma);
  rvec[o.beta + c] = static_cast<float>(dbeta);
}

// Keras Adam (beta1 0.9, beta2 0.999, epsilon 1e-7), lr_t = lr*sqrt(1-b2^t)/(1-b1^t).
// Entries in [skip_lo, skip_hi) are the BatchNorm moving statistics: not trained, moved towards the
// batch statistics with momentum 0.99 instead.
// kind 1: plain SGD (Keras default momentum 0): p -= lr g.  kind 2: RMSprop (rho 0.9, epsilon 1e-7, no momentum, not
// centered): v = rho v + (1 - rho) g^2, p -= lr g / (sqrt(v) + eps).  (libs/fit_model.py:71-74 accepts all three.)
static __global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ params, const float* __restrict__ grad, float* __restrict__ mom,
                                                          float* __restrict__ vel, long long n, float lr_t, float b1, float b2, float eps,
                                                          long long skip_lo, long long skip_hi, const BnBatch* __restrict__ stats,
                                                          float bn_momentum, int kind = 0) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (i >= skip_lo && i < skip_hi) {
    const long long k = i - skip_lo;
    const float batch = (k < kGcnChannelsDev) ? stats->mu[k] : stats->var[k - kGcnChannelsDev];
    params[i] = params[i] * bn_momentum + batch * (1.f - bn_momentum);
    return;
  }
  const float g = grad[i];
  if (kind == 1) {
    params[i] -= lr_t * g;
    return;
  }
  if (kind == 2) {
    const float v = 0.9f * vel[i] + 0.1f * g * g;
    vel[i] = v;
    params[i] -= lr_t * g / (sqrtf(v) + eps);
    return;
  }
  const float m = b1 * mom[i] + (1.f - b1) * g;
  const float v = b2 * vel[i] + (1.f - b2) * g * g;
  mom[i] = m;
  vel[i] = v;
  params[i] -= lr_t * m / (sqrtf(v) + eps);
}

// Device-side refresh of the packed / transposed weight copies after an optimizer step.
static __global__ void __launch_bounds__(256) repack_kernel(const float* __restrict__ W, const float* __restrict__ U, const float* __restrict__ bias,
                                                            int in, int h, float* __restrict__ packed, float* __restrict__ packed_t, int kpad) {
  const int K = in + h;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n_main = static_cast<long long>(K) * h * 4;
  if (i < n_main) {
    const int g = static_cast<int>(i & 3), j = static_cast<int>((i >> 2) % h), k = static_cast<int>((i >> 2) / h);
    const float v = (k < in) ? W[static_cast<long long>(k) * 4 * h + g * h + j] : U[static_cast<long long>(k - in) * 4 * h + g * h + j];
    packed[i] = v;
    packed_t[static_cast<long long>(g * h + j) * kpad + k] = v;
  } else if (i < n_main + 4 * h) {
    const int r = static_cast<int>(i - n_main), g = r & 3, j = r >> 2;
    packed[i] = bias[g * h + j];
  }
}

}  // namespace gnnqc
