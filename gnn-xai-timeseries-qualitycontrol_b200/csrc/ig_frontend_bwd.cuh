// Integrated gradients: backward through the CML graph front-end for a GROUP of interpolation points at once
// (xai/libs/integrated_gradients.py:955-1012,1185-1191 for the graph half of the model).
//
// For interpolation point k (alpha_k = k / m, trapezoid weight w_k) the LSTM backward leaves dx1_k = d(sum_b p_b)/d(x1 tile),
// and a feature row (t, n) of a window receives
//   attr[f] += w_k * x_f * coef_n * sum_c dx1_k[t][2+c] * prelu'(z_kc) * a[f][c],      z_kc = alpha_k * (x0 a[0][c] + x1 a[1][c]) + c_c.
// The first version ran one kernel per alpha: features read and attributions read-modify-written 51 times, ~210
// instructions per row and alpha (ncu: issue slots 73 %).  Here the dx1 tiles of a whole group of alphas are kept (HBM is
// large enough: 11.7 kB per window and alpha), and one pass handles them all:
//   * a CTA takes 8 consecutive windows at one timestep, so the dx1 values it stages are whole 32-byte sectors of the
//     [..][64 windows] tiles; per (window, k, c) it keeps G_kc = w_k dx1_k[t][2+c] a[f][c], split by the PReLU slope s_c into
//     base = sum_kc s_c G_kc (what every row gets) and Q_kc = (1 - s_c) G_kc (what rows with z_kc > 0 get on top);
//   * a row keeps u_c = x0 a[0][c] + x1 a[1][c] in registers, so per (k, c) it does one FMA, one compare and two
//     predicated adds; features are read once and attributions written once per group.
// Shared by the materialised-window path (ig.cu) and the resident-series path (series.cu), so both give the same bits.
#pragma once
#include "frontend.cuh"

namespace gnnqc {

constexpr int kIgWin = 8;         // windows per CTA (one 32-byte sector of a 64-window tile row)
constexpr int kIgMaxGroup = 51;   // alphas per pass (shared-memory staging: 8 x 51 x 16 float2 = 52 kB)

__host__ __device__ inline size_t ig_bwd_smem_bytes(int nk) {
  return size_t(kIgWin) * nk * kGcnChannelsDev * sizeof(float2) + size_t(kIgWin) * 4 * sizeof(float) * 9 + size_t(nk) * sizeof(float);
}

struct IgGroup {
  const float* dx;        // dx1 tiles of the group's first alpha: [tile][t_run][18][64]
  long long stride;       // floats between consecutive alphas
  int k0, nk, m_steps;    // alphas k0 .. k0 + nk - 1 of 0 .. m_steps
  int first;              // 1: assign (and zero the steps that cannot reach the score), 0: accumulate
  int scale_by_input;     // 0: the trapezoid mean of the raw gradients (gnnqc_ig_cml_raw)
};

struct IgSmem {
  float2* Q;      // [kIgWin][nk][16]: (1 - s_c) w_k dx1_k[2+c] (a[0][c], a[1][c])
  float* aux;     // [kIgWin][4]: base0, base1, and the target-series sums sum_k w_k dx1_k[0], [1]
  float* part;    // [8 warps][kIgWin][4]: per-warp partial sums of aux (summed in a fixed order: deterministic)
  float* alpha;   // [nk]
};
__device__ __forceinline__ IgSmem ig_smem(unsigned char* raw, int nk) {
  IgSmem s;
  s.Q = reinterpret_cast<float2*>(raw);
  s.aux = reinterpret_cast<float*>(s.Q + size_t(kIgWin) * nk * kGcnChannelsDev);
  s.part = s.aux + kIgWin * 4;
  s.alpha = s.part + 8 * kIgWin * 4;
  return s;
}

// Stages the group's dx1 values of timestep t for windows win0 .. win0 + 7 (win0 a multiple of 8).  256 threads.
__device__ __forceinline__ void ig_stage(const GcnFolded& p, const IgGroup& g, long long win0, int t, int t_run, const IgSmem& sm) {
  constexpr int CH = 2 + kGcnChannelsDev;
  const float m = static_cast<float>(g.m_steps);
  for (int k = threadIdx.x; k < g.nk; k += blockDim.x) sm.alpha[k] = static_cast<float>(g.k0 + k) / m;
  if (t >= t_run) return;                        // uniform over the CTA
  const long long tile = win0 / kTile;
  const int b0 = static_cast<int>(win0 % kTile);
  const int w_l = threadIdx.x & (kIgWin - 1);
  float b_0 = 0.f, b_1 = 0.f, ga0 = 0.f, ga1 = 0.f;
  for (int i = threadIdx.x >> 3; i < g.nk * kGcnChannelsDev; i += static_cast<int>(blockDim.x >> 3)) {
    const int k = i / kGcnChannelsDev, c = i % kGcnChannelsDev;
    const float* gp = g.dx + static_cast<long long>(k) * g.stride + ((tile * t_run + t) * CH) * kTile + b0 + w_l;
    const int kk = g.k0 + k;
    const float wk = ((kk == 0 || kk == g.m_steps) ? 0.5f : 1.0f) / m;
    const float gv = wk * __ldg(gp + (2 + c) * kTile);
    const float g0 = gv * p.a[0][c], g1 = gv * p.a[1][c], s = p.alpha[c];
    sm.Q[(w_l * g.nk + k) * kGcnChannelsDev + c] = make_float2(g0 * (1.f - s), g1 * (1.f - s));
    b_0 = fmaf(s, g0, b_0);
    b_1 = fmaf(s, g1, b_1);
    if (c == 0) {
      ga0 = fmaf(wk, __ldg(gp), ga0);
      ga1 = fmaf(wk, __ldg(gp + kTile), ga1);
    }
  }
  // lanes with the same window: 8 apart
  b_0 += __shfl_xor_sync(0xffffffffu, b_0, 8); b_0 += __shfl_xor_sync(0xffffffffu, b_0, 16);
  b_1 += __shfl_xor_sync(0xffffffffu, b_1, 8); b_1 += __shfl_xor_sync(0xffffffffu, b_1, 16);
  ga0 += __shfl_xor_sync(0xffffffffu, ga0, 8); ga0 += __shfl_xor_sync(0xffffffffu, ga0, 16);
  ga1 += __shfl_xor_sync(0xffffffffu, ga1, 8); ga1 += __shfl_xor_sync(0xffffffffu, ga1, 16);
  if ((threadIdx.x & 31) < kIgWin) {
    float* o = sm.part + ((threadIdx.x >> 5) * kIgWin + w_l) * 4;
    o[0] = b_0; o[1] = b_1; o[2] = ga0; o[3] = ga1;
  }
  __syncthreads();
  if (threadIdx.x < kIgWin * 4) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += sm.part[w * kIgWin * 4 + threadIdx.x];
    sm.aux[threadIdx.x] = v;
  }
}

// sum over the group's (k, c) of w_k dx1_k[2+c] prelu'(z_kc) a[f][c] for one row with features (x0, x1) of window slot w_l
__device__ __forceinline__ float2 ig_row(const GcnFolded& p, const IgGroup& g, float x0, float x1, const IgSmem& sm, int w_l) {
  float u[kGcnChannelsDev];
#pragma unroll
  for (int c = 0; c < kGcnChannelsDev; ++c) u[c] = fmaf(x0, p.a[0][c], x1 * p.a[1][c]);
  float sx[4] = {sm.aux[w_l * 4 + 0], 0.f, 0.f, 0.f}, sy[4] = {sm.aux[w_l * 4 + 1], 0.f, 0.f, 0.f};   // four chains: the adds are dependent otherwise
  const float2* q = sm.Q + size_t(w_l) * g.nk * kGcnChannelsDev;
  for (int k = 0; k < g.nk; ++k) {
    const float alpha = sm.alpha[k];
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) {
      const float2 qv = q[k * kGcnChannelsDev + c];
      if (fmaf(alpha, u[c], p.c[c]) > 0.f) {
        sx[c & 3] += qv.x;
        sy[c & 3] += qv.y;
      }
    }
  }
  return make_float2((sx[0] + sx[1]) + (sx[2] + sx[3]), (sy[0] + sy[1]) + (sy[2] + sy[3]));
}

// Materialised windows: feat rows (window, t, n), attr_feat the same layout.  Grid (ceil(windows / 8), T), 256 threads:
// thread = (window slot, node slot of 32); windows with more than 32 nodes loop.
static __global__ void __launch_bounds__(256) cml_frontend_bwd_multi_kernel(const float2* __restrict__ feat, const float2* __restrict__ anom,
                                                                            const float* __restrict__ coef, const int* __restrict__ node_off,
                                                                            const GcnFolded p, const IgGroup g, float2* __restrict__ attr_feat,
                                                                            float2* __restrict__ attr_anom, long long n_windows, int T, int t_run) {
  extern __shared__ __align__(16) unsigned char ig_raw[];
  const IgSmem sm = ig_smem(ig_raw, g.nk);
  const long long win0 = static_cast<long long>(blockIdx.x) * kIgWin;
  const int t = blockIdx.y;
  if (t >= t_run && !g.first) return;                  // dead steps are zeroed by the first group only
  ig_stage(p, g, win0, t, t_run, sm);
  __syncthreads();
  const int w_l = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long win = win0 + w_l;
  if (win >= n_windows) return;
  const int n0 = node_off[win], nn = node_off[win + 1] - n0;
  for (int n = lane; n < nn; n += 32) {
    const long long at = static_cast<long long>(n0) * T + static_cast<long long>(t) * nn + n;
    if (t >= t_run) {
      attr_feat[at] = make_float2(0.f, 0.f);
      continue;
    }
    const float2 v = __ldg(feat + at);
    const float cf = __ldg(coef + n0 + n);
    const float2 s = ig_row(p, g, v.x, v.y, sm, w_l);
    float2 a = g.first ? make_float2(0.f, 0.f) : attr_feat[at];
    a.x = fmaf(g.scale_by_input ? v.x * cf : cf, s.x, a.x);
    a.y = fmaf(g.scale_by_input ? v.y * cf : cf, s.y, a.y);
    attr_feat[at] = a;
  }
  if (lane == 0) {
    float2 a = make_float2(0.f, 0.f);
    if (t < t_run) {
      const float2 av = __ldg(anom + win * T + t);
      if (!g.first) a = attr_anom[win * T + t];
      a.x = fmaf(g.scale_by_input ? av.x : 1.f, sm.aux[w_l * 4 + 2], a.x);
      a.y = fmaf(g.scale_by_input ? av.y : 1.f, sm.aux[w_l * 4 + 3], a.y);
    }
    attr_anom[win * T + t] = a;
  }
}

}  // namespace gnnqc
