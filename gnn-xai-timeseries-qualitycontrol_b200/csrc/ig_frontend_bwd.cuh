// Integrated gradients: backward through the CML graph front-end for a GROUP of interpolation points at once
// (xai/libs/integrated_gradients.py:955-1012,1185-1191 for the graph half of the model).
//
// For interpolation point k (alpha_k = k / m, trapezoid weight w_k) the LSTM backward leaves dx1_k = d(sum_b p_b)/d(x1 tile),
// and a feature row (t, n) of a window receives
//   attr[f] += w_k * x_f * coef_n * sum_c dx1_k[t][2+c] * prelu'(z_kc) * a[f][c],      z_kc = alpha_k * (x0 a[0][c] + x1 a[1][c]) + c_c.
// The first version ran one kernel per alpha: features read and attributions read-modify-written 51 times, ~210
// instructions per row and alpha (ncu: issue slots 73 %).  Here the dx1 tiles of a whole group of alphas are kept (HBM is
// large enough: 11.7 kB per window and alpha), and one pass handles them all:
//   * a CTA takes 8 consecutive windows at one timestep, so the dx1 values it reads are whole 32-byte sectors of the
//     [..][64 windows] tiles.  One thread per (window, channel) walks the group's alphas and leaves the PREFIX SUMS
//     P_c[k] = sum_{k' < k} w_k' dx1_k'[c] in shared memory (18 channels: the two target-series channels only need the total);
//   * z_kc = alpha_k u_c + c_c is monotone in k, so the alphas with z_kc > 0 are a prefix (u_c < 0) or a suffix (u_c > 0) of
//     the group.  A row computes the crossing point per channel analytically and takes one difference of prefix sums:
//     ~20 instructions per channel where the version before did ~5 per (k, c) x 51.  (Where z_kc is within rounding of
//     zero the side is arbitrary - as it is between any two evaluations of z, the forward kernel's included.)
//     Features are read once and attributions written once per group.
// Shared by the materialised-window path (ig.cu) and the resident-series path (series.cu), so both give the same bits.
#pragma once
#include "frontend.cuh"

namespace gnnqc {

constexpr int kIgWin = 8;         // windows per CTA (one 32-byte sector of a 64-window tile row)
constexpr int kIgMaxGroup = 51;   // alphas per pass
constexpr int kIgCh = 2 + kGcnChannelsDev;   // channels of the x1 tile: target series (2), GCN output (16)

// prefix sums of one (window, channel): nk + 1 entries, padded to an odd count (bank spread of the per-thread walk)
__host__ __device__ inline int ig_row_len(int nk) { return (nk + 1) | 1; }
__host__ __device__ inline size_t ig_bwd_smem_bytes(int nk) {
  return (size_t(kIgWin) * kIgCh * ig_row_len(nk) + size_t(nk)) * sizeof(float);
}

struct IgGroup {
  const float* dx;        // dx1 tiles of the group's first alpha: [tile][t_run][18][64]
  long long stride;       // floats between consecutive alphas
  int k0, nk, m_steps;    // alphas k0 .. k0 + nk - 1 of 0 .. m_steps
  int first;              // 1: assign (and zero the steps that cannot reach the score), 0: accumulate
  int scale_by_input;     // 0: the trapezoid mean of the raw gradients (gnnqc_ig_cml_raw)
};

struct IgSmem {
  float* P;       // [kIgWin][18][row_len]: P[k] = sum_{k' < k} w_k' dx1_k'[channel]
  float* wk;      // [nk] trapezoid weights
};
__device__ __forceinline__ IgSmem ig_smem(unsigned char* raw, int nk) {
  IgSmem s;
  s.P = reinterpret_cast<float*>(raw);
  s.wk = s.P + size_t(kIgWin) * kIgCh * ig_row_len(nk);
  return s;
}

// Prefix sums of the group's dx1 values of timestep t for windows win0 .. win0 + 7 (win0 a multiple of 8).  256 threads;
// the caller synchronises before the rows read them.
__device__ __forceinline__ void ig_stage(const IgGroup& g, long long win0, int t, int t_run, const IgSmem& sm) {
  const float m = static_cast<float>(g.m_steps);
  for (int k = threadIdx.x; k < g.nk; k += blockDim.x) {
    const int kk = g.k0 + k;
    sm.wk[k] = ((kk == 0 || kk == g.m_steps) ? 0.5f : 1.0f) / m;
  }
  if (t >= t_run) return;                        // uniform over the CTA
  __syncthreads();
  if (threadIdx.x < kIgWin * kIgCh) {
    const int w_l = threadIdx.x & (kIgWin - 1), ch = threadIdx.x >> 3;
    const int L = ig_row_len(g.nk);
    const float* gp = g.dx + (((win0 / kTile) * t_run + t) * kIgCh + ch) * kTile + (win0 % kTile) + w_l;
    float* row = sm.P + (w_l * kIgCh + ch) * L;
    float run = 0.f;
    row[0] = 0.f;
    int k = 0;
    for (; k + 4 <= g.nk; k += 4) {              // the loads of four alphas in flight, the sum in order
      const float v0 = __ldg(gp + static_cast<long long>(k) * g.stride), v1 = __ldg(gp + static_cast<long long>(k + 1) * g.stride),
                  v2 = __ldg(gp + static_cast<long long>(k + 2) * g.stride), v3 = __ldg(gp + static_cast<long long>(k + 3) * g.stride);
      run = fmaf(sm.wk[k], v0, run); row[k + 1] = run;
      run = fmaf(sm.wk[k + 1], v1, run); row[k + 2] = run;
      run = fmaf(sm.wk[k + 2], v2, run); row[k + 3] = run;
      run = fmaf(sm.wk[k + 3], v3, run); row[k + 4] = run;
    }
    for (; k < g.nk; ++k) {
      run = fmaf(sm.wk[k], __ldg(gp + static_cast<long long>(k) * g.stride), run);
      row[k + 1] = run;
    }
  }
}

// sum_k w_k dx1_k[0], [1] of window slot w_l: the gradient with respect to the target series
__device__ __forceinline__ float2 ig_anom(const IgGroup& g, const IgSmem& sm, int w_l) {
  const int L = ig_row_len(g.nk);
  return make_float2(sm.P[(w_l * kIgCh + 0) * L + g.nk], sm.P[(w_l * kIgCh + 1) * L + g.nk]);
}

// sum over the group's (k, c) of w_k dx1_k[2+c] prelu'(z_kc) a[f][c] for one row with features (x0, x1) of window slot w_l
__device__ __forceinline__ float2 ig_row(const GcnFolded& p, const IgGroup& g, float x0, float x1, const IgSmem& sm, int w_l) {
  const int nk = g.nk, L = ig_row_len(nk);
  const float* pw = sm.P + (w_l * kIgCh + 2) * L;
  const float m = static_cast<float>(g.m_steps), half_k0 = 0.5f - static_cast<float>(g.k0), nkf = static_cast<float>(nk);
  float sx[2] = {0.f, 0.f}, sy[2] = {0.f, 0.f};
#pragma unroll
  for (int c = 0; c < kGcnChannelsDev; ++c) {
    const float u = fmaf(x0, p.a[0][c], x1 * p.a[1][c]), cc = p.c[c];
    const bool up = u > 0.f;
    // z_k = alpha_k u + c crosses zero at k* = -c m / u - k0: for u > 0 the alphas k > k* are on the positive side, for
    // u < 0 those with k < k*; either way the boundary index is ks = round(k* + 1/2), clamped to the group.  (u = 0 is
    // pushed to a tiny negative value: all or none, by the sign of c.)
    float ru;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ru) : "f"(up ? u : fminf(u, -1e-30f)));
    const float kc = fminf(fmaxf(fmaf(-cc * m, ru, half_k0), 0.f), nkf);
    const int ks = __float_as_int(kc + 12582912.f) - 0x4B400000;   // round to nearest through the 1.5 * 2^23 constant
    const float pk = pw[c * L + ks], pn = pw[c * L + nk];
    const float d = up ? pn - pk : pk;           // the alphas on the positive side: from ks on / before ks
    const float e = fmaf(p.alpha[c], pn - d, d); // slope 1 there, the PReLU slope on the others
    sx[c & 1] = fmaf(p.a[0][c], e, sx[c & 1]);
    sy[c & 1] = fmaf(p.a[1][c], e, sy[c & 1]);
  }
  return make_float2(sx[0] + sx[1], sy[0] + sy[1]);
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
  ig_stage(g, win0, t, t_run, sm);
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
      const float2 ga = ig_anom(g, sm, w_l);
      a.x = fmaf(g.scale_by_input ? av.x : 1.f, ga.x, a.x);
      a.y = fmaf(g.scale_by_input ? av.y : 1.f, ga.y, a.y);
    }
    attr_anom[win * T + t] = a;
  }
}

}  // namespace gnnqc
