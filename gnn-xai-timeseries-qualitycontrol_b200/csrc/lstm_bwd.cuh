// Input-gradient backward pass of one Keras LSTM layer over sequence tiles (BPTT without weight
// gradients): what tape.gradient(prediction, inputs) needs for integrated gradients
// (reference: xai/libs/integrated_gradients.py:967-987, which differentiates the whole model
// with tf.GradientTape; the layer arithmetic is Keras' LSTM as in lstm_fwd.cuh).
//
// Per step t (descending), with the forward's saved i, f, g, o, c_t (lstm_fwd.cuh, SAVE):
//   dh   = dy_t + dh_rec            (dy routed through MaxPooling1D's argmax for pooled layers)
//   dc  += dh * o * (1 - tanh(c_t)^2)
//   dz_i = dc*g*i(1-i)   dz_f = dc*c_{t-1}*f(1-f)   dz_g = dc*i*(1-g^2)   dz_o = dh*tanh(c_t)*o(1-o)
//   dc   = dc * f
//   [dx_t | dh_rec] = dz . [W | U]^T          <- the contraction, 4H long, (I+H) x 64 outputs
//
// A CTA owns one tile at a time.  Phase A (cell-wise, dc in registers) writes dz into a shared
// [4H][64] tile; phase B is a register-tiled contraction (thread = 4 output rows x 8 windows,
// packed FFMA2) against the transposed weights wT[col][Kpad], which sit in shared memory (or
// L1/L2 for the 128-unit layer).
#pragma once
#include <cstdint>

#include "lstm_fwd.cuh"

namespace gnnqc {

template <int I, int H, bool WGLOBAL>
struct BwdCfg {
  static constexpr int K = I + H;
  static constexpr int NW = (K + 15) / 16;   // warps; each covers 16 output rows x 64 windows
  static constexpr int NT = 32 * NW;
  static constexpr int KPAD = 16 * NW;
  static constexpr int COLS = 4 * H;
  static constexpr int GROUPS = H * (kTile / 4);          // float4 cell groups per step
  static constexpr int CPG = (GROUPS + NT - 1) / NT;      // groups owned by a thread
  static constexpr int CC = 32;                           // WGLOBAL: weight columns per streamed chunk
  static constexpr size_t W_FLOATS = WGLOBAL ? size_t(2) * CC * KPAD : size_t(COLS) * KPAD;
  static constexpr size_t SMEM_BYTES = (W_FLOATS + size_t(COLS) * kTile + size_t(H) * kTile) * 4;
};

// NSAVE: quantities per step in `save` (5 for attribution, 6 for training).  dz_out (training):
// when not null the gate pre-activation gradients are also written as dz_out[tile][t][4H][64] for
// the weight-gradient kernel.
template <int I, int H, bool POOL, bool LAST, bool WGLOBAL, int NSAVE = 5>
__global__ void __launch_bounds__((BwdCfg<I, H, WGLOBAL>::NT))
lstm_layer_bwd_kernel(const float* __restrict__ dy,    // grad wrt this layer's output (see header)
                      const float* __restrict__ save,  // [n_tiles][T][NSAVE][H][64]
                      const float* __restrict__ amax,  // POOL: [n_tiles][T/3][H][64]
                      const float* __restrict__ wT,    // [4H][KPAD]
                      float* __restrict__ dx,          // [n_tiles][T][I][64]
                      int T, int n_tiles, float* __restrict__ dz_out) {
  using C = BwdCfg<I, H, WGLOBAL>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* wsm = reinterpret_cast<float*>(smem_raw);
  float* dzs = wsm + C::W_FLOATS;
  float* dhs = dzs + C::COLS * kTile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (!WGLOBAL) {
    const float4* src = reinterpret_cast<const float4*>(wT);
    float4* dst = reinterpret_cast<float4*>(wsm);
    for (int i = tid; i < C::COLS * C::KPAD / 4; i += C::NT) dst[i] = __ldg(src + i);
  }
  const float4* w4 = reinterpret_cast<const float4*>(wsm);
  // phase-B coordinates
  const int rl = lane & 3, bl = lane >> 2;
  const int row0 = warp * 16 + rl * 4;
  const int t_dy = POOL ? T / 3 : T;

  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    float dc[C::CPG][4];
#pragma unroll
    for (int n = 0; n < C::CPG; ++n) dc[n][0] = dc[n][1] = dc[n][2] = dc[n][3] = 0.f;
    __syncthreads();  // weights loaded / previous tile finished with dzs, dhs
    for (int t = T - 1; t >= 0; --t) {
      // ---- phase A: cell-wise gradients ----------------------------------------------------
#pragma unroll
      for (int n = 0; n < C::CPG; ++n) {
        const int g = tid + n * C::NT;
        if (g < C::GROUPS) {
          const int u = g / (kTile / 4), c4 = g % (kTile / 4);
          constexpr int QS = H * kTile / 4;   // float4 stride between saved quantities
          const float4* sv = reinterpret_cast<const float4*>(save + (((tile * T + t) * NSAVE) * H + u) * kTile);
          const float4 ig = __ldg(sv + 0 * QS + c4), fg = __ldg(sv + 1 * QS + c4), gg = __ldg(sv + 2 * QS + c4),
                       og = __ldg(sv + 3 * QS + c4), ct = __ldg(sv + 4 * QS + c4);
          float4 cp = make_float4(0.f, 0.f, 0.f, 0.f);
          if (t > 0) cp = __ldg(reinterpret_cast<const float4*>(save + (((tile * T + t - 1) * NSAVE + 4) * H + u) * kTile) + c4);
          float4 dh = make_float4(0.f, 0.f, 0.f, 0.f);
          if (t < T - 1) dh = reinterpret_cast<const float4*>(dhs)[u * (kTile / 4) + c4];
          if (POOL) {
            const long long o = ((tile * t_dy + t / 3) * H + u) * (kTile / 4) + c4;
            const float4 a = __ldg(reinterpret_cast<const float4*>(amax) + o);
            const float4 d = __ldg(reinterpret_cast<const float4*>(dy) + o);
            const float ph = static_cast<float>(t % 3);
            dh.x += (a.x == ph) ? d.x : 0.f; dh.y += (a.y == ph) ? d.y : 0.f;
            dh.z += (a.z == ph) ? d.z : 0.f; dh.w += (a.w == ph) ? d.w : 0.f;
          } else if (LAST) {
            if (t == T - 1) {
              const float4 d = __ldg(reinterpret_cast<const float4*>(dy) + (tile * H + u) * (kTile / 4) + c4);
              dh.x += d.x; dh.y += d.y; dh.z += d.z; dh.w += d.w;
            }
          } else {
            const float4 d = __ldg(reinterpret_cast<const float4*>(dy) + ((tile * t_dy + t) * H + u) * (kTile / 4) + c4);
            dh.x += d.x; dh.y += d.y; dh.z += d.z; dh.w += d.w;
          }
          float4 zi, zf, zg, zo;
#define GNNQC_CELL_BWD(e, k)                                             \
          {                                                              \
            const float tc = tanh_fast(ct.e);                            \
            const float dcc = fmaf(dh.e * og.e, 1.f - tc * tc, dc[n][k]); \
            zi.e = dcc * gg.e * ig.e * (1.f - ig.e);                     \
            zf.e = dcc * cp.e * fg.e * (1.f - fg.e);                     \
            zg.e = dcc * ig.e * (1.f - gg.e * gg.e);                     \
            zo.e = dh.e * tc * og.e * (1.f - og.e);                      \
            dc[n][k] = dcc * fg.e;                                       \
          }
          GNNQC_CELL_BWD(x, 0) GNNQC_CELL_BWD(y, 1) GNNQC_CELL_BWD(z, 2) GNNQC_CELL_BWD(w, 3)
#undef GNNQC_CELL_BWD
          float4* dz4 = reinterpret_cast<float4*>(dzs);
          dz4[(0 * H + u) * (kTile / 4) + c4] = zi;
          dz4[(1 * H + u) * (kTile / 4) + c4] = zf;
          dz4[(2 * H + u) * (kTile / 4) + c4] = zg;
          dz4[(3 * H + u) * (kTile / 4) + c4] = zo;
          if (dz_out != nullptr) {
            float4* go = reinterpret_cast<float4*>(dz_out + (tile * T + t) * (C::COLS * kTile));
            go[(0 * H + u) * (kTile / 4) + c4] = zi;
            go[(1 * H + u) * (kTile / 4) + c4] = zf;
            go[(2 * H + u) * (kTile / 4) + c4] = zg;
            go[(3 * H + u) * (kTile / 4) + c4] = zo;
          }
        }
      }
      __syncthreads();
      // ---- phase B: [dx_t | dh_rec] = dz . [W|U]^T ---------------------------------------------
      unsigned long long acc[4][4];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[r][i] = 0ull;
      const ulonglong2* dz2 = reinterpret_cast<const ulonglong2*>(dzs);
      auto mul_cols = [&](const float4* __restrict__ wbase, int col0, int ncol) {
#pragma unroll 4
        for (int cc = 0; cc < ncol; ++cc) {
          const int col = col0 + cc;
          const float4 w = wbase[(cc * C::KPAD + row0) / 4];
          const ulonglong2 a = dz2[col * (kTile / 4) + bl], b = dz2[col * (kTile / 4) + 8 + bl];
          const unsigned long long w0 = dup2(w.x), w1 = dup2(w.y), w2 = dup2(w.z), w3 = dup2(w.w);
          ffma2(acc[0][0], a.x, w0); ffma2(acc[0][1], a.y, w0); ffma2(acc[0][2], b.x, w0); ffma2(acc[0][3], b.y, w0);
          ffma2(acc[1][0], a.x, w1); ffma2(acc[1][1], a.y, w1); ffma2(acc[1][2], b.x, w1); ffma2(acc[1][3], b.y, w1);
          ffma2(acc[2][0], a.x, w2); ffma2(acc[2][1], a.y, w2); ffma2(acc[2][2], b.x, w2); ffma2(acc[2][3], b.y, w2);
          ffma2(acc[3][0], a.x, w3); ffma2(acc[3][1], a.y, w3); ffma2(acc[3][2], b.x, w3); ffma2(acc[3][3], b.y, w3);
        }
      };
      if (WGLOBAL) {
        // Transposed weights streamed from L2 in 32-column chunks (double-buffered cp.async).
        constexpr int CHUNK_F4 = C::CC * C::KPAD / 4, N_CHUNKS = C::COLS / C::CC;
        auto issue_w = [&](int c) {
          float4* dst = reinterpret_cast<float4*>(wsm) + (c & 1) * CHUNK_F4;
          const float4* src = reinterpret_cast<const float4*>(wT) + static_cast<long long>(c) * CHUNK_F4;
          for (int i = tid; i < CHUNK_F4; i += C::NT) ptx::cp_async16(dst + i, src + i);
          ptx::cp_async_commit();
        };
        issue_w(0);
        for (int c = 0; c < N_CHUNKS; ++c) {
          ptx::cp_async_wait<0>();
          __syncthreads();
          if (c + 1 < N_CHUNKS) issue_w(c + 1);
          mul_cols(w4 + (c & 1) * CHUNK_F4, c * C::CC, C::CC);
        }
      } else {
        mul_cols(w4, 0, C::COLS);
      }
      // dhs was last read in phase A (before the barrier above), so it can be overwritten now.
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int row = row0 + r;
        if (row < C::K) {
          float v[8];
          unpack2(acc[r][0], v[0], v[1]); unpack2(acc[r][1], v[2], v[3]);
          unpack2(acc[r][2], v[4], v[5]); unpack2(acc[r][3], v[6], v[7]);
          float4* o4 = (row < I) ? reinterpret_cast<float4*>(dx + ((tile * T + t) * I + row) * kTile)
                                 : reinterpret_cast<float4*>(dhs + (row - I) * kTile);
          o4[bl] = make_float4(v[0], v[1], v[2], v[3]);
          o4[8 + bl] = make_float4(v[4], v[5], v[6], v[7]);
        }
      }
      __syncthreads();  // dh_rec visible to the next phase A; dzs free to be rewritten
    }
  }
}

}  // namespace gnnqc
