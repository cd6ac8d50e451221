// One Keras LSTM layer over a batch of sequence tiles, forward only.
//
// Arithmetic restated from keras.layers.LSTM as the reference configures it
// (libs/create_model.py:62-79: activation tanh, recurrent_activation sigmoid, use_bias,
// zero initial state):   z = x_t W + h_{t-1} U + b   with gate columns ordered i | f | c | o,
//   c_t = sig(z_f) c_{t-1} + sig(z_i) tanh(z_c),   h_t = sig(z_o) tanh(c_t).
// MaxPooling1D(pool 3, stride 3, 'valid') (create_model.py:68,76) is folded in when POOL.
//
// Data layout ("sequence tiles"): a tensor of n sequences x T steps x C channels is stored as
// [tile = n/64][t][c][64] floats, so everything a CTA needs for one step of one tile is one
// contiguous C*256-byte block -> one cp.async.bulk (TMA) per step into a shared-memory ring.
//
// Work split: a CTA owns one tile at a time (persistent, strided over tiles) and steps through
// time.  Thread (j-lane, b-lane) of warp (jw, bw) owns RJ hidden units x RB windows: all four
// gate pre-activations of those cells are accumulated in registers, the cell state c stays in
// registers for the whole sequence, h_t goes to a double-buffered shared tile for step t+1.
// Packed weights wpk[k][unit][gate] (k < I: kernel rows, k >= I: recurrent rows) followed by
// bias[unit][gate] are resident in shared memory (or read through L1/L2 when WGLOBAL, for the
// one layer whose 393 KB do not fit).
#pragma once
#include <cstdint>

#include "model_spec.h"
#include "ptx.cuh"

namespace gnnqc {

__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) { return 1.f - __fdividef(2.f, 1.f + __expf(2.f * x)); }

template <int I, int H, int RJ, int RB, int LJ, bool WGLOBAL>
struct LstmCfg {
  static constexpr int K = I + H;
  static constexpr int LB = 32 / LJ;                  // lanes along the window axis
  static constexpr int WJ = H / (LJ * RJ);            // warps along the unit axis
  static constexpr int WB = kTile / (LB * RB);        // warps along the window axis
  static constexpr int NT = 32 * WJ * WB;
  static constexpr int NQ = RB / 4;                   // float4 chunks of windows per thread
  static constexpr int XBYTES = I * kTile * 4;        // one step of input for one tile
  static constexpr int NST = (XBYTES >= 16384) ? 2 : 4;  // ring depth
  static constexpr size_t W_FLOATS = WGLOBAL ? 0 : size_t(K) * H * 4;
  static constexpr size_t SMEM_BYTES =
      (W_FLOATS + size_t(H) * 4 + size_t(NST) * I * kTile + 2 * size_t(H) * kTile) * 4 + NST * 8;
  static_assert(H % (LJ * RJ) == 0 && kTile % (LB * RB) == 0 && RB % 4 == 0, "bad tiling");
  static_assert(XBYTES % 16 == 0, "bulk copies need 16-byte multiples");
};

template <int I, int H, int RJ, int RB, int LJ, bool POOL, bool LAST, bool WGLOBAL>
__global__ void __launch_bounds__((LstmCfg<I, H, RJ, RB, LJ, WGLOBAL>::NT))
lstm_layer_kernel(const float* __restrict__ xin,   // [n_tiles][t_in_stride][I][64]
                  float* __restrict__ yout,        // [n_tiles][t_out_stride][H][64], LAST: [n_tiles][H][64]
                  const float* __restrict__ wpk,   // packed weights + bias
                  int T, int t_in_stride, int t_out_stride, int n_tiles) {
  using C = LstmCfg<I, H, RJ, RB, LJ, WGLOBAL>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* wsm = reinterpret_cast<float*>(smem_raw);
  float* bsm = wsm + C::W_FLOATS;
  float* xs = bsm + H * 4;
  float* hs = xs + C::NST * I * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(hs + 2 * H * kTile);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int jl = lane % LJ, bl = lane / LJ;
  const int jw = warp % C::WJ, bw = warp / C::WJ;
  const int unit0 = jw * (LJ * RJ) + jl;          // unit(r) = unit0 + r*LJ
  const int b4 = bw * (C::LB * C::NQ) + bl;       // float4 column; chunk q adds q*LB

  const int my_tiles = (n_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const long long total = static_cast<long long>(my_tiles) * T;
  if (total <= 0) return;

  if (tid == 0) {
    for (int s = 0; s < C::NST; ++s) ptx::mbar_init(&bars[s], 1);
    ptx::mbar_init_fence();
  }
  // Resident weights: plain vectorised copy, once per CTA.
  {
    const float4* src = reinterpret_cast<const float4*>(wpk);
    if (!WGLOBAL) {
      float4* dst = reinterpret_cast<float4*>(wsm);
      for (int i = tid; i < C::K * H; i += C::NT) dst[i] = __ldg(src + i);
    }
    float4* bdst = reinterpret_cast<float4*>(bsm);
    for (int i = tid; i < H; i += C::NT) bdst[i] = __ldg(src + C::K * H + i);
  }
  __syncthreads();

  auto issue_load = [&](long long q) {
    const int n = static_cast<int>(q / T), t = static_cast<int>(q % T);
    const long long tile = static_cast<long long>(blockIdx.x) + static_cast<long long>(n) * gridDim.x;
    const int s = static_cast<int>(q % C::NST);
    ptx::mbar_arrive_expect_tx(&bars[s], C::XBYTES);
    ptx::bulk_load(xs + s * I * kTile, xin + (tile * t_in_stride + t) * (I * kTile), C::XBYTES, &bars[s]);
  };
  if (tid == 0) {
    for (long long q = 0; q < C::NST && q < total; ++q) issue_load(q);
  }

  const float4* w4 = WGLOBAL ? reinterpret_cast<const float4*>(wpk) : reinterpret_cast<const float4*>(wsm);
  const float4* b4p = reinterpret_cast<const float4*>(bsm);

  float cst[RJ][RB];
  float pmax[RJ][RB];
  int n = 0, t = 0;
  for (long long q = 0; q < total; ++q) {
    const long long tile = static_cast<long long>(blockIdx.x) + static_cast<long long>(n) * gridDim.x;
    if (t == 0) {
#pragma unroll
      for (int r = 0; r < RJ; ++r)
#pragma unroll
        for (int i = 0; i < RB; ++i) cst[r][i] = 0.f;
    }
    float acc[RJ][4][RB];
#pragma unroll
    for (int r = 0; r < RJ; ++r) {
      const float4 bb = b4p[unit0 + r * LJ];
#pragma unroll
      for (int i = 0; i < RB; ++i) {
        acc[r][0][i] = bb.x; acc[r][1][i] = bb.y; acc[r][2][i] = bb.z; acc[r][3][i] = bb.w;
      }
    }
    auto contract = [&](const float4* __restrict__ act4, const float4* __restrict__ wrow4, int kn) {
#pragma unroll 2
      for (int k = 0; k < kn; ++k) {
        float4 w[RJ];
#pragma unroll
        for (int r = 0; r < RJ; ++r) w[r] = WGLOBAL ? __ldg(wrow4 + k * H + unit0 + r * LJ) : wrow4[k * H + unit0 + r * LJ];
        float xv[RB];
#pragma unroll
        for (int qq = 0; qq < C::NQ; ++qq) {
          const float4 v = act4[k * (kTile / 4) + b4 + qq * C::LB];
          xv[4 * qq + 0] = v.x; xv[4 * qq + 1] = v.y; xv[4 * qq + 2] = v.z; xv[4 * qq + 3] = v.w;
        }
#pragma unroll
        for (int r = 0; r < RJ; ++r)
#pragma unroll
          for (int i = 0; i < RB; ++i) {
            acc[r][0][i] = fmaf(w[r].x, xv[i], acc[r][0][i]);
            acc[r][1][i] = fmaf(w[r].y, xv[i], acc[r][1][i]);
            acc[r][2][i] = fmaf(w[r].z, xv[i], acc[r][2][i]);
            acc[r][3][i] = fmaf(w[r].w, xv[i], acc[r][3][i]);
          }
      }
    };
    // Input projection: needs only this step's x tile (TMA ring), not h.
    {
      const int s = static_cast<int>(q % C::NST);
      ptx::mbar_wait(&bars[s], static_cast<uint32_t>((q / C::NST) & 1));
      contract(reinterpret_cast<const float4*>(xs + s * I * kTile), w4, I);
    }
    // Recurrent part: h_{t-1} from the shared double buffer (zero state at t == 0).
    if (t > 0) contract(reinterpret_cast<const float4*>(hs + (t & 1) * H * kTile), w4 + I * H, H);

    float4* hnext = reinterpret_cast<float4*>(hs + ((t + 1) & 1) * H * kTile);
#pragma unroll
    for (int r = 0; r < RJ; ++r) {
      const int unit = unit0 + r * LJ;
      float hv[RB];
#pragma unroll
      for (int i = 0; i < RB; ++i) {
        const float ig = sigmoid_fast(acc[r][0][i]);
        const float fg = sigmoid_fast(acc[r][1][i]);
        const float gg = tanh_fast(acc[r][2][i]);
        const float og = sigmoid_fast(acc[r][3][i]);
        const float cn = fmaf(fg, cst[r][i], ig * gg);
        cst[r][i] = cn;
        hv[i] = og * tanh_fast(cn);
      }
#pragma unroll
      for (int qq = 0; qq < C::NQ; ++qq)
        hnext[unit * (kTile / 4) + b4 + qq * C::LB] = make_float4(hv[4 * qq], hv[4 * qq + 1], hv[4 * qq + 2], hv[4 * qq + 3]);
      if (POOL) {
        const int ph = t % 3;
#pragma unroll
        for (int i = 0; i < RB; ++i) pmax[r][i] = (ph == 0) ? hv[i] : fmaxf(pmax[r][i], hv[i]);
        if (ph == 2) {
          float4* yo = reinterpret_cast<float4*>(yout + ((tile * t_out_stride + t / 3) * H + unit) * kTile);
#pragma unroll
          for (int qq = 0; qq < C::NQ; ++qq)
            yo[b4 + qq * C::LB] = make_float4(pmax[r][4 * qq], pmax[r][4 * qq + 1], pmax[r][4 * qq + 2], pmax[r][4 * qq + 3]);
        }
      } else if (LAST) {
        if (t == T - 1) {
          float4* yo = reinterpret_cast<float4*>(yout + (tile * H + unit) * kTile);
#pragma unroll
          for (int qq = 0; qq < C::NQ; ++qq)
            yo[b4 + qq * C::LB] = make_float4(hv[4 * qq], hv[4 * qq + 1], hv[4 * qq + 2], hv[4 * qq + 3]);
        }
      } else {
        float4* yo = reinterpret_cast<float4*>(yout + ((tile * t_out_stride + t) * H + unit) * kTile);
#pragma unroll
        for (int qq = 0; qq < C::NQ; ++qq)
          yo[b4 + qq * C::LB] = make_float4(hv[4 * qq], hv[4 * qq + 1], hv[4 * qq + 2], hv[4 * qq + 3]);
      }
    }
    __syncthreads();  // h_t visible to all; everyone is done with this step's x slot
    if (tid == 0 && q + C::NST < total) issue_load(q + C::NST);
    if (++t == T) { t = 0; ++n; }
  }
}

}  // namespace gnnqc
