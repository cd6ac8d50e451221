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
// bias[unit][gate] are resident in shared memory; for the one layer whose 393 KB do not fit
// (WGLOBAL) they are streamed from L2 through a double-buffered 16-row chunk ring with cp.async.
#pragma once
#include <cstdint>

#include "model_spec.h"
#include "ptx.cuh"

namespace gnnqc {

__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) { return 1.f - __fdividef(2.f, 1.f + __expf(2.f * x)); }

// One LSTM cell update, plain form: 3 sigmoids + 2 tanh = 5 ex2 + 5 rcp on the MUFU pipe.
__device__ __forceinline__ float lstm_cell_plain(float zi, float zf, float zg, float zo, float& c) {
  const float cn = fmaf(sigmoid_fast(zf), c, sigmoid_fast(zi) * tanh_fast(zg));
  c = cn;
  return sigmoid_fast(zo) * tanh_fast(cn);
}

// Same update with the divisions merged (5 ex2 + 2 rcp): with a = e^-zi, f = e^-zf, g = e^-2zg,
//   c' = [c (1+a)(1+g) + (1-g)(1+f)] / [(1+f)(1+a)(1+g)],  h = (1-e^-2c') / [(1+e^-zo)(1+e^-2c')].
// Pre-activations are clamped to +-25 (sigmoid error < 2e-11) so the products stay finite.
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lstm_cell_fused(float zi, float zf, float zg, float zo, float& c) {
  constexpr float kL2E = 1.4426950408889634f;
  const float a = ex2_approx(-kL2E * fminf(fmaxf(zi, -25.f), 25.f));
  const float f = ex2_approx(-kL2E * fminf(fmaxf(zf, -25.f), 25.f));
  const float g = ex2_approx(-2.f * kL2E * fminf(fmaxf(zg, -12.5f), 12.5f));
  const float o = ex2_approx(-kL2E * fminf(fmaxf(zo, -25.f), 25.f));
  const float pa = (1.f + a) * (1.f + g);
  const float pf = 1.f + f;
  const float cn = __fdividef(fmaf(c, pa, (1.f - g) * pf), pf * pa);
  c = cn;
  const float e = ex2_approx(-2.f * kL2E * fminf(fmaxf(cn, -12.5f), 12.5f));
  return __fdividef(1.f - e, (1.f + o) * (1.f + e));
}

// Packed fp32x2 FMA (FFMA2 in SASS): one issue slot, two lanes of the fma pipe.
__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ unsigned long long dup2(float x) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

constexpr int kVarPacked = 1;     // accumulate with fma.rn.f32x2
constexpr int kVarFusedCell = 2;  // merged-reciprocal cell update
constexpr int kVarPipelined = 4;  // explicit register double-buffering of the k-loop fragments

template <int I, int H, int RJ, int RB, int LJ, bool WGLOBAL, int NST_ = 0>
struct LstmCfg {
  static constexpr int K = I + H;
  static constexpr int LB = 32 / LJ;                  // lanes along the window axis
  static constexpr int WJ = H / (LJ * RJ);            // warps along the unit axis
  static constexpr int WB = kTile / (LB * RB);        // warps along the window axis
  static constexpr int NT = 32 * WJ * WB;
  static constexpr int NQ = RB / 4;                   // float4 chunks of windows per thread
  static constexpr int HSTR = kTile + 4;              // padded row stride of the h tile (floats): kills the
                                                      // 8-way bank conflict of the per-unit float4 stores
  static constexpr int XBYTES = I * kTile * 4;        // one step of input for one tile
  static constexpr int NST = NST_ > 0 ? NST_ : ((XBYTES >= 16384) ? 2 : 4);  // ring depth
  static constexpr int KC = 16;                       // WGLOBAL: weight rows per streamed chunk
  static constexpr size_t W_FLOATS = WGLOBAL ? size_t(2) * KC * H * 4 : size_t(K) * H * 4;
  static_assert(!WGLOBAL || (I % KC == 0 && H % KC == 0), "streamed layers need chunk-aligned widths");
  static constexpr size_t SMEM_BYTES =
      (W_FLOATS + size_t(H) * 4 + size_t(NST) * I * kTile + 2 * size_t(H) * HSTR) * 4 + NST * 8;
  static_assert(H % (LJ * RJ) == 0 && kTile % (LB * RB) == 0 && RB % 4 == 0, "bad tiling");
  static_assert(XBYTES % 16 == 0, "bulk copies need 16-byte multiples");
};

// NSAVE > 0 (used by the attribution / training paths): additionally stores, per step, the gate
// activations i, f, g, o, the cell state c and (NSAVE == 6) the output h as
// save[tile][t][quantity][unit][64], and for pooled layers which of the three pooled steps held
// the maximum (amax[tile][t/3][unit][64], first maximum wins like TF's MaxPoolGrad).
template <int I, int H, int RJ, int RB, int LJ, bool POOL, bool LAST, bool WGLOBAL, int VAR = 0, int NST_ = 0, int NSAVE = 0>
__global__ void __launch_bounds__((LstmCfg<I, H, RJ, RB, LJ, WGLOBAL, NST_>::NT))
lstm_layer_kernel(const float* __restrict__ xin,   // [n_tiles][t_in_stride][I][64]
                  float* __restrict__ yout,        // [n_tiles][t_out_stride][H][64], LAST: [n_tiles][H][64]
                  const float* __restrict__ wpk,   // packed weights + bias
                  int T, int t_in_stride, int t_out_stride, int n_tiles,
                  float* __restrict__ save = nullptr, float* __restrict__ amax = nullptr) {
  using C = LstmCfg<I, H, RJ, RB, LJ, WGLOBAL, NST_>;
  constexpr bool PACKED = (VAR & kVarPacked) != 0;
  constexpr bool SAVE = NSAVE > 0;
  constexpr bool FUSED = (VAR & kVarFusedCell) != 0 && !SAVE;
  constexpr bool PIPE = (VAR & kVarPipelined) != 0;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* wsm = reinterpret_cast<float*>(smem_raw);
  float* bsm = wsm + C::W_FLOATS;
  float* xs = bsm + H * 4;
  float* hs = xs + C::NST * I * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(hs + 2 * H * C::HSTR);

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

  const float4* w4 = reinterpret_cast<const float4*>(wsm);
  const float4* b4p = reinterpret_cast<const float4*>(bsm);

  float cst[RJ][RB];
  float pmax[RJ][RB];
  float pidx[SAVE ? RJ : 1][SAVE ? RB : 1];
  int n = 0, t = 0;
  for (long long q = 0; q < total; ++q) {
    const long long tile = static_cast<long long>(blockIdx.x) + static_cast<long long>(n) * gridDim.x;
    if (t == 0) {
#pragma unroll
      for (int r = 0; r < RJ; ++r)
#pragma unroll
        for (int i = 0; i < RB; ++i) cst[r][i] = 0.f;
    }
    // Gate pre-activations of this thread's cells. Scalar form: acc[r][gate][window]; packed
    // form: acc2[r][gate][pair of adjacent windows] as one 64-bit register pair each.
    float acc[PACKED ? 1 : RJ][4][PACKED ? 1 : RB];
    unsigned long long acc2[PACKED ? RJ : 1][4][PACKED ? RB / 2 : 1];
#pragma unroll
    for (int r = 0; r < RJ; ++r) {
      const float4 bb = b4p[unit0 + r * LJ];
      if (PACKED) {
        const unsigned long long b0 = dup2(bb.x), b1 = dup2(bb.y), b2 = dup2(bb.z), b3 = dup2(bb.w);
#pragma unroll
        for (int i = 0; i < RB / 2; ++i) { acc2[r][0][i] = b0; acc2[r][1][i] = b1; acc2[r][2][i] = b2; acc2[r][3][i] = b3; }
      } else {
#pragma unroll
        for (int i = 0; i < RB; ++i) { acc[r][0][i] = bb.x; acc[r][1][i] = bb.y; acc[r][2][i] = bb.z; acc[r][3][i] = bb.w; }
      }
    }
    auto contract = [&](const float4* __restrict__ act4, const int astr, const float4* __restrict__ wrow4, const int kn) {
      if (PACKED && PIPE) {
        // Software-pipelined: fragments of row k+1 are in flight while row k is multiplied.
        const ulonglong2* act2 = reinterpret_cast<const ulonglong2*>(act4);
        auto load = [&](int k, float4 (&w)[RJ], unsigned long long (&xp)[RB / 2]) {
#pragma unroll
          for (int r = 0; r < RJ; ++r) w[r] = wrow4[k * H + unit0 + r * LJ];
#pragma unroll
          for (int qq = 0; qq < C::NQ; ++qq) {
            const ulonglong2 v = act2[k * astr + b4 + qq * C::LB];
            xp[2 * qq] = v.x; xp[2 * qq + 1] = v.y;
          }
        };
        auto mul = [&](const float4 (&w)[RJ], const unsigned long long (&xp)[RB / 2]) {
#pragma unroll
          for (int r = 0; r < RJ; ++r) {
            const unsigned long long w0 = dup2(w[r].x), w1 = dup2(w[r].y), w2 = dup2(w[r].z), w3 = dup2(w[r].w);
#pragma unroll
            for (int i = 0; i < RB / 2; ++i) {
              ffma2(acc2[r][0][i], xp[i], w0);
              ffma2(acc2[r][1][i], xp[i], w1);
              ffma2(acc2[r][2][i], xp[i], w2);
              ffma2(acc2[r][3][i], xp[i], w3);
            }
          }
        };
        float4 wa[RJ], wb[RJ];
        unsigned long long xa[RB / 2], xb[RB / 2];
        load(0, wa, xa);
        int k = 0;
#pragma unroll 1
        for (; k + 2 <= kn; k += 2) {
          load(k + 1, wb, xb);
          mul(wa, xa);
          if (k + 2 < kn) load(k + 2, wa, xa);
          mul(wb, xb);
        }
        if (k < kn) mul(wa, xa);
        return;
      }
#pragma unroll 2
      for (int k = 0; k < kn; ++k) {
        float4 w[RJ];
#pragma unroll
        for (int r = 0; r < RJ; ++r) w[r] = wrow4[k * H + unit0 + r * LJ];
        if (PACKED) {
          unsigned long long xp[RB / 2];
          const ulonglong2* act2 = reinterpret_cast<const ulonglong2*>(act4);
#pragma unroll
          for (int qq = 0; qq < C::NQ; ++qq) {
            const ulonglong2 v = act2[k * astr + b4 + qq * C::LB];
            xp[2 * qq] = v.x; xp[2 * qq + 1] = v.y;
          }
#pragma unroll
          for (int r = 0; r < RJ; ++r) {
            const unsigned long long w0 = dup2(w[r].x), w1 = dup2(w[r].y), w2 = dup2(w[r].z), w3 = dup2(w[r].w);
#pragma unroll
            for (int i = 0; i < RB / 2; ++i) {
              ffma2(acc2[r][0][i], xp[i], w0);
              ffma2(acc2[r][1][i], xp[i], w1);
              ffma2(acc2[r][2][i], xp[i], w2);
              ffma2(acc2[r][3][i], xp[i], w3);
            }
          }
        } else {
          float xv[RB];
#pragma unroll
          for (int qq = 0; qq < C::NQ; ++qq) {
            const float4 v = act4[k * astr + b4 + qq * C::LB];
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
      }
    };
    if (WGLOBAL) {
      // Weights streamed chunk by chunk; chunk c+1 is in flight while chunk c is multiplied.
      const int n_chunks = (t > 0 ? C::K : I) / C::KC;
      constexpr int CHUNK_F4 = C::KC * H;   // float4 per chunk
      auto issue_w = [&](int c) {
        float4* dst = reinterpret_cast<float4*>(wsm) + (c & 1) * CHUNK_F4;
        const float4* src = reinterpret_cast<const float4*>(wpk) + static_cast<long long>(c) * CHUNK_F4;
        for (int i = tid; i < CHUNK_F4; i += C::NT) ptx::cp_async16(dst + i, src + i);
        ptx::cp_async_commit();
      };
      issue_w(0);
      const int s = static_cast<int>(q % C::NST);
      ptx::mbar_wait(&bars[s], static_cast<uint32_t>((q / C::NST) & 1));
      for (int c = 0; c < n_chunks; ++c) {
        ptx::cp_async_wait<0>();
        __syncthreads();                       // chunk c landed for everyone; chunk c-1 fully consumed
        if (c + 1 < n_chunks) issue_w(c + 1);
        const int k0 = c * C::KC;
        const float4* wc = w4 + (c & 1) * CHUNK_F4;
        if (k0 < I) contract(reinterpret_cast<const float4*>(xs + s * I * kTile) + k0 * (kTile / 4), kTile / 4, wc, C::KC);
        else contract(reinterpret_cast<const float4*>(hs + (t & 1) * H * C::HSTR) + (k0 - I) * (C::HSTR / 4), C::HSTR / 4, wc, C::KC);
      }
    } else {
      // Input projection: needs only this step's x tile (TMA ring), not h.
      {
        const int s = static_cast<int>(q % C::NST);
        ptx::mbar_wait(&bars[s], static_cast<uint32_t>((q / C::NST) & 1));
        contract(reinterpret_cast<const float4*>(xs + s * I * kTile), kTile / 4, w4, I);
      }
      // Recurrent part: h_{t-1} from the shared double buffer (zero state at t == 0).
      if (t > 0) contract(reinterpret_cast<const float4*>(hs + (t & 1) * H * C::HSTR), C::HSTR / 4, w4 + I * H, H);
    }

    float4* hnext = reinterpret_cast<float4*>(hs + ((t + 1) & 1) * H * C::HSTR);
#pragma unroll
    for (int r = 0; r < RJ; ++r) {
      const int unit = unit0 + r * LJ;
      float hv[RB];
      float sv[SAVE ? NSAVE : 1][SAVE ? RB : 1];
#pragma unroll
      for (int i = 0; i < RB; ++i) {
        float zi, zf, zg, zo;
        if (PACKED) {
          float lo, hi;
          unpack2(acc2[r][0][i / 2], lo, hi); zi = (i & 1) ? hi : lo;
          unpack2(acc2[r][1][i / 2], lo, hi); zf = (i & 1) ? hi : lo;
          unpack2(acc2[r][2][i / 2], lo, hi); zg = (i & 1) ? hi : lo;
          unpack2(acc2[r][3][i / 2], lo, hi); zo = (i & 1) ? hi : lo;
        } else {
          zi = acc[r][0][i]; zf = acc[r][1][i]; zg = acc[r][2][i]; zo = acc[r][3][i];
        }
        if (SAVE) {
          const float ig = sigmoid_fast(zi), fg = sigmoid_fast(zf), gg = tanh_fast(zg), og = sigmoid_fast(zo);
          const float cn = fmaf(fg, cst[r][i], ig * gg);
          cst[r][i] = cn;
          hv[i] = og * tanh_fast(cn);
          sv[0][i] = ig; sv[1][i] = fg; sv[2][i] = gg; sv[3][i] = og; sv[4][i] = cn;
          if (NSAVE > 5) sv[NSAVE > 5 ? 5 : 0][i] = hv[i];
        } else {
          hv[i] = FUSED ? lstm_cell_fused(zi, zf, zg, zo, cst[r][i]) : lstm_cell_plain(zi, zf, zg, zo, cst[r][i]);
        }
      }
      if (SAVE) {
        float4* so = reinterpret_cast<float4*>(save + (((tile * T + t) * NSAVE) * H + unit) * kTile);
#pragma unroll
        for (int v = 0; v < NSAVE; ++v)
#pragma unroll
          for (int qq = 0; qq < C::NQ; ++qq)
            so[v * (H * kTile / 4) + b4 + qq * C::LB] = make_float4(sv[v][4 * qq], sv[v][4 * qq + 1], sv[v][4 * qq + 2], sv[v][4 * qq + 3]);
      }
#pragma unroll
      for (int qq = 0; qq < C::NQ; ++qq)
        hnext[unit * (C::HSTR / 4) + b4 + qq * C::LB] = make_float4(hv[4 * qq], hv[4 * qq + 1], hv[4 * qq + 2], hv[4 * qq + 3]);
      if (POOL) {
        const int ph = t % 3;
#pragma unroll
        for (int i = 0; i < RB; ++i) {
          if (SAVE) {
            if (ph == 0) pidx[r][i] = 0.f;
            else if (hv[i] > pmax[r][i]) pidx[r][i] = static_cast<float>(ph);
          }
          pmax[r][i] = (ph == 0) ? hv[i] : fmaxf(pmax[r][i], hv[i]);
        }
        if (ph == 2) {
          if (SAVE) {
            float4* ao = reinterpret_cast<float4*>(amax + ((tile * t_out_stride + t / 3) * H + unit) * kTile);
#pragma unroll
            for (int qq = 0; qq < C::NQ; ++qq)
              ao[b4 + qq * C::LB] = make_float4(pidx[r][4 * qq], pidx[r][4 * qq + 1], pidx[r][4 * qq + 2], pidx[r][4 * qq + 3]);
          }
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
