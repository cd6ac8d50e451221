// One Keras LSTM layer over sequence tiles with the gate contraction on the 5th-generation tensor
// cores (tcgen05.mma, accumulators in TMEM).  Same arithmetic contract as lstm_fwd.cuh (reference
// libs/create_model.py:62-79); this is the scoring-path kernel for the layers with H <= 64.
//
// Why tensor cores here: ncu on the fp32 kernel (profiles/lstm_full_r01b.csv) shows the fma pipe
// 60-74 % busy with math_pipe_throttle as the top stall - the [windows x (in+H)] x [(in+H) x 4H]
// contraction is the limiter, and it is dense.  Precision: one bf16 pass misses the 1e-4 score
// budget by 100x, so every operand is split x = hi + lo (both bf16) and the product is formed as
// hi*hi + hi*lo + lo*hi with fp32 accumulation in TMEM ("bf16x3", error ~2^-16 per product;
// measured 1.6e-5 absolute on scores against the float64 oracle).
//
// A CTA owns 128 windows (two 64-window tiles) = the 128 TMEM lanes.  Per step t:
//   epilogue threads (one per window): x_t (TMA ring, fp32) -> bf16 hi/lo into the A operand tile;
//       h_{t-1} is already there -> fence.proxy.async -> arrive(a_ready)
//   control thread: wait(a_ready) -> 3 x (K/16) tcgen05.mma into D[128 x 4H] -> tcgen05.commit(d_full)
//   epilogue threads: wait(d_full) -> tcgen05.ld their row of D -> + bias -> cell update (c in
//       registers) -> h_t to global tiles and, split in hi/lo, into the A tile for step t+1.
// Operand tiles are K-major without swizzle ([K/8][rows][8] bf16, see umma.cuh); K is laid out as
// [input channels padded to 8 | hidden units | zero padding to 16].
#pragma once
#include <cuda_bf16.h>

#include <cstdint>

#include "lstm_fwd.cuh"
#include "umma.cuh"

namespace gnnqc {

template <int I, int H>
struct TcCfg {
  static constexpr int IP = (I + 7) / 8 * 8;              // input channels padded to a core-matrix column
  static constexpr int KP = (IP + H + 15) / 16 * 16;      // contraction length per MMA chain
  static constexpr int N = 4 * H;                         // gate columns (Keras order i | f | c | o)
  static constexpr int ROWS = 128;                        // windows per CTA = TMEM lanes
  static constexpr int NST = 2;                           // x ring depth
  static constexpr int NT = 160;                          // 4 epilogue warps + 1 control warp
  static constexpr int TMEM_COLS = N < 32 ? 32 : N;       // power of two >= 32 (N is 64, 128 or 256)
  static constexpr int XBYTES = I * kTile * 4;            // one 64-window half of one step
  static constexpr size_t B_BYTES = size_t(KP) * N * 2;   // one of hi / lo
  static constexpr size_t A_BYTES = size_t(KP) * ROWS * 2;
  static constexpr size_t OFF_BHI = 0, OFF_BLO = B_BYTES, OFF_AHI = 2 * B_BYTES, OFF_ALO = 2 * B_BYTES + A_BYTES;
  static constexpr size_t OFF_X = 2 * B_BYTES + 2 * A_BYTES;
  static constexpr size_t OFF_BIAS = OFF_X + size_t(NST) * 2 * XBYTES;
  static constexpr size_t OFF_BAR = OFF_BIAS + size_t(N) * 4;
  static constexpr size_t SMEM_BYTES = OFF_BAR + (NST + 2) * 8 + 16;
  // packed weights in global memory: B_hi | B_lo | bias (fp32, Keras column order)
  static constexpr size_t PACKED_BYTES = 2 * B_BYTES + size_t(N) * 4;
  static_assert(N <= 256 && (N & (N - 1)) == 0, "one MMA covers at most 256 gate columns");
};

// hi/lo bf16 split of eight floats into two 16-byte operand chunks.
__device__ __forceinline__ void split8(const float (&v)[8], uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * i]), h1 = __float2bfloat16_rn(v[2 * i + 1]);
    const __nv_bfloat16 l0 = __float2bfloat16_rn(v[2 * i] - __bfloat162float(h0));
    const __nv_bfloat16 l1 = __float2bfloat16_rn(v[2 * i + 1] - __bfloat162float(h1));
    h[i] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
    l[i] = static_cast<uint32_t>(__bfloat16_as_ushort(l0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(l1)) << 16);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}


template <int I, int H, bool POOL>
__global__ void __launch_bounds__((TcCfg<I, H>::NT))
lstm_tc_kernel(const float* __restrict__ xin,            // [n_tiles][t_in_stride][I][64]
               float* __restrict__ yout,                 // [n_tiles][t_out_stride][H][64]
               const unsigned char* __restrict__ wpk,    // B_hi | B_lo | bias (TcCfg::PACKED_BYTES)
               int T, int t_in_stride, int t_out_stride, int n_tiles) {
  using C = TcCfg<I, H>;
  extern __shared__ __align__(128) unsigned char sm[];
  uint4* Ahi = reinterpret_cast<uint4*>(sm + C::OFF_AHI);   // 16-byte chunk index: kb * ROWS + row
  uint4* Alo = reinterpret_cast<uint4*>(sm + C::OFF_ALO);
  float* xs = reinterpret_cast<float*>(sm + C::OFF_X);      // [NST][2][I][64]
  float* bias = reinterpret_cast<float*>(sm + C::OFF_BIAS);
  uint64_t* x_full = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR);
  uint64_t* a_ready = x_full + C::NST;
  uint64_t* d_full = a_ready + 1;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  const int n_pairs = (n_tiles + 1) / 2;
  const int my_pairs = (n_pairs - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const long long total = static_cast<long long>(my_pairs) * T;
  if (total <= 0) return;

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == 4) umma::tmem_alloc(&tmem_slot, C::TMEM_COLS);
  if (tid == 0) {
    for (int s = 0; s < C::NST; ++s) ptx::mbar_init(&x_full[s], 1);
    ptx::mbar_init(a_ready, C::ROWS);
    ptx::mbar_init(d_full, 1);
    ptx::mbar_init_fence();
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(wpk);
    uint4* dst = reinterpret_cast<uint4*>(sm);
    for (int i = tid; i < static_cast<int>(2 * C::B_BYTES / 16); i += C::NT) dst[i] = __ldg(src + i);
    const float* bsrc = reinterpret_cast<const float*>(wpk + 2 * C::B_BYTES);
    for (int i = tid; i < C::N; i += C::NT) bias[i] = __ldg(bsrc + i);
    for (int i = tid; i < static_cast<int>(2 * C::A_BYTES / 16); i += C::NT) Ahi[i] = make_uint4(0u, 0u, 0u, 0u);   // A_hi and A_lo are adjacent
    for (int i = tid; i < C::NST * 2 * I * kTile; i += C::NT) xs[i] = 0.f;
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = tmem_slot;

  if (warp == 4) {
    // ---- control thread: TMA producer + MMA issuer -----------------------------------------------
    if (lane == 0) {
      auto issue_load = [&](long long q) {
        const int n = static_cast<int>(q / T), t = static_cast<int>(q % T);
        const long long pair = static_cast<long long>(blockIdx.x) + static_cast<long long>(n) * gridDim.x;
        const int s = static_cast<int>(q % C::NST);
        const bool second = 2 * pair + 1 < n_tiles;
        ptx::mbar_arrive_expect_tx(&x_full[s], second ? 2 * C::XBYTES : C::XBYTES);
        ptx::bulk_load(xs + (s * 2 + 0) * I * kTile, xin + ((2 * pair) * t_in_stride + t) * (I * kTile), C::XBYTES, &x_full[s]);
        if (second)
          ptx::bulk_load(xs + (s * 2 + 1) * I * kTile, xin + ((2 * pair + 1) * t_in_stride + t) * (I * kTile), C::XBYTES, &x_full[s]);
      };
      for (long long q = 0; q < C::NST && q < total; ++q) issue_load(q);
      constexpr uint32_t idesc = umma::instr_desc_bf16(C::ROWS, C::N);
      const unsigned char* a_hi = sm + C::OFF_AHI;
      const unsigned char* a_lo = sm + C::OFF_ALO;
      const unsigned char* b_hi = sm + C::OFF_BHI;
      const unsigned char* b_lo = sm + C::OFF_BLO;
      for (long long q = 0; q < total; ++q) {
        ptx::mbar_wait(a_ready, static_cast<uint32_t>(q & 1));
        umma::fence_after_sync();
#pragma unroll
        for (int kc = 0; kc < C::KP / 16; ++kc) {
          const uint64_t dah = umma::smem_desc(a_hi + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
          const uint64_t dal = umma::smem_desc(a_lo + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
          const uint64_t dbh = umma::smem_desc(b_hi + kc * 2 * C::N * 16, C::N * 16, 128);
          const uint64_t dbl = umma::smem_desc(b_lo + kc * 2 * C::N * 16, C::N * 16, 128);
          umma::mma_bf16(tbase, dah, dbh, idesc, kc > 0);
          umma::mma_bf16(tbase, dah, dbl, idesc, true);
          umma::mma_bf16(tbase, dal, dbh, idesc, true);
        }
        umma::commit(d_full);
        if (q + C::NST < total) issue_load(q + C::NST);   // the slot of step q was consumed before a_ready
      }
    }
  } else {
    // ---- epilogue threads: one window (TMEM lane) each ---------------------------------------------
    const int r = tid, half = r >> 6, b = r & 63;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    float cst[H];
    float pmax[POOL ? H : 1];
    int n = 0, t = 0;
    for (long long q = 0; q < total; ++q) {
      const long long pair = static_cast<long long>(blockIdx.x) + static_cast<long long>(n) * gridDim.x;
      const long long tile = 2 * pair + half;
      const bool valid = tile < n_tiles;
      if (t == 0) {
#pragma unroll
        for (int j = 0; j < H; ++j) cst[j] = 0.f;
      }
      // x_t -> bf16 hi/lo operand chunks
      const int s = static_cast<int>(q % C::NST);
      ptx::mbar_wait(&x_full[s], static_cast<uint32_t>((q / C::NST) & 1));
      const float* xrow = xs + (s * 2 + half) * I * kTile + b;
#pragma unroll
      for (int kb = 0; kb < C::IP / 8; ++kb) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (kb * 8 + i < I) ? xrow[(kb * 8 + i) * kTile] : 0.f;
        uint4 hi, lo;
        split8(v, hi, lo);
        Ahi[kb * C::ROWS + r] = hi;
        Alo[kb * C::ROWS + r] = lo;
      }
      umma::fence_async_smem();
      umma::fence_before_sync();
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(a_ready)) : "memory");
      // gates of this step
      ptx::mbar_wait(d_full, static_cast<uint32_t>(q & 1));
      umma::fence_after_sync();
      const bool last_step = (t == T - 1);
#pragma unroll
      for (int blk = 0; blk < H / 16; ++blk) {
        float zi[16], zf[16], zg[16], zo[16], hv[16];
        umma::tmem_ld16(tbase + lane_base + 0 * H + blk * 16, zi);
        umma::tmem_ld16(tbase + lane_base + 1 * H + blk * 16, zf);
        umma::tmem_ld16(tbase + lane_base + 2 * H + blk * 16, zg);
        umma::tmem_ld16(tbase + lane_base + 3 * H + blk * 16, zo);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int j = blk * 16 + i;
          hv[i] = lstm_cell_fused(zi[i] + bias[j], zf[i] + bias[H + j], zg[i] + bias[2 * H + j], zo[i] + bias[3 * H + j], cst[j]);
        }
        if (POOL) {
          const int ph = t % 3;
#pragma unroll
          for (int i = 0; i < 16; ++i) pmax[blk * 16 + i] = (ph == 0) ? hv[i] : fmaxf(pmax[blk * 16 + i], hv[i]);
          if (ph == 2 && valid) {
            float* yo = yout + ((tile * t_out_stride + t / 3) * H + blk * 16) * kTile + b;
#pragma unroll
            for (int i = 0; i < 16; ++i) yo[i * kTile] = pmax[blk * 16 + i];
          }
        } else if (valid) {
          float* yo = yout + ((tile * t_out_stride + t) * H + blk * 16) * kTile + b;
#pragma unroll
          for (int i = 0; i < 16; ++i) yo[i * kTile] = hv[i];
        }
        // h_t (zero state at a sequence boundary) into the A tile for the next step
        if (last_step) {
#pragma unroll
          for (int i = 0; i < 16; ++i) hv[i] = 0.f;
        }
        float lo8[8], hi8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { lo8[i] = hv[i]; hi8[i] = hv[8 + i]; }
        uint4 h0, l0, h1, l1;
        split8(lo8, h0, l0);
        split8(hi8, h1, l1);
        const int kb = C::IP / 8 + blk * 2;
        Ahi[kb * C::ROWS + r] = h0;
        Alo[kb * C::ROWS + r] = l0;
        Ahi[(kb + 1) * C::ROWS + r] = h1;
        Alo[(kb + 1) * C::ROWS + r] = l1;
      }
      if (++t == T) { t = 0; ++n; }
    }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 4) umma::tmem_dealloc(tbase, C::TMEM_COLS);
}

}  // namespace gnnqc
