// Tensor-core LSTM layer for the 64-unit layers of the TimeLayer stack (reference libs/create_model.py:62-79), second
// version.  Same operands and arithmetic as lstm_tc_big_kernel (fp16 hi/lo split products, fp32 accumulation in TMEM,
// fp32 cell update); the schedule is what changed.  For a 128-window step of a 64 -> 64 layer the tensor core needs
// ~3,100 cycles (3 passes x 8 K-slices x 2 chunks) and the cell update ~3,200 cycles of the CUDA cores, and
// lstm_tc_big_kernel ran them one after the other with only eight cell-update warps (ncu: 2.5 warps per scheduler,
// XU 27-37 %, 9,300 cycles per step).  Here:
//   * the INPUT half of the contraction (x_t W, K = I) does not depend on h_{t-1}: it is issued one step ahead, into a
//     second accumulator buffer (2 x 256 TMEM columns), while the cell update of the previous step runs - only the
//     recurrent half (h_{t-1} U, K = 64) stays on the step's critical path;
//   * sixteen cell-update warps instead of eight (four threads per window), so the CUDA-core part runs at the rate the
//     arithmetic allows (tools/micro/cell_rate.cu: ~50 cycles per warp-cell and scheduler with 4 warps);
//   * the gate columns stay cut into two chunks of 32 units with their own completion barriers, and every thread takes
//     8 units of EACH chunk: all sixteen warps update chunk 0 while the recurrent MMAs of chunk 1 are still running.
// One more warp issues the TMA loads and the MMAs (warp-uniform operands, one elected lane).
#pragma once
#include "lstm_tc_big.cuh"

namespace gnnqc {

template <int I, int H>
struct TcWideCfg {
  static constexpr int IP = I;                            // multiple of 16: the x slices and the h slices do not mix
  static constexpr int KP = I + H;
  static constexpr int KX = I / 16, KH = H / 16, KC = KP / 16;   // K = 16 slices: input part, recurrent part
  static constexpr int UB = 32, NC = 4 * UB, NCH = H / UB;      // chunks of 32 units x 4 gates
  static constexpr int ROWS = 128;
  static constexpr int CW = 16;                           // cell-update warps
  static constexpr int NT = (CW + 1) * 32;                // + the MMA / TMA warp
  static constexpr int DBUF = NCH * NC;                   // accumulator columns of one step
  static constexpr int TMEM_COLS = 2 * DBUF;              // two steps in flight
  static constexpr int PIECE_BYTES = 2 * 16 * NC * 2;     // hi | lo of one (chunk, K slice), as packed for lstm_tc_big_kernel
  static constexpr int N_PIECES = NCH * KC;
  static constexpr int XBYTES = I * kTile * 4;
  static constexpr size_t B_BYTES = size_t(N_PIECES) * PIECE_BYTES;
  static constexpr size_t A_BYTES = size_t(KP) * ROWS * 2;
  static constexpr size_t BT_BYTES = tc_bias_tile_bytes(NC);
  static constexpr size_t OFF_B = 0, OFF_AHI = B_BYTES, OFF_ALO = B_BYTES + A_BYTES, OFF_X = B_BYTES + 2 * A_BYTES;
  static constexpr size_t OFF_BIAS = OFF_X + 2 * size_t(XBYTES);
  static constexpr bool BIAS_MMA = OFF_BIAS + NCH * BT_BYTES + kOnesTileBytes + 1024 <= 227 * 1024;
  static constexpr size_t OFF_ONES = OFF_BIAS + (BIAS_MMA ? NCH * BT_BYTES : size_t(4 * H) * 4);
  static constexpr size_t OFF_BAR = OFF_ONES + (BIAS_MMA ? kOnesTileBytes : 0);
  static constexpr int N_BARS = 3 + NCH;                  // x_full, x_ready, a_ready, d_full[NCH]
  static constexpr size_t SMEM_BYTES = OFF_BAR + size_t(N_BARS) * 8 + 16;
  static_assert(H == 64 && I % 16 == 0 && TMEM_COLS == 512, "64-unit layers with 32 or 64 inputs");
  static_assert(SMEM_BYTES <= 227 * 1024, "operands do not fit in shared memory");
};

template <int I, int H, bool POOL, bool SAVE, bool COMPACT = false>
__global__ void __launch_bounds__((TcWideCfg<I, H>::NT), 1)
lstm_tc_wide_kernel(const float* __restrict__ xin,            // [n_tiles][t_in_stride][I][64]
                    float* __restrict__ yout,                 // [n_tiles][t_out_stride][H][64]
                    const unsigned char* __restrict__ wpk,    // pieces [chunk][K slice]{hi, lo} | bias tiles (lstm_tc.cu)
                    int T, int t_in_stride, int t_out_stride, int n_tiles,
                    float* __restrict__ save = nullptr,       // SAVE: [n_tiles][T][nsave][H][64] gates i, f, g, o, c (, h)
                    float* __restrict__ amax = nullptr,       // SAVE && POOL: [n_tiles][t_out_stride][H][64] argmax step
                    int nsave = 5) {
  using C = TcWideCfg<I, H>;
  extern __shared__ __align__(128) unsigned char sm[];
  uint4* Ahi = reinterpret_cast<uint4*>(sm + C::OFF_AHI);   // 16-byte chunk index: kb * ROWS + row
  uint4* Alo = reinterpret_cast<uint4*>(sm + C::OFF_ALO);
  float* xs = reinterpret_cast<float*>(sm + C::OFF_X);      // [2][I][64]
  uint64_t* x_full = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR);
  uint64_t* x_ready = x_full + 1;
  uint64_t* a_ready = x_full + 2;
  uint64_t* d_full = x_full + 3;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform (see lstm_tc_kernel)

  // contiguous, equal tile ranges per CTA, taken two tiles (128 windows) at a time
  const int n_cta = static_cast<int>(gridDim.x), cta = static_cast<int>(blockIdx.x);
  const int my_tiles = n_tiles / n_cta + (cta < n_tiles % n_cta ? 1 : 0);
  const long long tile0 = static_cast<long long>(cta) * (n_tiles / n_cta) + (cta < n_tiles % n_cta ? cta : n_tiles % n_cta);
  const int my_pairs = (my_tiles + 1) / 2;
  const long long total = static_cast<long long>(my_pairs) * T;
  if (total <= 0) return;

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == C::CW) umma::tmem_alloc(&tmem_slot, C::TMEM_COLS);
  if (tid == 0) {
    ptx::mbar_init(x_full, 1);
    ptx::mbar_init(x_ready, 32 * C::CW);
    ptx::mbar_init(a_ready, 32 * C::CW);
    for (int c = 0; c < C::NCH; ++c) ptx::mbar_init(&d_full[c], 1);
    ptx::mbar_init_fence();
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(wpk);
    uint4* dst = reinterpret_cast<uint4*>(sm + C::OFF_B);
    for (int i = tid; i < static_cast<int>(C::B_BYTES / 16); i += C::NT) dst[i] = __ldg(src + i);
    if (C::BIAS_MMA) {
      uint4* bdst = reinterpret_cast<uint4*>(sm + C::OFF_BIAS);
      for (int i = tid; i < static_cast<int>(C::NCH * C::BT_BYTES / 16); i += C::NT) bdst[i] = __ldg(src + C::B_BYTES / 16 + i);
      fill_ones_tile(sm + C::OFF_ONES, tid, C::NT);
    } else {
      const __half* bt = reinterpret_cast<const __half*>(wpk + C::B_BYTES);   // bias tile of chunk c, row gate * 32 + unit: {hi, lo, 0 ...}
      float* bias = reinterpret_cast<float*>(sm + C::OFF_BIAS);
      for (int i = tid; i < 4 * H; i += C::NT) {
        const int gate = i / H, unit = i % H, c = unit / C::UB, n = gate * C::UB + unit % C::UB;
        const __half* row = bt + (size_t(c) * C::BT_BYTES / 2) + n * 8;
        bias[i] = __half2float(row[0]) + __half2float(row[1]);
      }
    }
    for (int i = tid; i < static_cast<int>(2 * C::A_BYTES / 16); i += C::NT) Ahi[i] = make_uint4(0u, 0u, 0u, 0u);   // A_hi and A_lo are adjacent
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = __shfl_sync(0xffffffffu, tmem_slot, 0);

  if (warp == C::CW) {
    // ================= TMA loads + MMA issue (one elected lane, warp-uniform operands) ==================================
    if (umma::elect_one()) {
      const uint32_t sm_u = ptx::smem_addr(sm);
      auto issue_load = [&](long long q) {
        const int n = static_cast<int>(q / T), t = static_cast<int>(q % T);
        const long long tile_a = tile0 + 2 * n;
        const bool second = 2 * n + 1 < my_tiles;
        ptx::mbar_arrive_expect_tx(x_full, second ? 2 * C::XBYTES : C::XBYTES);
        ptx::bulk_load(xs, xin + (tile_a * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
        if (second) ptx::bulk_load(xs + I * kTile, xin + ((tile_a + 1) * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
      };
      constexpr uint32_t idesc = umma::instr_desc_f16(C::ROWS, C::NC);
      // one K = 16 slice of chunk c: three split products into D
      auto mma_slice = [&](uint32_t d, int c, int kc, bool accumulate) {
        const uint32_t bp = sm_u + C::OFF_B + static_cast<uint32_t>(c * C::KC + kc) * C::PIECE_BYTES;
        const uint64_t dah = umma::smem_desc_u(sm_u + C::OFF_AHI + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
        const uint64_t dal = umma::smem_desc_u(sm_u + C::OFF_ALO + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
        const uint64_t dbh = umma::smem_desc_u(bp, C::NC * 16, 128);
        const uint64_t dbl = umma::smem_desc_u(bp + C::PIECE_BYTES / 2, C::NC * 16, 128);
        umma::mma_bf16(d, dah, dbh, idesc, accumulate);
        umma::mma_bf16(d, dah, dbl, idesc, true);
        umma::mma_bf16(d, dal, dbh, idesc, true);
      };
      // the input half of step q: opens both chunks' accumulators of buffer q & 1
      auto x_part = [&](long long q) {
        const uint32_t d0 = tbase + static_cast<uint32_t>(q & 1) * C::DBUF;
        for (int c = 0; c < C::NCH; ++c) {
          if (C::BIAS_MMA)
            umma::mma_bf16(d0 + c * C::NC, umma::smem_desc_u(sm_u + C::OFF_ONES, C::ROWS * 16, 128),
                           umma::smem_desc_u(sm_u + C::OFF_BIAS + c * C::BT_BYTES, C::NC * 16, 128), idesc, false);
#pragma unroll
          for (int kc = 0; kc < C::KX; ++kc) mma_slice(d0 + c * C::NC, c, kc, C::BIAS_MMA || kc > 0);
        }
      };
      issue_load(0);
      ptx::mbar_wait(x_ready, 0u);            // x_0 converted
      umma::fence_after_sync();
      if (total > 1) issue_load(1);
      x_part(0);
      for (long long q = 0; q < total; ++q) {
        ptx::mbar_wait(a_ready, static_cast<uint32_t>(q & 1));     // h_{q-1} (or the zero state) is in the operand tile
        umma::fence_after_sync();
        const uint32_t d0 = tbase + static_cast<uint32_t>(q & 1) * C::DBUF;
        for (int c = 0; c < C::NCH; ++c) {
#pragma unroll
          for (int kc = C::KX; kc < C::KC; ++kc) mma_slice(d0 + c * C::NC, c, kc, true);
          umma::commit(&d_full[c]);
        }
        if (q + 1 < total) {
          ptx::mbar_wait(x_ready, static_cast<uint32_t>((q + 1) & 1));   // x_{q+1} converted (which also frees the staging slot)
          umma::fence_after_sync();
          if (q + 2 < total) issue_load(q + 2);
          x_part(q + 1);
        }
      }
    }
  } else {
    // ================= cell-update warps: one window (TMEM lane) and 16 units per thread ================================
    const int quarter = warp & 3, cs = warp >> 2;
    const int r = quarter * 32 + lane, half = quarter >> 1, b = (quarter & 1) * 32 + lane;
    // every thread takes 8 units of EACH chunk (units c * 32 + cs * 8 ..), so all sixteen warps work on chunk 0 while the
    // recurrent MMAs of chunk 1 are still running, instead of half of them waiting
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    f32x2 cst[8];
    float pmax[POOL ? 16 : 1];
    unsigned long long parg = 0ull;
    uint4 hold_hi[2], hold_lo[2];

    // this thread's share of x (fp32, TMA slot) -> fp16 hi/lo chunks of the A operand: I/8 column blocks over 4 threads per row
    auto convert_x = [&](bool valid) {
      if (!valid) return;
      const float* xrow = xs + half * I * kTile + b;
#pragma unroll
      for (int kk = 0; kk < I / 8 / 4; ++kk) {
        const int kb = cs + kk * 4;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = xrow[(kb * 8 + i) * kTile];
        uint4 hi, lo;
        split8(v, hi, lo);
        Ahi[kb * C::ROWS + r] = hi;
        Alo[kb * C::ROWS + r] = lo;
      }
    };
    auto publish = [&](uint64_t* bar) {
      umma::fence_async_smem();
      umma::fence_before_sync();
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(bar)) : "memory");
    };

    ptx::mbar_wait(x_full, 0u);
    convert_x(half < my_tiles);
    publish(x_ready);
    publish(a_ready);                                       // the zero state of the first step
    int n = 0, t = 0;
    for (long long q = 0; q < total; ++q) {
      const long long tile = tile0 + 2 * n + half;
      const bool valid = 2 * n + half < my_tiles;           // uniform over the warps of a half
      const uint32_t par = static_cast<uint32_t>(q & 1);
      if (t == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) cst[j] = 0ull;
      }
      const bool last_step = (t == T - 1);
      ptx::mbar_wait(&d_full[0], par);                      // chunk 0's accumulators; every earlier MMA (x part included) is done too
      umma::fence_after_sync();
      if (q + 1 < total) {                                  // x_{q+1}: the input half of the next step can start under this cell update
        ptx::mbar_wait(x_full, static_cast<uint32_t>((q + 1) & 1));
        const int n1 = (t + 1 == T) ? n + 1 : n;
        convert_x(2 * n1 + half < my_tiles);
        publish(x_ready);
      }
      {
#pragma unroll
        for (int b8 = 0; b8 < 2; ++b8) {                    // b8 = chunk
          if (b8 > 0) {
            ptx::mbar_wait(&d_full[b8], par);
            umma::fence_after_sync();
          }
          if (!valid) continue;
          float zi[8], zf[8], zg[8], zo[8], hv[8];
          const uint32_t col0 = tbase + par * C::DBUF + lane_base + static_cast<uint32_t>(b8 * C::NC + cs * 8);
          umma::tmem_ld8x4(col0 + 0 * C::UB, col0 + 1 * C::UB, col0 + 2 * C::UB, col0 + 3 * C::UB, zi, zf, zg, zo);
          const int j0 = b8 * C::UB + cs * 8;               // first unit of this block
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            f32x2 sv[5];
            f32x2 pi = pk2(zi[i], zi[i + 1]), pf = pk2(zf[i], zf[i + 1]), pg = pk2(zg[i], zg[i + 1]), po = pk2(zo[i], zo[i + 1]);
            if (!C::BIAS_MMA) {
              const float* bias = reinterpret_cast<const float*>(sm + C::OFF_BIAS);
              pi = add2(pi, *reinterpret_cast<const f32x2*>(bias + j0 + i));
              pf = add2(pf, *reinterpret_cast<const f32x2*>(bias + H + j0 + i));
              pg = add2(pg, *reinterpret_cast<const f32x2*>(bias + 2 * H + j0 + i));
              po = add2(po, *reinterpret_cast<const f32x2*>(bias + 3 * H + j0 + i));
            }
            up2(lstm_cell2<SAVE>(pi, pf, pg, po, cst[b8 * 4 + i / 2], sv), hv[i], hv[i + 1]);
            if (SAVE) {
              save_cells2<COMPACT>(save + (((tile * T + t) * nsave) * H + j0 + i) * kTile + b, nsave, H, sv, hv[i], hv[i + 1]);
            }
          }
          if (POOL) {
            const int ph = t % 3;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (SAVE) pool_arg_update(parg, b8 * 8 + i, ph, hv[i] > pmax[b8 * 8 + i]);
              pmax[b8 * 8 + i] = (ph == 0) ? hv[i] : fmaxf(pmax[b8 * 8 + i], hv[i]);
            }
            if (ph == 2) {
              float* yo = yout + ((tile * t_out_stride + t / 3) * H + j0) * kTile + b;
#pragma unroll
              for (int i = 0; i < 8; ++i) yo[i * kTile] = pmax[b8 * 8 + i];
              if (SAVE) {
                float* ao = amax + ((tile * t_out_stride + t / 3) * H + j0) * kTile + b;
#pragma unroll
                for (int i = 0; i < 8; ++i) ao[i * kTile] = static_cast<float>((parg >> (2 * (b8 * 8 + i))) & 3ull);
              }
            }
          } else {
            float* yo = yout + ((tile * t_out_stride + t) * H + j0) * kTile + b;
#pragma unroll
            for (int i = 0; i < 8; ++i) yo[i * kTile] = hv[i];
          }
          if (last_step) {   // the next step starts a new sequence from the zero state
#pragma unroll
            for (int i = 0; i < 8; ++i) hv[i] = 0.f;
          }
          split8(hv, hold_hi[b8], hold_lo[b8]);
        }
      }
      if (q + 1 < total) {
        // every recurrent MMA of this step is done with the operand tile (d_full[1] was waited for above): h_t goes in
        if (valid) {
#pragma unroll
          for (int b8 = 0; b8 < 2; ++b8) {
            const int kb = I / 8 + b8 * (C::UB / 8) + cs;
            Ahi[kb * C::ROWS + r] = hold_hi[b8];
            Alo[kb * C::ROWS + r] = hold_lo[b8];
          }
        }
        publish(a_ready);
      }
      if (++t == T) { t = 0; ++n; }
    }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == C::CW) umma::tmem_dealloc(tbase, C::TMEM_COLS);
}

}  // namespace gnnqc
