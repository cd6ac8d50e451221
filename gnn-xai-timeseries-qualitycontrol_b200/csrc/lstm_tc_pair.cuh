// Two consecutive narrow LSTM layers (I -> 16 -> 16 + MaxPooling1D(3)) in ONE tensor-core kernel: the first
// pair of the reference's TimeLayer stack (libs/create_model.py:62-79), 162 of its 528 dependent steps each.
// Same operands, arithmetic and tile dealing as lstm_tc_kernel; what changes is the schedule.  ncu on the
// single-layer kernel shows 36 % of the stall samples on the MMA round trip (arrive -> elected thread ->
// tcgen05.mma -> commit -> wake-up) that every step pays once.  Layer 2's step t only needs h1_t and
// h2_{t-1}, so here one round trip carries BOTH chains: in macro-step q the elected thread issues layer 1's
// step q and layer 2's step q-1, and every thread then updates the cells of both (32 units instead of 16 per
// hand-over).  Layer 1's output goes straight into layer 2's A operand (bf16 hi/lo, shared memory) and never
// touches HBM (680 MB written and read again per 65,536 windows otherwise).
#pragma once
#include "lstm_tc.cuh"

namespace gnnqc {

template <int I>
struct TcPairCfg {
  static constexpr int H = 16, N = 4 * H, ROWS = 128, GROUPS = 4, NT = GROUPS * ROWS;
  static constexpr int IP = (I + 7) / 8 * 8;
  static constexpr int KP1 = (IP + H + 15) / 16 * 16;     // layer 1: [x | h1 | pad]
  static constexpr int KP2 = 2 * H;                       // layer 2: [h1 | h2]
  static constexpr int XBYTES = I * kTile * 4;            // one 64-window half of one step
  static constexpr size_t B1_BYTES = size_t(KP1) * N * 2, B2_BYTES = size_t(KP2) * N * 2;   // one of hi / lo
  static constexpr size_t A1_BYTES = size_t(KP1) * ROWS * 2, A2_BYTES = size_t(KP2) * ROWS * 2;
  static constexpr size_t BT_BYTES = tc_bias_tile_bytes(N);
  static constexpr size_t OFF_B1 = 0, OFF_B2 = 2 * B1_BYTES, OFF_BIAS = OFF_B2 + 2 * B2_BYTES;   // bias tile 1 | bias tile 2 | ones tile
  static constexpr size_t OFF_ONES = OFF_BIAS + 2 * BT_BYTES;
  static constexpr size_t OFF_GROUPS = OFF_ONES + kOnesTileBytes;
  static constexpr size_t G_A1 = 0, G_A2 = 2 * A1_BYTES, G_X = G_A2 + 2 * A2_BYTES;          // hi then lo in each
  static constexpr size_t GROUP_BYTES = (G_X + 2 * size_t(XBYTES) + 127) / 128 * 128;
  static constexpr size_t OFF_BAR = OFF_GROUPS + GROUPS * GROUP_BYTES;
  static constexpr size_t SMEM_BYTES = OFF_BAR + size_t(GROUPS) * 3 * 8 + 16;               // x_full, a_ready, d_full per group
  static constexpr int TMEM_COLS = GROUPS * 2 * N;        // 512: D1 | D2 per group
  static_assert(SMEM_BYTES <= 227 * 1024, "operands do not fit in shared memory");
};

template <int I>
__global__ void __launch_bounds__((TcPairCfg<I>::NT), 1)
lstm_tc_pair_kernel(const float* __restrict__ xin,            // [n_tiles][t_in_stride][I][64]
                    float* __restrict__ yout,                 // layer 2, pooled: [n_tiles][t_out_stride][16][64]
                    const unsigned char* __restrict__ wpk1,   // TcCfg<I,16>::PACKED_BYTES
                    const unsigned char* __restrict__ wpk2,   // TcCfg<16,16>::PACKED_BYTES
                    int T, int t_in_stride, int t_out_stride, int n_tiles) {
  using C = TcPairCfg<I>;
  constexpr int H = C::H;
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ uint32_t tmem_slot;
  // The warp index goes through __shfl_sync so that the compiler knows everything derived from it (group, half,
  // tile, loop bounds, operand addresses) is warp-uniform: uniform branches, and the tcgen05.mma / cp.async.bulk
  // operands stay in uniform registers instead of being broadcast lane by lane around every instruction.
  const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int g = warp >> 2, quarter = warp & 3, half = quarter >> 1;
  const int r = quarter * 32 + lane, b = (quarter & 1) * 32 + lane;

  unsigned char* gsm = sm + C::OFF_GROUPS + size_t(g) * C::GROUP_BYTES;
  uint4* A1hi = reinterpret_cast<uint4*>(gsm + C::G_A1);
  uint4* A1lo = reinterpret_cast<uint4*>(gsm + C::G_A1 + C::A1_BYTES);
  uint4* A2hi = reinterpret_cast<uint4*>(gsm + C::G_A2);
  uint4* A2lo = reinterpret_cast<uint4*>(gsm + C::G_A2 + C::A2_BYTES);
  float* xs = reinterpret_cast<float*>(gsm + C::G_X);      // [2][I][64]
  uint64_t* x_full = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR) + g * 3;
  uint64_t* a_ready = x_full + 1;
  uint64_t* d_full = x_full + 2;

  // tile dealing as in lstm_tc_kernel
  const int n_cta = static_cast<int>(gridDim.x), cta = static_cast<int>(blockIdx.x);
  const int base = n_tiles / n_cta, rem = n_tiles % n_cta;
  const int my_tiles = base + (cta < rem ? 1 : 0);
  const long long tile0 = static_cast<long long>(cta) * base + (cta < rem ? cta : rem);
  const int R = (my_tiles + 2 * C::GROUPS - 1) / (2 * C::GROUPS);
  int rounds = my_tiles - 2 * g * R;
  rounds = rounds < 0 ? 0 : (rounds > R ? R : rounds);
  const int TM = T + 1;                                   // macro-steps per sequence (layer 2 lags one step)
  const long long total = static_cast<long long>(rounds) * TM;

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == 0) umma::tmem_alloc(&tmem_slot, C::TMEM_COLS);
  if (r == 0) {
    ptx::mbar_init(x_full, 1);
    ptx::mbar_init(a_ready, C::ROWS);
    ptx::mbar_init(d_full, 1);
    ptx::mbar_init_fence();
  }
  {
    const uint4* s1 = reinterpret_cast<const uint4*>(wpk1);
    const uint4* s2 = reinterpret_cast<const uint4*>(wpk2);
    uint4* d1 = reinterpret_cast<uint4*>(sm + C::OFF_B1);
    uint4* d2 = reinterpret_cast<uint4*>(sm + C::OFF_B2);
    for (int i = tid; i < static_cast<int>(2 * C::B1_BYTES / 16); i += C::NT) d1[i] = __ldg(s1 + i);
    for (int i = tid; i < static_cast<int>(2 * C::B2_BYTES / 16); i += C::NT) d2[i] = __ldg(s2 + i);
    uint4* bt = reinterpret_cast<uint4*>(sm + C::OFF_BIAS);
    for (int i = tid; i < static_cast<int>(C::BT_BYTES / 16); i += C::NT) {
      bt[i] = __ldg(s1 + 2 * C::B1_BYTES / 16 + i);
      bt[C::BT_BYTES / 16 + i] = __ldg(s2 + 2 * C::B2_BYTES / 16 + i);
    }
    fill_ones_tile(sm + C::OFF_ONES, tid, C::NT);
    uint4* gz = reinterpret_cast<uint4*>(gsm);
    for (int i = r; i < static_cast<int>(C::GROUP_BYTES / 16); i += C::ROWS) gz[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tb1 = tmem_slot + static_cast<uint32_t>(g * 2 * C::N), tb2 = tb1 + C::N;
  const bool issuer = quarter == 0;
  const int g_u = g;
  const uint32_t sm_u = ptx::smem_addr(sm);
  const uint32_t gsm_u = sm_u + static_cast<uint32_t>(C::OFF_GROUPS + size_t(g_u) * C::GROUP_BYTES);
  const uint32_t tb1_u = __shfl_sync(0xffffffffu, tmem_slot, 0) + static_cast<uint32_t>(g_u * 2 * C::N), tb2_u = tb1_u + C::N;
  const uint64_t d_ones = umma::smem_desc_u(sm_u + C::OFF_ONES, C::ROWS * 16, 128);
  const uint64_t d_bias1 = umma::smem_desc_u(sm_u + C::OFF_BIAS, C::N * 16, 128);
  const uint64_t d_bias2 = umma::smem_desc_u(sm_u + C::OFF_BIAS + C::BT_BYTES, C::N * 16, 128);

  // x of layer-1 step t of round rho (both halves through one barrier); `n` counts the loads issued
  auto issue_load = [&](int rho, int t) {
    const long long t_a = tile0 + static_cast<long long>(2 * g) * R + rho;
    const bool second = rho + (2 * g + 1) * R < my_tiles;
    ptx::mbar_arrive_expect_tx(x_full, second ? 2 * C::XBYTES : C::XBYTES);
    ptx::bulk_load(xs, xin + (t_a * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
    if (second) ptx::bulk_load(xs + I * kTile, xin + ((t_a + R) * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
  };
  if (r == 0 && total > 0) issue_load(0, 0);

  constexpr uint32_t idesc = umma::instr_desc_f16(C::ROWS, C::N);
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  f32x2 c1[H / 2], c2[H / 2];     // cell states, two units per register pair
  float pmax[H];
  int rho = 0, t = 0;          // t = macro-step inside the sequence: layer 1 runs step t (t < T), layer 2 step t-1 (t >= 1)
  uint32_t x_loads = 0;        // x loads consumed so far (barrier phase)
  for (long long q = 0; q < total; ++q) {
    const long long tile = tile0 + static_cast<long long>(2 * g + half) * R + rho;
    const bool valid = rho + (2 * g + half) * R < my_tiles;    // uniform over the two warps of a half
    if (t == 0) {
#pragma unroll
      for (int j = 0; j < H / 2; ++j) c1[j] = c2[j] = 0ull;
    }
    // x_t -> bf16 hi/lo operand chunks of layer 1
    if (t < T) {
      if (valid) {
        ptx::mbar_wait(x_full, x_loads & 1u);
        const float* xrow = xs + half * I * kTile + b;
#pragma unroll
        for (int kb = 0; kb < C::IP / 8; ++kb) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = (kb * 8 + i < I) ? xrow[(kb * 8 + i) * kTile] : 0.f;
          uint4 hi, lo;
          split8(v, hi, lo);
          A1hi[kb * C::ROWS + r] = hi;
          A1lo[kb * C::ROWS + r] = lo;
        }
        umma::fence_async_smem();
      }
      ++x_loads;
    }
    umma::fence_before_sync();
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(a_ready)) : "memory");
    if (issuer) {
      // ---- first warp of the group, one elected lane: layer 1 step t and layer 2 step t-1 behind one commit ----
      if (umma::elect_one()) {
        ptx::mbar_wait(a_ready, static_cast<uint32_t>(q & 1));
        umma::fence_after_sync();
        if (t < T) {
          umma::mma_bf16(tb1_u, d_ones, d_bias1, idesc, false);
#pragma unroll
          for (int kc = 0; kc < C::KP1 / 16; ++kc) {
            const uint64_t dah = umma::smem_desc_u(gsm_u + C::G_A1 + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
            const uint64_t dal = umma::smem_desc_u(gsm_u + C::G_A1 + C::A1_BYTES + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
            const uint64_t dbh = umma::smem_desc_u(sm_u + C::OFF_B1 + kc * 2 * C::N * 16, C::N * 16, 128);
            const uint64_t dbl = umma::smem_desc_u(sm_u + C::OFF_B1 + C::B1_BYTES + kc * 2 * C::N * 16, C::N * 16, 128);
            umma::mma_bf16(tb1_u, dah, dbh, idesc, true);
            umma::mma_bf16(tb1_u, dah, dbl, idesc, true);
            umma::mma_bf16(tb1_u, dal, dbh, idesc, true);
          }
        }
        if (t >= 1) {
          umma::mma_bf16(tb2_u, d_ones, d_bias2, idesc, false);
#pragma unroll
          for (int kc = 0; kc < C::KP2 / 16; ++kc) {
            const uint64_t dah = umma::smem_desc_u(gsm_u + C::G_A2 + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
            const uint64_t dal = umma::smem_desc_u(gsm_u + C::G_A2 + C::A2_BYTES + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
            const uint64_t dbh = umma::smem_desc_u(sm_u + C::OFF_B2 + kc * 2 * C::N * 16, C::N * 16, 128);
            const uint64_t dbl = umma::smem_desc_u(sm_u + C::OFF_B2 + C::B2_BYTES + kc * 2 * C::N * 16, C::N * 16, 128);
            umma::mma_bf16(tb2_u, dah, dbh, idesc, true);
            umma::mma_bf16(tb2_u, dah, dbl, idesc, true);
            umma::mma_bf16(tb2_u, dal, dbh, idesc, true);
          }
        }
        umma::commit(d_full);
        // the x slot of this macro-step has been converted by every thread: fetch the next layer-1 input
        if (t + 1 < T) issue_load(rho, t + 1);
        else if (t + 1 == TM && rho + 1 < rounds) issue_load(rho + 1, 0);
      }
      __syncwarp();
    }
    ptx::mbar_wait(d_full, static_cast<uint32_t>(q & 1));
    umma::fence_after_sync();
    if (valid) {
#pragma unroll
      for (int b8 = 0; b8 < H / 8; ++b8) {
        const int j0 = b8 * 8;
        if (t < T) {   // ---- layer 1, step t ------------------------------------------------------
          float zi[8], zf[8], zg[8], zo[8], hv[8];
          const uint32_t col0 = tb1 + lane_base + b8 * 8;
          umma::tmem_ld8x4(col0, col0 + H, col0 + 2 * H, col0 + 3 * H, zi, zf, zg, zo);
          f32x2 sv[5];
#pragma unroll
          for (int i = 0; i < 8; i += 2)
            up2(lstm_cell2<false>(pk2(zi[i], zi[i + 1]), pk2(zf[i], zf[i + 1]), pk2(zg[i], zg[i + 1]), pk2(zo[i], zo[i + 1]), c1[(j0 + i) / 2], sv),
                hv[i], hv[i + 1]);
          uint4 hi, lo;
          split8(hv, hi, lo);
          A2hi[b8 * C::ROWS + r] = hi;                       // layer 2's input of step t
          A2lo[b8 * C::ROWS + r] = lo;
          if (t == T - 1) hi = lo = make_uint4(0u, 0u, 0u, 0u);   // zero state for the next sequence
          A1hi[(C::IP / 8 + b8) * C::ROWS + r] = hi;
          A1lo[(C::IP / 8 + b8) * C::ROWS + r] = lo;
        }
        if (t >= 1) {  // ---- layer 2, step t-1 ----------------------------------------------------
          const int t2 = t - 1;
          float zi[8], zf[8], zg[8], zo[8], hv[8];
          const uint32_t col0 = tb2 + lane_base + b8 * 8;
          umma::tmem_ld8x4(col0, col0 + H, col0 + 2 * H, col0 + 3 * H, zi, zf, zg, zo);
          f32x2 sv[5];
#pragma unroll
          for (int i = 0; i < 8; i += 2)
            up2(lstm_cell2<false>(pk2(zi[i], zi[i + 1]), pk2(zf[i], zf[i + 1]), pk2(zg[i], zg[i + 1]), pk2(zo[i], zo[i + 1]), c2[(j0 + i) / 2], sv),
                hv[i], hv[i + 1]);
          const int ph = t2 % 3;
#pragma unroll
          for (int i = 0; i < 8; ++i) pmax[j0 + i] = (ph == 0) ? hv[i] : fmaxf(pmax[j0 + i], hv[i]);
          if (ph == 2) {
            float* yo = yout + ((tile * t_out_stride + t2 / 3) * H + j0) * kTile + b;
#pragma unroll
            for (int i = 0; i < 8; ++i) yo[i * kTile] = pmax[j0 + i];
          }
          uint4 hi, lo;
          if (t2 == T - 1) hi = lo = make_uint4(0u, 0u, 0u, 0u);
          else split8(hv, hi, lo);
          A2hi[(H / 8 + b8) * C::ROWS + r] = hi;
          A2lo[(H / 8 + b8) * C::ROWS + r] = lo;
        }
      }
    }
    if (++t == TM) { t = 0; ++rho; }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem_slot, C::TMEM_COLS);
}

}  // namespace gnnqc
