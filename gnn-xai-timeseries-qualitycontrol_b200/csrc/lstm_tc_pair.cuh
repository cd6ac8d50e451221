// Two consecutive LSTM layers (I -> H -> H + MaxPooling1D(3)) in ONE tensor-core kernel: the pairs (1, 2) and (3, 4)
// of the reference's TimeLayer stack (libs/create_model.py:62-79).  Same operands and arithmetic as lstm_tc_kernel
// (fp16 hi/lo split products, fp32 accumulation in TMEM, fp32 cell update); what changes is the schedule.
//
// A single layer's step is a serial chain: cell update -> h_t into the A operand -> tcgen05.mma -> wait -> cell update,
// and ncu on the one-layer kernels shows the warps spending a third of their time in that wait (all window groups of an
// SM drift into lockstep: they compute together, then queue on the one tensor core together).  With two layers in one
// kernel the chain can be software-pipelined inside every window group, because the two layers' MMAs need each other's
// results one step apart:
//   P(m): wait D1(m) -> layer-1 cells of step m -> h1_m into A1, x_{m+1} into A1 -> MMA1(m+1) is issued
//   Q(m): wait D2(m-1) -> layer-2 cells of step m-1 -> h1_m and h2_{m-1} into A2 -> MMA2(m) is issued
// so MMA1(m+1) runs under Q(m) and MMA2(m) under P(m+1): by the time a warp comes back for an accumulator it has long
// been committed, and the tensor core works while the CUDA cores do.  h1 never touches HBM (it goes from registers
// into layer 2's operand tile).  The MMA chain is issued by whichever warp of the group arrives LAST (a shared-memory
// counter), so no warp waits for the others just to issue.
//
// H = 16: four groups of 128 threads (one window per thread).  H = 32: two groups of 256 threads, two threads per
// window, each taking 16 units of both layers - the accumulators of a group (2 x 4H columns) fill half of TMEM either way.
#pragma once
#include "lstm_tc.cuh"

namespace gnnqc {

template <int I, int H>
struct TcPairCfg {
  static constexpr int CS = H / 16;                       // threads per window (TMEM lane), 16 units of each layer per thread
  static constexpr int GROUPS = 4 / CS;
  static constexpr int GT = 128 * CS;                     // threads per group
  static constexpr int WPG = GT / 32;                     // warps per group
  static constexpr int NT = GROUPS * GT;                  // 512
  static constexpr int N = 4 * H, ROWS = 128;
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
  static constexpr size_t BAR_BYTES = 32;                 // per group: x_full, d1_full, d2_full (mbarriers), cnt1, cnt2 (u32)
  static constexpr size_t SMEM_BYTES = OFF_BAR + size_t(GROUPS) * BAR_BYTES + 16;
  static constexpr int TMEM_COLS = GROUPS * 2 * N;        // 512: D1 | D2 per group
  static_assert(H == 16 || H == 32, "pairs of 16- or 32-unit layers");
  static_assert((IP / 8) % CS == 0, "input columns are split evenly over the threads of a window");
  static_assert(SMEM_BYTES <= 227 * 1024, "operands do not fit in shared memory");
};

template <int I, int H>
__global__ void __launch_bounds__((TcPairCfg<I, H>::NT), 1)
lstm_tc_pair_kernel(const float* __restrict__ xin,            // [n_tiles][t_in_stride][I][64]
                    float* __restrict__ yout,                 // layer 2, pooled: [n_tiles][t_out_stride][H][64]
                    const unsigned char* __restrict__ wpk1,   // TcCfg<I,H>::PACKED_BYTES
                    const unsigned char* __restrict__ wpk2,   // TcCfg<H,H>::PACKED_BYTES
                    int T, int t_in_stride, int t_out_stride, int n_tiles) {
  using C = TcPairCfg<I, H>;
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ uint32_t tmem_slot;
  // warp index through __shfl_sync: everything derived from it is provably warp-uniform (see lstm_tc_kernel)
  const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int g = warp / C::WPG, wl = warp % C::WPG, cs = wl >> 2, quarter = wl & 3, half = quarter >> 1;
  const int r = quarter * 32 + lane, b = (quarter & 1) * 32 + lane;   // TMEM lane = operand row, window in its 64-window tile
  const int u0 = cs * 16;                                              // first of this thread's 16 units

  unsigned char* gsm = sm + C::OFF_GROUPS + size_t(g) * C::GROUP_BYTES;
  uint4* A1hi = reinterpret_cast<uint4*>(gsm + C::G_A1);               // 16-byte chunk index: kb * ROWS + row
  uint4* A1lo = reinterpret_cast<uint4*>(gsm + C::G_A1 + C::A1_BYTES);
  uint4* A2hi = reinterpret_cast<uint4*>(gsm + C::G_A2);
  uint4* A2lo = reinterpret_cast<uint4*>(gsm + C::G_A2 + C::A2_BYTES);
  float* xs = reinterpret_cast<float*>(gsm + C::G_X);                  // [2][I][64]
  uint64_t* x_full = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR + size_t(g) * C::BAR_BYTES);
  uint64_t* d1_full = x_full + 1;
  uint64_t* d2_full = x_full + 2;
  uint32_t* cnt1 = reinterpret_cast<uint32_t*>(x_full + 3);
  uint32_t* cnt2 = cnt1 + 1;

  // tile dealing as in lstm_tc_kernel: contiguous range per CTA, half-group slots filled round by round
  const int n_cta = static_cast<int>(gridDim.x), cta = static_cast<int>(blockIdx.x);
  const int base = n_tiles / n_cta, rem = n_tiles % n_cta;
  const int my_tiles = base + (cta < rem ? 1 : 0);
  const long long tile0 = static_cast<long long>(cta) * base + (cta < rem ? cta : rem);
  const int R = (my_tiles + 2 * C::GROUPS - 1) / (2 * C::GROUPS);
  int rounds = my_tiles - 2 * g * R;
  rounds = rounds < 0 ? 0 : (rounds > R ? R : rounds);

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == 0) umma::tmem_alloc(&tmem_slot, C::TMEM_COLS);
  if (wl == 0 && lane == 0) {
    ptx::mbar_init(x_full, 1);
    ptx::mbar_init(d1_full, 1);
    ptx::mbar_init(d2_full, 1);
    *cnt1 = 0u;
    *cnt2 = 0u;
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
    for (int i = wl * 32 + lane; i < static_cast<int>(C::GROUP_BYTES / 16); i += C::GT) gz[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  if (rounds == 0) {           // this group has no tile at all (small batches)
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem_slot, C::TMEM_COLS);
    return;
  }
  const uint32_t tmem0 = __shfl_sync(0xffffffffu, tmem_slot, 0);
  const uint32_t tb1 = tmem0 + static_cast<uint32_t>(g * 2 * C::N), tb2 = tb1 + C::N;
  const uint32_t sm_u = ptx::smem_addr(sm);
  const uint32_t gsm_u = sm_u + static_cast<uint32_t>(C::OFF_GROUPS + size_t(g) * C::GROUP_BYTES);
  const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
  constexpr uint32_t idesc = umma::instr_desc_f16(C::ROWS, C::N);

  // x of layer-1 step t of round rho (both halves through one barrier)
  auto issue_load = [&](int rho, int t) {
    const long long t_a = tile0 + static_cast<long long>(2 * g) * R + rho;
    const bool second = rho + (2 * g + 1) * R < my_tiles;
    ptx::mbar_arrive_expect_tx(x_full, second ? 2 * C::XBYTES : C::XBYTES);
    ptx::bulk_load(xs, xin + (t_a * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
    if (second) ptx::bulk_load(xs + I * kTile, xin + ((t_a + R) * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
  };
  // Publishes this warp's operand writes and tells whether it was the group's last warp to do so (that warp issues the MMAs).
  auto arrive_last = [&](uint32_t* cnt) -> bool {
    umma::fence_async_smem();
    umma::fence_before_sync();
    __syncwarp();
    uint32_t old = 0;
    if (lane == 0)   // one acquire-release atomic: this warp's operand writes are published, the last arriver sees everyone's
      asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(ptx::smem_addr(cnt)) : "memory");
    old = __shfl_sync(0xffffffffu, old, 0);
    return (old + 1u) % C::WPG == 0u;
  };
  auto mma_chain = [&](uint32_t d_tmem, uint32_t a_off, size_t a_bytes, uint32_t b_off, size_t b_bytes, uint32_t bias_off, int kslices) {
    umma::mma_bf16(d_tmem, umma::smem_desc_u(sm_u + C::OFF_ONES, C::ROWS * 16, 128), umma::smem_desc_u(sm_u + bias_off, C::N * 16, 128), idesc, false);
    for (int kc = 0; kc < kslices; ++kc) {
      const uint64_t dah = umma::smem_desc_u(gsm_u + a_off + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
      const uint64_t dal = umma::smem_desc_u(gsm_u + a_off + static_cast<uint32_t>(a_bytes) + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
      const uint64_t dbh = umma::smem_desc_u(sm_u + b_off + kc * 2 * C::N * 16, C::N * 16, 128);
      const uint64_t dbl = umma::smem_desc_u(sm_u + b_off + static_cast<uint32_t>(b_bytes) + kc * 2 * C::N * 16, C::N * 16, 128);
      umma::mma_bf16(d_tmem, dah, dbh, idesc, true);
      umma::mma_bf16(d_tmem, dah, dbl, idesc, true);
      umma::mma_bf16(d_tmem, dal, dbh, idesc, true);
    }
  };
  // this thread's share of the staged x (fp32) -> fp16 hi/lo chunks of layer 1's operand
  auto convert_x = [&](bool valid) {
    if (!valid) return;
    const float* xrow = xs + half * I * kTile + b;
#pragma unroll
    for (int kk = 0; kk < C::IP / 8 / C::CS; ++kk) {
      const int kb = cs + kk * C::CS;
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = (kb * 8 + i < I) ? xrow[(kb * 8 + i) * kTile] : 0.f;
      uint4 hi, lo;
      split8(v, hi, lo);
      A1hi[kb * C::ROWS + r] = hi;
      A1lo[kb * C::ROWS + r] = lo;
    }
  };

  uint32_t n_x = 0, n_d1 = 0, n_d2 = 0;    // barrier phases consumed so far
  // ---- prologue: x of the first step, MMA1 of the first step --------------------------------------------------------
  if (wl == 0 && umma::elect_one()) issue_load(0, 0);
  __syncwarp();
  {
    const bool valid0 = (2 * g + half) * R < my_tiles;
    ptx::mbar_wait(x_full, n_x & 1u);
    ++n_x;
    convert_x(valid0);
    if (arrive_last(cnt1)) {
      if (umma::elect_one()) {
        umma::fence_after_sync();
        mma_chain(tb1, C::G_A1, C::A1_BYTES, C::OFF_B1, C::B1_BYTES, C::OFF_BIAS, C::KP1 / 16);
        umma::commit(d1_full);
        if (T > 1) issue_load(0, 1);
        else if (rounds > 1) issue_load(1, 0);
      }
      __syncwarp();
    }
  }

  f32x2 c1[8], c2[8];            // cell states of this thread's 16 units per layer, two units per register pair
  float pmax[16];
  uint4 hold_hi[2], hold_lo[2];  // h1 of the current step, already split, until layer 2's operand tile is free
  for (int rho = 0; rho < rounds; ++rho) {
    const long long tile = tile0 + static_cast<long long>(2 * g + half) * R + rho;
    const bool valid = rho + (2 * g + half) * R < my_tiles;            // uniform over the warps of a half
    const bool next_valid = (rho + 1) + (2 * g + half) * R < my_tiles;
    for (int m = 0; m <= T; ++m) {
      // ================= P(m): layer 1, step m =======================================================================
      if (m < T) {
        if (m == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) c1[j] = 0ull;
        }
        ptx::mbar_wait(d1_full, n_d1 & 1u);
        ++n_d1;
        umma::fence_after_sync();
        if (valid) {
#pragma unroll
          for (int b8 = 0; b8 < 2; ++b8) {
            float zi[8], zf[8], zg[8], zo[8], hv[8];
            const uint32_t col0 = tb1 + lane_base + u0 + b8 * 8;
            umma::tmem_ld8x4(col0, col0 + H, col0 + 2 * H, col0 + 3 * H, zi, zf, zg, zo);
            f32x2 sv[5];
#pragma unroll
            for (int i = 0; i < 8; i += 2)
              up2(lstm_cell2<false>(pk2(zi[i], zi[i + 1]), pk2(zf[i], zf[i + 1]), pk2(zg[i], zg[i + 1]), pk2(zo[i], zo[i + 1]), c1[b8 * 4 + i / 2], sv),
                  hv[i], hv[i + 1]);
            split8(hv, hold_hi[b8], hold_lo[b8]);
            const int kb = C::IP / 8 + cs * 2 + b8;
            const bool last_step = (m == T - 1);                      // zero state for the next sequence
            A1hi[kb * C::ROWS + r] = last_step ? make_uint4(0u, 0u, 0u, 0u) : hold_hi[b8];
            A1lo[kb * C::ROWS + r] = last_step ? make_uint4(0u, 0u, 0u, 0u) : hold_lo[b8];
          }
        }
        // layer 1's next step: m + 1 of this round, or step 0 of the next round
        const bool more = (m + 1 < T) || (rho + 1 < rounds);
        if (more) {
          ptx::mbar_wait(x_full, n_x & 1u);
          ++n_x;
          convert_x(m + 1 < T ? valid : next_valid);
          if (arrive_last(cnt1)) {
            if (umma::elect_one()) {
              umma::fence_after_sync();
              mma_chain(tb1, C::G_A1, C::A1_BYTES, C::OFF_B1, C::B1_BYTES, C::OFF_BIAS, C::KP1 / 16);
              umma::commit(d1_full);
              // every thread of the group has converted the staged x: fetch the one after
              if (m + 2 < T) issue_load(rho, m + 2);
              else if (m + 2 == T) { if (rho + 1 < rounds) issue_load(rho + 1, 0); }
              else if (T > 1) issue_load(rho + 1, 1);                  // m + 1 == T: the step just issued was step 0 of round rho + 1
              else if (rho + 2 < rounds) issue_load(rho + 2, 0);
            }
            __syncwarp();
          }
        }
      }
      // ================= Q(m): layer 2, step m - 1; then its operand for step m =======================================
      if (m >= 1) {
        const int t2 = m - 1;
        if (t2 == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) c2[j] = 0ull;
        }
        ptx::mbar_wait(d2_full, n_d2 & 1u);
        ++n_d2;
        umma::fence_after_sync();
        if (valid) {
#pragma unroll
          for (int b8 = 0; b8 < 2; ++b8) {
            float zi[8], zf[8], zg[8], zo[8], hv[8];
            const uint32_t col0 = tb2 + lane_base + u0 + b8 * 8;
            umma::tmem_ld8x4(col0, col0 + H, col0 + 2 * H, col0 + 3 * H, zi, zf, zg, zo);
            f32x2 sv[5];
#pragma unroll
            for (int i = 0; i < 8; i += 2)
              up2(lstm_cell2<false>(pk2(zi[i], zi[i + 1]), pk2(zf[i], zf[i + 1]), pk2(zg[i], zg[i + 1]), pk2(zo[i], zo[i + 1]), c2[b8 * 4 + i / 2], sv),
                  hv[i], hv[i + 1]);
            const int ph = t2 % 3;
#pragma unroll
            for (int i = 0; i < 8; ++i) pmax[b8 * 8 + i] = (ph == 0) ? hv[i] : fmaxf(pmax[b8 * 8 + i], hv[i]);
            if (ph == 2) {
              float* yo = yout + ((tile * t_out_stride + t2 / 3) * H + u0 + b8 * 8) * kTile + b;
#pragma unroll
              for (int i = 0; i < 8; ++i) yo[i * kTile] = pmax[b8 * 8 + i];
            }
            uint4 hi, lo;
            if (m == T) hi = lo = make_uint4(0u, 0u, 0u, 0u);         // zero state for the next sequence
            else split8(hv, hi, lo);
            const int kb = H / 8 + cs * 2 + b8;
            A2hi[kb * C::ROWS + r] = hi;
            A2lo[kb * C::ROWS + r] = lo;
          }
        }
      }
      if (m < T) {
        if (valid) {
#pragma unroll
          for (int b8 = 0; b8 < 2; ++b8) {                             // h1_m: layer 2's input of step m
            A2hi[(cs * 2 + b8) * C::ROWS + r] = hold_hi[b8];
            A2lo[(cs * 2 + b8) * C::ROWS + r] = hold_lo[b8];
          }
        }
        if (arrive_last(cnt2)) {
          if (umma::elect_one()) {
            umma::fence_after_sync();
            mma_chain(tb2, C::G_A2, C::A2_BYTES, C::OFF_B2, C::B2_BYTES, C::OFF_BIAS + C::BT_BYTES, C::KP2 / 16);
            umma::commit(d2_full);
          }
          __syncwarp();
        }
      }
    }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem_slot, C::TMEM_COLS);
}

}  // namespace gnnqc
