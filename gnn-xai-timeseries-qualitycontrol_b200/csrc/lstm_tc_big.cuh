// Tensor-core LSTM layer for the wide layers of the TimeLayer stack (H = 64 and H = 128; reference
// libs/create_model.py:62-79).  Same arithmetic contract as lstm_tc.cuh (bf16x3 split products,
// fp32 accumulation in TMEM, fp32 cell update); what changes is how the work is cut so that the
// operands of a [128 windows x (I+H)] x [(I+H) x 4H] step fit next to each other on one SM:
//
//  * the 4H gate columns are cut into chunks of 32 units x 4 gates = 128 columns; each chunk is its
//    own MMA chain into its own 128 TMEM columns with its own completion barrier, so the cell update
//    of chunk c overlaps the MMAs of chunk c+1;
//  * two groups of 128 cell-update threads (warps 0-3 and 4-7; a warp reaches the TMEM lanes
//    32*(warp%4)..+31) take alternate chunks, so a thread carries the cell state of H/2 units;
//  * the weight operand is stored in 8 KB pieces (one chunk x one K=16 slice, hi then lo).  For
//    H = 64 all pieces stay resident in shared memory; for H = 128 (384 KB of weights) a producer
//    thread streams them from L2 through a ring of pieces with TMA, every step.
//  * h_t can only be written into the A operand once every MMA of step t has read it, so a thread
//    keeps its h values (already split in bf16 hi/lo) in registers until the last chunk completes.
#pragma once
#include <cuda_bf16.h>

#include <cstdint>

#include "lstm_tc.cuh"

namespace gnnqc {

template <int I, int H, bool STREAM>
struct TcBigCfg {
  static constexpr int IP = (I + 7) / 8 * 8;
  static constexpr int KP = (IP + H + 15) / 16 * 16;
  static constexpr int KC = KP / 16;                      // K = 16 slices per chain
  static constexpr int UB = 32;                           // units per chunk
  static constexpr int NC = 4 * UB;                       // gate columns per chunk
  static constexpr int NCH = H / UB;                      // chunks
  static constexpr int G = 2;                             // cell-update thread groups
  static constexpr int CPG = NCH / G;                     // chunks per group
  static constexpr int ROWS = 128;
  static constexpr int NT = (4 * G + 2) * 32;             // + MMA/x warp + weight producer warp
  static constexpr int TMEM_COLS = NCH * NC;              // 256 or 512
  static constexpr int PIECE_BYTES = 2 * 16 * NC * 2;     // hi | lo of one (chunk, K slice)
  static constexpr int N_PIECES = NCH * KC;
  static constexpr int RING = STREAM ? 8 : N_PIECES;
  static constexpr int XBYTES = I * kTile * 4;            // one 64-window half of one step
  static constexpr size_t B_BYTES = size_t(RING) * PIECE_BYTES;
  static constexpr size_t A_BYTES = size_t(KP) * ROWS * 2;
  static constexpr size_t OFF_B = 0, OFF_AHI = B_BYTES, OFF_ALO = B_BYTES + A_BYTES, OFF_X = B_BYTES + 2 * A_BYTES;
  static constexpr size_t BT_BYTES = tc_bias_tile_bytes(NC);       // one bias tile per chunk (resident also when the weights stream)
  static constexpr size_t OFF_BIAS = OFF_X + 2 * size_t(XBYTES);
  // 64 -> 64 units leaves no room for the bias / ones tiles next to the resident weights: fp32 biases in shared memory instead
  static constexpr bool BIAS_MMA = OFF_BIAS + NCH * BT_BYTES + kOnesTileBytes + 1024 <= 227 * 1024;
  static constexpr size_t OFF_ONES = OFF_BIAS + (BIAS_MMA ? NCH * BT_BYTES : size_t(4 * H) * 4);
  static constexpr size_t OFF_BAR = OFF_ONES + (BIAS_MMA ? kOnesTileBytes : 0);
  static constexpr int N_BARS = 2 + NCH + 2 * RING;       // x_full, a_ready, d_full[NCH], b_full[RING], b_empty[RING]
  static constexpr size_t SMEM_BYTES = OFF_BAR + size_t(N_BARS) * 8 + 16;
  static constexpr size_t PACKED_BYTES = size_t(N_PIECES) * PIECE_BYTES + NCH * BT_BYTES;
  static_assert(I % 8 == 0 && H % (UB * G) == 0 && TMEM_COLS <= 512, "shape not covered");
  static_assert(SMEM_BYTES <= 227 * 1024, "operands do not fit in shared memory");
};

template <int I, int H, bool POOL, bool LAST, bool STREAM, bool SAVE = false, bool COMPACT = false>
__global__ void __launch_bounds__((TcBigCfg<I, H, STREAM>::NT), 1)
lstm_tc_big_kernel(const float* __restrict__ xin,            // [n_tiles][t_in_stride][I][64]
                   float* __restrict__ yout,                 // [n_tiles][t_out_stride][H][64], LAST: [n_tiles][H][64]
                   const unsigned char* __restrict__ wpk,    // pieces [chunk][K slice]{hi, lo} | bias
                   int T, int t_in_stride, int t_out_stride, int n_tiles,
                   float* __restrict__ save = nullptr,       // SAVE: [n_tiles][T][5][H][64] gates i, f, g, o and c
                   float* __restrict__ amax = nullptr,       // SAVE && POOL: [n_tiles][t_out_stride][H][64] argmax step
               int nsave = 5) {                          // SAVE: 5, 6 (h_t as well: training) or kSaveCompact (save_cells2)
  using C = TcBigCfg<I, H, STREAM>;
  extern __shared__ __align__(128) unsigned char sm[];
  uint4* Ahi = reinterpret_cast<uint4*>(sm + C::OFF_AHI);   // 16-byte chunk index: kb * ROWS + row
  uint4* Alo = reinterpret_cast<uint4*>(sm + C::OFF_ALO);
  float* xs = reinterpret_cast<float*>(sm + C::OFF_X);      // [2][I][64]
  uint64_t* x_full = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR);
  uint64_t* a_ready = x_full + 1;
  uint64_t* d_full = a_ready + 1;
  uint64_t* b_full = d_full + C::NCH;
  uint64_t* b_empty = b_full + C::RING;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int kMmaWarp = 4 * C::G, kLoadWarp = 4 * C::G + 1;

  // Tiles are dealt to CTAs in contiguous, equal ranges (6 or 7 of 1,024 tiles on 148 SMs) and taken two
  // at a time; in a CTA's last round the second half may be empty, and its two warps then skip the
  // cell work.
  const int n_cta = static_cast<int>(gridDim.x), cta = static_cast<int>(blockIdx.x);
  const int my_tiles = n_tiles / n_cta + (cta < n_tiles % n_cta ? 1 : 0);
  const long long tile0 = static_cast<long long>(cta) * (n_tiles / n_cta) + (cta < n_tiles % n_cta ? cta : n_tiles % n_cta);
  const int my_pairs = (my_tiles + 1) / 2;
  const long long total = static_cast<long long>(my_pairs) * T;
  if (total <= 0) return;

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == kMmaWarp) umma::tmem_alloc(&tmem_slot, C::TMEM_COLS);
  if (tid == 0) {
    ptx::mbar_init(x_full, 1);
    ptx::mbar_init(a_ready, 128 * C::G);
    for (int c = 0; c < C::NCH; ++c) ptx::mbar_init(&d_full[c], 1);
    for (int s = 0; s < C::RING; ++s) {
      ptx::mbar_init(&b_full[s], 1);
      ptx::mbar_init(&b_empty[s], 1);
    }
    ptx::mbar_init_fence();
  }
  {
    if (!STREAM) {
      const uint4* src = reinterpret_cast<const uint4*>(wpk);
      uint4* dst = reinterpret_cast<uint4*>(sm + C::OFF_B);
      for (int i = tid; i < static_cast<int>(C::B_BYTES / 16); i += C::NT) dst[i] = __ldg(src + i);
    }
    if (C::BIAS_MMA) {
      const uint4* bsrc = reinterpret_cast<const uint4*>(wpk + size_t(C::N_PIECES) * C::PIECE_BYTES);
      uint4* bdst = reinterpret_cast<uint4*>(sm + C::OFF_BIAS);
      for (int i = tid; i < static_cast<int>(C::NCH * C::BT_BYTES / 16); i += C::NT) bdst[i] = __ldg(bsrc + i);
      fill_ones_tile(sm + C::OFF_ONES, tid, C::NT);
    } else {
      // bias tile of chunk c, row n = gate * 32 + unit-in-chunk: {hi, lo, 0 ...}  ->  bias[gate * H + c * 32 + unit]
      const __half* bt = reinterpret_cast<const __half*>(wpk + size_t(C::N_PIECES) * C::PIECE_BYTES);
      float* bias = reinterpret_cast<float*>(sm + C::OFF_BIAS);
      for (int i = tid; i < 4 * H; i += C::NT) {
        const int gate = i / H, unit = i % H, c = unit / C::UB, n = gate * C::UB + unit % C::UB;
        const __half* row = bt + (size_t(c) * C::BT_BYTES / 2) + n * 8;
        bias[i] = __half2float(row[0]) + __half2float(row[1]);
      }
    }
    for (int i = tid; i < static_cast<int>(2 * C::A_BYTES / 16); i += C::NT) Ahi[i] = make_uint4(0u, 0u, 0u, 0u);   // A_hi and A_lo are adjacent
    for (int i = tid; i < 2 * I * kTile; i += C::NT) xs[i] = 0.f;
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = tmem_slot;

  const int warp_u = __shfl_sync(0xffffffffu, warp, 0);     // warp-uniform (see umma::elect_one)
  if (warp_u == kMmaWarp) {
    // ---- x loads + MMA issue (one elected lane, warp-uniform operands) ------------------------------
    const uint32_t sm_u = ptx::smem_addr(sm);
    const uint32_t tbase_u = __shfl_sync(0xffffffffu, tmem_slot, 0);
    if (umma::elect_one()) {
      auto issue_load = [&](long long q) {
        const int n = static_cast<int>(q / T), t = static_cast<int>(q % T);
        const long long tile_a = tile0 + 2 * n;
        const bool second = 2 * n + 1 < my_tiles;
        ptx::mbar_arrive_expect_tx(x_full, second ? 2 * C::XBYTES : C::XBYTES);
        ptx::bulk_load(xs, xin + (tile_a * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
        if (second) ptx::bulk_load(xs + I * kTile, xin + ((tile_a + 1) * t_in_stride + t) * (I * kTile), C::XBYTES, x_full);
      };
      issue_load(0);
      constexpr uint32_t idesc = umma::instr_desc_f16(C::ROWS, C::NC);
      const uint64_t d_ones = umma::smem_desc_u(sm_u + C::OFF_ONES, C::ROWS * 16, 128);
      const uint32_t a_hi = sm_u + C::OFF_AHI, a_lo = sm_u + C::OFF_ALO, bsm = sm_u + C::OFF_B;
      long long pc = 0;   // pieces consumed (STREAM)
      for (long long q = 0; q < total; ++q) {
        ptx::mbar_wait(a_ready, static_cast<uint32_t>(q & 1));
        umma::fence_after_sync();
        if (q + 1 < total) issue_load(q + 1);   // every thread has converted x_q by now
#pragma unroll 1
        for (int c = 0; c < C::NCH; ++c) {
          if (C::BIAS_MMA) umma::mma_bf16(tbase_u + c * C::NC, d_ones, umma::smem_desc_u(sm_u + C::OFF_BIAS + c * C::BT_BYTES, C::NC * 16, 128), idesc, false);
#pragma unroll
          for (int kc = 0; kc < C::KC; ++kc) {
            int slot = c * C::KC + kc;
            if (STREAM) {
              slot = static_cast<int>(pc % C::RING);
              ptx::mbar_wait(&b_full[slot], static_cast<uint32_t>((pc / C::RING) & 1));
              umma::fence_after_sync();
            }
            const uint32_t bp = bsm + static_cast<uint32_t>(slot) * C::PIECE_BYTES;
            const uint64_t dah = umma::smem_desc_u(a_hi + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
            const uint64_t dal = umma::smem_desc_u(a_lo + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
            const uint64_t dbh = umma::smem_desc_u(bp, C::NC * 16, 128);
            const uint64_t dbl = umma::smem_desc_u(bp + C::PIECE_BYTES / 2, C::NC * 16, 128);
            umma::mma_bf16(tbase_u + c * C::NC, dah, dbh, idesc, C::BIAS_MMA || kc > 0);
            umma::mma_bf16(tbase_u + c * C::NC, dah, dbl, idesc, true);
            umma::mma_bf16(tbase_u + c * C::NC, dal, dbh, idesc, true);
            if (STREAM) {
              umma::commit(&b_empty[slot]);
              ++pc;
            }
          }
          umma::commit(&d_full[c]);
        }
      }
    }
  } else if (warp_u == kLoadWarp) {
    // ---- weight pieces through the ring (H = 128 only) ---------------------------------------------
    if (STREAM && lane == 0) {
      const long long n_loads = total * C::N_PIECES;
      int piece = 0;
      for (long long pc = 0; pc < n_loads; ++pc) {
        const int slot = static_cast<int>(pc % C::RING);
        const long long round = pc / C::RING;
        if (round > 0) ptx::mbar_wait(&b_empty[slot], static_cast<uint32_t>((round - 1) & 1));
        ptx::mbar_arrive_expect_tx(&b_full[slot], C::PIECE_BYTES);
        ptx::bulk_load(sm + C::OFF_B + size_t(slot) * C::PIECE_BYTES, wpk + size_t(piece) * C::PIECE_BYTES, C::PIECE_BYTES, &b_full[slot]);
        if (++piece == C::N_PIECES) piece = 0;
      }
    }
  } else {
    // ---- cell-update threads: one window (TMEM lane) and every other chunk of units each ----------
    const int quarter = warp & 3, g = warp >> 2;
    const int r = quarter * 32 + lane, half = r >> 6, b = r & 63;
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    f32x2 cst[C::CPG * C::UB / 2];                   // cell states, two units per register pair
    float pmax[POOL ? C::CPG * C::UB : 1];
    unsigned long long parg = 0ull;
    uint4 hold_hi[C::CPG][C::UB / 8], hold_lo[C::CPG][C::UB / 8];
#pragma unroll
    for (int ci = 0; ci < C::CPG; ++ci)
#pragma unroll
      for (int b8 = 0; b8 < C::UB / 8; ++b8) hold_hi[ci][b8] = hold_lo[ci][b8] = make_uint4(0u, 0u, 0u, 0u);

    // this thread's share of x_q (fp32, TMA slot) -> bf16 hi/lo chunks of the A operand
    auto convert_x = [&](long long q) {
      ptx::mbar_wait(x_full, static_cast<uint32_t>(q & 1));
      const float* xrow = xs + half * I * kTile + b;
#pragma unroll
      for (int kk = 0; kk < C::IP / 8 / C::G; ++kk) {
        const int kb = g + kk * C::G;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = xrow[(kb * 8 + i) * kTile];
        uint4 hi, lo;
        split8(v, hi, lo);
        Ahi[kb * C::ROWS + r] = hi;
        Alo[kb * C::ROWS + r] = lo;
      }
    };
    auto publish_a = [&]() {
      umma::fence_async_smem();
      umma::fence_before_sync();
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(a_ready)) : "memory");
    };

    convert_x(0);
    publish_a();
    int n = 0, t = 0;
    for (long long q = 0; q < total; ++q) {
      const long long tile = tile0 + 2 * n + half;
      const bool valid = 2 * n + half < my_tiles;     // uniform over the two warps of a half
      const uint32_t par = static_cast<uint32_t>(q & 1);
      if (t == 0) {
#pragma unroll
        for (int j = 0; j < C::CPG * C::UB / 2; ++j) cst[j] = 0ull;
      }
      const bool last_step = (t == T - 1);
#pragma unroll
      for (int ci = 0; ci < C::CPG; ++ci) {
        const int c = g + ci * C::G;
        ptx::mbar_wait(&d_full[c], par);
        umma::fence_after_sync();
        const uint32_t col0 = tbase + lane_base + static_cast<uint32_t>(c * C::NC);
        if (!valid) continue;
#pragma unroll
        for (int b8 = 0; b8 < C::UB / 8; ++b8) {
          float zi[8], zf[8], zg[8], zo[8], hv[8];
          umma::tmem_ld8x4(col0 + 0 * C::UB + b8 * 8, col0 + 1 * C::UB + b8 * 8, col0 + 2 * C::UB + b8 * 8, col0 + 3 * C::UB + b8 * 8,
                           zi, zf, zg, zo);
          const int j0 = c * C::UB + b8 * 8;       // first unit of this block
          const int l0 = ci * C::UB + b8 * 8;      // its slot in cst / pmax
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
            up2(lstm_cell2<SAVE>(pi, pf, pg, po, cst[(l0 + i) / 2], sv), hv[i], hv[i + 1]);
            if (SAVE && valid) {
              save_cells2<COMPACT>(save + (((tile * T + t) * nsave) * H + j0 + i) * kTile + b, nsave, H, sv, hv[i], hv[i + 1]);
            }
          }
          if (POOL) {
            const int ph = t % 3;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (SAVE) pool_arg_update(parg, l0 + i, ph, hv[i] > pmax[l0 + i]);
              pmax[l0 + i] = (ph == 0) ? hv[i] : fmaxf(pmax[l0 + i], hv[i]);
            }
            if (ph == 2 && valid) {
              float* yo = yout + ((tile * t_out_stride + t / 3) * H + j0) * kTile + b;
#pragma unroll
              for (int i = 0; i < 8; ++i) yo[i * kTile] = pmax[l0 + i];
              if (SAVE) {
                float* ao = amax + ((tile * t_out_stride + t / 3) * H + j0) * kTile + b;
#pragma unroll
                for (int i = 0; i < 8; ++i) ao[i * kTile] = static_cast<float>((parg >> (2 * (l0 + i))) & 3ull);
              }
            }
          } else if (LAST) {
            if (last_step && valid) {
              float* yo = yout + (tile * H + j0) * kTile + b;
#pragma unroll
              for (int i = 0; i < 8; ++i) yo[i * kTile] = hv[i];
            }
          } else if (valid) {
            float* yo = yout + ((tile * t_out_stride + t) * H + j0) * kTile + b;
#pragma unroll
            for (int i = 0; i < 8; ++i) yo[i * kTile] = hv[i];
          }
          if (last_step) {   // the next step starts a new sequence from the zero state
#pragma unroll
            for (int i = 0; i < 8; ++i) hv[i] = 0.f;
          }
          split8(hv, hold_hi[ci][b8], hold_lo[ci][b8]);
        }
      }
      // every MMA of this step has to be done with the A operand before h_t and x_{t+1} go in
      ptx::mbar_wait(&d_full[C::NCH - 1], par);
#pragma unroll
      for (int ci = 0; ci < C::CPG; ++ci) {
        const int c = g + ci * C::G;
#pragma unroll
        for (int b8 = 0; b8 < C::UB / 8; ++b8) {
          const int kb = C::IP / 8 + c * (C::UB / 8) + b8;
          Ahi[kb * C::ROWS + r] = hold_hi[ci][b8];
          Alo[kb * C::ROWS + r] = hold_lo[ci][b8];
        }
      }
      if (q + 1 < total) {
        convert_x(q + 1);
        publish_a();
      }
      if (++t == T) { t = 0; ++n; }
    }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) umma::tmem_dealloc(tbase, C::TMEM_COLS);
}

}  // namespace gnnqc
