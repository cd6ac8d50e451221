// Input-gradient backward pass of one Keras LSTM layer (BPTT without weight gradients) with the
// [windows x 4H] x [4H x (I+H)] contraction on the tensor cores.  Same contract as lstm_bwd.cuh
// (reference xai/libs/integrated_gradients.py:967-987; tape.gradient through the Keras LSTM):
// per step t, descending, from the forward's saved i, f, g, o, c_t
//   dh   = dy_t + dh_rec          (dy routed through MaxPooling1D's argmax for pooled layers)
//   dc  += dh * o * (1 - tanh(c_t)^2)
//   dz_i = dc*g*i(1-i)   dz_f = dc*c_{t-1}*f(1-f)   dz_g = dc*i*(1-g^2)   dz_o = dh*tanh(c_t)*o(1-o)
//   dc   = dc * f
//   [dx_t | dh_rec] = dz . [W | U]^T
//
// Mapping.  A group of 128 threads owns 128 windows (two tiles) = the 128 TMEM lanes, one window
// per thread, dc in registers.  The contraction runs over K = 4H in chunks of 8 units x 4 gates:
// the threads compute a chunk's dz (fp32), split it in bf16 hi/lo and store it as the A operand of
// a two-buffer ring; one MMA thread per group multiplies each chunk with the matching rows of
// [W | U]^T (B operand, bf16 hi/lo, hi*hi + hi*lo + lo*hi) and accumulates D[128 x (I+H)] in TMEM,
// so the cell arithmetic of chunk c+1 overlaps the MMAs of chunk c.  D is double-buffered across
// steps: while step t accumulates into D[t&1], the threads read dh_rec (their own lane, 8 columns per
// chunk) and flush dx_{t+1} from D[(t+1)&1].  For H = 128 the 384 KB of weights are streamed from
// L2 through a ring by a producer thread (TMA).
// SPLIT = 2 deals the chunks of a window over two threads (two warps share each 32 lanes, a thread carries dc for half
// the units): the step's serial chain gets shorter (~4 us instead of ~6 us for a 16-unit layer) at half the windows in
// flight per SM.  lstm_tc_bwd.cu picks the shape per launch: always split for H >= 64, split for H <= 32 when all tile
// pairs fit in one wave of it.
#pragma once
#include <cuda_bf16.h>

#include <cstdint>

#include "lstm_tc.cuh"

namespace gnnqc {

template <int I, int H, int GRP, int SPLIT, bool STREAM>
struct TcBwdCfg {
  static constexpr int IP = (I + 15) / 16 * 16;           // dx columns padded so the MMA N is a multiple of 16
  static constexpr int N = IP + H;                        // output columns: dx | pad | dh_rec
  static constexpr int UB = 8;                            // units per chunk
  static constexpr int NCHUNK = H / UB;
  static constexpr int CPS = NCHUNK / SPLIT;              // chunks per thread and step
  static constexpr int NBUF = 2;                          // A ring depth
  static constexpr int A_BUF = 2 * 4 * 128 * 16;          // hi | lo, each [4 kb][128 rows][8] bf16
  static constexpr int PIECE = 128 * N;                   // B rows of one chunk: [slice][hi|lo][2 kb][N][8]
  static constexpr int RING = STREAM ? 4 : NCHUNK;
  static constexpr size_t B_BYTES = size_t(RING) * PIECE;
  static constexpr size_t OFF_B = 0, OFF_A = B_BYTES;
  static constexpr size_t OFF_BAR = OFF_A + size_t(GRP) * NBUF * A_BUF;
  static constexpr int BARS_PER_GROUP = 2 * NBUF + 2;     // a_ready[NBUF], a_free[NBUF], d_full[2]
  static constexpr int N_BARS = GRP * BARS_PER_GROUP + 2 * RING;   // + b_full[RING], b_empty[RING]
  static constexpr size_t SMEM_BYTES = OFF_BAR + size_t(N_BARS) * 8 + 16;
  static constexpr int CT = GRP * SPLIT * 128;            // compute threads
  static constexpr int NT = CT + 32 * GRP + (STREAM ? 32 : 0);
  static constexpr int TMEM_GROUP = 2 * N;                // D[2] per group
  static constexpr size_t PACKED_BYTES = size_t(NCHUNK) * PIECE;
  static_assert(H % (UB * SPLIT) == 0 && GRP * TMEM_GROUP <= 512 && N <= 256 && (!STREAM || GRP == 1), "shape not covered");
  // a thread's successive chunks must use successive phases of its ring buffer's barriers (parity waits cannot tell
  // phases two apart): one split walks both buffers in turn, two splits own one buffer each - four splits would alias
  static_assert(SPLIT == 1 || SPLIT == NBUF, "ring depth must match the unit split");
  static_assert(SMEM_BYTES <= 227 * 1024, "operands do not fit in shared memory");
};

// TRAIN: the forward saved six quantities per step (h_t too) and dz is also written for the weight-gradient kernel.
// COMPACT: the forward saved the gates as 16-bit fixed point + c (three planes per step, save_cells2 in lstm_tc.cuh).
template <int I, int H, bool POOL, bool LAST, int GRP, int SPLIT, bool STREAM, bool TRAIN = false, bool COMPACT = false>
__global__ void __launch_bounds__((TcBwdCfg<I, H, GRP, SPLIT, STREAM>::NT), 1)
lstm_tc_bwd_kernel(const float* __restrict__ dy,            // grad wrt this layer's output (layouts of lstm_bwd.cuh)
                   const float* __restrict__ save,          // [n_tiles][T][5 | 6 | 3][H][64]
                   const float* __restrict__ amax,          // POOL: [n_tiles][T/3][H][64]
                   const unsigned char* __restrict__ wpk,   // B pieces per chunk (TcBwdCfg::PACKED_BYTES)
                   float* __restrict__ dx,                  // [n_tiles][T][I][64]
                   int T, int n_tiles,
                   float* __restrict__ dz_out = nullptr) {  // TRAIN: [n_tiles][T][4H][64] gate pre-activation gradients
  using C = TcBwdCfg<I, H, GRP, SPLIT, STREAM>;
  constexpr int nsave = TRAIN ? 6 : (COMPACT ? kSaveCompact : 5);
  static_assert(!(TRAIN && COMPACT), "training keeps fp32 gates");
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int kMmaWarp0 = C::CT / 32, kLoadWarp = C::CT / 32 + GRP;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR);
  uint64_t* b_full = bars + GRP * C::BARS_PER_GROUP;
  uint64_t* b_empty = b_full + C::RING;

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == kMmaWarp0) umma::tmem_alloc(&tmem_slot, 512);
  if (tid == 0) {
    for (int g = 0; g < GRP; ++g) {
      uint64_t* gb = bars + g * C::BARS_PER_GROUP;
      for (int s = 0; s < C::NBUF; ++s) {
        ptx::mbar_init(&gb[s], 128);             // a_ready
        ptx::mbar_init(&gb[C::NBUF + s], 1);     // a_free
      }
      ptx::mbar_init(&gb[2 * C::NBUF + 0], 1);   // d_full
      ptx::mbar_init(&gb[2 * C::NBUF + 1], 1);
    }
    for (int s = 0; s < C::RING; ++s) {
      ptx::mbar_init(&b_full[s], 1);
      ptx::mbar_init(&b_empty[s], 1);
    }
    ptx::mbar_init_fence();
  }
  if (!STREAM) {
    const uint4* src = reinterpret_cast<const uint4*>(wpk);
    uint4* dst = reinterpret_cast<uint4*>(sm + C::OFF_B);
    for (int i = tid; i < static_cast<int>(C::B_BYTES / 16); i += C::NT) dst[i] = __ldg(src + i);
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();

  const int n_pairs = (n_tiles + 1) / 2;
  const int n_slots = static_cast<int>(gridDim.x) * GRP;
  auto slot_steps = [&](int g) -> long long {
    const int slot = g * static_cast<int>(gridDim.x) + static_cast<int>(blockIdx.x);   // group-major: every SM gets 3-4 of 512 pairs
    const int mine = slot < n_pairs ? (n_pairs - slot + n_slots - 1) / n_slots : 0;
    return static_cast<long long>(mine) * T;
  };

  if (warp >= kMmaWarp0 && warp < kMmaWarp0 + GRP) {
    // ---- MMA issue: one thread per group -----------------------------------------------------------
    if (lane == 0) {
      const int g = warp - kMmaWarp0;
      uint64_t* a_ready = bars + g * C::BARS_PER_GROUP;
      uint64_t* a_free = a_ready + C::NBUF;
      uint64_t* d_full = a_ready + 2 * C::NBUF;
      const uint32_t tbase = tmem_slot + static_cast<uint32_t>(g * C::TMEM_GROUP);
      const unsigned char* abase = sm + C::OFF_A + size_t(g) * C::NBUF * C::A_BUF;
      constexpr uint32_t idesc = umma::instr_desc_bf16(128, C::N);
      const long long total = slot_steps(g);
      long long q = 0;   // chunks consumed
      for (long long si = 0; si < total; ++si) {
        const uint32_t dcol = tbase + static_cast<uint32_t>((si & 1) * C::N);
#pragma unroll 1
        for (int c = 0; c < C::NCHUNK; ++c, ++q) {
          const int buf = static_cast<int>(q % C::NBUF);
          ptx::mbar_wait(&a_ready[buf], static_cast<uint32_t>((q / C::NBUF) & 1));
          int slot = c;
          if (STREAM) {
            slot = static_cast<int>(q % C::RING);
            ptx::mbar_wait(&b_full[slot], static_cast<uint32_t>((q / C::RING) & 1));
          }
          umma::fence_after_sync();
          const unsigned char* ap = abase + size_t(buf) * C::A_BUF;
          const unsigned char* bp = sm + C::OFF_B + size_t(slot) * C::PIECE;
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            const uint64_t dah = umma::smem_desc(ap + s * 2 * 128 * 16, 128 * 16, 128);
            const uint64_t dal = umma::smem_desc(ap + C::A_BUF / 2 + s * 2 * 128 * 16, 128 * 16, 128);
            const uint64_t dbh = umma::smem_desc(bp + (s * 2 + 0) * 2 * C::N * 16, C::N * 16, 128);
            const uint64_t dbl = umma::smem_desc(bp + (s * 2 + 1) * 2 * C::N * 16, C::N * 16, 128);
            umma::mma_bf16(dcol, dah, dbh, idesc, c > 0 || s > 0);
            umma::mma_bf16(dcol, dah, dbl, idesc, true);
            umma::mma_bf16(dcol, dal, dbh, idesc, true);
          }
          umma::commit(&a_free[buf]);
          if (STREAM) umma::commit(&b_empty[slot]);
        }
        umma::commit(&d_full[si & 1]);
      }
    }
  } else if (STREAM && warp == kLoadWarp) {
    // ---- weight pieces through the ring (H = 128) ---------------------------------------------------
    if (lane == 0) {
      const long long n_loads = slot_steps(0) * C::NCHUNK;
      int piece = 0;
      for (long long pc = 0; pc < n_loads; ++pc) {
        const int slot = static_cast<int>(pc % C::RING);
        const long long round = pc / C::RING;
        if (round > 0) ptx::mbar_wait(&b_empty[slot], static_cast<uint32_t>((round - 1) & 1));
        ptx::mbar_arrive_expect_tx(&b_full[slot], C::PIECE);
        ptx::bulk_load(sm + C::OFF_B + size_t(slot) * C::PIECE, wpk + size_t(piece) * C::PIECE, C::PIECE, &b_full[slot]);
        if (++piece == C::NCHUNK) piece = 0;
      }
    }
  } else if (tid < C::CT) {
    // ---- cell gradients: one window (TMEM lane) per thread ------------------------------------------
    const int g = tid / (SPLIT * 128), sp = (tid / 128) % SPLIT, r = tid & 127, half = r >> 6, b = r & 63;
    uint64_t* a_ready = bars + g * C::BARS_PER_GROUP;
    uint64_t* a_free = a_ready + C::NBUF;
    uint64_t* d_full = a_ready + 2 * C::NBUF;
    const uint32_t tlane = tmem_slot + static_cast<uint32_t>(g * C::TMEM_GROUP) + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    unsigned char* abase = sm + C::OFF_A + size_t(g) * C::NBUF * C::A_BUF;
    const int slot = g * static_cast<int>(gridDim.x) + static_cast<int>(blockIdx.x);
    const long long total = slot_steps(g);
    const int t_dy = POOL ? T / 3 : T;
    constexpr int NBLK = (I + 7) / 8;   // 8-column blocks of dx
    float dcv[C::CPS * C::UB];

    // dx of step (tile_f, t_f) sits in D[s]: this thread stores its window's columns
    auto flush_dx = [&](int s, long long tile_f, int t_f, bool valid_f) {
#pragma unroll
      for (int blk = 0; blk < NBLK; ++blk) {
        if (blk % SPLIT != sp) continue;
        float v[8];
        umma::tmem_ld8(tlane + s * C::N + blk * 8, v);
        if (valid_f) {
          float* o = dx + ((tile_f * T + t_f) * I + blk * 8) * kTile + b;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (blk * 8 + i < I) o[i * kTile] = v[i];
        }
      }
    };

    long long prev_tile = 0;
    int prev_t = 0;
    bool prev_valid = false;
    for (long long si = 0; si < total; ++si) {
      const int n = static_cast<int>(si / T), t = T - 1 - static_cast<int>(si % T);
      const long long tile = 2 * (static_cast<long long>(slot) + static_cast<long long>(n) * n_slots) + half;
      const bool valid = tile < n_tiles;          // uniform over the two warps of a half
      const bool new_seq = (t == T - 1);
      if (new_seq) {
#pragma unroll
        for (int j = 0; j < C::CPS * C::UB; ++j) dcv[j] = 0.f;
      }
      const int sprev = static_cast<int>((si + 1) & 1);   // D buffer of step si - 1
      if (si > 0) {
        ptx::mbar_wait(&d_full[sprev], static_cast<uint32_t>(((si - 1) >> 1) & 1));
        umma::fence_after_sync();
        flush_dx(sprev, prev_tile, prev_t, prev_valid);
      }
      const float ph = static_cast<float>(t % 3);
#pragma unroll
      for (int ci = 0; ci < C::CPS; ++ci) {   // unrolled: dcv stays in registers
        const int c = sp + ci * SPLIT, j0 = c * C::UB;
        float dh[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) dh[i] = 0.f;
        if (!new_seq) umma::tmem_ld8(tlane + sprev * C::N + C::IP + j0, dh);
        float zi[8], zf[8], zg[8], zo[8];
        if (valid) {
          const float* sv = save + (((tile * T + t) * nsave) * H + j0) * kTile + b;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float ig, fg, gg, og, ct, cp;
            if (COMPACT) {
              const uint32_t* su = reinterpret_cast<const uint32_t*>(sv);
              unpack_gates(__ldg(su + (0 * H + i) * kTile), __ldg(su + (1 * H + i) * kTile), ig, fg, gg, og);
              ct = __ldg(sv + (2 * H + i) * kTile);
              cp = (t > 0) ? __ldg(sv + (i - H) * kTile) : 0.f;   // c of step t-1: plane 2 of the previous step's block
            } else {
              ig = __ldg(sv + (0 * H + i) * kTile); fg = __ldg(sv + (1 * H + i) * kTile); gg = __ldg(sv + (2 * H + i) * kTile);
              og = __ldg(sv + (3 * H + i) * kTile); ct = __ldg(sv + (4 * H + i) * kTile);
              cp = (t > 0) ? __ldg(sv + (i - (nsave - 4) * H) * kTile) : 0.f;           // c of step t-1: quantity 4 of the previous step's block
            }
            float d = dh[i];
            if (POOL) {
              const long long o = ((tile * t_dy + t / 3) * H + j0 + i) * kTile + b;
              d += (__ldg(amax + o) == ph) ? __ldg(dy + o) : 0.f;
            } else if (LAST) {
              if (new_seq) d += __ldg(dy + (tile * H + j0 + i) * kTile + b);
            } else {
              d += __ldg(dy + ((tile * t_dy + t) * H + j0 + i) * kTile + b);
            }
            const float e = ex2_approx(fminf(ct * tc_gate_scale(2), 36.f));
            const float th = fmaf(-e, rcp_approx(1.f + e), rcp_approx(1.f + e));   // tanh(c_t)
            const float dcc = fmaf(d * og, 1.f - th * th, dcv[ci * C::UB + i]);
            zi[i] = dcc * gg * ig * (1.f - ig);
            zf[i] = dcc * cp * fg * (1.f - fg);
            zg[i] = dcc * ig * (1.f - gg * gg);
            zo[i] = d * th * og * (1.f - og);
            dcv[ci * C::UB + i] = dcc * fg;
          }
          if (TRAIN) {
            float* dzo = dz_out + ((tile * T + t) * 4 * H + j0) * kTile + b;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              dzo[(0 * H + i) * kTile] = zi[i];
              dzo[(1 * H + i) * kTile] = zf[i];
              dzo[(2 * H + i) * kTile] = zg[i];
              dzo[(3 * H + i) * kTile] = zo[i];
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) zi[i] = zf[i] = zg[i] = zo[i] = 0.f;
        }
        const long long q = si * C::NCHUNK + c;
        const int buf = static_cast<int>(q % C::NBUF);
        const long long use = q / C::NBUF;
        if (use > 0) ptx::mbar_wait(&a_free[buf], static_cast<uint32_t>((use - 1) & 1));
        uint4* Ahi = reinterpret_cast<uint4*>(abase + size_t(buf) * C::A_BUF);
        uint4* Alo = reinterpret_cast<uint4*>(abase + size_t(buf) * C::A_BUF + C::A_BUF / 2);
        uint4 hi, lo;
        split8_bf16(zi, hi, lo); Ahi[0 * 128 + r] = hi; Alo[0 * 128 + r] = lo;
        split8_bf16(zf, hi, lo); Ahi[1 * 128 + r] = hi; Alo[1 * 128 + r] = lo;
        split8_bf16(zg, hi, lo); Ahi[2 * 128 + r] = hi; Alo[2 * 128 + r] = lo;
        split8_bf16(zo, hi, lo); Ahi[3 * 128 + r] = hi; Alo[3 * 128 + r] = lo;
        umma::fence_async_smem();
        umma::fence_before_sync();
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(&a_ready[buf])) : "memory");
      }
      prev_tile = tile;
      prev_t = t;
      prev_valid = valid;
    }
    if (total > 0) {
      const int s = static_cast<int>((total - 1) & 1);
      ptx::mbar_wait(&d_full[s], static_cast<uint32_t>(((total - 1) >> 1) & 1));
      umma::fence_after_sync();
      flush_dx(s, prev_tile, prev_t, prev_valid);
    }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp0) umma::tmem_dealloc(tmem_slot, 512);
}

}  // namespace gnnqc
