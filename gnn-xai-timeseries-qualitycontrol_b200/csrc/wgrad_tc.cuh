// Weight gradients of one LSTM layer on the tensor cores (training step, reference libs/fit_model.py:61-112: what
// tape.gradient leaves for kernel, recurrent_kernel and bias of a Keras LSTM):
//   dW[a][c] = sum_{window, t} x_t[a] dz_t[c]      dU[a][c] = sum_{window, t >= 1} h_{t-1}[a] dz_t[c]      db[c] = sum dz_t[c]
// This is the one large-K contraction of the repository (K = windows x steps); train_kernels.cuh's wgrad_kernel ran
// it on the CUDA cores (packed FFMA2) in two launches per layer.  Here, per layer, ONE launch forms
//   D[c][n] = sum_items sum_{b < 64} dz_item[c][b] * R_item[n][b],      R = [x_t ; h_{t-1} ; 1]   (n < I + H + 1)
// with item = (tile of 64 windows, step t): dz is the M operand (a chunk of 128 gate columns per CTA), R the N
// operand, the 64 windows of the tile the K extent of the item (4 MMAs of K = 16).  Both blocks are K-major in HBM
// already ([row][64 windows] fp32); eight staging warps split them into bf16 hi/lo (hi*hi + hi*lo + lo*hi, fp32
// accumulation in TMEM - bf16 because dz goes down to 1e-10) and write the core-matrix layout of umma.cuh, one warp
// per 8-window K group so that both the 32-byte global reads (whole sectors) and the 16-byte shared stores
// (consecutive rows) are conflict-free; a ninth warp issues the MMAs, two to four stages deep.
// Deterministic: a CTA owns a fixed slice of the items and leaves D in partial[slice][n][c]; wgrad_tc_reduce_kernel
// adds the slices in a fixed order.
#pragma once
#include <algorithm>

#include "lstm_tc.cuh"

namespace gnnqc {

struct WgradTcArgs {
  const float* x;           // layer input tiles: x + tile * x_tile + t * x_step = [I][64]
  const float* h;           // saved h_t:         h + tile * h_tile + t * h_step = [H][64]
  const float* dz;          // dz + tile * dz_tile + t * dz_step = [4H][64]
  long long x_tile, x_step, h_tile, h_step, dz_tile, dz_step;
  int I, H, T;
  long long n_items;        // tiles * T
  int n16;                  // I + H + 1 rounded up to a multiple of 16 (the MMA N)
  int n_stages;
  float* partial;           // [slices][I + H + 1][4H]
};

constexpr int kWgTcStageThreads = 256, kWgTcThreads = kWgTcStageThreads + 32;
__host__ __device__ inline size_t wgrad_tc_stage_bytes(int n16) { return size_t(2) * 128 * 128 + size_t(2) * n16 * 128; }
__host__ __device__ inline size_t wgrad_tc_smem_bytes(int n16, int n_stages) { return n_stages * wgrad_tc_stage_bytes(n16) + 128; }
inline int wgrad_tc_stages(int n16) {
  const size_t room = size_t(200) * 1024;
  return static_cast<int>(std::max<size_t>(2, std::min<size_t>(4, room / wgrad_tc_stage_bytes(n16))));
}

// grid (slices, ceil(4H / 128)), kWgTcThreads threads
static __global__ void __launch_bounds__(kWgTcThreads, 1) wgrad_tc_kernel(const WgradTcArgs g) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int NS = g.n_stages, n16 = g.n16, NC = 4 * g.H, rows_out = g.I + g.H + 1;
  const size_t stage_bytes = wgrad_tc_stage_bytes(n16);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + size_t(NS) * stage_bytes);   // full[NS] | free[NS] | done
  uint64_t* full = bars;
  uint64_t* freeb = bars + NS;
  uint64_t* done = bars + 2 * NS;
  const int c0 = static_cast<int>(blockIdx.y) * 128;
  const int n_slices = static_cast<int>(gridDim.x), slice = static_cast<int>(blockIdx.x);
  const long long per = (g.n_items + n_slices - 1) / n_slices;
  const long long i_begin = slice * per, i_end = (i_begin + per < g.n_items) ? i_begin + per : g.n_items;
  const long long mine = i_end > i_begin ? i_end - i_begin : 0;

  // ---- set-up: zero every stage (rows that are never staged must read as 0), the ones row, barriers, TMEM ----
  for (size_t i = tid; i < size_t(NS) * stage_bytes / 16; i += kWgTcThreads) reinterpret_cast<uint4*>(sm)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  for (int i = tid; i < NS * 8; i += kWgTcThreads) {   // R row I + H = 1.0 (bf16 0x3F80) in the hi half of every stage
    unsigned char* bhi = sm + size_t(i / 8) * stage_bytes + 2 * 128 * 128;
    *reinterpret_cast<uint4*>(bhi + (size_t(i % 8) * n16 + (g.I + g.H)) * 16) = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
  }
  if (warp == 8) umma::tmem_alloc(&tmem_slot, 256);
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) {
      ptx::mbar_init(&full[s], kWgTcStageThreads);
      ptx::mbar_init(&freeb[s], 1);
    }
    ptx::mbar_init(done, 1);
    ptx::mbar_init_fence();
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = tmem_slot;

  if (warp == 8) {
    // ---- MMA issue: warp-uniform control flow, one elected lane ------------------------------------------------
    const uint32_t idesc = umma::instr_desc_bf16(128, n16);
    const uint32_t sm0 = ptx::smem_addr(sm);
    for (long long it = 0; it < mine; ++it) {
      const int s = static_cast<int>(it % NS);
      ptx::mbar_wait(&full[s], static_cast<uint32_t>((it / NS) & 1));
      umma::fence_after_sync();
      const uint32_t a_hi = sm0 + static_cast<uint32_t>(s * stage_bytes), a_lo = a_hi + 128 * 128;
      const uint32_t b_hi = a_lo + 128 * 128, b_lo = b_hi + static_cast<uint32_t>(n16) * 128;
      if (umma::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {   // K = 16 windows each: two 8-window core-matrix columns
          const uint64_t dah = umma::smem_desc_u(a_hi + ks * 2 * 128 * 16, 128 * 16, 128);
          const uint64_t dal = umma::smem_desc_u(a_lo + ks * 2 * 128 * 16, 128 * 16, 128);
          const uint64_t dbh = umma::smem_desc_u(b_hi + ks * 2 * n16 * 16, n16 * 16, 128);
          const uint64_t dbl = umma::smem_desc_u(b_lo + ks * 2 * n16 * 16, n16 * 16, 128);
          umma::mma_bf16(tbase, dah, dbh, idesc, it > 0 || ks > 0);
          umma::mma_bf16(tbase, dah, dbl, idesc, true);
          umma::mma_bf16(tbase, dal, dbh, idesc, true);
        }
        umma::commit(&freeb[s]);
        if (it + 1 == mine) umma::commit(done);
      }
      __syncwarp();
    }
  } else {
    // ---- staging: warp = 8-window K group, lanes = rows ---------------------------------------------------------
    // The global loads of item i + 1 are issued before item i is converted and stored (two register sets), so a
    // staging thread never sits on an HBM round trip.
    const int kb = warp;
    constexpr int kRowBlocksB = 6;               // (I + H) / 32 rounded up: at most 192 rows of R come from memory
    const int nb_b = (g.I + g.H + 31) / 32;
    auto load = [&](long long it, float4 (&a)[4][2], float4 (&b)[kRowBlocksB][2]) {
      const long long item = i_begin + it, tile = item / g.T;
      const int t = static_cast<int>(item % g.T);
      const float* dz = g.dz + tile * g.dz_tile + t * g.dz_step + kb * 8;
      const float* xt = g.x + tile * g.x_tile + t * g.x_step + kb * 8;
      const float* hp = g.h + tile * g.h_tile + (t - 1) * g.h_step + kb * 8;
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = c0 + lane + 32 * j;
        const float4* src = reinterpret_cast<const float4*>(dz + static_cast<long long>(r) * kTile);
        a[j][0] = r < NC ? __ldg(src) : zero;
        a[j][1] = r < NC ? __ldg(src + 1) : zero;
      }
#pragma unroll
      for (int j = 0; j < kRowBlocksB; ++j) {
        if (j >= nb_b) break;
        const int n = lane + 32 * j;
        const float* row = n < g.I ? xt + static_cast<long long>(n) * kTile : hp + static_cast<long long>(n - g.I) * kTile;
        const bool live = n < g.I || (n < g.I + g.H && t > 0);     // no recurrent input at the first step: zeros
        b[j][0] = live ? __ldg(reinterpret_cast<const float4*>(row)) : zero;
        b[j][1] = live ? __ldg(reinterpret_cast<const float4*>(row) + 1) : zero;
      }
    };
    auto put = [&](unsigned char* hi_base, int rows_total, int row, const float4& v0, const float4& v1) {
      const float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
      uint4 hi, lo;
      split8_bf16(v, hi, lo);
      const size_t at = (size_t(kb) * rows_total + row) * 16;
      *reinterpret_cast<uint4*>(hi_base + at) = hi;
      *reinterpret_cast<uint4*>(hi_base + size_t(rows_total) * 128 + at) = lo;
    };
    auto store = [&](long long it, const float4 (&a)[4][2], const float4 (&b)[kRowBlocksB][2]) {
      const int s = static_cast<int>(it % NS);
      if (it >= NS) ptx::mbar_wait(&freeb[s], static_cast<uint32_t>((it / NS - 1) & 1));
      unsigned char* a_hi = sm + size_t(s) * stage_bytes;
      unsigned char* b_hi = a_hi + 2 * 128 * 128;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (c0 + 32 * j < NC) put(a_hi, 128, lane + 32 * j, a[j][0], a[j][1]);   // rows past 4H stay zero from the set-up
#pragma unroll
      for (int j = 0; j < kRowBlocksB; ++j) {
        if (j >= nb_b) break;
        if (lane + 32 * j < g.I + g.H) put(b_hi, n16, lane + 32 * j, b[j][0], b[j][1]);
      }
      umma::fence_async_smem();
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(&full[s])) : "memory");
    };
    float4 a0[4][2], b0[kRowBlocksB][2], a1[4][2], b1[kRowBlocksB][2];
    if (mine > 0) load(0, a0, b0);
    for (long long it = 0; it < mine; it += 2) {
      if (it + 1 < mine) load(it + 1, a1, b1);
      store(it, a0, b0);
      if (it + 1 < mine) {
        if (it + 2 < mine) load(it + 2, a0, b0);
        store(it + 1, a1, b1);
      }
    }
    // ---- epilogue: warps 0-3 own the 128 TMEM lanes = gate columns c0 .. c0 + 127 --------------------------------
    if (warp < 4) {
      const int c = c0 + warp * 32 + lane;
      float* out = g.partial + static_cast<long long>(slice) * rows_out * NC + c;
      if (mine > 0) {
        ptx::mbar_wait(done, 0);
        umma::fence_after_sync();
        const uint32_t tl = tbase + (static_cast<uint32_t>(warp * 32) << 16);
        for (int n0 = 0; n0 < rows_out; n0 += 8) {
          float v[8];
          umma::tmem_ld8(tl + n0, v);
          if (c < NC) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (n0 + i < rows_out) out[static_cast<long long>(n0 + i) * NC] = v[i];
          }
        }
      } else if (c < NC) {
        for (int n = 0; n < rows_out; ++n) out[static_cast<long long>(n) * NC] = 0.f;
      }
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 8) umma::tmem_dealloc(tbase, 256);
}

// dW, dU, db = sum over the slices, in a fixed order: one warp per element, lane l adds slices l, l + 32, ... and the
// 32 partial sums go through a fixed shuffle tree (as wgrad_reduce_kernel).
static __global__ void __launch_bounds__(256) wgrad_tc_reduce_kernel(const float* __restrict__ partial, int slices, int I, int H,
                                                                     float* __restrict__ out_w, float* __restrict__ out_u, float* __restrict__ out_b) {
  const int NC = 4 * H, rows = I + H + 1;
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= rows * NC) return;
  float s = 0.f;
  for (int k = lane; k < slices; k += 32) s += partial[static_cast<long long>(k) * rows * NC + i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if (lane != 0) return;
  const int n = i / NC, c = i % NC;
  if (n < I) out_w[n * NC + c] = s;
  else if (n < I + H) out_u[(n - I) * NC + c] = s;
  else out_b[c] = s;
}

}  // namespace gnnqc
