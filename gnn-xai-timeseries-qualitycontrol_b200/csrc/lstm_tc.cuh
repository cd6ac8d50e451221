// One Keras LSTM layer over sequence tiles with the gate contraction on the 5th-generation tensor
// cores (tcgen05.mma, accumulators in TMEM).  Same arithmetic contract as lstm_fwd.cuh (reference
// libs/create_model.py:62-79); this is the scoring-path kernel for the narrow layers (H = 16, 32);
// lstm_tc_big.cuh covers H = 64 and 128.
//
// Why tensor cores here: ncu on the fp32 kernel (profiles/lstm_full_r01b.csv) shows the fma pipe
// 60-74 % busy with math_pipe_throttle as the top stall - the [windows x (in+H)] x [(in+H) x 4H]
// contraction is the limiter, and it is dense.  Precision: one bf16 pass misses the 1e-4 score
// budget by 100x, so every operand is split x = hi + lo (both bf16) and the product is formed as
// hi*hi + hi*lo + lo*hi with fp32 accumulation in TMEM ("bf16x3", error ~2^-16 per product;
// measured 1.6e-5 absolute on scores against the float64 oracle).
//
// One persistent CTA per SM holds the weight operand once and runs FOUR independent groups of 128
// threads.  A group owns 128 windows (two 64-window tiles) = the 128 TMEM lanes of its own
// accumulator columns, its own A operand tile, x ring and barriers; the groups only share the SM,
// so the latency chain of one group's step (below) is covered by the other three.  With the cell
// update left on the CUDA cores the step is bound by the transcendental unit (7 MUFU per unit and
// step), which is why several groups per SM matter and more tensor throughput would not.
// Per step t, in a group:
//   all threads (one per window): x_t (TMA ring, fp32) -> bf16 hi/lo into the A operand tile;
//       h_{t-1} is already there -> fence.proxy.async -> arrive(a_ready)
//   elected thread: wait(a_ready) -> 3 x (K/16) tcgen05.mma into D[128 x 4H] -> tcgen05.commit(d_full)
//       -> TMA for a later step
//   all threads: wait(d_full) -> tcgen05.ld their row of D -> + bias -> cell update (c in
//       registers) -> h_t to global tiles and, split in hi/lo, into the A tile for step t+1.
// Tiles are dealt to CTAs in contiguous, equal ranges and to the 8 half-group slots of a CTA round
// by round, so 1024 tiles on 148 SMs are 7 (or 6) tiles per SM in ONE round; the two warps of an
// unused half skip the cell work.
// Operand tiles are K-major without swizzle ([K/8][rows][8] bf16, see umma.cuh); K is laid out as
// [input channels padded to 8 | hidden units | zero padding to 16].
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdint>

#include "lstm_fwd.cuh"
#include "umma.cuh"

namespace gnnqc {

// Bias through the tensor core: one extra K = 16 MMA per step opens each chain with accumulate = false,
//   A = the constant "ones" tile (k = 0, 1 hold 1.0), B = the bias tile (k = 0: bias_hi, k = 1: bias_lo),
// so the accumulators arrive with the bias already added (no shared-memory bias reads, no adds in the cell update).
constexpr int kOnesTileBytes = 2 * 128 * 16;            // [2 core-matrix columns][128 rows][8] fp16
__device__ __forceinline__ void fill_ones_tile(unsigned char* dst, int tid, int nthreads) {
  uint4* d = reinterpret_cast<uint4*>(dst);
  for (int i = tid; i < 256; i += nthreads) d[i] = (i < 128) ? make_uint4(0x3C003C00u, 0u, 0u, 0u) : make_uint4(0u, 0u, 0u, 0u);
}
__host__ __device__ constexpr size_t tc_bias_tile_bytes(int n) { return size_t(2) * n * 16; }   // [2][n][8] fp16

// SPLIT = threads per window: 1 (four groups of 128 windows per CTA, every thread all H units) or 2 (two groups, a thread
// takes every other block of 8 units: half the cell arithmetic on the per-step chain, half the windows in flight).
template <int I, int H, int SPLIT = 1>
struct TcCfg {
  static constexpr int IP = (I + 7) / 8 * 8;              // input channels padded to a core-matrix column
  static constexpr int KP = (IP + H + 15) / 16 * 16;      // contraction length per MMA chain
  static constexpr int N = 4 * H;                         // gate columns (Keras order i | f | c | o)
  static constexpr int ROWS = 128;                        // windows per group = TMEM lanes
  static constexpr int GROUPS = 4 / SPLIT;
  static constexpr int NT = GROUPS * SPLIT * ROWS;
  static constexpr int UPT = H / SPLIT;                   // units per thread
  static constexpr int NST = (I >= 32 && H >= 32) ? 1 : 2;   // x ring depth per group
  static constexpr int TMEM_COLS = GROUPS * N;            // 256 or 512
  static constexpr int XBYTES = I * kTile * 4;            // one 64-window half of one step
  static constexpr size_t B_BYTES = size_t(KP) * N * 2;   // one of hi / lo
  static constexpr size_t A_BYTES = size_t(KP) * ROWS * 2;
  static constexpr size_t BT_BYTES = tc_bias_tile_bytes(N);
  static constexpr size_t G_AHI = 0, G_ALO = A_BYTES, G_X = 2 * A_BYTES;          // within a group's block
  static constexpr size_t GROUP_BYTES = (2 * A_BYTES + size_t(NST) * 2 * XBYTES + 127) / 128 * 128;
  // 32 -> 32 units is the one shape whose operands leave no room for the two extra tiles: it keeps fp32 biases in
  // shared memory (rebuilt from the packed hi + lo rows) and adds them in the cell update.
  // (decided for the four-group shape and kept for the two-group one, so that both give the same bits)
  static constexpr bool BIAS_MMA = 2 * B_BYTES + BT_BYTES + kOnesTileBytes + 4 * GROUP_BYTES + 256 <= 227 * 1024;
  static constexpr size_t OFF_BHI = 0, OFF_BLO = B_BYTES, OFF_BIAS = 2 * B_BYTES;
  static constexpr size_t OFF_ONES = OFF_BIAS + (BIAS_MMA ? BT_BYTES : size_t(N) * 4);
  static constexpr size_t OFF_GROUPS = OFF_ONES + (BIAS_MMA ? kOnesTileBytes : 0);
  static constexpr size_t OFF_BAR = OFF_GROUPS + GROUPS * GROUP_BYTES;
  static constexpr int BARS_PER_GROUP = NST + 2;          // x_full[NST], a_ready, d_full
  static constexpr size_t SMEM_BYTES = OFF_BAR + size_t(GROUPS) * BARS_PER_GROUP * 8 + 16;
  // packed weights in global memory: B_hi | B_lo | bias tile (fp16 hi / lo rows, Keras column order)
  static constexpr size_t PACKED_BYTES = 2 * B_BYTES + BT_BYTES;
  static_assert(N <= 256 && (N & (N - 1)) == 0 && TMEM_COLS <= 512, "one MMA covers at most 256 gate columns");
  static_assert((SPLIT == 1 || SPLIT == 2) && H % (8 * SPLIT) == 0, "unit split not covered");
  static_assert(SMEM_BYTES <= 227 * 1024, "operands do not fit in shared memory");
};

// hi/lo bf16 split of eight floats into two 16-byte operand chunks (two values per cvt).  bf16 keeps fp32's
// exponent range, which the BACKWARD kernels need (their operand is dz, down to 1e-10); the forward kernels use
// the fp16 split below.
__device__ __forceinline__ void split8_bf16(const float (&v)[8], uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 hp = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);   // .x = low half-word
    h[i] = *reinterpret_cast<const uint32_t*>(&hp);
    const float r0 = v[2 * i] - __uint_as_float(h[i] << 16);
    const float r1 = v[2 * i + 1] - __uint_as_float(h[i] & 0xFFFF0000u);
    const __nv_bfloat162 lp = __floats2bfloat162_rn(r0, r1);
    l[i] = *reinterpret_cast<const uint32_t*>(&lp);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// ---- packed fp32x2 arithmetic (FADD2 / FMUL2 / FFMA2: one issue slot for two lanes) -----------------------------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up2(f32x2 p, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(p)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// hi/lo fp16 split (forward kernels): x = hi + lo with 11 + 11 mantissa bits, so hi*hi + hi*lo + lo*hi carries ~2^-22
// per product instead of bf16's ~2^-16 - measured on the scores: 1e-4 -> ~1e-6 against the float64 oracle.  fp16's
// range is enough here: hidden states are in [-1, 1], weights O(1), inputs are normalised sensor values; the
// conversion saturates (no inf can reach the tensor core), so a raw input beyond +-65504 is clipped.  Values below
// 2^-24 ~ 6e-8 in magnitude flush in the lo part: an ABSOLUTE error of 3e-8 on an O(1) operand.
__device__ __forceinline__ uint32_t f16x2_sat(float lo_elem, float hi_elem) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_elem), "f"(lo_elem));
  return r;
}
__device__ __forceinline__ void split8(const float (&v)[8], uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    h[i] = f16x2_sat(v[2 * i], v[2 * i + 1]);
    const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
    float r0, r1;
    up2(sub2(pk2(v[2 * i], v[2 * i + 1]), pk2(hf.x, hf.y)), r0, r1);
    l[i] = f16x2_sat(r0, r1);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// ---- cell update ------------------------------------------------------------------------------------------------
// The packed weights and biases of the tensor-core layers carry the factors of the exponentials (-log2 e for the
// i, f, o columns, -2 log2 e for the candidate column, tc_gate_scale), so the accumulator IS the ex2 argument:
//   a = e^-zi, f = e^-zf, g = e^-2zg, o = e^-zo,
//   c' = [c (1+a)(1+g) + (1-g)(1+f)] / [(1+f)(1+a)(1+g)],  h = (1-e^-2c') / [(1+o)(1+e^-2c')].
// Only large POSITIVE arguments can overflow the products, so one min per exponential is enough (2^36 ~ e^25:
// sigmoid error < 2e-11).  Two units are updated at a time in packed fp32x2 arithmetic.  The step is bound by the
// transcendental unit (MUFU: 16 lanes per clock and SM) with the FMA pipe mostly idle, so kPolyExps of the five
// exponentials per unit are evaluated on the FMA pipe instead: round-to-nearest range reduction with the 1.5*2^23
// constant, a degree-5 minimax polynomial for 2^f on [-0.5, 0.5] (2.3e-7 relative, the same 2 ulp as ex2.approx),
// exponent inserted with an integer add.
__host__ __device__ constexpr float tc_gate_scale(int gate) { return gate == 2 ? -2.8853900817779268f : -1.4426950408889634f; }
#ifndef GNNQC_POLY_EXPS
#define GNNQC_POLY_EXPS 1
#endif
constexpr int kPolyExps = GNNQC_POLY_EXPS;

__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ f32x2 ex2_mufu2(f32x2 z) {
  float z0, z1;
  up2(z, z0, z1);
  return pk2(ex2_approx(fminf(z0, 36.f)), ex2_approx(fminf(z1, 36.f)));
}
__device__ __forceinline__ f32x2 ex2_poly2(f32x2 z) {
  float z0, z1;
  up2(z, z0, z1);
  const f32x2 x = pk2(fmaxf(fminf(z0, 36.f), -126.f), fmaxf(fminf(z1, 36.f), -126.f));
  const f32x2 t = add2(x, pk2(12582912.f, 12582912.f));          // low mantissa bits of t = round(x)
  const f32x2 fr = sub2(x, add2(t, pk2(-12582912.f, -12582912.f)));
  f32x2 p = fma2(pk2(0.0013276472454890609f, 0.0013276472454890609f), fr, pk2(0.009675540961325169f, 0.009675540961325169f));
  p = fma2(p, fr, pk2(0.05550713092088699f, 0.05550713092088699f));
  p = fma2(p, fr, pk2(0.24022120237350464f, 0.24022120237350464f));
  p = fma2(p, fr, pk2(0.6931469440460205f, 0.6931469440460205f));
  p = fma2(p, fr, pk2(1.0000001192092896f, 1.0000001192092896f));
  float p0, p1, t0, t1;
  up2(p, p0, p1);
  up2(t, t0, t1);
  return pk2(__int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23)), __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23)));
}
template <int SLOT>   // SLOT = which of the five exponentials (0: forget, 1: output, 2: cell, 3: input, 4: candidate)
__device__ __forceinline__ f32x2 ex2_sel2(f32x2 z) { return SLOT < kPolyExps ? ex2_poly2(z) : ex2_mufu2(z); }

// Two units at once.  SAVE also returns the gates i, f, g, o and the new cell state for the backward pass (layout of
// lstm_fwd.cuh's SAVE mode); they come out of the two merged reciprocals by multiplication.
template <bool SAVE>
__device__ __forceinline__ f32x2 lstm_cell2(f32x2 zi, f32x2 zf, f32x2 zg, f32x2 zo, f32x2& c, f32x2 (&sv)[5]) {
  const f32x2 one = pk2(1.f, 1.f);
  const f32x2 a = ex2_sel2<3>(zi);
  const f32x2 f = ex2_sel2<0>(zf);
  const f32x2 g = ex2_sel2<4>(zg);
  const f32x2 o = ex2_sel2<1>(zo);
  const f32x2 g1 = add2(g, one), pf = add2(f, one);
  f32x2 pa, a1;
  if (SAVE) {
    a1 = add2(a, one);
    pa = mul2(a1, g1);
  } else {
    pa = fma2(a, g1, g1);                       // (1+a)(1+g)
  }
  const f32x2 tg = mul2(sub2(one, g), pf);      // (1-g)(1+f)
  float d0, d1;
  up2(mul2(pf, pa), d0, d1);
  const f32x2 r1 = pk2(rcp_approx(d0), rcp_approx(d1));
  const f32x2 cn = mul2(fma2(c, pa, tg), r1);
  c = cn;
  if (SAVE) {
    sv[0] = mul2(r1, mul2(pf, g1));             // 1/(1+a)
    sv[1] = mul2(r1, pa);                       // 1/(1+f)
    sv[2] = mul2(r1, mul2(tg, a1));             // (1-g)/(1+g)
    sv[4] = cn;
  }
  const f32x2 e = ex2_sel2<2>(mul2(cn, pk2(tc_gate_scale(2), tc_gate_scale(2))));
  const f32x2 e1 = add2(e, one);
  up2(fma2(o, e1, e1), d0, d1);                 // (1+o)(1+e)
  const f32x2 r2 = pk2(rcp_approx(d0), rcp_approx(d1));
  if (SAVE) sv[3] = mul2(r2, e1);               // 1/(1+o)
  return mul2(sub2(one, e), r2);
}

// What the gate-saving forward keeps of a unit pair (j, j + 1) for the backward pass; so = the pair's slot in plane 0 of
// its (tile, step) block, planes H * 64 floats apart.
//   nsave 5 / 6: fp32 planes i, f, g, o, c (and h_t: training, whose weight gradients need it).
//   nsave 3 (integrated gradients): 12 instead of 20 bytes per cell - the saved-gate stream is what bounds an
//     interpolation step (written once and read once per alpha, 83 % of HBM peak in the first layers' forward).  The gates
//     are bounded, so they are kept as 16-bit FIXED point: plane 0 = i | f << 16 (unorm16, absolute error 7.6e-6),
//     plane 1 = g (snorm16, 1.5e-5) | o << 16; c stays fp32 in plane 2.  (fp16 would have 2.4e-4 in [0.5, 1].)
constexpr int kSaveCompact = 3;
__device__ __forceinline__ uint32_t pack_gates_uu(float lo, float hi) {
  return __byte_perm(__float2uint_rn(lo * 65535.f), __float2uint_rn(hi * 65535.f), 0x5410);
}
__device__ __forceinline__ uint32_t pack_gates_su(float lo, float hi) {
  return __byte_perm(static_cast<uint32_t>(__float2int_rn(lo * 32767.f)), __float2uint_rn(hi * 65535.f), 0x5410);
}
__device__ __forceinline__ void unpack_gates(uint32_t w0, uint32_t w1, float& ig, float& fg, float& gg, float& og) {
  ig = __uint2float_rn(w0 & 0xffffu) * (1.f / 65535.f);
  fg = __uint2float_rn(w0 >> 16) * (1.f / 65535.f);
  gg = __int2float_rn(static_cast<int>(static_cast<short>(w1 & 0xffffu))) * (1.f / 32767.f);
  og = __uint2float_rn(w1 >> 16) * (1.f / 65535.f);
}
template <bool COMPACT>
__device__ __forceinline__ void save_cells2(float* so, int nsave, int H, const f32x2 (&sv)[5], float h0, float h1) {
  if (COMPACT) {
    float i0, i1, f0, f1, g0, g1, o0, o1;
    up2(sv[0], i0, i1); up2(sv[1], f0, f1); up2(sv[2], g0, g1); up2(sv[3], o0, o1);
    uint32_t* su = reinterpret_cast<uint32_t*>(so);
    su[0] = pack_gates_uu(i0, f0);
    su[kTile] = pack_gates_uu(i1, f1);
    su[H * kTile] = pack_gates_su(g0, o0);
    su[(H + 1) * kTile] = pack_gates_su(g1, o1);
    up2(sv[4], so[2 * H * kTile], so[(2 * H + 1) * kTile]);
  } else {
#pragma unroll
    for (int v = 0; v < 5; ++v) up2(sv[v], so[v * H * kTile], so[(v * H + 1) * kTile]);
    if (nsave > 5) {
      so[5 * H * kTile] = h0;
      so[(5 * H + 1) * kTile] = h1;
    }
  }
}

// Which of the three pooled steps held the maximum (first one on ties, as lstm_fwd.cuh), two bits per unit.
__device__ __forceinline__ void pool_arg_update(unsigned long long& bits, int unit, int ph, bool greater) {
  const unsigned long long m = 3ull << (2 * unit);
  if (ph == 0) bits &= ~m;
  else if (greater) bits = (bits & ~m) | (static_cast<unsigned long long>(ph) << (2 * unit));
}

template <int I, int H, bool POOL, bool SAVE = false, bool COMPACT = false, int SPLIT = 1>   // COMPACT: nsave == kSaveCompact
__global__ void __launch_bounds__((TcCfg<I, H, SPLIT>::NT), 1)
lstm_tc_kernel(const float* __restrict__ xin,            // [n_tiles][t_in_stride][I][64]
               float* __restrict__ yout,                 // [n_tiles][t_out_stride][H][64]
               const unsigned char* __restrict__ wpk,    // B_hi | B_lo | bias (TcCfg::PACKED_BYTES)
               int T, int t_in_stride, int t_out_stride, int n_tiles,
               float* __restrict__ save = nullptr,       // SAVE: [n_tiles][T][5][H][64] gates i, f, g, o and c
               float* __restrict__ amax = nullptr,       // SAVE && POOL: [n_tiles][t_out_stride][H][64] argmax step
               int nsave = 5) {                          // SAVE: 5, 6 (h_t as well: training) or kSaveCompact (save_cells2)
  using C = TcCfg<I, H, SPLIT>;
  extern __shared__ __align__(128) unsigned char sm[];
  float* bias = reinterpret_cast<float*>(sm + C::OFF_BIAS);
  __shared__ uint32_t tmem_slot;
  // The warp index goes through __shfl_sync so that the compiler knows everything derived from it (group, half,
  // tile, loop bounds, operand addresses) is warp-uniform: uniform branches, and the tcgen05.mma / cp.async.bulk
  // operands stay in uniform registers instead of being broadcast lane by lane around every instruction.
  const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int g = warp / (4 * SPLIT), sp = (warp >> 2) % SPLIT, quarter = warp & 3, half = quarter >> 1;
  const int r = quarter * 32 + lane, b = (quarter & 1) * 32 + lane;   // group, unit split, TMEM lane, tile of the pair, window in tile

  unsigned char* gsm = sm + C::OFF_GROUPS + size_t(g) * C::GROUP_BYTES;
  uint4* Ahi = reinterpret_cast<uint4*>(gsm + C::G_AHI);   // 16-byte chunk index: kb * ROWS + row
  uint4* Alo = reinterpret_cast<uint4*>(gsm + C::G_ALO);
  float* xs = reinterpret_cast<float*>(gsm + C::G_X);      // [NST][2][I][64]
  uint64_t* x_full = reinterpret_cast<uint64_t*>(sm + C::OFF_BAR) + g * C::BARS_PER_GROUP;
  uint64_t* a_ready = x_full + C::NST;
  uint64_t* d_full = a_ready + 1;

  // This CTA's contiguous range of tiles, dealt to the 8 half-group slots round by round:
  // slot s takes local tiles s*R + rho (rho = round), so every round is equally full.
  const int n_cta = static_cast<int>(gridDim.x), cta = static_cast<int>(blockIdx.x);
  const int base = n_tiles / n_cta, rem = n_tiles % n_cta;
  const int my_tiles = base + (cta < rem ? 1 : 0);
  const long long tile0 = static_cast<long long>(cta) * base + (cta < rem ? cta : rem);
  const int R = (my_tiles + 2 * C::GROUPS - 1) / (2 * C::GROUPS);
  // rounds in which this group's first half has a tile (the second half has one in a prefix of them)
  int rounds = my_tiles - 2 * g * R;
  rounds = rounds < 0 ? 0 : (rounds > R ? R : rounds);
  const long long total = static_cast<long long>(rounds) * T;

  // ---- set-up --------------------------------------------------------------------------------
  if (warp == 0) umma::tmem_alloc(&tmem_slot, C::TMEM_COLS);
  if (r == 0 && sp == 0) {
    for (int s = 0; s < C::NST; ++s) ptx::mbar_init(&x_full[s], 1);
    ptx::mbar_init(a_ready, C::ROWS * SPLIT);
    ptx::mbar_init(d_full, 1);
    ptx::mbar_init_fence();
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(wpk);
    uint4* dst = reinterpret_cast<uint4*>(sm);
    for (int i = tid; i < static_cast<int>(2 * C::B_BYTES / 16); i += C::NT) dst[i] = __ldg(src + i);
    if (C::BIAS_MMA) {
      for (int i = tid; i < static_cast<int>(C::BT_BYTES / 16); i += C::NT) dst[2 * C::B_BYTES / 16 + i] = __ldg(src + 2 * C::B_BYTES / 16 + i);
      fill_ones_tile(sm + C::OFF_ONES, tid, C::NT);
    } else {
      const __half* bt = reinterpret_cast<const __half*>(wpk + 2 * C::B_BYTES);   // row n of the bias tile: {hi, lo, 0 ...}
      for (int i = tid; i < C::N; i += C::NT) bias[i] = __half2float(bt[i * 8]) + __half2float(bt[i * 8 + 1]);
    }
    uint4* gz = reinterpret_cast<uint4*>(gsm);
    for (int i = r + sp * C::ROWS; i < static_cast<int>(C::GROUP_BYTES / 16); i += C::ROWS * SPLIT) gz[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  umma::fence_async_smem();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = tmem_slot + static_cast<uint32_t>(g * C::N);
  const bool issuer = quarter == 0 && sp == 0;
  const int g_u = g;
  const uint32_t sm_u = ptx::smem_addr(sm);
  const uint32_t gsm_u = sm_u + static_cast<uint32_t>(C::OFF_GROUPS + size_t(g_u) * C::GROUP_BYTES);
  const uint32_t tbase_u = __shfl_sync(0xffffffffu, tmem_slot, 0) + static_cast<uint32_t>(g_u * C::N);
  const uint64_t d_ones = umma::smem_desc_u(sm_u + C::OFF_ONES, C::ROWS * 16, 128);
  const uint64_t d_bias = umma::smem_desc_u(sm_u + C::OFF_BIAS, C::N * 16, 128);

  // x_q of this group: both halves of one step through one barrier
  auto issue_load = [&](long long q) {
    const int rho = static_cast<int>(q / T), t = static_cast<int>(q % T);
    const long long t_a = tile0 + static_cast<long long>(2 * g) * R + rho;
    const bool second = rho + (2 * g + 1) * R < my_tiles;
    const int s = static_cast<int>(q % C::NST);
    ptx::mbar_arrive_expect_tx(&x_full[s], second ? 2 * C::XBYTES : C::XBYTES);
    ptx::bulk_load(xs + (s * 2 + 0) * I * kTile, xin + (t_a * t_in_stride + t) * (I * kTile), C::XBYTES, &x_full[s]);
    if (second) ptx::bulk_load(xs + (s * 2 + 1) * I * kTile, xin + ((t_a + R) * t_in_stride + t) * (I * kTile), C::XBYTES, &x_full[s]);
  };
  if (r == 0 && sp == 0)
    for (long long q = 0; q < C::NST && q < total; ++q) issue_load(q);

  constexpr uint32_t idesc = umma::instr_desc_f16(C::ROWS, C::N);
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  f32x2 cst[C::UPT / 2];                                      // cell states, two units per register pair
  float pmax[POOL ? C::UPT : 1];
  unsigned long long parg = 0ull;
  int rho = 0, t = 0;
  for (long long q = 0; q < total; ++q) {
    const long long tile = tile0 + static_cast<long long>(2 * g + half) * R + rho;
    const bool valid = rho + (2 * g + half) * R < my_tiles;    // uniform over the two warps of a half
    if (t == 0) {
#pragma unroll
      for (int j = 0; j < C::UPT / 2; ++j) cst[j] = 0ull;
    }
    // x_t -> fp16 hi/lo operand chunks (the column blocks dealt over the threads of a window)
    const int s = static_cast<int>(q % C::NST);
    if (valid) {
      ptx::mbar_wait(&x_full[s], static_cast<uint32_t>((q / C::NST) & 1));
      const float* xrow = xs + (s * 2 + half) * I * kTile + b;
#pragma unroll
      for (int kb = 0; kb < C::IP / 8; ++kb) {
        if (SPLIT > 1 && kb % SPLIT != sp) continue;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (kb * 8 + i < I) ? xrow[(kb * 8 + i) * kTile] : 0.f;
        uint4 hi, lo;
        split8(v, hi, lo);
        Ahi[kb * C::ROWS + r] = hi;
        Alo[kb * C::ROWS + r] = lo;
      }
      umma::fence_async_smem();
    }
    umma::fence_before_sync();
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(a_ready)) : "memory");
    if (issuer) {
      // ---- first warp of the group, one elected lane: MMA chain of this step, then the x load that reuses the freed slot ----
      if (umma::elect_one()) {
        ptx::mbar_wait(a_ready, static_cast<uint32_t>(q & 1));
        umma::fence_after_sync();
        if (C::BIAS_MMA) umma::mma_bf16(tbase_u, d_ones, d_bias, idesc, false);
#pragma unroll
        for (int kc = 0; kc < C::KP / 16; ++kc) {
          const uint64_t dah = umma::smem_desc_u(gsm_u + C::G_AHI + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
          const uint64_t dal = umma::smem_desc_u(gsm_u + C::G_ALO + kc * 2 * C::ROWS * 16, C::ROWS * 16, 128);
          const uint64_t dbh = umma::smem_desc_u(sm_u + C::OFF_BHI + kc * 2 * C::N * 16, C::N * 16, 128);
          const uint64_t dbl = umma::smem_desc_u(sm_u + C::OFF_BLO + kc * 2 * C::N * 16, C::N * 16, 128);
          umma::mma_bf16(tbase_u, dah, dbh, idesc, C::BIAS_MMA || kc > 0);
          umma::mma_bf16(tbase_u, dah, dbl, idesc, true);
          umma::mma_bf16(tbase_u, dal, dbh, idesc, true);
        }
        umma::commit(d_full);
        if (q + C::NST < total) issue_load(q + C::NST);
      }
      __syncwarp();
    }
    // gates of this step
    ptx::mbar_wait(d_full, static_cast<uint32_t>(q & 1));
    umma::fence_after_sync();
    if (valid) {
      const bool last_step = (t == T - 1);
#pragma unroll
      for (int lb = 0; lb < C::UPT / 8; ++lb) {
        const int b8 = lb * SPLIT + sp, l0 = lb * 8;     // unit block of this thread, its slot in cst / pmax
        float zi[8], zf[8], zg[8], zo[8], hv[8];
        const uint32_t col0 = tbase + lane_base + b8 * 8;
        umma::tmem_ld8x4(col0, col0 + H, col0 + 2 * H, col0 + 3 * H, zi, zf, zg, zo);
        const int j0 = b8 * 8;
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          f32x2 pi = pk2(zi[i], zi[i + 1]), pf = pk2(zf[i], zf[i + 1]), pg = pk2(zg[i], zg[i + 1]), po = pk2(zo[i], zo[i + 1]);
          if (!C::BIAS_MMA) {
            pi = add2(pi, *reinterpret_cast<const f32x2*>(bias + j0 + i));
            pf = add2(pf, *reinterpret_cast<const f32x2*>(bias + H + j0 + i));
            pg = add2(pg, *reinterpret_cast<const f32x2*>(bias + 2 * H + j0 + i));
            po = add2(po, *reinterpret_cast<const f32x2*>(bias + 3 * H + j0 + i));
          }
          f32x2 sv[5];
          up2(lstm_cell2<SAVE>(pi, pf, pg, po, cst[(l0 + i) / 2], sv), hv[i], hv[i + 1]);
          if (SAVE) {
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
          if (ph == 2) {
            float* yo = yout + ((tile * t_out_stride + t / 3) * H + j0) * kTile + b;
#pragma unroll
            for (int i = 0; i < 8; ++i) yo[i * kTile] = pmax[l0 + i];
            if (SAVE) {
              float* ao = amax + ((tile * t_out_stride + t / 3) * H + j0) * kTile + b;
#pragma unroll
              for (int i = 0; i < 8; ++i) ao[i * kTile] = static_cast<float>((parg >> (2 * (l0 + i))) & 3ull);
            }
          }
        } else {
          float* yo = yout + ((tile * t_out_stride + t) * H + j0) * kTile + b;
#pragma unroll
          for (int i = 0; i < 8; ++i) yo[i * kTile] = hv[i];
        }
        // h_t (zero state at a sequence boundary) into the A tile for the next step
        if (last_step) {
#pragma unroll
          for (int i = 0; i < 8; ++i) hv[i] = 0.f;
        }
        uint4 hi, lo;
        split8(hv, hi, lo);
        const int kb = C::IP / 8 + b8;
        Ahi[kb * C::ROWS + r] = hi;
        Alo[kb * C::ROWS + r] = lo;
      }
    }
    if (++t == T) { t = 0; ++rho; }
  }
  // ---- tear-down ----------------------------------------------------------------------------------
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem_slot, C::TMEM_COLS);
}

}  // namespace gnnqc
