// Host side of the tensor-core LSTM layers (lstm_tc.cuh): weight packing (bf16 hi/lo split, UMMA
// K-major layout), kernel configuration and launch.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "device_model.cuh"
#include "lstm_tc.cuh"
#include "lstm_tc_big.cuh"
#include "lstm_tc_pair.cuh"
#include "lstm_tc_wide.cuh"

namespace gnnqc {
namespace {

// Whether the gate-saving narrow kernel runs with two threads per window (filled in from measurements).
bool tc_fwd_split2(int hid, long long n_tiles, int sm_count) {
  (void)hid;
  return (n_tiles + 1) / 2 <= 2LL * sm_count;   // all tile pairs in one wave of the two-group shape
}

template <int I, int H, bool POOL, bool SAVE>
struct TcLaunch {
  using Cfg = TcCfg<I, H>;
  using Cfg2 = TcCfg<I, H, 2>;   // two threads per window (gate-saving kernels only: IG and training forward)
  static auto kernel() { return lstm_tc_kernel<I, H, POOL, SAVE>; }
  static auto kernel_compact() { return lstm_tc_kernel<I, H, POOL, SAVE, SAVE>; }   // 16-bit fixed-point gates (save_cells2)
  static auto kernel2() { return lstm_tc_kernel<I, H, POOL, SAVE, false, SAVE ? 2 : 1>; }
  static auto kernel2_compact() { return lstm_tc_kernel<I, H, POOL, SAVE, SAVE, SAVE ? 2 : 1>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
    if (SAVE) {
      GNNQC_CUDA(cudaFuncSetAttribute(kernel_compact(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
      GNNQC_CUDA(cudaFuncSetAttribute(kernel2(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg2::SMEM_BYTES)));
      GNNQC_CUDA(cudaFuncSetAttribute(kernel2_compact(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg2::SMEM_BYTES)));
    }
  }
  static void launch(const float* x, float* y, const unsigned char* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count, cudaStream_t st,
                     float* save, float* amax, int nsave) {
    // One persistent CTA per SM (window groups inside, lstm_tc.cuh); tiles are dealt evenly.
    if (std::getenv("GNNQC_DEBUG")) {
      cudaFuncAttributes fa;
      GNNQC_CUDA(cudaFuncGetAttributes(&fa, kernel()));
      std::fprintf(stderr, "[gnnqc] tc I=%d H=%d regs=%d local=%zu smem=%zu tmem_cols=%d\n", I, H, fa.numRegs, fa.localSizeBytes, size_t(Cfg::SMEM_BYTES), Cfg::TMEM_COLS);
    }
    const long long grid = std::min<long long>(n_tiles, sm_count);
    // Two threads per window shorten the per-step chain and halve the windows in flight: GNNQC_TC_FWD_SPLIT2 = 0 / 1
    // forces a shape, otherwise the rule measured for this layer shape applies (tc_fwd_split2).
    static const int force = [] { const char* e = std::getenv("GNNQC_TC_FWD_SPLIT2"); return e ? std::atoi(e) : -1; }();
    const bool split2 = SAVE && (force >= 0 ? force != 0 : tc_fwd_split2(H, n_tiles, sm_count));
    if (split2)
      (nsave == kSaveCompact ? kernel2_compact() : kernel2())<<<static_cast<unsigned>(grid), Cfg2::NT, Cfg2::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), save, amax, nsave);
    else
      (SAVE && nsave == kSaveCompact ? kernel_compact() : kernel())<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), save, amax, nsave);
    GNNQC_CUDA(cudaGetLastError());
  }
};

// Wide layers (lstm_tc_big.cuh): one CTA per SM, weights resident (H = 64) or streamed (H = 128).
template <int I, int H, bool POOL, bool LAST, bool SAVE>
struct TcBigLaunch {
  static constexpr bool STREAM = (H > 64);
  using Cfg = TcBigCfg<I, H, STREAM>;
  static auto kernel() { return lstm_tc_big_kernel<I, H, POOL, LAST, STREAM, SAVE>; }
  static auto kernel_compact() { return lstm_tc_big_kernel<I, H, POOL, LAST, STREAM, SAVE, SAVE>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
    if (SAVE) GNNQC_CUDA(cudaFuncSetAttribute(kernel_compact(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  }
  static void launch(const float* x, float* y, const unsigned char* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count, cudaStream_t st,
                     float* save, float* amax, int nsave) {
    const long long n_pairs = (n_tiles + 1) / 2;
    const long long grid = std::min<long long>(n_pairs, sm_count);
    if (std::getenv("GNNQC_DEBUG")) {
      cudaFuncAttributes fa;
      GNNQC_CUDA(cudaFuncGetAttributes(&fa, kernel()));
      std::fprintf(stderr, "[gnnqc] tc-big I=%d H=%d regs=%d local=%zu smem=%zu tmem_cols=%d\n", I, H, fa.numRegs, fa.localSizeBytes, size_t(Cfg::SMEM_BYTES), Cfg::TMEM_COLS);
    }
    (SAVE && nsave == kSaveCompact ? kernel_compact() : kernel())<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), save, amax, nsave);
    GNNQC_CUDA(cudaGetLastError());
  }
};

// 64-unit layers, second version (lstm_tc_wide.cuh): input projection one step ahead, sixteen cell-update warps.
template <int I, int H, bool POOL, bool SAVE>
struct TcWideLaunch {
  using Cfg = TcWideCfg<I, H>;
  static auto kernel() { return lstm_tc_wide_kernel<I, H, POOL, SAVE>; }
  static auto kernel_compact() { return lstm_tc_wide_kernel<I, H, POOL, SAVE, SAVE>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
    if (SAVE) GNNQC_CUDA(cudaFuncSetAttribute(kernel_compact(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  }
  static void launch(const float* x, float* y, const unsigned char* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count, cudaStream_t st,
                     float* save, float* amax, int nsave) {
    const long long grid = std::min<long long>((n_tiles + 1) / 2, sm_count);
    (SAVE && nsave == kSaveCompact ? kernel_compact() : kernel())<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), save, amax, nsave);
    GNNQC_CUDA(cudaGetLastError());
  }
};

enum class TcOp { Query, Configure, Launch };

struct TcArgs {
  const float* x = nullptr;
  float* y = nullptr;
  const unsigned char* w = nullptr;
  int T = 0, t_in = 0, t_out = 0, sm_count = 0;
  long long n_tiles = 0;
  cudaStream_t st = nullptr;
  float* save = nullptr;   // non-null: the gate-saving variant (integrated gradients forward)
  float* amax = nullptr;
  int nsave = 5;
};

template <int I, int H, bool POOL>
bool tc_case(TcOp op, const TcArgs& a) {
  if (op == TcOp::Configure) {
    TcLaunch<I, H, POOL, false>::configure();
    TcLaunch<I, H, POOL, true>::configure();
  } else if (op == TcOp::Launch) {
    if (a.save) TcLaunch<I, H, POOL, true>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, a.save, a.amax, a.nsave);
    else TcLaunch<I, H, POOL, false>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, nullptr, nullptr, 5);
  }
  return true;
}

template <int I, int H, bool POOL, bool LAST>
bool tc_big_case(TcOp op, const TcArgs& a) {
  if (op == TcOp::Configure) {
    TcBigLaunch<I, H, POOL, LAST, false>::configure();
    TcBigLaunch<I, H, POOL, LAST, true>::configure();
  } else if (op == TcOp::Launch) {
    if (a.save) TcBigLaunch<I, H, POOL, LAST, true>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, a.save, a.amax, a.nsave);
    else TcBigLaunch<I, H, POOL, LAST, false>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, nullptr, nullptr, 5);
  }
  return true;
}

bool tc_is_big(int hid) { return hid >= 64; }

bool tc_wide_enabled() {   // GNNQC_TC_WIDE=0: the first-version kernel for the 64-unit layers (A/B)
  static const bool on = [] { const char* e = std::getenv("GNNQC_TC_WIDE"); return !(e && std::strcmp(e, "0") == 0); }();
  return on;
}

template <int I, int H, bool POOL>
bool tc_wide_case(TcOp op, const TcArgs& a) {
  if (op == TcOp::Configure) {
    TcWideLaunch<I, H, POOL, false>::configure();
    TcWideLaunch<I, H, POOL, true>::configure();
  } else if (op == TcOp::Launch) {
    if (a.save) TcWideLaunch<I, H, POOL, true>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, a.save, a.amax, a.nsave);
    else TcWideLaunch<I, H, POOL, false>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, nullptr, nullptr, 5);
  }
  return true;
}

bool tc_dispatch(TcOp op, int in, int hid, bool pool, bool last, const TcArgs& a) {
  if (in == 32 && hid == 64 && !pool && !last) {
    if (op == TcOp::Configure) tc_wide_case<32, 64, false>(op, a);
    if (op != TcOp::Configure && tc_wide_enabled()) return tc_wide_case<32, 64, false>(op, a);
    return tc_big_case<32, 64, false, false>(op, a);
  }
  if (in == 64 && hid == 64 && pool && !last) {
    if (op == TcOp::Configure) tc_wide_case<64, 64, true>(op, a);
    if (op != TcOp::Configure && tc_wide_enabled()) return tc_wide_case<64, 64, true>(op, a);
    return tc_big_case<64, 64, true, false>(op, a);
  }
  if (in == 64 && hid == 128 && !pool && last) return tc_big_case<64, 128, false, true>(op, a);
  if (last) return false;
#define GNNQC_TC_CASE(I_, H_, POOL_) \
  if (in == I_ && hid == H_ && pool == POOL_) return tc_case<I_, H_, POOL_>(op, a);
  GNNQC_TC_CASE(2, 16, false)
  GNNQC_TC_CASE(3, 16, false)
  GNNQC_TC_CASE(18, 16, false)
  GNNQC_TC_CASE(19, 16, false)
  GNNQC_TC_CASE(16, 16, true)
  GNNQC_TC_CASE(16, 32, false)
  GNNQC_TC_CASE(32, 32, true)
#undef GNNQC_TC_CASE
  return false;
}

int tc_max_hidden() {
  static const int v = [] {
    const char* e = std::getenv("GNNQC_TC");
    if (e && std::strcmp(e, "0") == 0) return 0;
    const char* m = std::getenv("GNNQC_TC_MAXH");
    return m ? std::atoi(m) : 128;
  }();
  return v;
}

}  // namespace

namespace {
template <int I, int H>
bool pair_case(TcOp op, const TcArgs& a, const unsigned char* w2) {
  using Cfg = TcPairCfg<I, H>;
  if (op == TcOp::Configure) {
    GNNQC_CUDA(cudaFuncSetAttribute(lstm_tc_pair_kernel<I, H>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  } else if (op == TcOp::Launch) {
    const long long grid = std::min<long long>(a.n_tiles, a.sm_count);
    lstm_tc_pair_kernel<I, H><<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, a.st>>>(a.x, a.y, a.w, w2, a.T, a.t_in, a.t_out,
                                                                                               static_cast<int>(a.n_tiles));
    GNNQC_CUDA(cudaGetLastError());
  }
  return true;
}
bool pair_dispatch(TcOp op, int in, int hid, const TcArgs& a, const unsigned char* w2) {
  if (hid == 16) {
    if (in == 2) return pair_case<2, 16>(op, a, w2);
    if (in == 3) return pair_case<3, 16>(op, a, w2);
    if (in == 18) return pair_case<18, 16>(op, a, w2);
    if (in == 19) return pair_case<19, 16>(op, a, w2);
  }
  if (hid == 32 && in == 16) return pair_case<16, 32>(op, a, w2);
  return false;
}
}  // namespace

// Two consecutive layers in -> hid -> hid (+ pool) as one software-pipelined kernel (lstm_tc_pair.cuh): layers 1+2 and 3+4
// of the stack.  GNNQC_TC_PAIR=0: one launch per layer (A/B).
bool tc_pair_supported(int in0, int hid0, int in1, int hid1) {
  static const bool off = [] { const char* e = std::getenv("GNNQC_TC_PAIR"); return e && std::strcmp(e, "0") == 0; }();
  return !off && tc_max_hidden() >= hid0 && in1 == hid0 && hid1 == hid0 && pair_dispatch(TcOp::Query, in0, hid0, TcArgs(), nullptr);
}
void tc_pair_configure(int in0, int hid0) { pair_dispatch(TcOp::Configure, in0, hid0, TcArgs(), nullptr); }
void tc_pair_launch(int in0, int hid0, const float* x, float* y, const unsigned char* w1, const unsigned char* w2, int T, int t_in, int t_out,
                    long long n_tiles, int sm_count, cudaStream_t st) {
  TcArgs a;
  a.x = x; a.y = y; a.w = w1; a.T = T; a.t_in = t_in; a.t_out = t_out; a.n_tiles = n_tiles; a.sm_count = sm_count; a.st = st;
  if (!pair_dispatch(TcOp::Launch, in0, hid0, a, w2)) throw std::invalid_argument("no fused tensor-core kernel for this pair of layers");
}

bool tc_layer_supported(int in, int hid, bool pool, bool last) {
  if (hid > tc_max_hidden()) return false;
  return tc_dispatch(TcOp::Query, in, hid, pool, last, TcArgs());
}

size_t tc_packed_bytes(int in, int hid) {
  if (tc_is_big(hid)) {
    const int kp = ((in + 7) / 8 * 8 + hid + 15) / 16 * 16;
    return size_t(hid / 32) * (kp / 16) * 8192 + size_t(hid / 32) * tc_bias_tile_bytes(128);
  }
  const int ip = (in + 7) / 8 * 8, kp = (ip + hid + 15) / 16 * 16, n = 4 * hid;
  return size_t(2) * kp * n * 2 + tc_bias_tile_bytes(n);
}

namespace {
// One packed value: where the hi and lo halves of the fp16 split go (element index inside the blob, in halfs).
// Shared by the host packer and its device twin so that the two are bit-identical by construction.
struct TcSlot { size_t hi, lo; };
__host__ __device__ inline TcSlot tc_weight_slot(int k, int col, int hid, int kp) {   // col: Keras gate column
  const int n = 4 * hid;
  if (hid >= 64) {   // pieces [chunk of 32 units][K slice of 16]{hi, lo}, each half [2][128][8], operand row = gate * 32 + unit-in-chunk
    const int gate = col / hid, unit = col % hid, c = unit / 32, nrow = gate * 32 + unit % 32, kc = k / 16, kk = k % 16;
    const size_t base = (size_t(c) * (kp / 16) + kc) * 4096 + (size_t(kk / 8) * 128 + nrow) * 8 + (kk % 8);
    return {base, base + 2048};
  }
  const size_t at = (size_t(k / 8) * n + col) * 8 + (k % 8);   // B_hi | B_lo, each [KP/8][N][8]
  return {at, size_t(kp) * n + at};
}
__host__ __device__ inline TcSlot tc_bias_slot(int col, int hid, int kp) {            // row of the bias tile: k = 0 hi, k = 1 lo
  const int n = 4 * hid;
  if (hid >= 64) {
    const int gate = col / hid, unit = col % hid, c = unit / 32, nrow = gate * 32 + unit % 32;
    const size_t base = size_t(hid / 32) * (kp / 16) * 4096 + size_t(c) * (tc_bias_tile_bytes(128) / 2) + size_t(nrow) * 8;
    return {base, base + 1};
  }
  const size_t base = size_t(2) * kp * n + size_t(col) * 8;
  return {base, base + 1};
}
__host__ __device__ inline float tc_weight_value(const float* W, const float* U, int k, int col, int in, int hid) {
  const int ip = (in + 7) / 8 * 8, n = 4 * hid;
  float v = 0.f;
  if (k < in) v = W[size_t(k) * n + col];
  else if (k >= ip && k < ip + hid) v = U[size_t(k - ip) * n + col];
  return v * tc_gate_scale(col / hid);
}
}  // namespace

// Gate columns are pre-multiplied by the factor of their exponential (lstm_cell2).  Weights: fp16 hi / lo of
// [W ; U] in the UMMA K-major layout (k = [inputs | pad | hidden | pad]); biases: one more operand tile (lstm_tc.cuh).
void tc_pack_host(const float* W, const float* U, const float* bias, int in, int hid, unsigned char* out) {
  const int ip = (in + 7) / 8 * 8, kp = (ip + hid + 15) / 16 * 16, n = 4 * hid;
  std::memset(out, 0, tc_packed_bytes(in, hid));
  __half* o = reinterpret_cast<__half*>(out);
  for (int k = 0; k < kp; ++k)
    for (int col = 0; col < n; ++col) {
      const float v = tc_weight_value(W, U, k, col, in, hid);
      const __half h = __float2half_rn(v);
      const TcSlot s = tc_weight_slot(k, col, hid, kp);
      o[s.hi] = h;
      o[s.lo] = __float2half_rn(v - __half2float(h));
    }
  for (int col = 0; col < n; ++col) {
    const float v = bias[col] * tc_gate_scale(col / hid);
    const __half h = __float2half_rn(v);
    const TcSlot s = tc_bias_slot(col, hid, kp);
    o[s.hi] = h;
    o[s.lo] = __float2half_rn(v - __half2float(h));
  }
}

namespace {
// Device twin of tc_pack_host (same slot maps, same fp32 scaling and fp16 rounding), for the refresh after an optimizer step.
__global__ void __launch_bounds__(256) tc_repack_kernel(const float* __restrict__ W, const float* __restrict__ U, const float* __restrict__ bias,
                                                        int in, int hid, unsigned char* __restrict__ out) {
  const int ip = (in + 7) / 8 * 8, kp = (ip + hid + 15) / 16 * 16, n = 4 * hid;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n_main = static_cast<long long>(kp) * n;
  __half* o = reinterpret_cast<__half*>(out);
  if (i < n_main) {
    const int k = static_cast<int>(i / n), col = static_cast<int>(i % n);
    const float v = tc_weight_value(W, U, k, col, in, hid);
    const __half h = __float2half_rn(v);
    const TcSlot s = tc_weight_slot(k, col, hid, kp);
    o[s.hi] = h;
    o[s.lo] = __float2half_rn(v - __half2float(h));
  } else if (i < n_main + n) {
    const int col = static_cast<int>(i - n_main);
    const float v = bias[col] * tc_gate_scale(col / hid);
    const __half h = __float2half_rn(v);
    const TcSlot s = tc_bias_slot(col, hid, kp);
    o[s.hi] = h;
    o[s.lo] = __float2half_rn(v - __half2float(h));   // the other entries of the bias tile were zeroed when the blob was first packed
  }
}
}  // namespace

void tc_repack_device(const float* W, const float* U, const float* bias, int in, int hid, unsigned char* out, cudaStream_t st) {
  const int ip = (in + 7) / 8 * 8, kp = (ip + hid + 15) / 16 * 16, n = 4 * hid;
  const long long total = static_cast<long long>(kp) * n + n;
  tc_repack_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(W, U, bias, in, hid, out);
  GNNQC_CUDA(cudaGetLastError());
}

void tc_configure(int in, int hid, bool pool, bool last) { tc_dispatch(TcOp::Configure, in, hid, pool, last, TcArgs()); }

void tc_launch(int in, int hid, bool pool, bool last, const float* x, float* y, const unsigned char* w, int T, int t_in, int t_out, long long n_tiles,
               int sm_count, cudaStream_t st, float* save, float* amax, int nsave) {
  TcArgs a;
  a.save = save; a.amax = amax; a.nsave = nsave;
  a.x = x; a.y = y; a.w = w; a.T = T; a.t_in = t_in; a.t_out = t_out; a.n_tiles = n_tiles; a.sm_count = sm_count; a.st = st;
  if (!tc_dispatch(TcOp::Launch, in, hid, pool, last, a)) throw std::invalid_argument("no tensor-core kernel for this layer shape");
}

}  // namespace gnnqc
