// Host side of the tensor-core LSTM backward kernels (lstm_tc_bwd.cuh): [W | U]^T operand packing
// (bf16 hi/lo split, UMMA K-major layout, one piece per chunk of 8 units), configuration and launch.
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "device_model.cuh"
#include "lstm_tc_bwd.cuh"

namespace gnnqc {
namespace {

template <int I, int H, bool POOL, bool LAST, int GRP, int SPLIT>
struct TcBwdLaunch {
  static constexpr bool STREAM = (H > 64);
  using Cfg = TcBwdCfg<I, H, GRP, SPLIT, STREAM>;
  static auto kernel() { return lstm_tc_bwd_kernel<I, H, POOL, LAST, GRP, SPLIT, STREAM, false>; }
  static auto kernel_train() { return lstm_tc_bwd_kernel<I, H, POOL, LAST, GRP, SPLIT, STREAM, true>; }
  static auto kernel_compact() { return lstm_tc_bwd_kernel<I, H, POOL, LAST, GRP, SPLIT, STREAM, false, true>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
    GNNQC_CUDA(cudaFuncSetAttribute(kernel_train(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
    GNNQC_CUDA(cudaFuncSetAttribute(kernel_compact(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  }
  static void launch(const float* dy, const float* save, const float* amax, const unsigned char* wpk, float* dx, int T, long long n_tiles,
                     int sm_count, cudaStream_t st, int nsave, float* dz_out) {
    const long long n_pairs = (n_tiles + 1) / 2;
    const long long grid = std::min<long long>(n_pairs, sm_count);
    if (std::getenv("GNNQC_DEBUG")) {
      cudaFuncAttributes fa;
      GNNQC_CUDA(cudaFuncGetAttributes(&fa, kernel()));
      std::fprintf(stderr, "[gnnqc] tc-bwd I=%d H=%d regs=%d local=%zu smem=%zu threads=%d\n", I, H, fa.numRegs, fa.localSizeBytes, size_t(Cfg::SMEM_BYTES), Cfg::NT);
    }
    if (nsave == 6 && dz_out)
      kernel_train()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(dy, save, amax, wpk, dx, T, static_cast<int>(n_tiles), dz_out);
    else if (nsave == 5 && !dz_out)
      kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(dy, save, amax, wpk, dx, T, static_cast<int>(n_tiles), nullptr);
    else if (nsave == kSaveCompact && !dz_out)
      kernel_compact()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(dy, save, amax, wpk, dx, T, static_cast<int>(n_tiles), nullptr);
    else
      throw std::invalid_argument("tensor-core backward: nsave/dz_out combination not built");
    GNNQC_CUDA(cudaGetLastError());
  }
};

enum class BwdOp { Query, Configure, Launch };

struct BwdArgs {
  const float* dy = nullptr;
  const float* save = nullptr;
  const float* amax = nullptr;
  const unsigned char* w = nullptr;
  float* dx = nullptr;
  int T = 0, sm_count = 0;
  long long n_tiles = 0;
  cudaStream_t st = nullptr;
  int nsave = 5;
  float* dz_out = nullptr;
};

template <int I, int H, bool POOL, bool LAST, int GRP, int SPLIT>
bool bwd_case(BwdOp op, const BwdArgs& a) {
  using L = TcBwdLaunch<I, H, POOL, LAST, GRP, SPLIT>;
  if (op == BwdOp::Configure) L::configure();
  else if (op == BwdOp::Launch) L::launch(a.dy, a.save, a.amax, a.w, a.dx, a.T, a.n_tiles, a.sm_count, a.st, a.nsave, a.dz_out);
  return true;
}

bool bwd_dispatch(BwdOp op, int in, int hid, bool pool, bool last, const BwdArgs& a) {
  // Narrow layers come in two shapes: four groups of 128 threads per SM (one thread per window, all units), which
  // keeps the most windows in flight, and two groups of 256 threads (two threads per window, half the units each),
  // whose step is shorter (~4 us against ~6 us).  The second is faster as long as all tile pairs fit in one wave of
  // it (measured: training steps of 128 and 4,096 windows; a 65,536-window IG chunk is 25 % slower with it).
  // GNNQC_TC_BWD_SPLIT2=0 / 1 forces one of them (A/B).
  static const int force = [] { const char* e = std::getenv("GNNQC_TC_BWD_SPLIT2"); return e ? std::atoi(e) : -1; }();
  const bool split2 = force >= 0 ? force != 0 : (a.n_tiles + 1) / 2 <= 2LL * a.sm_count;
#define GNNQC_NARROW(I_, H_, POOL_)                                                    \
  if (in == I_ && hid == H_ && pool == POOL_ && !last) {                               \
    if (op == BwdOp::Configure) return bwd_case<I_, H_, POOL_, false, 4, 1>(op, a) && bwd_case<I_, H_, POOL_, false, 2, 2>(op, a); \
    return split2 ? bwd_case<I_, H_, POOL_, false, 2, 2>(op, a) : bwd_case<I_, H_, POOL_, false, 4, 1>(op, a);                   \
  }
  GNNQC_NARROW(2, 16, false)
  GNNQC_NARROW(18, 16, false)
  GNNQC_NARROW(16, 16, true)
  GNNQC_NARROW(16, 32, false)
  GNNQC_NARROW(32, 32, true)
#undef GNNQC_NARROW
  // 64-unit layers: two threads per window as well (32 units each; 96 instead of 168 registers, twice the warps per
  // scheduler).  IG backward 26.6 -> 21.4 ms (layer 5) and 32.5 -> 25.2 ms (layer 6) per 51 x 65,536 windows.
  // GNNQC_TC_BWD_WIDE2=0: one thread per window (A/B).
  static const bool wide2 = [] { const char* e = std::getenv("GNNQC_TC_BWD_WIDE2"); return !(e && std::atoi(e) == 0); }();
  if (in == 32 && hid == 64 && !pool && !last) {
    if (op == BwdOp::Configure) return bwd_case<32, 64, false, false, 2, 1>(op, a) && bwd_case<32, 64, false, false, 2, 2>(op, a);
    return wide2 ? bwd_case<32, 64, false, false, 2, 2>(op, a) : bwd_case<32, 64, false, false, 2, 1>(op, a);
  }
  if (in == 64 && hid == 64 && pool && !last) {
    if (op == BwdOp::Configure) return bwd_case<64, 64, true, false, 2, 1>(op, a) && bwd_case<64, 64, true, false, 2, 2>(op, a);
    return wide2 ? bwd_case<64, 64, true, false, 2, 2>(op, a) : bwd_case<64, 64, true, false, 2, 1>(op, a);
  }
  if (in == 64 && hid == 128 && !pool && last) return bwd_case<64, 128, false, true, 1, 2>(op, a);
  return false;
}

}  // namespace

bool tc_bwd_supported(int in, int hid, bool pool, bool last) {
  static const bool off = [] { const char* e = std::getenv("GNNQC_TC"); return e && std::strcmp(e, "0") == 0; }();
  return !off && bwd_dispatch(BwdOp::Query, in, hid, pool, last, BwdArgs());
}

size_t tc_bwd_packed_bytes(int in, int hid) {
  const int n = (in + 15) / 16 * 16 + hid;
  return size_t(hid / 8) * 128 * n;
}

// Piece of chunk c: [slice s = 0, 1][hi | lo][kb = 0, 1][n < N][8] bf16 with
//   k = c*32 + (2 s + kb) * 8 + u  <->  Keras column (2 s + kb) * H + 8 c + u   (gate-major inside a chunk)
//   n < I: kernel row n;  IP <= n < IP + H: recurrent row n - IP;  zero otherwise.
void tc_bwd_pack_host(const float* W, const float* U, int in, int hid, unsigned char* out) {
  const int ip = (in + 15) / 16 * 16, n_out = ip + hid, n4 = 4 * hid;
  for (int c = 0; c < hid / 8; ++c) {
    unsigned char* piece = out + size_t(c) * 128 * n_out;
    for (int s = 0; s < 2; ++s)
      for (int kb = 0; kb < 2; ++kb)
        for (int n = 0; n < n_out; ++n)
          for (int u = 0; u < 8; ++u) {
            const int col = (2 * s + kb) * hid + 8 * c + u;
            float v = 0.f;
            if (n < in) v = W[size_t(n) * n4 + col];
            else if (n >= ip) v = U[size_t(n - ip) * n4 + col];
            const __nv_bfloat16 h = __float2bfloat16_rn(v);
            const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
            const size_t at = (size_t(kb) * n_out + n) * 8 + u;
            reinterpret_cast<__nv_bfloat16*>(piece + size_t(s * 2 + 0) * 2 * n_out * 16)[at] = h;
            reinterpret_cast<__nv_bfloat16*>(piece + size_t(s * 2 + 1) * 2 * n_out * 16)[at] = l;
          }
  }
}

namespace {
// Device twin of tc_bwd_pack_host.
__global__ void __launch_bounds__(256) tc_bwd_repack_kernel(const float* __restrict__ W, const float* __restrict__ U, int in, int hid,
                                                            unsigned char* __restrict__ out) {
  const int ip = (in + 15) / 16 * 16, n_out = ip + hid, n4 = 4 * hid;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(n_out) * n4) return;
  const int n = static_cast<int>(i / n4), col = static_cast<int>(i % n4);
  float v = 0.f;
  if (n < in) v = W[static_cast<long long>(n) * n4 + col];
  else if (n >= ip) v = U[static_cast<long long>(n - ip) * n4 + col];
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
  const int gate = col / hid, unit = col % hid, c = unit / 8, u = unit % 8, s = gate / 2, kb = gate % 2;
  unsigned char* piece = out + static_cast<size_t>(c) * 128 * n_out;
  const size_t at = (static_cast<size_t>(kb) * n_out + n) * 8 + u;
  reinterpret_cast<__nv_bfloat16*>(piece + static_cast<size_t>(s * 2 + 0) * 2 * n_out * 16)[at] = h;
  reinterpret_cast<__nv_bfloat16*>(piece + static_cast<size_t>(s * 2 + 1) * 2 * n_out * 16)[at] = l;
}
}  // namespace

void tc_bwd_repack_device(const float* W, const float* U, int in, int hid, unsigned char* out, cudaStream_t st) {
  const long long total = static_cast<long long>((in + 15) / 16 * 16 + hid) * 4 * hid;
  tc_bwd_repack_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(W, U, in, hid, out);
  GNNQC_CUDA(cudaGetLastError());
}

void tc_bwd_configure(int in, int hid, bool pool, bool last) { bwd_dispatch(BwdOp::Configure, in, hid, pool, last, BwdArgs()); }

void tc_bwd_launch(int in, int hid, bool pool, bool last, const float* dy, const float* save, const float* amax, const unsigned char* w,
                   float* dx, int T, long long n_tiles, int sm_count, cudaStream_t st, int nsave, float* dz_out) {
  BwdArgs a;
  a.nsave = nsave; a.dz_out = dz_out;
  a.dy = dy; a.save = save; a.amax = amax; a.w = w; a.dx = dx; a.T = T; a.n_tiles = n_tiles; a.sm_count = sm_count; a.st = st;
  if (!bwd_dispatch(BwdOp::Launch, in, hid, pool, last, a)) throw std::invalid_argument("no tensor-core backward kernel for this layer shape");
}

}  // namespace gnnqc
