// Host side of the tensor-core LSTM layers (lstm_tc.cuh): weight packing (bf16 hi/lo split, UMMA
// K-major layout), kernel configuration and launch.
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "device_model.cuh"
#include "lstm_tc.cuh"

namespace gnnqc {
namespace {

template <int I, int H, bool POOL>
struct TcLaunch {
  using Cfg = TcCfg<I, H>;
  static auto kernel() { return lstm_tc_kernel<I, H, POOL>; }
  static void configure() { GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES))); }
  static void launch(const float* x, float* y, const unsigned char* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count, cudaStream_t st) {
    // The occupancy API answers 1 for kernels that allocate tensor memory, although CTAs do share an
    // SM as long as their TMEM columns, shared memory and registers fit; size the persistent grid
    // from those three limits directly.
    cudaFuncAttributes fa;
    GNNQC_CUDA(cudaFuncGetAttributes(&fa, kernel()));
    const int regs_per_cta = ((fa.numRegs + 7) / 8 * 8) * ((Cfg::NT + 31) / 32 * 32);
    int per_sm = std::min({512 / Cfg::TMEM_COLS, static_cast<int>((227 * 1024) / (Cfg::SMEM_BYTES + fa.sharedSizeBytes + 1024)), 65536 / regs_per_cta});
    if (const char* e = std::getenv("GNNQC_TC_CTAS")) per_sm = std::atoi(e);
    per_sm = std::max(1, per_sm);
    if (std::getenv("GNNQC_DEBUG")) std::fprintf(stderr, "[gnnqc] tc I=%d H=%d ctas/sm=%d regs=%d smem=%zu tmem_cols=%d\n", I, H, per_sm, fa.numRegs, size_t(Cfg::SMEM_BYTES), Cfg::TMEM_COLS);
    const long long n_pairs = (n_tiles + 1) / 2;
    const long long grid = std::min<long long>(n_pairs, static_cast<long long>(sm_count) * per_sm);
    kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles));
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
};

template <int I, int H, bool POOL>
bool tc_case(TcOp op, const TcArgs& a) {
  if (op == TcOp::Configure) TcLaunch<I, H, POOL>::configure();
  else if (op == TcOp::Launch) TcLaunch<I, H, POOL>::launch(a.x, a.y, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st);
  return true;
}

bool tc_dispatch(TcOp op, int in, int hid, bool pool, const TcArgs& a) {
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
    return m ? std::atoi(m) : 32;
  }();
  return v;
}

}  // namespace

bool tc_layer_supported(int in, int hid, bool pool, bool last) {
  if (last || hid > tc_max_hidden()) return false;
  return tc_dispatch(TcOp::Query, in, hid, pool, TcArgs());
}

size_t tc_packed_bytes(int in, int hid) {
  const int ip = (in + 7) / 8 * 8, kp = (ip + hid + 15) / 16 * 16, n = 4 * hid;
  return size_t(2) * kp * n * 2 + size_t(n) * 4;
}

// B_hi | B_lo in [KP/8][N][8] bf16 (row n = Keras gate column, k = [inputs | pad | hidden | pad]), then bias.
void tc_pack_host(const float* W, const float* U, const float* bias, int in, int hid, unsigned char* out) {
  const int ip = (in + 7) / 8 * 8, kp = (ip + hid + 15) / 16 * 16, n = 4 * hid;
  __nv_bfloat16* bhi = reinterpret_cast<__nv_bfloat16*>(out);
  __nv_bfloat16* blo = bhi + size_t(kp) * n;
  for (int k = 0; k < kp; ++k)
    for (int col = 0; col < n; ++col) {
      float v = 0.f;
      if (k < in) v = W[size_t(k) * n + col];
      else if (k >= ip && k < ip + hid) v = U[size_t(k - ip) * n + col];
      const __nv_bfloat16 h = __float2bfloat16_rn(v);
      const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
      const size_t at = (size_t(k / 8) * n + col) * 8 + (k % 8);
      bhi[at] = h;
      blo[at] = l;
    }
  std::memcpy(out + size_t(2) * kp * n * 2, bias, size_t(n) * 4);
}

void tc_configure(int in, int hid, bool pool) { tc_dispatch(TcOp::Configure, in, hid, pool, TcArgs()); }

void tc_launch(int in, int hid, bool pool, const float* x, float* y, const unsigned char* w, int T, int t_in, int t_out, long long n_tiles,
               int sm_count, cudaStream_t st) {
  TcArgs a;
  a.x = x; a.y = y; a.w = w; a.T = T; a.t_in = t_in; a.t_out = t_out; a.n_tiles = n_tiles; a.sm_count = sm_count; a.st = st;
  if (!tc_dispatch(TcOp::Launch, in, hid, pool, a)) throw std::invalid_argument("no tensor-core kernel for this layer shape");
}

}  // namespace gnnqc
