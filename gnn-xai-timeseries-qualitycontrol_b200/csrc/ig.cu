// Integrated gradients for the CML GCN model on the GPU (gnnqc_ig_cml).
//
// Restates xai/libs/integrated_gradients.py (reference):
//   :898-917   zero baseline
//   :919-942   alphas = linspace(0, 1, m_steps+1); both inputs interpolated with the same alpha
//   :955-1004  per alpha: forward of the whole batch, gradient of the (summed) prediction w.r.t.
//              the interpolated node features and target series
//   :1006-1012 trapezoid: mean over k of (g_k + g_{k+1})/2
//   :1185-1191 scale by (x - baseline)
//   :1229      prediction of the un-interpolated batch
// Per alpha the device runs: front-end (inputs scaled by alpha) -> 7 x LSTM forward with saved
// gates -> head forward+gradient -> 7 x LSTM input-gradient backward -> front-end backward, which
// accumulates w_k * x * grad straight into the caller's attribution buffers, so nothing but the
// inputs is read from and nothing but the attributions is written to caller memory.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <cmath>

#include <functional>

#include "device_model.cuh"
#include "frontend.cuh"
#include "ig_frontend_bwd.cuh"
#include "lstm_bwd.cuh"
#include "lstm_fwd.cuh"

namespace gnnqc {
namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

size_t ig_scratch_budget_bytes() {
  const char* env = std::getenv("GNNQC_IG_SCRATCH_MB");
  size_t mb = env ? std::strtoull(env, nullptr, 10) : 24576;
  return std::max<size_t>(mb, 64) << 20;
}

// Forward with SAVE: packed FFMA2 accumulation, plain cell (the gates are needed individually).
template <int I, int H, bool POOL, bool LAST>
struct FwdSave {
  static constexpr bool WGLOBAL = (size_t(I + H) * H * 16 > 160 * 1024);
  static constexpr int RJ = 2, RB = 8, LJ = 8, NST = 2, VAR = kVarPacked;
  using Cfg = LstmCfg<I, H, RJ, RB, LJ, WGLOBAL, NST>;
  static auto kernel() { return lstm_layer_kernel<I, H, RJ, RB, LJ, POOL, LAST, WGLOBAL, VAR, NST, 5>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  }
  static void launch(const float* x, float* y, const float* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count,
                     float* save, float* amax, cudaStream_t st) {
    int per_sm = 1;
    GNNQC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel(), Cfg::NT, Cfg::SMEM_BYTES));
    if (per_sm < 1) throw CudaError("lstm_layer_kernel<SAVE> does not fit on an SM");
    long long grid = std::min<long long>(n_tiles, static_cast<long long>(sm_count) * per_sm);
    kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), save, amax);
    GNNQC_CUDA(cudaGetLastError());
  }
};

template <int I, int H, bool POOL, bool LAST>
struct Bwd {
  static constexpr bool WGLOBAL = (size_t(4 * H) * ((I + H + 15) / 16 * 16) * 4 > 150 * 1024);
  using Cfg = BwdCfg<I, H, WGLOBAL>;
  static auto kernel() { return lstm_layer_bwd_kernel<I, H, POOL, LAST, WGLOBAL>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  }
  static void launch(const float* dy, const float* save, const float* amax, const float* wT, float* dx, int T, long long n_tiles,
                     int sm_count, cudaStream_t st) {
    int per_sm = 1;
    GNNQC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel(), Cfg::NT, Cfg::SMEM_BYTES));
    if (per_sm < 1) throw CudaError("lstm_layer_bwd_kernel does not fit on an SM");
    long long grid = std::min<long long>(n_tiles, static_cast<long long>(sm_count) * per_sm);
    kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(dy, save, amax, wT, dx, T, static_cast<int>(n_tiles), nullptr);
    GNNQC_CUDA(cudaGetLastError());
  }
};

enum class Op { Configure, Forward, Backward };

struct LayerArgs {
  const float* in = nullptr;   // forward: x tile; backward: dy
  float* out = nullptr;        // forward: y tile; backward: dx
  const float* w = nullptr;    // forward: packed; backward: transposed
  const unsigned char* wtc = nullptr;   // tensor-core operand tiles of this direction (null: fp32 kernel)
  float* save = nullptr;
  float* amax = nullptr;
  int nsave = 5;               // kSaveCompact: 16-bit fixed-point gates (tensor-core kernels on both sides only)
  int T = 0, t_in = 0, t_out = 0;
  long long n_tiles = 0;
  int sm_count = 0;
  cudaStream_t st = nullptr;
};

template <int I, int H, bool POOL, bool LAST>
void layer_op(Op op, const LayerArgs& a) {
  if (op == Op::Configure) {
    FwdSave<I, H, POOL, LAST>::configure();
    Bwd<I, H, POOL, LAST>::configure();
    if (tc_bwd_supported(I, H, POOL, LAST)) tc_bwd_configure(I, H, POOL, LAST);
  } else if (op == Op::Forward) {
    if (a.wtc && tc_layer_supported(I, H, POOL, LAST))   // gate contraction on the tensor cores (lstm_tc.cuh, SAVE mode)
      return tc_launch(I, H, POOL, LAST, a.in, a.out, a.wtc, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, a.save, a.amax, a.nsave);
    FwdSave<I, H, POOL, LAST>::launch(a.in, a.out, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.save, a.amax, a.st);
  } else {
    if (a.wtc && tc_bwd_supported(I, H, POOL, LAST))   // contraction on the tensor cores (lstm_tc_bwd.cuh)
      return tc_bwd_launch(I, H, POOL, LAST, a.in, a.save, a.amax, a.wtc, a.out, a.T, a.n_tiles, a.sm_count, a.st, a.nsave);
    Bwd<I, H, POOL, LAST>::launch(a.in, a.save, a.amax, a.w, a.out, a.T, a.n_tiles, a.sm_count, a.st);
  }
}

void dispatch(Op op, int in, int hid, bool pool, bool last, const LayerArgs& a) {
#define GNNQC_CASE(I_, H_, POOL_, LAST_) \
  if (in == I_ && hid == H_ && pool == POOL_ && last == LAST_) return layer_op<I_, H_, POOL_, LAST_>(op, a);
  GNNQC_CASE(18, 16, false, false)
  GNNQC_CASE(16, 16, true, false)
  GNNQC_CASE(16, 32, false, false)
  GNNQC_CASE(32, 32, true, false)
  GNNQC_CASE(32, 64, false, false)
  GNNQC_CASE(64, 64, true, false)
  GNNQC_CASE(64, 128, false, true)
#undef GNNQC_CASE
  throw std::invalid_argument("integrated gradients: no kernel for this LSTM layer shape");
}


}  // namespace

void DeviceModel::configure_ig_kernels() {
  const ModelSpec& s = host.spec;
  if (!s.has_gcn || s.soilnet) return;
  configure_cml_frontend();
  GNNQC_CUDA(cudaFuncSetAttribute(cml_frontend_bwd_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ig_bwd_smem_bytes(kIgMaxGroup))));
  GNNQC_CUDA(cudaGetLastError());
  LayerArgs none;
  for (int l = 0; l < s.n_lstm; ++l) {
    const bool last = (l == s.n_lstm - 1);
    dispatch(Op::Configure, s.lstm_in[l], s.lstm_hid[l], !last && (l % 2 == 1), last, none);
  }
}

// Integrated gradients over windows whatever their source: `fwd(alpha, win0, nw, nt, x1)` builds the x1 tiles of a chunk
// at the interpolation point alpha * x, `bwd(alpha, wk, first, win0, nw, dx1)` turns d(sum p)/d(x1) into attributions
// (trapezoid weight wk; `first` assigns instead of accumulating).  `extra_bytes` of scratch come first and are handed
// to `prepare` (node coefficients / per-window node lists).
void DeviceModel::ig_core(long long B, int T, int m_steps, float* scores, cudaStream_t st, size_t extra_bytes,
                          const std::function<void(char*)>& prepare,
                          const std::function<void(float, long long, long long, long long, float*)>& fwd,
                          const std::function<void(const IgGroup&, long long, long long)>& bwd, bool scale_by_input) {
  if (B <= 0) return;
  const ModelSpec& s = host.spec;
  GNNQC_CUDA(cudaSetDevice(device));
  refresh_tc();
  // GNNQC_IG_TC=0: forward passes on the fp32 CUDA-core kernels (A/B)
  static const bool ig_tc = [] { const char* e = std::getenv("GNNQC_IG_TC"); return !(e && std::strcmp(e, "0") == 0); }();
  static const bool ig_tc_bwd = [] { const char* e = std::getenv("GNNQC_IG_TC_BWD"); return ig_tc && !(e && std::strcmp(e, "0") == 0); }();
  StackPlan plan(s, T);
  const int L = s.n_lstm;
  const long long n_tiles = (B + kTile - 1) / kTile;
  // Saved gates as 16-bit fixed point + fp32 c (12 instead of 20 bytes per cell, lstm_tc.cuh: save_cells2) when both the
  // forward and the backward of every layer run on the tensor-core kernels; GNNQC_IG_COMPACT=0 keeps fp32 gates (A/B).
  static const bool ig_compact_env = [] { const char* e = std::getenv("GNNQC_IG_COMPACT"); return !(e && std::strcmp(e, "0") == 0); }();
  bool compact = ig_compact_env && ig_tc && ig_tc_bwd && d_packed_tc && d_packed_tcb;
  for (int l = 0; l < L && compact; ++l) {
    const bool last = (l == L - 1), pool = !last && (l % 2 == 1);
    compact = tc_layer_supported(s.lstm_in[l], s.lstm_hid[l], pool, last) && tc_bwd_supported(s.lstm_in[l], s.lstm_hid[l], pool, last);
  }
  const int nsave = compact ? 3 : 5;

  // Per-tile scratch: forward ping/pong, gradient ping/pong, dy7, saved gates and pool argmax.
  const size_t sa = align_up(plan.bytes_a, 256), sb = align_up(plan.bytes_b, 256);
  size_t sg = 0;
  for (int l = 0; l < L; ++l) sg = std::max(sg, size_t(plan.t_run[l]) * s.lstm_in[l] * kTile * sizeof(float));
  sg = align_up(sg, 256);
  size_t save_off[kMaxLstmLayers], amax_off[kMaxLstmLayers], per_tile_save = 0;
  for (int l = 0; l < L; ++l) {
    save_off[l] = per_tile_save;
    per_tile_save += size_t(plan.t_run[l]) * s.lstm_hid[l] * nsave * kTile * sizeof(float);
    amax_off[l] = per_tile_save;
    if (l % 2 == 1 && l != L - 1) per_tile_save += size_t(plan.t_out[l]) * s.lstm_hid[l] * kTile * sizeof(float);
  }
  const size_t s7 = size_t(s.lstm_hid[L - 1]) * kTile * sizeof(float);
  const size_t per_tile = sa + sb + 2 * sg + s7 + per_tile_save;
  extra_bytes = align_up(extra_bytes, 256);
  const long long chunk = balanced_chunk(n_tiles, static_cast<long long>(ig_scratch_budget_bytes() / per_tile));
  // dx1 of a whole group of interpolation points stays resident, so the front-end backward reads the features and writes
  // the attributions once per group instead of once per alpha (ig_frontend_bwd.cuh); GNNQC_IG_DX_MB bounds it (default 48 GiB: 51 alphas of a 75,776-window wave)
  static const size_t dx_budget = [] { const char* e = std::getenv("GNNQC_IG_DX_MB"); return (e ? std::strtoull(e, nullptr, 10) : size_t(49152)) << 20; }();
  const long long group = std::max<long long>(1, std::min<long long>(std::min<long long>(m_steps + 1, kIgMaxGroup), static_cast<long long>(dx_budget / (size_t(chunk) * sg))));
  const size_t dx_bytes = size_t(group) * size_t(chunk) * sg;
  char* base = static_cast<char*>(scratch(extra_bytes + size_t(chunk) * per_tile + dx_bytes));
  float* dx_all = reinterpret_cast<float*>(base + extra_bytes + size_t(chunk) * per_tile);
  const long long dx_stride = static_cast<long long>(size_t(chunk) * sg / sizeof(float));
  char* p = base + extra_bytes;
  float* buf_a = reinterpret_cast<float*>(p); p += size_t(chunk) * sa;
  float* buf_b = reinterpret_cast<float*>(p); p += size_t(chunk) * sb;
  float* g0 = reinterpret_cast<float*>(p); p += size_t(chunk) * sg;
  float* g1 = reinterpret_cast<float*>(p); p += size_t(chunk) * sg;
  float* dy7 = reinterpret_cast<float*>(p); p += size_t(chunk) * s7;
  char* saves = p;   // layer l of the chunk: saves + chunk*save_off[l] (layer-major, then tile)

  prepare(base);
  const float* prm = d_params;
  static const char* const kFwd[kMaxLstmLayers] = {"ig_fwd_l1", "ig_fwd_l2", "ig_fwd_l3", "ig_fwd_l4", "ig_fwd_l5", "ig_fwd_l6", "ig_fwd_l7"};
  static const char* const kBwd[kMaxLstmLayers] = {"ig_bwd_l1", "ig_bwd_l2", "ig_bwd_l3", "ig_bwd_l4", "ig_bwd_l5", "ig_bwd_l6", "ig_bwd_l7"};

  for (long long t0 = 0; t0 < n_tiles; t0 += chunk) {
    const long long nt = std::min(chunk, n_tiles - t0);
    const long long win0 = t0 * kTile, nw = std::min<long long>(B - win0, nt * kTile);
    for (int k = 0; k <= m_steps; ++k) {
      const float alpha = static_cast<float>(k) / static_cast<float>(m_steps);
      const float wk = ((k == 0 || k == m_steps) ? 0.5f : 1.0f) / static_cast<float>(m_steps);
      mark_begin("ig_frontend", st);
      fwd(alpha, win0, nw, nt, buf_a);
      GNNQC_CUDA(cudaGetLastError());
      mark_end(st);
      // forward, saving gates
      const float* in = buf_a;
      int t_in = plan.t_run[0];
      for (int l = 0; l < L; ++l) {
        const bool last = (l == L - 1), pool = !last && (l % 2 == 1);
        LayerArgs a;
        a.in = in; a.out = (l % 2 == 0) ? buf_b : buf_a; a.w = d_packed + packed_off[l];
        a.wtc = ig_tc ? d_packed_tc + packed_tc_off[l] : nullptr;
        a.save = reinterpret_cast<float*>(saves + size_t(chunk) * save_off[l]);
        a.amax = pool ? reinterpret_cast<float*>(saves + size_t(chunk) * amax_off[l]) : nullptr;
        a.nsave = nsave;
        a.T = plan.t_run[l]; a.t_in = t_in; a.t_out = plan.t_out[l]; a.n_tiles = nt; a.sm_count = sm_count; a.st = st;
        mark_begin(kFwd[l], st);
        dispatch(Op::Forward, s.lstm_in[l], s.lstm_hid[l], pool, last, a);
        mark_end(st);
        in = a.out;
        t_in = plan.t_out[l];
      }
      if ((s.off_d1_w | s.off_d2_w) & 3) throw std::invalid_argument("integrated gradients: dense kernels must start on a 16-byte boundary of the parameter vector");
      mark_begin("ig_head", st);
      head_grad_kernel<128, 64><<<static_cast<unsigned>(nt), 256, 0, st>>>(in, prm + s.off_d1_w, prm + s.off_d1_b, prm + s.off_d2_w,
                                                                            prm + s.off_d2_b, prm + s.off_out_w, prm + s.off_out_b,
                                                                            s.leaky_alpha, dy7, (k == m_steps) ? scores + win0 : nullptr, nw);
      GNNQC_CUDA(cudaGetLastError());
      mark_end(st);
      // backward
      const float* dy = dy7;
      for (int l = L - 1; l >= 0; --l) {
        const bool last = (l == L - 1), pool = !last && (l % 2 == 1);
        LayerArgs a;
        a.in = dy; a.out = (l % 2 == 0) ? g0 : g1; a.w = d_packed_t + packed_t_off[l];
        if (l == 0) a.out = dx_all + static_cast<long long>(k % group) * dx_stride;   // kept until its group is complete
        a.wtc = ig_tc_bwd ? d_packed_tcb + packed_tcb_off[l] : nullptr;
        a.save = reinterpret_cast<float*>(saves + size_t(chunk) * save_off[l]);
        a.amax = pool ? reinterpret_cast<float*>(saves + size_t(chunk) * amax_off[l]) : nullptr;
        a.nsave = nsave;
        a.T = plan.t_run[l]; a.n_tiles = nt; a.sm_count = sm_count; a.st = st;
        mark_begin(kBwd[l], st);
        dispatch(Op::Backward, s.lstm_in[l], s.lstm_hid[l], pool, last, a);
        mark_end(st);
        dy = a.out;
      }
      (void)wk;
      if ((k + 1) % group == 0 || k == m_steps) {
        IgGroup g;
        g.k0 = static_cast<int>(k / group * group);
        g.nk = k - g.k0 + 1;
        g.m_steps = m_steps;
        g.dx = dx_all;
        g.stride = dx_stride;
        g.first = g.k0 == 0 ? 1 : 0;
        g.scale_by_input = scale_by_input ? 1 : 0;
        mark_begin("ig_frontend_bwd", st);
        bwd(g, win0, nw);
        GNNQC_CUDA(cudaGetLastError());
        mark_end(st);
      }
    }
  }
}

void DeviceModel::ig_cml(const float* feat, const float* anom, const long long* edges, const int* node_off,
                         const long long* edge_off, long long B, int T, long long n_nodes, int m_steps, float* attr_feat,
                         float* attr_anom, float* scores, cudaStream_t st, bool scale_by_input) {
  if (B <= 0) return;
  GNNQC_CUDA(cudaSetDevice(device));
  const StackPlan plan(host.spec, T);
  const size_t coef_bytes = align_up(size_t(n_nodes) * sizeof(float), 256);
  const GcnFolded folded = this->folded(st);
  const float2* feat2 = reinterpret_cast<const float2*>(feat);
  float* coef = nullptr;
  ig_core(B, T, m_steps, scores, st, 2 * coef_bytes,
          [&](char* extra) {
            int* deg = reinterpret_cast<int*>(extra);
            coef = reinterpret_cast<float*>(extra + coef_bytes);
            GNNQC_CUDA(cudaMemsetAsync(deg, 0, size_t(n_nodes) * sizeof(int), st));
            mark_begin("cml_node_coef", st);
            cml_node_coef_kernel<<<static_cast<unsigned>((B * 32 + 255) / 256), 256, 0, st>>>(edges, node_off, edge_off, deg, coef, B);
            GNNQC_CUDA(cudaGetLastError());
            mark_end(st);
          },
          [&](float alpha, long long win0, long long nw, long long nt, float* x1) {
            launch_cml_frontend(feat2, n_nodes * T, reinterpret_cast<const float2*>(anom) + win0 * T, coef, node_off + win0, folded, x1, nw, T,
                                plan.t_run[0], nt, alpha, sm_count, st);
          },
          [&](const IgGroup& g, long long win0, long long nw) {
            const dim3 grid(static_cast<unsigned>((nw + kIgWin - 1) / kIgWin), static_cast<unsigned>(T));
            cml_frontend_bwd_multi_kernel<<<grid, 256, ig_bwd_smem_bytes(g.nk), st>>>(feat2, reinterpret_cast<const float2*>(anom) + win0 * T, coef,
                                                                                     node_off + win0, folded, g, reinterpret_cast<float2*>(attr_feat),
                                                                                     reinterpret_cast<float2*>(attr_anom) + win0 * T, nw, T, plan.t_run[0]);
          },
          scale_by_input);
}

}  // namespace gnnqc
