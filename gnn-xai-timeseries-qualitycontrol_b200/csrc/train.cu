// Training step on the GPU (gnnqc_train_*), restating libs/fit_model.py:61-112 as Keras executes it:
//   forward in training mode (BatchNormalization of the GeneralConv uses the batch statistics over
//   all node rows), loss = sum_b class_weight[y_b] * BCE(logit_b, y_b) / batch_size, backward through
//   head, the seven LSTM layers (BPTT), max-pooling, node pooling and the GeneralConv, Adam update,
//   moving-average update of the BN statistics.
// Split in three calls so that data-parallel ranks can sum across GPUs in between (the only
// exchange steps of the whole path):
//   gnnqc_train_moments_cml   local feature moments (6 doubles)     -> all-reduce (sum)
//   gnnqc_train_grads_*       local reduction vector (grads, raw GCN sums, loss) -> all-reduce (sum)
//   gnnqc_train_apply         BN-backward closure, Adam, BN moving statistics, weight re-packing
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>

#include "device_model.cuh"
#include "lstm_bwd.cuh"
#include "soil_kernels.cuh"
#include "train_kernels.cuh"
#include "wgrad_tc.cuh"

namespace gnnqc {
namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <int I, int H, bool POOL, bool LAST>
struct TrainFwd {
  static constexpr bool WGLOBAL = (size_t(I + H) * H * 16 > 160 * 1024);
  static constexpr int RJ = 2, RB = 8, LJ = 8, NST = 2, VAR = kVarPacked;
  using Cfg = LstmCfg<I, H, RJ, RB, LJ, WGLOBAL, NST>;
  static auto kernel() { return lstm_layer_kernel<I, H, RJ, RB, LJ, POOL, LAST, WGLOBAL, VAR, NST, 6>; }
  static void configure() { GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES))); }
  static void launch(const float* x, float* y, const float* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count, float* save,
                     float* amax, cudaStream_t st) {
    int per_sm = 1;
    GNNQC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel(), Cfg::NT, Cfg::SMEM_BYTES));
    long long grid = std::min<long long>(n_tiles, static_cast<long long>(sm_count) * std::max(per_sm, 1));
    kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), save, amax);
    GNNQC_CUDA(cudaGetLastError());
  }
};

template <int I, int H, bool POOL, bool LAST>
struct TrainBwd {
  static constexpr bool WGLOBAL = (size_t(4 * H) * ((I + H + 15) / 16 * 16) * 4 > 150 * 1024);
  using Cfg = BwdCfg<I, H, WGLOBAL>;
  static auto kernel() { return lstm_layer_bwd_kernel<I, H, POOL, LAST, WGLOBAL, 6>; }
  static void configure() { GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES))); }
  static void launch(const float* dy, const float* save, const float* amax, const float* wT, float* dx, int T, long long n_tiles,
                     int sm_count, float* dz_out, cudaStream_t st) {
    int per_sm = 1;
    GNNQC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel(), Cfg::NT, Cfg::SMEM_BYTES));
    long long grid = std::min<long long>(n_tiles, static_cast<long long>(sm_count) * std::max(per_sm, 1));
    kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(dy, save, amax, wT, dx, T, static_cast<int>(n_tiles), dz_out);
    GNNQC_CUDA(cudaGetLastError());
  }
};

enum class Op { Configure, Forward, Backward };
struct LayerArgs {
  const float* in = nullptr;
  float* out = nullptr;
  const float* w = nullptr;
  const unsigned char* wtc = nullptr;   // tensor-core operand tiles of this direction (null: fp32 kernel)
  float* save = nullptr;
  float* amax = nullptr;
  float* dz = nullptr;
  int T = 0, t_in = 0, t_out = 0;
  long long n_tiles = 0;
  int sm_count = 0;
  cudaStream_t st = nullptr;
};

template <int I, int H, bool POOL, bool LAST>
void layer_op(Op op, const LayerArgs& a) {
  if (op == Op::Configure) {
    TrainFwd<I, H, POOL, LAST>::configure();
    TrainBwd<I, H, POOL, LAST>::configure();
    if (tc_bwd_supported(I, H, POOL, LAST)) tc_bwd_configure(I, H, POOL, LAST);
  } else if (op == Op::Forward) {
    if (a.wtc && tc_layer_supported(I, H, POOL, LAST))   // lstm_tc*.cuh in SAVE mode with h_t as sixth quantity
      return tc_launch(I, H, POOL, LAST, a.in, a.out, a.wtc, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.st, a.save, a.amax, 6);
    TrainFwd<I, H, POOL, LAST>::launch(a.in, a.out, a.w, a.T, a.t_in, a.t_out, a.n_tiles, a.sm_count, a.save, a.amax, a.st);
  } else {
    if (a.wtc && tc_bwd_supported(I, H, POOL, LAST))     // lstm_tc_bwd.cuh, also writing dz for the weight-gradient kernel
      return tc_bwd_launch(I, H, POOL, LAST, a.in, a.save, a.amax, a.wtc, a.out, a.T, a.n_tiles, a.sm_count, a.st, 6, a.dz);
    TrainBwd<I, H, POOL, LAST>::launch(a.in, a.save, a.amax, a.w, a.out, a.T, a.n_tiles, a.sm_count, a.dz, a.st);
  }
}

void dispatch(Op op, int in, int hid, bool pool, bool last, const LayerArgs& a) {
#define GNNQC_CASE(I_, H_, POOL_, LAST_) \
  if (in == I_ && hid == H_ && pool == POOL_ && last == LAST_) return layer_op<I_, H_, POOL_, LAST_>(op, a);
  GNNQC_CASE(2, 16, false, false)
  GNNQC_CASE(3, 16, false, false)
  GNNQC_CASE(18, 16, false, false)
  GNNQC_CASE(19, 16, false, false)
  GNNQC_CASE(16, 16, true, false)
  GNNQC_CASE(16, 32, false, false)
  GNNQC_CASE(32, 32, true, false)
  GNNQC_CASE(32, 64, false, false)
  GNNQC_CASE(64, 64, true, false)
  GNNQC_CASE(64, 128, false, true)
#undef GNNQC_CASE
  throw std::invalid_argument("training: no kernel for this LSTM layer shape");
}

constexpr size_t kPartialBytes = size_t(64) << 20;

}  // namespace

struct TrainState {
  float* mom = nullptr;
  float* vel = nullptr;
  GcnFolded* folded = nullptr;
  BnBatch* stats = nullptr;
  long long step = 0;
  int kind = 0;            // 0 adam | 1 sgd | 2 rmsprop
  float last_lr = 0.f;
  // the weight gradients of layer l run on a side stream while the backward pass goes on with layer l - 1
  cudaStream_t side = nullptr;
  cudaEvent_t ev_layer[kMaxLstmLayers] = {};
  cudaEvent_t ev_done = nullptr;
};

void DeviceModel::configure_train_kernels() {
  const ModelSpec& s = host.spec;
  LayerArgs none;
  for (int l = 0; l < s.n_lstm; ++l) {
    const bool last = (l == s.n_lstm - 1);
    dispatch(Op::Configure, s.lstm_in[l], s.lstm_hid[l], !last && (l % 2 == 1), last, none);
  }
  GNNQC_CUDA(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kWgradSmem)));
  GNNQC_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024));
}

TrainState* DeviceModel::train_state() {
  if (!train) {
    train = new TrainState();
    const size_t n = size_t(host.spec.n_params);
    GNNQC_CUDA(cudaSetDevice(device));
    GNNQC_CUDA(cudaMalloc(&train->mom, n * sizeof(float)));
    GNNQC_CUDA(cudaMalloc(&train->vel, n * sizeof(float)));
    GNNQC_CUDA(cudaMemset(train->mom, 0, n * sizeof(float)));
    GNNQC_CUDA(cudaMemset(train->vel, 0, n * sizeof(float)));
    // a checkpoint written by model.save carries the optimizer state: resume where it stopped
    train->kind = host.optimizer;
    train->step = host.opt_iterations;
    train->last_lr = host.opt_learning_rate;
    if (host.opt_m.size() == n) GNNQC_CUDA(cudaMemcpy(train->mom, host.opt_m.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    if (host.opt_v.size() == n) GNNQC_CUDA(cudaMemcpy(train->vel, host.opt_v.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMalloc(&train->folded, sizeof(GcnFolded)));
    GNNQC_CUDA(cudaMalloc(&train->stats, sizeof(BnBatch)));
  }
  return train;
}

void DeviceModel::free_train_state() {
  if (!train) return;
  cudaFree(train->mom);
  cudaFree(train->vel);
  cudaFree(train->folded);
  cudaFree(train->stats);
  if (train->side) cudaStreamDestroy(train->side);
  for (cudaEvent_t e : train->ev_layer)
    if (e) cudaEventDestroy(e);
  if (train->ev_done) cudaEventDestroy(train->ev_done);
  delete train;
  train = nullptr;
}

void DeviceModel::set_optimizer(int kind) {
  if (kind < 0 || kind > 2) throw std::invalid_argument("optimizer must be 0 (adam), 1 (sgd) or 2 (rmsprop)");
  TrainState* ts = train_state();
  if (ts->kind != kind) {   // slots of another rule mean nothing to this one
    const size_t n = size_t(host.spec.n_params);
    GNNQC_CUDA(cudaMemset(ts->mom, 0, n * sizeof(float)));
    GNNQC_CUDA(cudaMemset(ts->vel, 0, n * sizeof(float)));
    ts->step = 0;
  }
  ts->kind = kind;
  host.optimizer = kind;
}

long long DeviceModel::train_iterations() const { return train ? train->step : host.opt_iterations; }

// Brings the optimizer state to the host copy (for gnnqc_model_save).
void DeviceModel::sync_host_optimizer() {
  if (!train || device < 0) return;
  GNNQC_CUDA(cudaSetDevice(device));
  GNNQC_CUDA(cudaDeviceSynchronize());
  const size_t n = size_t(host.spec.n_params);
  host.optimizer = train->kind;
  host.opt_iterations = train->step;
  host.opt_learning_rate = train->last_lr;
  host.opt_m.assign(train->kind == 0 ? n : 0, 0.f);
  host.opt_v.assign(train->kind != 1 ? n : 0, 0.f);
  if (train->kind == 0) GNNQC_CUDA(cudaMemcpy(host.opt_m.data(), train->mom, n * sizeof(float), cudaMemcpyDeviceToHost));
  if (train->kind != 1) GNNQC_CUDA(cudaMemcpy(host.opt_v.data(), train->vel, n * sizeof(float), cudaMemcpyDeviceToHost));
}

void DeviceModel::sync_host_params() {
  if (!params_dirty || device < 0) return;
  GNNQC_CUDA(cudaSetDevice(device));
  GNNQC_CUDA(cudaDeviceSynchronize());
  GNNQC_CUDA(cudaMemcpy(host.params.data(), d_params, size_t(host.spec.n_params) * sizeof(float), cudaMemcpyDeviceToHost));
  params_dirty = false;
}

void DeviceModel::train_moments(const float* feat, long long n_rows, double* moments, cudaStream_t st) {
  GNNQC_CUDA(cudaSetDevice(device));
  const bool soil = host.spec.soilnet;       // three input channels: ten moments instead of six
  const int n_mom = soil ? kSoilMoments : 6;
  GNNQC_CUDA(cudaMemsetAsync(moments, 0, n_mom * sizeof(double), st));
  mark_begin("train_moments", st);
  if (soil) feat_moments3_kernel<<<sm_count * 4, 256, 0, st>>>(feat, n_rows, moments);
  else feat_moments_kernel<<<sm_count * 4, 256, 0, st>>>(reinterpret_cast<const float2*>(feat), n_rows, moments);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);
  const double n = static_cast<double>(n_rows);
  GNNQC_CUDA(cudaMemcpyAsync(moments + n_mom - 1, &n, sizeof(double), cudaMemcpyHostToDevice, st));
  GNNQC_CUDA(cudaStreamSynchronize(st));  // `n` lives on this stack frame
}

// Shared by the GCN and baseline paths.  x_first: the first layer's input tiles (already built).
void DeviceModel::train_grads(const float* feat, const float* anom, const long long* edges, const int* node_off,
                              const long long* edge_off, const float* x_seq, const long long* labels, long long B, int T,
                              long long n_nodes, const double* moments, float cw0, float cw1, long long global_batch, float* rvec,
                              cudaStream_t st, long long n_edges) {
  const ModelSpec& s = host.spec;
  const bool gcn = s.has_gcn;
  const bool soil = gcn && s.soilnet;        // every sensor of every window is a sequence (create_model.py:224-231)
  const long long n_win = B;
  if (soil) B = n_nodes;                     // from here on B counts sequences
  GNNQC_CUDA(cudaSetDevice(device));
  refresh_tc();   // no-op inside a training loop: train_apply refreshes the operand tiles on the device
  TrainState* ts = train_state();
  // GNNQC_TRAIN_TC=0: forward/backward on the fp32 CUDA-core kernels (A/B)
  static const bool train_tc = [] { const char* e = std::getenv("GNNQC_TRAIN_TC"); return !(e && std::strcmp(e, "0") == 0); }();
  // GNNQC_TRAIN_TC_BWD_MIN_TILES: smallest step (in 64-window tiles) whose backward runs on the tensor-core kernels
  // (default 2 per SM); read per call so a test can send a small batch down that path
  long long tc_bwd_min_tiles = 1;
  if (const char* e = std::getenv("GNNQC_TRAIN_TC_BWD_MIN_TILES")) tc_bwd_min_tiles = std::max(1LL, std::atoll(e));
  StackPlan plan(s, T);
  const int L = s.n_lstm;
  const long long nt = (B + kTile - 1) / kTile;
  const long long np = s.n_params;

  // ---- scratch layout (whole local batch at once; training batches are small) -------------------
  size_t off = 0;
  auto take = [&off](size_t bytes) { size_t at = off; off += align_up(bytes, 256); return at; };
  const size_t o_deg = take(size_t(std::max<long long>(n_nodes, 1)) * 4), o_coef = take(size_t(std::max<long long>(n_nodes, 1)) * 4);
  // SoilNet: CSR by receiving node (forward) and by sending node (backward)
  const size_t nb = size_t(n_nodes + 1) * sizeof(int);
  size_t scan_bytes = 0;
  size_t o_csr[7] = {0}, o_scan = 0, o_nbr = 0, o_nbrT = 0;
  if (soil) {
    GNNQC_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, static_cast<int*>(nullptr), static_cast<int*>(nullptr), static_cast<int>(n_nodes + 1), st));
    for (int k = 0; k < 7; ++k) o_csr[k] = take(nb);      // deg, cursor, row_ptr, seq_win, degT, cursorT, rowT_ptr
    o_scan = take(scan_bytes);
    o_nbr = take(size_t(std::max<long long>(n_edges, 1)) * sizeof(int));
    o_nbrT = take(size_t(std::max<long long>(n_edges, 1)) * sizeof(int));
  }
  const size_t o_x1 = take(size_t(nt) * plan.t_run[0] * s.lstm_in[0] * kTile * 4);
  size_t o_y[kMaxLstmLayers], o_save[kMaxLstmLayers], o_amax[kMaxLstmLayers], o_dz[kMaxLstmLayers];
  size_t g_bytes = 0;
  for (int l = 0; l < L; ++l) {
    o_y[l] = take(size_t(nt) * plan.t_out[l] * s.lstm_hid[l] * kTile * 4);
    o_save[l] = take(size_t(nt) * plan.t_run[l] * 6 * s.lstm_hid[l] * kTile * 4);
    o_amax[l] = (l % 2 == 1 && l != L - 1) ? take(size_t(nt) * plan.t_out[l] * s.lstm_hid[l] * kTile * 4) : 0;
    o_dz[l] = take(size_t(nt) * plan.t_run[l] * 4 * s.lstm_hid[l] * kTile * 4);
    g_bytes = std::max(g_bytes, size_t(nt) * plan.t_run[l] * s.lstm_in[l] * kTile * 4);
  }
  const size_t o_g0 = take(g_bytes), o_g1 = take(g_bytes);
  const size_t o_dy7 = take(size_t(nt) * s.lstm_hid[L - 1] * kTile * 4);
  const size_t o_head = take(size_t(nt) * kHeadTrainFloats * 4);
  const size_t o_lossp = take(size_t(nt) * 4);
  const dim3 grid_fe(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
  const dim3 grid_feb(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + kFeBwdSteps - 1) / kFeBwdSteps));
  const size_t o_gcnp = take(size_t(grid_feb.x) * grid_feb.y * kRvecGcn * 4);
  const size_t o_part = take(kPartialBytes);
  char* base = static_cast<char*>(scratch(off));
  auto F = [base](size_t o) { return reinterpret_cast<float*>(base + o); };

  GNNQC_CUDA(cudaMemsetAsync(rvec, 0, size_t(np + kRvecGcn + 1) * sizeof(float), st));
  auto I32 = [base](size_t o) { return reinterpret_cast<int*>(base + o); };

  // ---- forward ----------------------------------------------------------------------------------
  const float* prm = d_params;
  if (soil) {
    int *deg = I32(o_csr[0]), *cursor = I32(o_csr[1]), *row_ptr = I32(o_csr[2]), *seq_win = I32(o_csr[3]);
    int *degT = I32(o_csr[4]), *cursorT = I32(o_csr[5]), *rowT_ptr = I32(o_csr[6]);
    GNNQC_CUDA(cudaMemsetAsync(deg, 0, nb, st));
    GNNQC_CUDA(cudaMemsetAsync(cursor, 0, nb, st));
    GNNQC_CUDA(cudaMemsetAsync(degT, 0, nb, st));
    GNNQC_CUDA(cudaMemsetAsync(cursorT, 0, nb, st));
    const unsigned wblocks = static_cast<unsigned>((n_win * 32 + 255) / 256), nblocks = static_cast<unsigned>((n_nodes + 255) / 256);
    mark_begin("soil_csr", st);
    soil_count_kernel<<<wblocks, 256, 0, st>>>(edges, node_off, edge_off, deg, seq_win, n_win, 0);
    soil_count_kernel<<<wblocks, 256, 0, st>>>(edges, node_off, edge_off, degT, seq_win, n_win, 1);
    GNNQC_CUDA(cub::DeviceScan::ExclusiveSum(base + o_scan, scan_bytes, deg, row_ptr, static_cast<int>(n_nodes + 1), st));
    GNNQC_CUDA(cub::DeviceScan::ExclusiveSum(base + o_scan, scan_bytes, degT, rowT_ptr, static_cast<int>(n_nodes + 1), st));
    soil_fill_kernel<<<wblocks, 256, 0, st>>>(edges, node_off, edge_off, row_ptr, cursor, I32(o_nbr), n_win, 0);
    soil_fill_kernel<<<wblocks, 256, 0, st>>>(edges, node_off, edge_off, rowT_ptr, cursorT, I32(o_nbrT), n_win, 1);
    soil_sort_kernel<<<nblocks, 256, 0, st>>>(row_ptr, deg, I32(o_nbr), n_nodes);
    soil_sort_kernel<<<nblocks, 256, 0, st>>>(rowT_ptr, degT, I32(o_nbrT), n_nodes);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    bn_prepare3_kernel<<<1, 32, 0, st>>>(prm + s.off_gcn_kernel, prm + s.off_gcn_bias, prm + s.off_prelu, prm + s.off_bn_gamma,
                                         prm + s.off_bn_beta, moments, s.bn_eps, ts->folded, ts->stats);
    GNNQC_CUDA(cudaGetLastError());
    mark_begin("train_frontend", st);
    soil_frontend_devparam_kernel<<<grid_fe, 256, 0, st>>>(feat, node_off, seq_win, row_ptr, deg, I32(o_nbr), ts->folded, F(o_x1), 0, B, T,
                                                          plan.t_run[0]);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
  } else if (gcn) {
    GNNQC_CUDA(cudaMemsetAsync(base + o_deg, 0, size_t(n_nodes) * 4, st));
    mark_begin("cml_node_coef", st);
    cml_node_coef_kernel<<<static_cast<unsigned>((B * 32 + 255) / 256), 256, 0, st>>>(edges, node_off, edge_off, reinterpret_cast<int*>(base + o_deg), F(o_coef), B);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    bn_prepare_kernel<<<1, 32, 0, st>>>(prm + s.off_gcn_kernel, prm + s.off_gcn_bias, prm + s.off_prelu, prm + s.off_bn_gamma,
                                        prm + s.off_bn_beta, moments, s.bn_eps, ts->folded, ts->stats);
    GNNQC_CUDA(cudaGetLastError());
    mark_begin("train_frontend", st);
    cml_frontend_devparam_kernel<<<grid_fe, 256, 0, st>>>(reinterpret_cast<const float2*>(feat), reinterpret_cast<const float2*>(anom), F(o_coef),
                                                          node_off, ts->folded, F(o_x1), B, T, plan.t_run[0]);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
  } else {
    mark_begin("seq_to_tiles", st);
    if (s.lstm_in[0] == 3) seq_to_tiles_kernel<3><<<grid_fe, 256, 0, st>>>(x_seq, F(o_x1), B, T, plan.t_run[0]);
    else seq_to_tiles_kernel<2><<<grid_fe, 256, 0, st>>>(x_seq, F(o_x1), B, T, plan.t_run[0]);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
  }
  static const char* const kFwd[kMaxLstmLayers] = {"train_fwd_l1", "train_fwd_l2", "train_fwd_l3", "train_fwd_l4", "train_fwd_l5", "train_fwd_l6", "train_fwd_l7"};
  static const char* const kBwd[kMaxLstmLayers] = {"train_bwd_l1", "train_bwd_l2", "train_bwd_l3", "train_bwd_l4", "train_bwd_l5", "train_bwd_l6", "train_bwd_l7"};
  const float* in = F(o_x1);
  int t_in = plan.t_run[0];
  for (int l = 0; l < L; ++l) {
    const bool last = (l == L - 1), pool = !last && (l % 2 == 1);
    LayerArgs a;
    a.in = in; a.out = F(o_y[l]); a.w = d_packed + packed_off[l]; a.save = F(o_save[l]); a.amax = pool ? F(o_amax[l]) : nullptr;
    a.wtc = train_tc ? d_packed_tc + packed_tc_off[l] : nullptr;
    a.T = plan.t_run[l]; a.t_in = t_in; a.t_out = plan.t_out[l]; a.n_tiles = nt; a.sm_count = sm_count; a.st = st;
    mark_begin(kFwd[l], st);
    dispatch(Op::Forward, s.lstm_in[l], s.lstm_hid[l], pool, last, a);
    mark_end(st);
    in = a.out;
    t_in = plan.t_out[l];
  }
  mark_begin("train_head", st);
  head_train_kernel<128, 64><<<static_cast<unsigned>(nt), 256, 0, st>>>(in, prm + s.off_d1_w, prm + s.off_d1_b, prm + s.off_d2_w, prm + s.off_d2_b,
                                                                        prm + s.off_out_w, prm + s.off_out_b, s.leaky_alpha, labels, cw0, cw1,
                                                                        1.f / static_cast<float>(global_batch), F(o_dy7), F(o_head), F(o_lossp),
                                                                        nullptr, B);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);

  // ---- weight gradients ------------------------------------------------------------------------------
  auto wgrad = [&](const char* name, const float* A, long long a_tile, long long a_step, int NA, const float* Bm, long long b_tile,
                   long long b_step, int NC, int n_steps, int bias, float* out_main, float* out_bias) {
    if (n_steps <= 0) return;
    WgradArgs g;
    g.A = A; g.B = Bm; g.a_tile = a_tile; g.a_step = a_step; g.b_tile = b_tile; g.b_step = b_step;
    g.NA = NA; g.NC = NC; g.n_steps = n_steps; g.n_items = nt * n_steps; g.bias = bias; g.partial = F(o_part);
    const int rows = NA + bias;
    const unsigned gx = (rows + kWgA - 1) / kWgA, gy = (NC + kWgC - 1) / kWgC;
    long long slices = std::max<long long>(1, (static_cast<long long>(sm_count) * 16) / (gx * gy));
    slices = std::min<long long>(slices, g.n_items);
    slices = std::min<long long>(slices, 512);
    slices = std::min<long long>(slices, static_cast<long long>(kPartialBytes / (size_t(rows) * NC * 4)));
    mark_begin(name, st);
    wgrad_kernel<<<dim3(gx, gy, static_cast<unsigned>(slices)), 64, kWgradSmem, st>>>(g);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    mark_begin("wgrad_reduce", st);
    wgrad_reduce_kernel<<<(rows * NC * 32 + 255) / 256, 256, 0, st>>>(F(o_part), static_cast<int>(slices), rows, NC, out_main, NA, out_bias);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
  };
  // GNNQC_TRAIN_TC_WGRAD=0: the LSTM weight gradients on the CUDA-core kernel (A/B)
  static const bool wgrad_tc = [] { const char* e = std::getenv("GNNQC_TRAIN_TC_WGRAD"); return !(e && std::strcmp(e, "0") == 0); }();
  auto layer_wgrad = [&](int l) {
    const int I = s.lstm_in[l], H = s.lstm_hid[l], Tl = plan.t_run[l];
    const float* xin = (l == 0) ? F(o_x1) : F(o_y[l - 1]);
    const int t_in_l = (l == 0) ? plan.t_run[0] : plan.t_out[l - 1];
    const long long dz_tile = static_cast<long long>(Tl) * 4 * H * kTile, dz_step = 4LL * H * kTile;
    if (train_tc && wgrad_tc && I + H + 1 <= 256 && Tl > 0) {   // dW, dU, db of the layer in one tensor-core launch (wgrad_tc.cuh)
      WgradTcArgs g;
      g.x = xin; g.x_tile = static_cast<long long>(t_in_l) * I * kTile; g.x_step = static_cast<long long>(I) * kTile;
      g.h = F(o_save[l]) + 5LL * H * kTile; g.h_tile = static_cast<long long>(Tl) * 6 * H * kTile; g.h_step = 6LL * H * kTile;
      g.dz = F(o_dz[l]); g.dz_tile = dz_tile; g.dz_step = dz_step;
      g.I = I; g.H = H; g.T = Tl; g.n_items = nt * Tl;
      g.n16 = (I + H + 1 + 15) / 16 * 16;
      g.n_stages = wgrad_tc_stages(g.n16);
      g.partial = F(o_part);
      const int m_chunks = (4 * H + 127) / 128, rows = I + H + 1;
      long long slices = std::max<long long>(1, sm_count / m_chunks);
      slices = std::min<long long>(slices, g.n_items);
      slices = std::min<long long>(slices, static_cast<long long>(kPartialBytes / (size_t(rows) * 4 * H * 4)));
      mark_begin("wgrad_tc", st);
      wgrad_tc_kernel<<<dim3(static_cast<unsigned>(slices), m_chunks), kWgTcThreads, wgrad_tc_smem_bytes(g.n16, g.n_stages), st>>>(g);
      GNNQC_CUDA(cudaGetLastError());
      mark_end(st);
      mark_begin("wgrad_reduce", st);
      wgrad_tc_reduce_kernel<<<(rows * 4 * H * 32 + 255) / 256, 256, 0, st>>>(F(o_part), static_cast<int>(slices), I, H, rvec + s.off_lstm_w[l],
                                                                             rvec + s.off_lstm_u[l], rvec + s.off_lstm_b[l]);
      GNNQC_CUDA(cudaGetLastError());
      mark_end(st);
      return;
    }
    wgrad("wgrad_w", xin, static_cast<long long>(t_in_l) * I * kTile, static_cast<long long>(I) * kTile, I, F(o_dz[l]), dz_tile, dz_step, 4 * H, Tl,
          1, rvec + s.off_lstm_w[l], rvec + s.off_lstm_b[l]);
    // recurrent kernel: h_{t-1} (saved quantity 5) against dz_t, t = 1..T-1
    wgrad("wgrad_u", F(o_save[l]) + 5LL * H * kTile, static_cast<long long>(Tl) * 6 * H * kTile, 6LL * H * kTile, H, F(o_dz[l]) + dz_step, dz_tile,
          dz_step, 4 * H, Tl - 1, 0, rvec + s.off_lstm_u[l], nullptr);
  };

  // ---- backward through the LSTM stack -------------------------------------------------------------
  // GNNQC_TRAIN_OVERLAP=0: everything on the caller's stream (A/B).  The partial-sum scratch is shared by the layers'
  // weight-gradient launches, which stay in order on the one side stream.
  static const bool overlap_env = [] { const char* e = std::getenv("GNNQC_TRAIN_OVERLAP"); return !(e && std::strcmp(e, "0") == 0); }();
  const bool overlap = overlap_env;
  const cudaStream_t main_st = st;
  if (overlap && !ts->side) {
    GNNQC_CUDA(cudaStreamCreateWithFlags(&ts->side, cudaStreamNonBlocking));
    for (int l = 0; l < kMaxLstmLayers; ++l) GNNQC_CUDA(cudaEventCreateWithFlags(&ts->ev_layer[l], cudaEventDisableTiming));
    GNNQC_CUDA(cudaEventCreateWithFlags(&ts->ev_done, cudaEventDisableTiming));
  }
  const float* dy = F(o_dy7);
  for (int l = L - 1; l >= 0; --l) {
    const bool last = (l == L - 1), pool = !last && (l % 2 == 1);
    LayerArgs a;
    a.in = dy; a.out = (l % 2 == 0) ? F(o_g0) : F(o_g1); a.w = d_packed_t + packed_t_off[l]; a.save = F(o_save[l]);
    // tensor-core backward at every batch size (lstm_tc_bwd.cu picks the two-threads-per-window shape of the narrow
    // layers for small steps); GNNQC_TRAIN_TC_BWD_MIN_TILES=<n> sends smaller steps of the narrow layers to the fp32 kernel
    a.wtc = (train_tc && (nt >= tc_bwd_min_tiles || s.lstm_hid[l] >= 64)) ? d_packed_tcb + packed_tcb_off[l] : nullptr;
    a.amax = pool ? F(o_amax[l]) : nullptr; a.dz = F(o_dz[l]);
    a.T = plan.t_run[l]; a.n_tiles = nt; a.sm_count = sm_count; a.st = st;
    mark_begin(kBwd[l], st);
    dispatch(Op::Backward, s.lstm_in[l], s.lstm_hid[l], pool, last, a);
    mark_end(st);
    dy = a.out;
    if (overlap) {   // dz of layer l is complete: its weight gradients start on the side stream
      GNNQC_CUDA(cudaEventRecord(ts->ev_layer[l], main_st));
      GNNQC_CUDA(cudaStreamWaitEvent(ts->side, ts->ev_layer[l], 0));
      st = ts->side;
      layer_wgrad(l);
      st = main_st;
    }
  }
  const float* dx1 = dy;
  if (!overlap)
    for (int l = 0; l < L; ++l) layer_wgrad(l);

  {
    if (overlap) st = ts->side;   // the dense layers' gradients share the partial-sum scratch: same stream as the layers'
    const long long ht = kHeadTrainFloats;
    const int D = s.dense_units, HL = s.lstm_hid[L - 1];
    wgrad("wgrad_dense", F(o_y[L - 1]), static_cast<long long>(HL) * kTile, 0, HL, F(o_head) + 2LL * D * kTile, ht, 0, D, 1, 1, rvec + s.off_d1_w,
          rvec + s.off_d1_b);
    wgrad("wgrad_dense", F(o_head), ht, 0, D, F(o_head) + 3LL * D * kTile, ht, 0, D, 1, 1, rvec + s.off_d2_w, rvec + s.off_d2_b);
    wgrad("wgrad_dense", F(o_head) + 1LL * D * kTile, ht, 0, D, F(o_head) + 4LL * D * kTile, ht, 0, 1, 1, 1, rvec + s.off_out_w, rvec + s.off_out_b);
    st = main_st;
  }
  if (soil) {
    mark_begin("train_frontend_bwd", st);
    soil_frontend_train_bwd_kernel<<<grid_feb, 256, 0, st>>>(feat, node_off, I32(o_csr[3]), I32(o_csr[6]), I32(o_csr[4]), I32(o_nbrT), I32(o_csr[0]),
                                                            prm + s.off_gcn_kernel, prm + s.off_gcn_bias, ts->folded, ts->stats, s.bn_eps, dx1,
                                                            F(o_gcnp), B, T, plan.t_run[0]);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    sum_partials_kernel<<<1, 128, 0, st>>>(F(o_gcnp), static_cast<long long>(grid_feb.x) * grid_feb.y, kSoilGcnSums, rvec + np);
    GNNQC_CUDA(cudaGetLastError());
  } else if (gcn) {
    mark_begin("train_frontend_bwd", st);
    cml_frontend_train_bwd_kernel<<<grid_feb, 256, 0, st>>>(reinterpret_cast<const float2*>(feat), F(o_coef), node_off, prm + s.off_gcn_kernel,
                                                           prm + s.off_gcn_bias, ts->folded, ts->stats, s.bn_eps, dx1, F(o_gcnp), B, T,
                                                           plan.t_run[0]);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    sum_partials_kernel<<<1, 128, 0, st>>>(F(o_gcnp), static_cast<long long>(grid_feb.x) * grid_feb.y, kGcnSums, rvec + np);
    GNNQC_CUDA(cudaGetLastError());
  }
  sum_partials_kernel<<<1, 128, 0, st>>>(F(o_lossp), nt, 1, rvec + np + kRvecGcn);
  GNNQC_CUDA(cudaGetLastError());
  if (overlap) {   // the caller's stream owns the reduction vector again only when the side stream is done with it
    GNNQC_CUDA(cudaEventRecord(ts->ev_done, ts->side));
    GNNQC_CUDA(cudaStreamWaitEvent(main_st, ts->ev_done, 0));
  }
}

void DeviceModel::train_apply(float* rvec, const double* moments, float lr, cudaStream_t st) {
  const ModelSpec& s = host.spec;
  GNNQC_CUDA(cudaSetDevice(device));
  TrainState* ts = train_state();
  const long long np = s.n_params;
  long long skip_lo = 0, skip_hi = 0;
  if (s.has_gcn) {
    if (!moments) throw std::invalid_argument("gnnqc_train_apply: a GCN model needs the (global) feature moments");
    GcnOffsets o{s.off_gcn_kernel, s.off_gcn_bias, s.off_prelu, s.off_bn_gamma, s.off_bn_beta, s.off_bn_mean, s.off_bn_var};
    if (s.soilnet) gcn_finalize3_kernel<<<1, 32, 0, st>>>(rvec, np, d_params, o, moments, ts->stats, s.bn_eps);
    else gcn_finalize_kernel<<<1, 32, 0, st>>>(rvec, np, d_params, o, moments, ts->stats, s.bn_eps);
    GNNQC_CUDA(cudaGetLastError());
    skip_lo = s.off_bn_mean;
    skip_hi = s.off_bn_var + kGcnChannels;
  }
  ts->step += 1;
  ts->last_lr = lr;
  const double b1 = 0.9, b2 = 0.999;
  const float lr_t = ts->kind == 0 ? static_cast<float>(lr * std::sqrt(1.0 - std::pow(b2, double(ts->step))) / (1.0 - std::pow(b1, double(ts->step)))) : lr;
  mark_begin("adam", st);
  adam_kernel<<<static_cast<unsigned>((np + 255) / 256), 256, 0, st>>>(d_params, rvec, ts->mom, ts->vel, np, lr_t, 0.9f, 0.999f, 1e-7f, skip_lo,
                                                                       skip_hi, ts->stats, 0.99f, ts->kind);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);
  for (int l = 0; l < s.n_lstm; ++l) {
    const int in = s.lstm_in[l], h = s.lstm_hid[l], kpad = (in + h + 15) / 16 * 16;
    const long long n = static_cast<long long>(in + h + 1) * h * 4;
    repack_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(d_params + s.off_lstm_w[l], d_params + s.off_lstm_u[l],
                                                                          d_params + s.off_lstm_b[l], in, h, d_packed + packed_off[l],
                                                                          d_packed_t + packed_t_off[l], kpad);
    GNNQC_CUDA(cudaGetLastError());
    // the tensor-core operand tiles too, so neither the next training step nor the next scoring call goes through the host
    tc_repack_device(d_params + s.off_lstm_w[l], d_params + s.off_lstm_u[l], d_params + s.off_lstm_b[l], in, h, d_packed_tc + packed_tc_off[l], st);
    tc_bwd_repack_device(d_params + s.off_lstm_w[l], d_params + s.off_lstm_u[l], in, h, d_packed_tcb + packed_tcb_off[l], st);
  }
  params_dirty = true;
  tc_stale = false;
  folded_valid = false;
}

}  // namespace gnnqc
