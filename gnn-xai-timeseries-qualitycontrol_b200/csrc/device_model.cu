#include "device_model.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "frontend.cuh"
#include "lstm_fwd.cuh"

namespace gnnqc {

// ---------------------------------------------------------------------------------------------
StackPlan::StackPlan(const ModelSpec& s, int T_) : T(T_) {
  const int n_pairs = (s.n_lstm - 1) / 2;  // LSTM pairs, each followed by MaxPooling1D(3)
  int t_last = T;
  for (int p = 0; p < n_pairs; ++p) t_last /= 3;
  if (t_last < 1) throw std::invalid_argument("sequence length " + std::to_string(T) + " is too short for this model's pooling stack");
  // Walk back up: the pair feeding a stage with n steps must run 3n steps.
  int need = t_last;
  t_run[s.n_lstm - 1] = need;
  t_out[s.n_lstm - 1] = 1;
  for (int p = n_pairs - 1; p >= 0; --p) {
    need *= 3;
    t_run[2 * p] = t_run[2 * p + 1] = need;
    t_out[2 * p] = need;          // full sequence to its partner
    t_out[2 * p + 1] = need / 3;  // pooled
  }
  // Ping/pong: every launch reads one buffer and writes the other.  Which layers land in which buffer depends on the
  // schedule (fused layer pairs skip a hand-over), so both are sized for the largest tensor that can land in them.
  size_t b = 0;
  for (int l = 0; l < s.n_lstm; ++l) b = std::max(b, size_t(t_out[l]) * s.lstm_hid[l]);
  const size_t a = std::max(b, size_t(t_run[0]) * s.lstm_in[0]);
  bytes_a = a * kTile * sizeof(float);
  bytes_b = b * kTile * sizeof(float);
}

// ---------------------------------------------------------------------------------------------
void DeviceModel::mark_begin(const char* name, cudaStream_t st) {
  ++launches;
  if (!profiling) return;
  auto get = [&]() {
    if (!event_pool.empty()) { cudaEvent_t e = event_pool.back(); event_pool.pop_back(); return e; }
    cudaEvent_t e;
    GNNQC_CUDA(cudaEventCreate(&e));
    return e;
  };
  LaunchRecord r{name, get(), get()};
  GNNQC_CUDA(cudaEventRecord(r.start, st));
  records.push_back(r);
}

void DeviceModel::mark_end(cudaStream_t st) {
  if (!profiling) return;
  GNNQC_CUDA(cudaEventRecord(records.back().stop, st));
}

void DeviceModel::profile_reset() {
  for (LaunchRecord& r : records) { event_pool.push_back(r.start); event_pool.push_back(r.stop); }
  records.clear();
}

DeviceModel::~DeviceModel() {
  if (device >= 0) {
    cudaSetDevice(device);
    free_train_state();
    free_host_stage(host_stage);
    profile_reset();
    for (cudaEvent_t e : event_pool) cudaEventDestroy(e);
    cudaFree(d_params);
    cudaFree(d_packed);
    cudaFree(d_packed_t);
    cudaFree(d_packed_tc);
    cudaFree(d_packed_tcb);
    cudaFree(d_folded);
    cudaFree(d_scratch);
  }
}

void DeviceModel::upload() {
  if (device < 0) return;
  const ModelSpec& s = host.spec;
  GNNQC_CUDA(cudaSetDevice(device));
  if (!d_params) GNNQC_CUDA(cudaMalloc(&d_params, size_t(s.n_params) * sizeof(float)));
  GNNQC_CUDA(cudaMemcpy(d_params, host.params.data(), size_t(s.n_params) * sizeof(float), cudaMemcpyHostToDevice));
  // Packed per-layer copy: wpk[k][unit][gate], k < in: kernel rows, then recurrent rows, then bias.
  size_t total = 0;
  for (int l = 0; l < s.n_lstm; ++l) {
    packed_off[l] = total;
    total += size_t(s.lstm_in[l] + s.lstm_hid[l] + 1) * s.lstm_hid[l] * 4;
  }
  std::vector<float> packed(total);
  for (int l = 0; l < s.n_lstm; ++l) {
    const int in = s.lstm_in[l], h = s.lstm_hid[l];
    const float* W = host.params.data() + s.off_lstm_w[l];
    const float* U = host.params.data() + s.off_lstm_u[l];
    const float* bias = host.params.data() + s.off_lstm_b[l];
    float* out = packed.data() + packed_off[l];
    for (int k = 0; k < in + h; ++k) {
      const float* row = (k < in) ? W + size_t(k) * 4 * h : U + size_t(k - in) * 4 * h;
      for (int j = 0; j < h; ++j)
        for (int g = 0; g < 4; ++g) out[(size_t(k) * h + j) * 4 + g] = row[g * h + j];
    }
    float* bo = out + size_t(in + h) * h * 4;
    for (int j = 0; j < h; ++j)
      for (int g = 0; g < 4; ++g) bo[j * 4 + g] = bias[g * h + j];
  }
  if (!d_packed) GNNQC_CUDA(cudaMalloc(&d_packed, total * sizeof(float)));
  GNNQC_CUDA(cudaMemcpy(d_packed, packed.data(), total * sizeof(float), cudaMemcpyHostToDevice));

  // Transposed copy for the backward contraction: wT[col][Kpad], col = gate*H + unit (Keras
  // column order), row k < in: kernel, then recurrent, zero padded to a multiple of 16 rows.
  size_t total_t = 0;
  for (int l = 0; l < s.n_lstm; ++l) {
    packed_t_off[l] = total_t;
    const int kpad = (s.lstm_in[l] + s.lstm_hid[l] + 15) / 16 * 16;
    total_t += size_t(4) * s.lstm_hid[l] * kpad;
  }
  std::vector<float> packed_t(total_t, 0.f);
  for (int l = 0; l < s.n_lstm; ++l) {
    const int in = s.lstm_in[l], h = s.lstm_hid[l], kpad = (in + h + 15) / 16 * 16;
    const float* W = host.params.data() + s.off_lstm_w[l];
    const float* U = host.params.data() + s.off_lstm_u[l];
    float* out = packed_t.data() + packed_t_off[l];
    for (int col = 0; col < 4 * h; ++col)
      for (int k = 0; k < in + h; ++k)
        out[size_t(col) * kpad + k] = (k < in) ? W[size_t(k) * 4 * h + col] : U[size_t(k - in) * 4 * h + col];
  }
  if (!d_packed_t) GNNQC_CUDA(cudaMalloc(&d_packed_t, total_t * sizeof(float)));
  GNNQC_CUDA(cudaMemcpy(d_packed_t, packed_t.data(), total_t * sizeof(float), cudaMemcpyHostToDevice));

  // bf16 hi/lo operand tiles for the layers that run on the tensor cores.
  size_t total_tc = 0;
  for (int l = 0; l < s.n_lstm; ++l) {
    packed_tc_off[l] = total_tc;
    total_tc += (tc_packed_bytes(s.lstm_in[l], s.lstm_hid[l]) + 255) / 256 * 256;
  }
  std::vector<unsigned char> packed_tc(total_tc, 0);
  for (int l = 0; l < s.n_lstm; ++l)
    tc_pack_host(host.params.data() + s.off_lstm_w[l], host.params.data() + s.off_lstm_u[l], host.params.data() + s.off_lstm_b[l],
                 s.lstm_in[l], s.lstm_hid[l], packed_tc.data() + packed_tc_off[l]);
  if (!d_packed_tc) GNNQC_CUDA(cudaMalloc(&d_packed_tc, total_tc));
  GNNQC_CUDA(cudaMemcpy(d_packed_tc, packed_tc.data(), total_tc, cudaMemcpyHostToDevice));
  size_t total_tcb = 0;
  for (int l = 0; l < s.n_lstm; ++l) {
    packed_tcb_off[l] = total_tcb;
    total_tcb += (tc_bwd_packed_bytes(s.lstm_in[l], s.lstm_hid[l]) + 255) / 256 * 256;
  }
  std::vector<unsigned char> packed_tcb(total_tcb, 0);
  for (int l = 0; l < s.n_lstm; ++l)
    tc_bwd_pack_host(host.params.data() + s.off_lstm_w[l], host.params.data() + s.off_lstm_u[l], s.lstm_in[l], s.lstm_hid[l],
                     packed_tcb.data() + packed_tcb_off[l]);
  if (!d_packed_tcb) GNNQC_CUDA(cudaMalloc(&d_packed_tcb, total_tcb));
  GNNQC_CUDA(cudaMemcpy(d_packed_tcb, packed_tcb.data(), total_tcb, cudaMemcpyHostToDevice));
  tc_stale = false;
  folded_valid = false;
}

void DeviceModel::refresh_tc() {
  if (!tc_stale) return;
  sync_host_params();
  upload();
}

void* DeviceModel::scratch(size_t bytes) {
  if (bytes > scratch_bytes) {
    GNNQC_CUDA(cudaSetDevice(device));
    if (d_scratch) {
      GNNQC_CUDA(cudaDeviceSynchronize());
      GNNQC_CUDA(cudaFree(d_scratch));
      d_scratch = nullptr;
      scratch_bytes = 0;
    }
    GNNQC_CUDA(cudaMalloc(&d_scratch, bytes));
    scratch_bytes = bytes;
  }
  return d_scratch;
}

// ---------------------------------------------------------------------------------------------
// Per-layer tiling choice: RJ units x RB windows per thread, LJ lanes along the unit axis.
namespace {

template <int I, int H, bool POOL, bool LAST, int VAR>
struct LayerLaunch {
  static constexpr bool WGLOBAL = (size_t(I + H) * H * 16 > 160 * 1024);
  static constexpr int RJ = 2, RB = 8, LJ = (VAR & 8) ? 4 : 8, NST = 2;
  using Cfg = LstmCfg<I, H, RJ, RB, LJ, WGLOBAL, NST>;
  static auto kernel() { return lstm_layer_kernel<I, H, RJ, RB, LJ, POOL, LAST, WGLOBAL, (VAR & 7), NST>; }
  static void configure() {
    GNNQC_CUDA(cudaFuncSetAttribute(kernel(), cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES)));
  }
  static void launch(const float* x, float* y, const float* wpk, int T, int t_in, int t_out, long long n_tiles,
                     int sm_count, cudaStream_t st) {
    int per_sm = 1;
    GNNQC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel(), Cfg::NT, Cfg::SMEM_BYTES));
    if (per_sm < 1) throw CudaError("lstm_layer_kernel does not fit on an SM");
    long long grid = std::min<long long>(n_tiles, static_cast<long long>(sm_count) * per_sm);
    kernel()<<<static_cast<unsigned>(grid), Cfg::NT, Cfg::SMEM_BYTES, st>>>(x, y, wpk, T, t_in, t_out, static_cast<int>(n_tiles), nullptr, nullptr);
    GNNQC_CUDA(cudaGetLastError());
  }
};

// Kernel variant (bit 0: packed FFMA2 accumulation, bit 1: merged-reciprocal cell update,
// bit 2: software-pipelined k-loop, bit 3: 4x8 instead of 8x4 lane layout).
// GNNQC_LSTM_VARIANT (0, 3, 7, 11, 15) overrides the default for A/B measurements.
int lstm_variant() {
  static const int v = [] {
    const char* e = std::getenv("GNNQC_LSTM_VARIANT");
    return e ? (std::atoi(e) & 15) : -1;
  }();
  return v;
}

template <bool CONFIGURE_ONLY, int I, int H, bool POOL, bool LAST>
void layer_op(const float* x, float* y, const float* wpk, int T, int t_in, int t_out, long long n_tiles, int sm_count,
              cudaStream_t st) {
  if (CONFIGURE_ONLY) {
    LayerLaunch<I, H, POOL, LAST, 0>::configure();
    LayerLaunch<I, H, POOL, LAST, 3>::configure();
    LayerLaunch<I, H, POOL, LAST, 7>::configure();
    LayerLaunch<I, H, POOL, LAST, 11>::configure();
    LayerLaunch<I, H, POOL, LAST, 15>::configure();
    return;
  }
  // Default (no override): measured best per role on B200 - pooled layers like the software-
  // pipelined loop (variant 7), the others the 4x8 lane layout (variant 11); profiles/README.md.
  const int variant = lstm_variant() >= 0 ? lstm_variant() : (POOL ? 7 : 11);
  switch (variant) {
    case 0: LayerLaunch<I, H, POOL, LAST, 0>::launch(x, y, wpk, T, t_in, t_out, n_tiles, sm_count, st); break;
    case 7: LayerLaunch<I, H, POOL, LAST, 7>::launch(x, y, wpk, T, t_in, t_out, n_tiles, sm_count, st); break;
    case 11: LayerLaunch<I, H, POOL, LAST, 11>::launch(x, y, wpk, T, t_in, t_out, n_tiles, sm_count, st); break;
    case 15: LayerLaunch<I, H, POOL, LAST, 15>::launch(x, y, wpk, T, t_in, t_out, n_tiles, sm_count, st); break;
    default: LayerLaunch<I, H, POOL, LAST, 3>::launch(x, y, wpk, T, t_in, t_out, n_tiles, sm_count, st); break;
  }
}

// Dispatch on the (input width, hidden width, role) combinations the reference's models use.
template <bool CFG>
void layer_dispatch(int in, int hid, bool pool, bool last, const float* x, float* y, const float* wpk, int T, int t_in,
                    int t_out, long long n_tiles, int sm_count, cudaStream_t st) {
#define GNNQC_CASE(I_, H_, POOL_, LAST_)                                                            \
  if (in == I_ && hid == H_ && pool == POOL_ && last == LAST_) {                                    \
    layer_op<CFG, I_, H_, POOL_, LAST_>(x, y, wpk, T, t_in, t_out, n_tiles, sm_count, st);          \
    return;                                                                                         \
  }
  GNNQC_CASE(2, 16, false, false)    // model_cml_baseline time1
  GNNQC_CASE(3, 16, false, false)    // model_soilnet_baseline time1
  GNNQC_CASE(18, 16, false, false)   // model_cml time1
  GNNQC_CASE(19, 16, false, false)   // model_soilnet time1
  GNNQC_CASE(16, 16, true, false)    // time2 + pool
  GNNQC_CASE(16, 32, false, false)   // time_layers[0]
  GNNQC_CASE(32, 32, true, false)    // time_layers[1] + pool
  GNNQC_CASE(32, 64, false, false)   // time_layers[2]
  GNNQC_CASE(64, 64, true, false)    // time_layers[3] + pool
  GNNQC_CASE(64, 128, false, true)   // time4 (last state only)
#undef GNNQC_CASE
  throw std::invalid_argument("no kernel for an LSTM layer with input width " + std::to_string(in) + " and " +
                              std::to_string(hid) + " units (supported: the filter_1_size=16, n_stacks=2 stack)");
}

}  // namespace

void DeviceModel::configure_kernels() {
  const ModelSpec& s = host.spec;
  for (int l = 0; l < s.n_lstm; ++l) {
    const bool last = (l == s.n_lstm - 1);
    const bool pool = !last && (l % 2 == 1);
    layer_dispatch<true>(s.lstm_in[l], s.lstm_hid[l], pool, last, nullptr, nullptr, nullptr, 0, 0, 0, 0, 0, nullptr);
    if (tc_layer_supported(s.lstm_in[l], s.lstm_hid[l], pool, last)) tc_configure(s.lstm_in[l], s.lstm_hid[l], pool, last);
  }
  for (int l = 0; l + 2 < s.n_lstm; l += 2)
    if (tc_pair_supported(s.lstm_in[l], s.lstm_hid[l], s.lstm_in[l + 1], s.lstm_hid[l + 1])) tc_pair_configure(s.lstm_in[l], s.lstm_hid[l]);
  configure_cml_frontend();
  GNNQC_CUDA(cudaGetLastError());
  if (s.lstm_hid[s.n_lstm - 1] != 128 || s.dense_units != 64)
    throw std::invalid_argument("head kernel supports a 128 -> 64 -> 64 -> 1 head only");
}

void DeviceModel::run_stack(const StackPlan& plan, float* buf_a, float* buf_b, long long n_tiles, long long n_seq,
                            float* scores, cudaStream_t st) {
  const ModelSpec& s = host.spec;
  const float* in = buf_a;
  float* bufs[2] = {buf_a, buf_b};
  int cur = 0;                       // `in` lives in bufs[cur]; every layer writes the other buffer
  int t_in = plan.t_run[0];
  static const char* const kNames[kMaxLstmLayers] = {"lstm_l1", "lstm_l2", "lstm_l3", "lstm_l4", "lstm_l5", "lstm_l6", "lstm_l7"};
  static const char* const kPairNames[kMaxLstmLayers] = {"lstm_l1_l2", "", "lstm_l3_l4", "", "lstm_l5_l6", "", ""};
  for (int l = 0; l < s.n_lstm;) {
    const bool last = (l == s.n_lstm - 1);
    const bool pool = !last && (l % 2 == 1);
    float* out = bufs[cur ^ 1];
    // An (un-pooled, pooled) pair of equally wide layers as one software-pipelined kernel when the shapes allow it
    // (lstm_tc_pair.cuh); the first layer's output then never leaves the SM.
    if (l % 2 == 0 && l + 2 < s.n_lstm && plan.t_run[l] == plan.t_run[l + 1] &&
        tc_pair_supported(s.lstm_in[l], s.lstm_hid[l], s.lstm_in[l + 1], s.lstm_hid[l + 1])) {
      mark_begin(kPairNames[l], st);
      tc_pair_launch(s.lstm_in[l], s.lstm_hid[l], in, out, d_packed_tc + packed_tc_off[l], d_packed_tc + packed_tc_off[l + 1], plan.t_run[l], t_in,
                     plan.t_out[l + 1], n_tiles, sm_count, st);
      mark_end(st);
      in = out;
      cur ^= 1;
      t_in = plan.t_out[l + 1];
      l += 2;
      continue;
    }
    mark_begin(kNames[l], st);
    if (tc_layer_supported(s.lstm_in[l], s.lstm_hid[l], pool, last))
      tc_launch(s.lstm_in[l], s.lstm_hid[l], pool, last, in, out, d_packed_tc + packed_tc_off[l], plan.t_run[l], t_in, plan.t_out[l], n_tiles,
                sm_count, st);
    else
      layer_dispatch<false>(s.lstm_in[l], s.lstm_hid[l], pool, last, in, out, d_packed + packed_off[l], plan.t_run[l], t_in,
                            plan.t_out[l], n_tiles, sm_count, st);
    mark_end(st);
    in = out;
    cur ^= 1;
    t_in = plan.t_out[l];
    ++l;
  }
  const float* p = d_params;
  mark_begin("head", st);
  head_kernel<128, 64><<<static_cast<unsigned>(n_tiles), 128, 0, st>>>(in, p + s.off_d1_w, p + s.off_d1_b, p + s.off_d2_w,
                                                                       p + s.off_d2_b, p + s.off_out_w, p + s.off_out_b,
                                                                       s.leaky_alpha, scores, n_seq);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);
}

static size_t scratch_budget_bytes() {
  const char* env = std::getenv("GNNQC_SCRATCH_MB");
  size_t mb = env ? std::strtoull(env, nullptr, 10) : 8192;
  return std::max<size_t>(mb, 16) << 20;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

void DeviceModel::score_seq(const float* x, long long n_seq, int T, float* scores, cudaStream_t st, const int* node_off, int n_windows) {
  if (n_seq <= 0) return;
  const ModelSpec& s = host.spec;
  GNNQC_CUDA(cudaSetDevice(device));
  StackPlan plan(s, T);
  const long long n_tiles = (n_seq + kTile - 1) / kTile;
  const size_t sa = align_up(plan.bytes_a, 256), sb = align_up(plan.bytes_b, 256);
  const long long chunk = balanced_chunk(n_tiles, static_cast<long long>(scratch_budget_bytes() / (sa + sb)));
  char* base = static_cast<char*>(scratch(size_t(chunk) * (sa + sb)));
  float* buf_a = reinterpret_cast<float*>(base);
  float* buf_b = reinterpret_cast<float*>(base + size_t(chunk) * sa);
  const int c_in = s.lstm_in[0];
  for (long long t0 = 0; t0 < n_tiles; t0 += chunk) {
    const long long nt = std::min(chunk, n_tiles - t0);
    const long long seq0 = t0 * kTile, ns = std::min<long long>(n_seq - seq0, nt * kTile);
    dim3 grid(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
    const float* xs = x + seq0 * T * c_in;
    mark_begin("seq_to_tiles", st);
    if (node_off) {   // rows in (window, t, node) order: graph_reshape happens in the tile builder
      if (c_in == 3) rows_to_tiles_kernel<3><<<grid, 256, 0, st>>>(x, node_off, n_windows, buf_a, seq0, n_seq, T, plan.t_run[0]);
      else if (c_in == 2) rows_to_tiles_kernel<2><<<grid, 256, 0, st>>>(x, node_off, n_windows, buf_a, seq0, n_seq, T, plan.t_run[0]);
      else throw std::invalid_argument("row-layout scoring is for the baseline models (2 or 3 input channels)");
    } else if (c_in == 2) seq_to_tiles_kernel<2><<<grid, 256, 0, st>>>(xs, buf_a, ns, T, plan.t_run[0]);
    else if (c_in == 3) seq_to_tiles_kernel<3><<<grid, 256, 0, st>>>(xs, buf_a, ns, T, plan.t_run[0]);
    else if (c_in == 18) seq_to_tiles_kernel<18><<<grid, 256, 0, st>>>(xs, buf_a, ns, T, plan.t_run[0]);
    else if (c_in == 19) seq_to_tiles_kernel<19><<<grid, 256, 0, st>>>(xs, buf_a, ns, T, plan.t_run[0]);
    else throw std::invalid_argument("unsupported sequence width");
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    run_stack(plan, buf_a, buf_b, nt, ns, scores + seq0, st);
  }
}

// Same arithmetic on the host (parameters loaded or set from the host) and on the device (after an optimizer step).
struct FoldOffsets { long long kernel, bias, prelu, gamma, beta, mean, var; int n_features; float eps; };
__host__ __device__ inline void fold_channel(const float* p, const FoldOffsets& o, int c, GcnFolded* f) {
  const float scale = p[o.gamma + c] / sqrtf(p[o.var + c] + o.eps);
  for (int k = 0; k < 3; ++k) f->a[k][c] = k < o.n_features ? p[o.kernel + k * kGcnChannels + c] * scale : 0.f;
  f->c[c] = fmaf(p[o.bias + c] - p[o.mean + c], scale, p[o.beta + c]);   // explicit fma: the same bits on the host and on the device
  f->alpha[c] = p[o.prelu + c];
}
static __global__ void gcn_fold_kernel(const float* __restrict__ params, FoldOffsets o, GcnFolded* __restrict__ out) {
  if (threadIdx.x < kGcnChannels) fold_channel(params, o, threadIdx.x, out);
}

const GcnFolded& DeviceModel::folded(cudaStream_t st) {
  if (folded_valid) return folded_cache;
  const ModelSpec& s = host.spec;
  const FoldOffsets o{s.off_gcn_kernel, s.off_gcn_bias, s.off_prelu, s.off_bn_gamma, s.off_bn_beta, s.off_bn_mean, s.off_bn_var, s.n_features, s.bn_eps};
  if (params_dirty && device >= 0) {
    // the device copy is authoritative (train_apply): fold there, 320 bytes back, once per training -> scoring transition
    if (!d_folded) GNNQC_CUDA(cudaMalloc(&d_folded, sizeof(GcnFolded)));
    gcn_fold_kernel<<<1, 32, 0, st>>>(d_params, o, d_folded);
    GNNQC_CUDA(cudaGetLastError());
    GNNQC_CUDA(cudaMemcpyAsync(&folded_cache, d_folded, sizeof(GcnFolded), cudaMemcpyDeviceToHost, st));
    GNNQC_CUDA(cudaStreamSynchronize(st));
  } else {
    for (int c = 0; c < kGcnChannels; ++c) fold_channel(host.params.data(), o, c, &folded_cache);
  }
  folded_valid = true;
  return folded_cache;
}

void DeviceModel::score_cml(const float* feat, const float* anom, const long long* edges, const int* node_off,
                            const long long* edge_off, long long B, int T, long long n_nodes, float* scores,
                            cudaStream_t st) {
  if (B <= 0) return;
  const ModelSpec& s = host.spec;
  GNNQC_CUDA(cudaSetDevice(device));
  StackPlan plan(s, T);
  const long long n_tiles = (B + kTile - 1) / kTile;
  const size_t sa = align_up(plan.bytes_a, 256), sb = align_up(plan.bytes_b, 256);
  const size_t coef_bytes = align_up(size_t(n_nodes) * sizeof(float), 256);
  const long long chunk = balanced_chunk(n_tiles, static_cast<long long>(scratch_budget_bytes() / (sa + sb)));
  char* base = static_cast<char*>(scratch(2 * coef_bytes + size_t(chunk) * (sa + sb)));
  int* deg = reinterpret_cast<int*>(base);
  float* coef = reinterpret_cast<float*>(base + coef_bytes);
  float* buf_a = reinterpret_cast<float*>(base + 2 * coef_bytes);
  float* buf_b = reinterpret_cast<float*>(base + 2 * coef_bytes + size_t(chunk) * sa);

  GNNQC_CUDA(cudaMemsetAsync(deg, 0, size_t(n_nodes) * sizeof(int), st));
  {
    const long long threads = B * 32;
    mark_begin("cml_node_coef", st);
    cml_node_coef_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, st>>>(edges, node_off, edge_off, deg, coef, B);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
  }
  const GcnFolded folded = this->folded(st);
  for (long long t0 = 0; t0 < n_tiles; t0 += chunk) {
    const long long nt = std::min(chunk, n_tiles - t0);
    const long long win0 = t0 * kTile, nw = std::min<long long>(B - win0, nt * kTile);
    dim3 grid(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
    mark_begin("cml_frontend", st);
    static const bool plain_frontend = [] { const char* e = std::getenv("GNNQC_FRONTEND"); return e && std::strcmp(e, "plain") == 0; }();
    if (plain_frontend)   // A/B: direct loads, one CTA per (tile, 32 steps)
      cml_frontend_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float2*>(feat), reinterpret_cast<const float2*>(anom) + win0 * T,
                                                coef, node_off + win0, folded, buf_a, nw, T, plan.t_run[0]);
    else
      launch_cml_frontend(reinterpret_cast<const float2*>(feat), n_nodes * T, reinterpret_cast<const float2*>(anom) + win0 * T, coef,
                          node_off + win0, folded, buf_a, nw, T, plan.t_run[0], nt, 1.f, sm_count, st);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    run_stack(plan, buf_a, buf_b, nt, nw, scores + win0, st);
  }
}

}  // namespace gnnqc
