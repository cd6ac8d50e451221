// Device-resident model: uploaded weights (flat + per-layer packed copies), scratch arena,
// and the host-side launch plan of the scoring pipeline.
#pragma once
#include <cuda_runtime.h>

#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "model_spec.h"

namespace gnnqc {

struct CudaError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

#define GNNQC_CUDA(expr)                                                                        \
  do {                                                                                          \
    cudaError_t err__ = (expr);                                                                 \
    if (err__ != cudaSuccess)                                                                   \
      throw ::gnnqc::CudaError(std::string(#expr) + ": " + cudaGetErrorString(err__));          \
  } while (0)

// Number of steps each stage of the TimeLayer actually has to run for a T-step input.
// MaxPooling1D(3, 'valid') floors, and only the last state of the final LSTM is used
// (libs/create_model.py:68,76,78-79), so steps past 27*T7 (first pair), 9*T7 (second pair)
// and 3*T7 (third pair) cannot influence the score and are not computed.
struct StackPlan {
  int T = 0;
  int t_run[kMaxLstmLayers] = {0};  // steps executed per layer
  int t_out[kMaxLstmLayers] = {0};  // time extent of each layer's stored output
  size_t bytes_a = 0, bytes_b = 0;  // per-tile ping/pong buffer sizes
  explicit StackPlan(const ModelSpec& s, int T);
  size_t bytes_per_tile() const { return bytes_a + bytes_b; }
};

// Optional per-launch timing (CUDA events on the launching stream) + an always-on launch counter.
struct LaunchRecord {
  const char* name;
  cudaEvent_t start, stop;
};

// Tiles per launch when a batch does not fit the scratch budget: as few chunks as needed, all the same size
// (a short last chunk would leave most SMs idle for a full sequence length).
inline long long balanced_chunk(long long n_tiles, long long max_tiles) {
  if (max_tiles < 1) max_tiles = 1;
  if (n_tiles <= max_tiles) return n_tiles < 1 ? 1 : n_tiles;
  const long long parts = (n_tiles + max_tiles - 1) / max_tiles;
  return (n_tiles + parts - 1) / parts;
}

struct TrainState;
struct IgGroup;     // a group of interpolation points for the front-end backward (csrc/ig_frontend_bwd.cuh)
struct HostStage;   // staging slots and streams of gnnqc_score_cml_host (csrc/host_scoring.cu)
void free_host_stage(HostStage* h);

// Tensor-core LSTM layers (csrc/lstm_tc.cu).
bool tc_layer_supported(int in, int hid, bool pool, bool last);
size_t tc_packed_bytes(int in, int hid);
void tc_pack_host(const float* W, const float* U, const float* bias, int in, int hid, unsigned char* out);
void tc_configure(int in, int hid, bool pool, bool last);
void tc_launch(int in, int hid, bool pool, bool last, const float* x, float* y, const unsigned char* w, int T, int t_in, int t_out, long long n_tiles,
               int sm_count, cudaStream_t st, float* save = nullptr, float* amax = nullptr, int nsave = 5);   // save != nullptr: also store gates (IG / training forward)
bool tc_pair_supported(int in0, int hid0, int in1, int hid1);
void tc_pair_configure(int in0, int hid0);
void tc_pair_launch(int in0, int hid0, const float* x, float* y, const unsigned char* w1, const unsigned char* w2, int T, int t_in, int t_out,
                    long long n_tiles, int sm_count, cudaStream_t st);
void tc_repack_device(const float* W, const float* U, const float* bias, int in, int hid, unsigned char* out, cudaStream_t st);
// Tensor-core LSTM backward (csrc/lstm_tc_bwd.cu).
bool tc_bwd_supported(int in, int hid, bool pool, bool last);
size_t tc_bwd_packed_bytes(int in, int hid);
void tc_bwd_pack_host(const float* W, const float* U, int in, int hid, unsigned char* out);
void tc_bwd_configure(int in, int hid, bool pool, bool last);
void tc_bwd_launch(int in, int hid, bool pool, bool last, const float* dy, const float* save, const float* amax, const unsigned char* w,
                   float* dx, int T, long long n_tiles, int sm_count, cudaStream_t st, int nsave = 5, float* dz_out = nullptr);
void tc_bwd_repack_device(const float* W, const float* U, int in, int hid, unsigned char* out, cudaStream_t st);
// optimizer slots and per-step BN state (csrc/train.cu)

struct DeviceModel {
  HostModel host;
  int device = -1;              // < 0: host-only handle
  int sm_count = 0;
  float* d_params = nullptr;    // flat vector, layout of ModelSpec
  float* d_packed = nullptr;    // per-layer [k][unit][gate] + bias[unit][gate]
  size_t packed_off[kMaxLstmLayers] = {0};
  float* d_packed_t = nullptr;  // per-layer transposed copy wT[col][Kpad] for the backward kernels
  size_t packed_t_off[kMaxLstmLayers] = {0};
  unsigned char* d_packed_tc = nullptr;   // bf16 hi/lo operand tiles for the tensor-core layers
  size_t packed_tc_off[kMaxLstmLayers] = {0};
  unsigned char* d_packed_tcb = nullptr;  // [W | U]^T operand pieces for the tensor-core backward
  size_t packed_tcb_off[kMaxLstmLayers] = {0};
  bool tc_stale = false;                   // set by train_apply; refreshed lazily before scoring
  void refresh_tc();
  void* d_scratch = nullptr;
  size_t scratch_bytes = 0;

  long long launches = 0;       // kernels enqueued by this handle since creation
  bool profiling = false;
  std::vector<LaunchRecord> records;
  std::vector<cudaEvent_t> event_pool;
  void mark_begin(const char* name, cudaStream_t st);
  void mark_end(cudaStream_t st);
  void profile_reset();

  ~DeviceModel();
  void upload();                // host.params -> d_params, d_packed
  void* scratch(size_t bytes);  // grow-only arena
  void configure_kernels();     // opt-in shared memory sizes, once per handle

  // x1: sequence tiles [n_tiles][plan.t_run[0]][lstm_in[0]][64] already in `buf_a`.
  void run_stack(const StackPlan& plan, float* buf_a, float* buf_b, long long n_tiles, long long n_seq,
                 float* scores, cudaStream_t stream);
  void score_seq(const float* x, long long n_seq, int T, float* scores, cudaStream_t stream, const int* node_off = nullptr, int n_windows = 0);
  // Training (csrc/train.cu).  After train_apply the device copy of the parameters is the
  // authoritative one; sync_host_params() brings host.params up to date on demand.
  HostStage* host_stage = nullptr;
  TrainState* train = nullptr;
  bool params_dirty = false;
  // BatchNorm (inference) folded into the GeneralConv projection, as the front-end kernels take it by value.  After a
  // training step the device copy of the parameters is authoritative: the fold then runs on the device and 320 bytes
  // come back once (not the whole parameter vector, and not on every scoring call).
  GcnFolded folded_cache;
  bool folded_valid = false;
  GcnFolded* d_folded = nullptr;
  const GcnFolded& folded(cudaStream_t stream);
  void configure_train_kernels();
  TrainState* train_state();
  void free_train_state();
  void sync_host_params();
  void sync_host_optimizer();
  long long train_iterations() const;
  void set_optimizer(int kind);      // 0 adam | 1 sgd | 2 rmsprop
  void train_moments(const float* feat, long long n_rows, double* moments, cudaStream_t stream);
  void train_grads(const float* feat, const float* anom, const long long* edges, const int* node_off, const long long* edge_off,
                   const float* x_seq, const long long* labels, long long B, int T, long long n_nodes, const double* moments,
                   float cw0, float cw1, long long global_batch, float* rvec, cudaStream_t stream, long long n_edges = 0);
  void train_apply(float* rvec, const double* moments, float lr, cudaStream_t stream);
  // SoilNet GCN (csrc/soilnet.cu): scores [n_nodes], one per (window, sensor).
  void score_soilnet(const float* feat, const long long* edges, const int* node_off, const long long* edge_off, long long B,
                     int T, long long n_nodes, long long n_edges, float* scores, cudaStream_t stream);
  // Integrated gradients for the CML GCN model (csrc/ig.cu).
  void configure_ig_kernels();
  void ig_core(long long B, int T, int m_steps, float* scores, cudaStream_t stream, size_t extra_bytes,
               const std::function<void(char*)>& prepare,
               const std::function<void(float, long long, long long, long long, float*)>& fwd,
               const std::function<void(const IgGroup&, long long, long long)>& bwd, bool scale_by_input = true);
  void ig_cml(const float* feat, const float* anom, const long long* edges, const int* node_off,
              const long long* edge_off, long long B, int T, long long n_nodes, int m_steps, float* attr_feat,
              float* attr_anom, float* scores, cudaStream_t stream, bool scale_by_input = true);
  void score_cml(const float* feat, const float* anom, const long long* edges, const int* node_off,
                 const long long* edge_off, long long B, int T, long long n_nodes, float* scores,
                 cudaStream_t stream);
};

}  // namespace gnnqc
