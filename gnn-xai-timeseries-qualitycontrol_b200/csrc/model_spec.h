// Host-side description of one anomaly model (GCN or LSTM-baseline) and its weights.
//
// Mirrors what the reference builds in libs/create_model.py:
//   GCNClassifier      (:140-240)  GeneralConv -> pool / graph_reshape -> TimeLayer -> 3x Dense
//   BaselineClassifier (:261-377)  the same TimeLayer + Dense head straight on the series
//   TimeLayer          (:43-123)   LSTM,LSTM,pool3,(LSTM,LSTM,pool3) x n_stacks, LSTM(last state)
// Weights come from the Keras SavedModel checkpoint (variables/variables.*), resolved through
// the checkpoint's object graph by attribute path (gcn_layer/kernel, time_layer/time1/cell/...).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace gnnqc {

constexpr int kGcnChannels = 16;   // graph_convolution.units in every shipped model
constexpr int kMaxLstmLayers = 7;  // 2 + 2*n_stacks + 1 with n_stacks = 2
constexpr int kTile = 64;          // sequences per device-side tile (see lstm_fwd.cuh)

// BatchNorm (inference) folded into the GeneralConv projection: z = prelu(x0*a0 + x1*a1 (+ x2*a2) + c).
struct GcnFolded {
  float a[3][kGcnChannels];
  float c[kGcnChannels];
  float alpha[kGcnChannels];
};

struct ModelSpec {
  bool has_gcn = false;     // GCNClassifier vs BaselineClassifier
  bool soilnet = false;     // ds_type: 'soilnet' (F=3, every node is a sequence) or 'cml' (F=2)
  int n_features = 2;       // F
  int n_lstm = 7;
  int lstm_in[kMaxLstmLayers] = {0};
  int lstm_hid[kMaxLstmLayers] = {0};
  int dense_units = 64;
  float leaky_alpha = 0.3f;  // LeakyReLU(alpha) after dense and dense2 (create_model.py:210,212)
  float bn_eps = 1e-3f;      // Keras BatchNormalization default, recorded in keras_metadata.pb

  // Flat parameter vector layout (floats). GCN block first (if any), then per LSTM layer
  // kernel[in,4H], recurrent_kernel[H,4H], bias[4H] (Keras gate order i,f,c,o), then the head.
  int64_t off_gcn_kernel = -1, off_gcn_bias = -1, off_prelu = -1;
  int64_t off_bn_gamma = -1, off_bn_beta = -1, off_bn_mean = -1, off_bn_var = -1;
  int64_t off_lstm_w[kMaxLstmLayers] = {0}, off_lstm_u[kMaxLstmLayers] = {0}, off_lstm_b[kMaxLstmLayers] = {0};
  int64_t off_d1_w = 0, off_d1_b = 0, off_d2_w = 0, off_d2_b = 0, off_out_w = 0, off_out_b = 0;
  int64_t n_params = 0;      // total floats, BN moving statistics included
  int64_t n_trainable = 0;   // without bn_mean / bn_var

  // Fills lstm_in/lstm_hid for the reference's TimeLayer given filter_1_size and n_stacks,
  // then assigns all offsets.
  void layout(bool gcn, bool soil, int features, int filter_1_size, int n_stacks, int dense);
};

struct HostModel {
  ModelSpec spec;
  std::vector<float> params;
  std::vector<int32_t> model_info;  // [timestep_before, timestep_after, batch_size(, freq)]
  std::string ds_type;              // "cml" | "soilnet"
  std::string normalization;        // "rolling_median" | "scale_range" | ...
  // Optimizer state as model.save keeps it (Keras 2.11 optimizer: optimizer/_iterations, optimizer/_learning_rate,
  // optimizer/_variables/1.. = the slots of every trainable variable in order: Adam m, v interleaved; RMSprop one
  // velocity each; plain SGD none).  Flat, in parameter-vector layout (moving statistics unused); empty = none saved.
  int optimizer = 0;                // 0 adam | 1 sgd | 2 rmsprop  (libs/fit_model.py:71-74)
  long long opt_iterations = 0;
  float opt_learning_rate = 0.f;
  std::vector<float> opt_m, opt_v;
};

// `dir` is a SavedModel directory (containing variables/variables.index) or the
// variables prefix itself. Throws std::runtime_error.
HostModel load_saved_model(const std::string& dir);

// Writes variables/variables.{index,data-00000-of-00001} under `dir` with the same keys and
// object-graph paths the reference checkpoints use, so the file round-trips through
// load_saved_model (and is shaped for Keras' object-based restore).
void save_saved_model(const HostModel& m, const std::string& dir);

// Name of each flat-vector segment, in order, e.g. "time_layer/time1/cell/kernel".
struct ParamSegment {
  std::string path;
  int64_t offset;
  std::vector<int64_t> shape;
  bool trainable;
};
std::vector<ParamSegment> param_segments(const ModelSpec& s);
std::vector<ParamSegment> checkpoint_order(const ModelSpec& s);   // the order of `variables/N` in the reference's checkpoints

}  // namespace gnnqc
