#include "model_spec.h"

#include <filesystem>
#include <stdexcept>
#include <sys/stat.h>

#include "tf_bundle.h"

namespace gnnqc {
namespace {

bool file_exists(const std::string& p) {
  struct stat st;
  return ::stat(p.c_str(), &st) == 0;
}

// Attribute path of LSTM layer `l` (execution order) below the object that owns the stack.
std::string lstm_attr(const ModelSpec& s, int l) {
  if (l == 0) return "time1";
  if (l == 1) return "time2";
  if (l == s.n_lstm - 1) return "time4";
  return "time_layers/" + std::to_string(l - 2);
}

const char* kVarSuffix = "/.ATTRIBUTES/VARIABLE_VALUE";

}  // namespace

void ModelSpec::layout(bool gcn, bool soil, int features, int filter_1_size, int n_stacks, int dense) {
  has_gcn = gcn;
  soilnet = soil;
  n_features = features;
  dense_units = dense;
  n_lstm = 3 + 2 * n_stacks;
  if (n_lstm > kMaxLstmLayers) throw std::runtime_error("model_spec: n_stacks > 2 is not supported");
  // TimeLayer (create_model.py:61-79): time1/time2 at filter_1_size, stack i at
  // filter_1_size*2^(i+1) twice, time4 at filter_1_size*2^(n_stacks+1).
  int first_in = gcn ? (soil ? kGcnChannels + features : features + kGcnChannels) : features;
  int h = filter_1_size;
  lstm_in[0] = first_in; lstm_hid[0] = h;
  lstm_in[1] = h;        lstm_hid[1] = h;
  int prev = h;
  for (int i = 0; i < n_stacks; ++i) {
    int hh = filter_1_size << (i + 1);
    lstm_in[2 + 2 * i] = prev; lstm_hid[2 + 2 * i] = hh;
    lstm_in[3 + 2 * i] = hh;   lstm_hid[3 + 2 * i] = hh;
    prev = hh;
  }
  lstm_in[n_lstm - 1] = prev;
  lstm_hid[n_lstm - 1] = filter_1_size << (n_stacks + 1);

  int64_t o = 0;
  auto take = [&o](int64_t n) { int64_t at = o; o += n; return at; };
  if (gcn) {
    off_gcn_kernel = take(int64_t(features) * kGcnChannels);
    off_gcn_bias = take(kGcnChannels);
    off_prelu = take(kGcnChannels);
    off_bn_gamma = take(kGcnChannels);
    off_bn_beta = take(kGcnChannels);
    off_bn_mean = take(kGcnChannels);
    off_bn_var = take(kGcnChannels);
  }
  for (int l = 0; l < n_lstm; ++l) {
    off_lstm_w[l] = take(int64_t(lstm_in[l]) * 4 * lstm_hid[l]);
    off_lstm_u[l] = take(int64_t(lstm_hid[l]) * 4 * lstm_hid[l]);
    off_lstm_b[l] = take(4 * lstm_hid[l]);
  }
  int last = lstm_hid[n_lstm - 1];
  off_d1_w = take(int64_t(last) * dense);
  off_d1_b = take(dense);
  off_d2_w = take(int64_t(dense) * dense);
  off_d2_b = take(dense);
  off_out_w = take(dense);
  off_out_b = take(1);
  n_params = o;
  n_trainable = o - (gcn ? 2 * kGcnChannels : 0);
}

std::vector<ParamSegment> param_segments(const ModelSpec& s) {
  std::vector<ParamSegment> v;
  if (s.has_gcn) {
    v.push_back({"gcn_layer/kernel", s.off_gcn_kernel, {s.n_features, kGcnChannels}, true});
    v.push_back({"gcn_layer/bias", s.off_gcn_bias, {kGcnChannels}, true});
    v.push_back({"gcn_layer/activation/alpha", s.off_prelu, {kGcnChannels}, true});
    v.push_back({"gcn_layer/batch_norm/gamma", s.off_bn_gamma, {kGcnChannels}, true});
    v.push_back({"gcn_layer/batch_norm/beta", s.off_bn_beta, {kGcnChannels}, true});
    v.push_back({"gcn_layer/batch_norm/moving_mean", s.off_bn_mean, {kGcnChannels}, false});
    v.push_back({"gcn_layer/batch_norm/moving_variance", s.off_bn_var, {kGcnChannels}, false});
  }
  const std::string stack = s.has_gcn ? "time_layer/" : "";
  for (int l = 0; l < s.n_lstm; ++l) {
    std::string base = stack + lstm_attr(s, l) + "/cell/";
    int64_t in = s.lstm_in[l], h = s.lstm_hid[l];
    v.push_back({base + "kernel", s.off_lstm_w[l], {in, 4 * h}, true});
    v.push_back({base + "recurrent_kernel", s.off_lstm_u[l], {h, 4 * h}, true});
    v.push_back({base + "bias", s.off_lstm_b[l], {4 * h}, true});
  }
  const std::string d1 = s.has_gcn ? "dense" : "dense1";
  int64_t last = s.lstm_hid[s.n_lstm - 1], d = s.dense_units;
  v.push_back({d1 + "/kernel", s.off_d1_w, {last, d}, true});
  v.push_back({d1 + "/bias", s.off_d1_b, {d}, true});
  v.push_back({"dense2/kernel", s.off_d2_w, {d, d}, true});
  v.push_back({"dense2/bias", s.off_d2_b, {d}, true});
  v.push_back({"dense_out/kernel", s.off_out_w, {d, 1}, true});
  v.push_back({"dense_out/bias", s.off_out_b, {1}, true});
  return v;
}

HostModel load_saved_model(const std::string& dir) {
  std::string prefix = dir;
  if (file_exists(dir + "/variables/variables.index")) prefix = dir + "/variables/variables";
  else if (file_exists(dir + "/variables.index")) prefix = dir + "/variables";
  else if (!file_exists(dir + ".index"))
    throw std::runtime_error("load_saved_model: no variables/variables.index under '" + dir + "'");
  BundleReader r(prefix);
  if (r.objects().empty()) throw std::runtime_error("load_saved_model: checkpoint has no object graph");

  HostModel m;
  const bool gcn = !r.key_for_path("gcn_layer/kernel").empty();
  const std::string stack = gcn ? "time_layer/" : "";
  std::string k1 = r.key_for_path(stack + "time1/cell/kernel");
  std::string u1 = r.key_for_path(stack + "time1/cell/recurrent_kernel");
  if (k1.empty() || u1.empty())
    throw std::runtime_error("load_saved_model: not a GCNClassifier/BaselineClassifier checkpoint (time1 LSTM missing); "
                             "only sequence_layer.algorithm='lstm' models are supported");
  const BundleEntry& e_k1 = r.entry(k1);
  const BundleEntry& e_u1 = r.entry(u1);
  if (e_k1.shape.size() != 2 || e_u1.shape.size() != 2) throw std::runtime_error("load_saved_model: bad LSTM kernel rank");
  int filter = static_cast<int>(e_u1.shape[0]);
  int first_in = static_cast<int>(e_k1.shape[0]);
  int n_stacks = 0;
  while (!r.key_for_path(stack + "time_layers/" + std::to_string(2 * n_stacks) + "/cell/kernel").empty()) ++n_stacks;
  int features = first_in;
  if (gcn) {
    const BundleEntry& gk = r.entry(r.key_for_path("gcn_layer/kernel"));
    if (gk.shape.size() != 2 || gk.shape[1] != kGcnChannels)
      throw std::runtime_error("load_saved_model: GeneralConv with channels != 16 is not supported");
    features = static_cast<int>(gk.shape[0]);
  }
  std::string d1 = gcn ? "dense" : "dense1";
  const BundleEntry& e_d1 = r.entry(r.key_for_path(d1 + "/kernel"));
  int dense = static_cast<int>(e_d1.shape.at(1));

  // Metadata variables (create_model.py:159-165, :279-281).
  if (r.has(std::string("model_info") + kVarSuffix)) m.model_info = r.read_i32(std::string("model_info") + kVarSuffix);
  if (r.has(std::string("model_type") + kVarSuffix)) m.ds_type = r.read_scalar_string(std::string("model_type") + kVarSuffix);
  if (r.has(std::string("model_normalization") + kVarSuffix))
    m.normalization = r.read_scalar_string(std::string("model_normalization") + kVarSuffix);
  else if (r.has(std::string("normalization") + kVarSuffix))
    m.normalization = r.read_scalar_string(std::string("normalization") + kVarSuffix);
  if (m.ds_type.empty()) m.ds_type = (features == 3) ? "soilnet" : "cml";
  const bool soil = (m.ds_type == "soilnet");
  if ((soil && features != 3) || (!soil && features != 2))
    throw std::runtime_error("load_saved_model: feature count " + std::to_string(features) + " does not match ds_type '" + m.ds_type + "'");

  m.spec.layout(gcn, soil, features, filter, n_stacks, dense);
  if (m.spec.lstm_in[0] != first_in)
    throw std::runtime_error("load_saved_model: first LSTM input width " + std::to_string(first_in) + " unexpected");
  m.params.assign(static_cast<size_t>(m.spec.n_params), 0.f);
  for (const ParamSegment& seg : param_segments(m.spec)) {
    std::string key = r.key_for_path(seg.path);
    if (key.empty()) throw std::runtime_error("load_saved_model: variable '" + seg.path + "' missing from object graph");
    const BundleEntry& e = r.entry(key);
    if (e.shape != seg.shape) throw std::runtime_error("load_saved_model: variable '" + seg.path + "' has an unexpected shape");
    std::vector<float> v = r.read_f32(key);
    std::copy(v.begin(), v.end(), m.params.begin() + seg.offset);
  }
  // Optimizer slots (absent from checkpoints written without them, e.g. tests/golden/models): the k-th trainable variable
  // in checkpoint order owns optimizer/_variables/(2k+1) (m) and (2k+2) (v) for Adam, or /(k+1) for RMSprop.
  const std::string it_key = std::string("optimizer/_iterations") + kVarSuffix;
  if (r.has(it_key)) {
    std::vector<int64_t> it = r.read_i64(it_key);
    m.opt_iterations = it.empty() ? 0 : it[0];
    const std::string lr_key = std::string("optimizer/_learning_rate") + kVarSuffix;
    if (r.has(lr_key)) { std::vector<float> lr = r.read_f32(lr_key); m.opt_learning_rate = lr.empty() ? 0.f : lr[0]; }
    std::vector<ParamSegment> trainable;
    for (const ParamSegment& seg : checkpoint_order(m.spec))
      if (seg.trainable) trainable.push_back(seg);
    const int n_tr = static_cast<int>(trainable.size());
    auto slot_key = [](int n) { return "optimizer/_variables/" + std::to_string(n) + kVarSuffix; };
    const bool adam = r.has(slot_key(2 * n_tr)) && !r.has(slot_key(2 * n_tr + 1));
    const bool rms = !adam && r.has(slot_key(n_tr)) && !r.has(slot_key(n_tr + 1));
    if (adam || rms) {
      m.optimizer = adam ? 0 : 2;
      m.opt_v.assign(static_cast<size_t>(m.spec.n_params), 0.f);
      if (adam) m.opt_m.assign(static_cast<size_t>(m.spec.n_params), 0.f);
      for (int k = 0; k < n_tr; ++k) {
        const ParamSegment& seg = trainable[k];
        for (int which = 0; which < (adam ? 2 : 1); ++which) {
          const std::string key = slot_key(adam ? 2 * k + 1 + which : k + 1);
          if (r.entry(key).shape != seg.shape) throw std::runtime_error("load_saved_model: optimizer slot '" + key + "' does not match '" + seg.path + "'");
          std::vector<float> v = r.read_f32(key);
          std::vector<float>& dst = (adam && which == 0) ? m.opt_m : m.opt_v;
          std::copy(v.begin(), v.end(), dst.begin() + seg.offset);
        }
      }
    } else {
      m.optimizer = 1;   // iteration counter without slots: plain SGD
    }
  }
  return m;
}

// The reference checkpoints number `variables/N` as Keras lists them: GCN block, then the stacked layers (time_layers),
// then time1, time2, time4, then the head.
std::vector<ParamSegment> checkpoint_order(const ModelSpec& s) {
  std::vector<ParamSegment> segs = param_segments(s), out;
  auto add_if = [&](const std::string& needle) {
    for (const ParamSegment& seg : segs)
      if (seg.path.find(needle) != std::string::npos) out.push_back(seg);
  };
  add_if("gcn_layer/");
  add_if("time_layers/");
  add_if("time1/");
  add_if("time2/");
  add_if("time4/");
  add_if("dense");
  return out;
}

void save_saved_model(const HostModel& m, const std::string& dir) {
  std::error_code ec;
  std::filesystem::create_directories(dir + "/variables", ec);
  if (ec) throw std::runtime_error("save_saved_model: cannot create " + dir + "/variables: " + ec.message());
  const ModelSpec& s = m.spec;
  const std::vector<ParamSegment> order = checkpoint_order(s);
  BundleWriter w;
  const std::string cls = s.has_gcn ? "gcn_classifier/" : "baseline_classifier/";
  for (int n = 0; n < static_cast<int>(order.size()); ++n) {
    const ParamSegment& seg = order[n];
    std::string key = "variables/" + std::to_string(n) + kVarSuffix;
    w.add_f32(key, seg.shape, m.params.data() + seg.offset);
    w.track(seg.path, key, cls + seg.path);
  }
  // Optimizer state, when a training run produced some (model.save keeps it: SURVEY.md section 5, checkpoint row).
  if (m.opt_iterations > 0) {
    const std::string it_key = std::string("optimizer/_iterations") + kVarSuffix, lr_key = std::string("optimizer/_learning_rate") + kVarSuffix;
    const int64_t it = m.opt_iterations;
    w.add_i64(it_key, {}, &it);
    w.track("optimizer/_iterations", it_key, "iteration");
    w.add_f32(lr_key, {}, &m.opt_learning_rate);
    w.track("optimizer/_learning_rate", lr_key, "learning_rate");
    const char* const kind = m.optimizer == 0 ? "Adam" : "RMSprop";
    int n = 0;
    for (const ParamSegment& seg : order) {
      if (!seg.trainable || m.optimizer == 1) continue;
      for (int which = 0; which < (m.optimizer == 0 ? 2 : 1); ++which) {
        const std::vector<float>& src = (m.optimizer == 0 && which == 0) ? m.opt_m : m.opt_v;
        if (src.size() != m.params.size()) continue;
        const std::string path = "optimizer/_variables/" + std::to_string(++n);
        const std::string key = path + kVarSuffix;
        w.add_f32(key, seg.shape, src.data() + seg.offset);
        w.track(path, key, std::string(kind) + (m.optimizer == 0 ? (which == 0 ? "/m/" : "/v/") : "/velocity/") + cls + seg.path);
      }
    }
  }
  if (!m.model_info.empty()) {
    std::string key = std::string("model_info") + kVarSuffix;
    w.add_i32(key, {static_cast<int64_t>(m.model_info.size())}, m.model_info.data());
    w.track("model_info", key, "Variable");
  }
  if (s.has_gcn) {
    std::string kt = std::string("model_type") + kVarSuffix, kn = std::string("model_normalization") + kVarSuffix;
    w.add_scalar_string(kt, m.ds_type);
    w.track("model_type", kt, "Variable");
    w.add_scalar_string(kn, m.normalization);
    w.track("model_normalization", kn, "Variable");
  } else {
    std::string kn = std::string("normalization") + kVarSuffix;
    w.add_scalar_string(kn, m.normalization);
    w.track("normalization", kn, "Variable");
  }
  w.finish(dir + "/variables/variables");
}

}  // namespace gnnqc
