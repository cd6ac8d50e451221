// extern "C" surface of libgnnqc.so (declarations and reference citations: include/gnnqc.h).
#include <cstring>
#include <exception>
#include <new>
#include <string>

#include "../../include/gnnqc.h"
#include "capi_internal.h"
#include "device_model.cuh"
#include "model_spec.h"

using gnnqc::DeviceModel;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

// Runs `fn`, translating exceptions into error codes + gnnqc_last_error().
template <typename F>
int guarded(F&& fn) {
  try {
    fn();
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::invalid_argument& e) {
    return fail(GNNQC_ERR_ARG, e.what());
  } catch (const std::bad_alloc&) {
    return fail(GNNQC_ERR_STATE, "out of host memory");
  } catch (const std::exception& e) {
    return fail(GNNQC_ERR_CKPT, e.what());
  }
}

void attach_device(DeviceModel& dm, int device) {
  dm.device = device;
  if (device < 0) return;
  int count = 0;
  cudaError_t err = cudaGetDeviceCount(&count);
  if (err != cudaSuccess || count == 0)
    throw gnnqc::CudaError(std::string("no usable CUDA device (") + cudaGetErrorString(err) + "); libgnnqc has no CPU fallback");
  if (device >= count) throw std::invalid_argument("device index " + std::to_string(device) + " out of range");
  GNNQC_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  GNNQC_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    throw gnnqc::CudaError(std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major * 10 + prop.minor) +
                           "; libgnnqc is built for sm_100a (B200) only");
  dm.sm_count = prop.multiProcessorCount;
  dm.configure_kernels();
  dm.configure_ig_kernels();
  dm.configure_train_kernels();
  dm.upload();
}

int need_device(const gnnqc_model* m, const char* what) {
  if (!m) return fail(GNNQC_ERR_ARG, std::string(what) + ": null model handle");
  if (m->dm.device < 0) return fail(GNNQC_ERR_STATE, std::string(what) + ": handle was opened host-only (device < 0); there is no CPU path");
  return GNNQC_OK;
}

}  // namespace

int gnnqc_internal_fail(int code, const std::string& msg) { return fail(code, msg); }

extern "C" {

int gnnqc_abi_version(void) { return GNNQC_ABI_VERSION; }

const char* gnnqc_last_error(void) { return g_last_error.c_str(); }

int gnnqc_model_load(const char* savedmodel_dir, int device, gnnqc_model** out) {
  if (!savedmodel_dir || !out) return fail(GNNQC_ERR_ARG, "gnnqc_model_load: null argument");
  *out = nullptr;
  gnnqc_model* m = nullptr;
  int rc = guarded([&] {
    m = new gnnqc_model();
    m->dm.host = gnnqc::load_saved_model(savedmodel_dir);
    attach_device(m->dm, device);
  });
  if (rc != GNNQC_OK) {
    delete m;
    return rc;
  }
  *out = m;
  return GNNQC_OK;
}

int gnnqc_model_create(int has_gcn, int soilnet, int filter_1_size, int n_stacks, int dense_units,
                       const float* params_host, int64_t n_params, const int32_t* model_info, int n_model_info,
                       const char* normalization, int device, gnnqc_model** out) {
  if (!out) return fail(GNNQC_ERR_ARG, "gnnqc_model_create: null out");
  *out = nullptr;
  gnnqc_model* m = nullptr;
  int rc = guarded([&] {
    m = new gnnqc_model();
    gnnqc::HostModel& h = m->dm.host;
    h.spec.layout(has_gcn != 0, soilnet != 0, soilnet ? 3 : 2, filter_1_size, n_stacks, dense_units);
    if (params_host && n_params != h.spec.n_params)
      throw std::invalid_argument("gnnqc_model_create: expected " + std::to_string(h.spec.n_params) + " parameters, got " +
                                  std::to_string(n_params));
    h.params.assign(static_cast<size_t>(h.spec.n_params), 0.f);
    if (params_host) std::memcpy(h.params.data(), params_host, sizeof(float) * h.spec.n_params);
    if (model_info && n_model_info > 0) h.model_info.assign(model_info, model_info + n_model_info);
    h.ds_type = soilnet ? "soilnet" : "cml";
    h.normalization = normalization ? normalization : "";
    attach_device(m->dm, device);
  });
  if (rc != GNNQC_OK) {
    delete m;
    return rc;
  }
  *out = m;
  return GNNQC_OK;
}

int gnnqc_model_describe(const gnnqc_model* m, gnnqc_model_desc* out) {
  if (!m || !out) return fail(GNNQC_ERR_ARG, "gnnqc_model_describe: null argument");
  const gnnqc::HostModel& h = m->dm.host;
  std::memset(out, 0, sizeof(*out));
  out->has_gcn = h.spec.has_gcn;
  out->soilnet = h.spec.soilnet;
  out->n_features = h.spec.n_features;
  out->n_lstm = h.spec.n_lstm;
  for (int l = 0; l < h.spec.n_lstm; ++l) {
    out->lstm_in[l] = h.spec.lstm_in[l];
    out->lstm_hid[l] = h.spec.lstm_hid[l];
  }
  out->dense_units = h.spec.dense_units;
  out->n_model_info = static_cast<int32_t>(std::min<size_t>(h.model_info.size(), 4));
  for (int i = 0; i < out->n_model_info; ++i) out->model_info[i] = h.model_info[i];
  if (out->n_model_info >= 2) {
    // sequence_length = int((timestep_before + timestep_after)/freq + 1)  (create_model.py:12,
    // :245); freq defaults to 1 for CML and 15 for SoilNet when model_info has 3 entries.
    int freq = out->n_model_info >= 4 ? out->model_info[3] : (h.spec.soilnet ? 15 : 1);
    if (freq < 1) freq = 1;
    out->seq_len = (out->model_info[0] + out->model_info[1]) / freq + 1;
  }
  out->n_params = h.spec.n_params;
  out->n_trainable = h.spec.n_trainable;
  std::strncpy(out->ds_type, h.ds_type.c_str(), sizeof(out->ds_type) - 1);
  std::strncpy(out->normalization, h.normalization.c_str(), sizeof(out->normalization) - 1);
  return GNNQC_OK;
}

int gnnqc_model_get_params(const gnnqc_model* m, float* out_host, int64_t n) {
  if (!m || !out_host) return fail(GNNQC_ERR_ARG, "gnnqc_model_get_params: null argument");
  if (n != m->dm.host.spec.n_params) return fail(GNNQC_ERR_ARG, "gnnqc_model_get_params: wrong length");
  if (int rc = guarded([&] { const_cast<gnnqc_model*>(m)->dm.sync_host_params(); })) return rc;
  std::memcpy(out_host, m->dm.host.params.data(), sizeof(float) * n);
  return GNNQC_OK;
}

int gnnqc_model_set_params(gnnqc_model* m, const float* in_host, int64_t n) {
  if (!m || !in_host) return fail(GNNQC_ERR_ARG, "gnnqc_model_set_params: null argument");
  if (n != m->dm.host.spec.n_params) return fail(GNNQC_ERR_ARG, "gnnqc_model_set_params: wrong length");
  return guarded([&] {
    std::memcpy(m->dm.host.params.data(), in_host, sizeof(float) * n);
    m->dm.params_dirty = false;      // the host copy is authoritative again
    m->dm.upload();
  });
}

int gnnqc_model_segment(const gnnqc_model* m, int idx, char path[128], int64_t* offset, int64_t shape[4], int32_t* rank,
                        int32_t* trainable) {
  if (!m) return fail(GNNQC_ERR_ARG, "gnnqc_model_segment: null model");
  std::vector<gnnqc::ParamSegment> segs = gnnqc::param_segments(m->dm.host.spec);
  if (idx < 0 || idx >= static_cast<int>(segs.size())) return fail(GNNQC_ERR_ARG, "gnnqc_model_segment: index past the last segment");
  const gnnqc::ParamSegment& s = segs[idx];
  if (path) {
    std::memset(path, 0, 128);
    std::strncpy(path, s.path.c_str(), 127);
  }
  if (offset) *offset = s.offset;
  if (rank) *rank = static_cast<int32_t>(s.shape.size());
  if (shape)
    for (size_t i = 0; i < 4; ++i) shape[i] = i < s.shape.size() ? s.shape[i] : 1;
  if (trainable) *trainable = s.trainable;
  return GNNQC_OK;
}

int gnnqc_model_save(const gnnqc_model* m, const char* dir) {
  if (!m || !dir) return fail(GNNQC_ERR_ARG, "gnnqc_model_save: null argument");
  return guarded([&] {
    const_cast<gnnqc_model*>(m)->dm.sync_host_params();
    const_cast<gnnqc_model*>(m)->dm.sync_host_optimizer();
    gnnqc::save_saved_model(m->dm.host, dir);
  });
}

void gnnqc_model_free(gnnqc_model* m) { delete m; }

int gnnqc_score_seq(gnnqc_model* m, const float* x_dev, int64_t n_seq, int32_t T, float* scores_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_score_seq")) return rc;
  if (n_seq < 0 || T <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_score_seq: bad n_seq / T");
  if (n_seq > 0 && (!x_dev || !scores_dev)) return fail(GNNQC_ERR_ARG, "gnnqc_score_seq: null buffer");
  return guarded([&] {
    m->dm.refresh_tc();
    m->dm.score_seq(x_dev, n_seq, T, scores_dev, static_cast<cudaStream_t>(stream));
  });
}

int gnnqc_score_rows(gnnqc_model* m, const float* features_dev, const int32_t* node_off_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                     float* scores_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_score_rows")) return rc;
  if (m->dm.host.spec.has_gcn) return fail(GNNQC_ERR_STATE, "gnnqc_score_rows: model is not a BaselineClassifier");
  if (B < 0 || T <= 0 || n_nodes_total < 0 || B >= (int64_t(1) << 31)) return fail(GNNQC_ERR_ARG, "gnnqc_score_rows: bad B / T / n_nodes_total");
  if (n_nodes_total > 0 && (!features_dev || !node_off_dev || !scores_dev)) return fail(GNNQC_ERR_ARG, "gnnqc_score_rows: null buffer");
  return guarded([&] {
    m->dm.refresh_tc();
    m->dm.score_seq(features_dev, n_nodes_total, T, scores_dev, static_cast<cudaStream_t>(stream), node_off_dev, static_cast<int>(B));
  });
}

int gnnqc_score_cml(gnnqc_model* m, const float* features_dev, const float* anom_dev, const int64_t* edges_dev,
                    const int32_t* node_off_dev, const int64_t* edge_off_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                    float* scores_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_score_cml")) return rc;
  if (!m->dm.host.spec.has_gcn || m->dm.host.spec.soilnet)
    return fail(GNNQC_ERR_STATE, "gnnqc_score_cml: model is not a CML GCNClassifier");
  if (B < 0 || T <= 0 || n_nodes_total < 0) return fail(GNNQC_ERR_ARG, "gnnqc_score_cml: bad B / T / n_nodes_total");
  if (B > 0 && (!features_dev || !anom_dev || !edges_dev || !node_off_dev || !edge_off_dev || !scores_dev))
    return fail(GNNQC_ERR_ARG, "gnnqc_score_cml: null buffer");
  return guarded([&] {
    m->dm.refresh_tc();
    m->dm.score_cml(features_dev, anom_dev, reinterpret_cast<const long long*>(edges_dev), node_off_dev,
                    reinterpret_cast<const long long*>(edge_off_dev), B, T, n_nodes_total, scores_dev,
                    static_cast<cudaStream_t>(stream));
  });
}

int gnnqc_score_soilnet(gnnqc_model* m, const float* features_dev, const int64_t* edges_dev, const int32_t* node_off_dev,
                        const int64_t* edge_off_dev, int64_t B, int32_t T, int64_t n_nodes_total, int64_t n_edges_total,
                        float* scores_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_score_soilnet")) return rc;
  if (!m->dm.host.spec.has_gcn || !m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_score_soilnet: model is not a SoilNet GCNClassifier");
  if (B < 0 || T <= 0 || n_nodes_total < 0 || n_edges_total < 0) return fail(GNNQC_ERR_ARG, "gnnqc_score_soilnet: bad B / T / totals");
  if (n_nodes_total >= (int64_t(1) << 31) - 1 || n_edges_total >= (int64_t(1) << 31) - 1)
    return fail(GNNQC_ERR_ARG, "gnnqc_score_soilnet: more than 2^31 nodes or edges in one call; split the batch");
  if (B > 0 && (!features_dev || !edges_dev || !node_off_dev || !edge_off_dev || !scores_dev))
    return fail(GNNQC_ERR_ARG, "gnnqc_score_soilnet: null buffer");
  return guarded([&] {
    m->dm.refresh_tc();
    m->dm.score_soilnet(features_dev, reinterpret_cast<const long long*>(edges_dev), node_off_dev,
                        reinterpret_cast<const long long*>(edge_off_dev), B, T, n_nodes_total, n_edges_total, scores_dev,
                        static_cast<cudaStream_t>(stream));
  });
}

int gnnqc_ig_cml(gnnqc_model* m, const float* features_dev, const float* anom_dev, const int64_t* edges_dev,
                 const int32_t* node_off_dev, const int64_t* edge_off_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                 int32_t m_steps, float* attr_features_dev, float* attr_anom_dev, float* scores_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_ig_cml")) return rc;
  if (!m->dm.host.spec.has_gcn || m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_ig_cml: model is not a CML GCNClassifier");
  if (B < 0 || T <= 0 || n_nodes_total < 0 || m_steps < 1) return fail(GNNQC_ERR_ARG, "gnnqc_ig_cml: bad B / T / n_nodes_total / m_steps");
  if (B > 0 && (!features_dev || !anom_dev || !edges_dev || !node_off_dev || !edge_off_dev || !attr_features_dev || !attr_anom_dev || !scores_dev))
    return fail(GNNQC_ERR_ARG, "gnnqc_ig_cml: null buffer");
  return guarded([&] {
    m->dm.refresh_tc();
    m->dm.ig_cml(features_dev, anom_dev, reinterpret_cast<const long long*>(edges_dev), node_off_dev,
                 reinterpret_cast<const long long*>(edge_off_dev), B, T, n_nodes_total, m_steps, attr_features_dev, attr_anom_dev,
                 scores_dev, static_cast<cudaStream_t>(stream));
  });
}

int gnnqc_ig_cml_raw(gnnqc_model* m, const float* features_dev, const float* anom_dev, const int64_t* edges_dev,
                 const int32_t* node_off_dev, const int64_t* edge_off_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                 int32_t m_steps, float* attr_features_dev, float* attr_anom_dev, float* scores_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_ig_cml_raw")) return rc;
  if (!m->dm.host.spec.has_gcn || m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_ig_cml_raw: model is not a CML GCNClassifier");
  if (B < 0 || T <= 0 || n_nodes_total < 0 || m_steps < 1) return fail(GNNQC_ERR_ARG, "gnnqc_ig_cml_raw: bad B / T / n_nodes_total / m_steps");
  if (B > 0 && (!features_dev || !anom_dev || !edges_dev || !node_off_dev || !edge_off_dev || !attr_features_dev || !attr_anom_dev || !scores_dev))
    return fail(GNNQC_ERR_ARG, "gnnqc_ig_cml_raw: null buffer");
  return guarded([&] {
    m->dm.refresh_tc();
    m->dm.ig_cml(features_dev, anom_dev, reinterpret_cast<const long long*>(edges_dev), node_off_dev,
                 reinterpret_cast<const long long*>(edge_off_dev), B, T, n_nodes_total, m_steps, attr_features_dev, attr_anom_dev,
                 scores_dev, static_cast<cudaStream_t>(stream), false);
  });
}

int64_t gnnqc_train_vector_len(const gnnqc_model* m) { return m ? m->dm.host.spec.n_params + 97 : 0; }

int gnnqc_train_moments_cml(gnnqc_model* m, const float* features_dev, int64_t n_rows, double* moments_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_train_moments_cml")) return rc;
  if (!m->dm.host.spec.has_gcn || m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_train_moments_cml: model is not a CML GCNClassifier");
  if (!features_dev || !moments_dev || n_rows <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_train_moments_cml: bad argument");
  return guarded([&] { m->dm.train_moments(features_dev, n_rows, moments_dev, static_cast<cudaStream_t>(stream)); });
}

int gnnqc_train_grads_cml(gnnqc_model* m, const float* features_dev, const float* anom_dev, const int64_t* edges_dev,
                          const int32_t* node_off_dev, const int64_t* edge_off_dev, const int64_t* labels_dev, int64_t B, int32_t T,
                          int64_t n_nodes_total, const double* moments_dev, float class_weight_0, float class_weight_1,
                          int64_t global_batch, float* rvec_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_train_grads_cml")) return rc;
  if (!m->dm.host.spec.has_gcn || m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_train_grads_cml: model is not a CML GCNClassifier");
  if (B <= 0 || T <= 0 || n_nodes_total <= 0 || global_batch <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_train_grads_cml: bad sizes");
  if (!features_dev || !anom_dev || !edges_dev || !node_off_dev || !edge_off_dev || !labels_dev || !moments_dev || !rvec_dev)
    return fail(GNNQC_ERR_ARG, "gnnqc_train_grads_cml: null buffer");
  return guarded([&] {
    m->dm.train_grads(features_dev, anom_dev, reinterpret_cast<const long long*>(edges_dev), node_off_dev,
                      reinterpret_cast<const long long*>(edge_off_dev), nullptr, reinterpret_cast<const long long*>(labels_dev), B, T,
                      n_nodes_total, moments_dev, class_weight_0, class_weight_1, global_batch, rvec_dev, static_cast<cudaStream_t>(stream));
  });
}

int gnnqc_train_moments_soilnet(gnnqc_model* m, const float* features_dev, int64_t n_rows, double* moments_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_train_moments_soilnet")) return rc;
  if (!m->dm.host.spec.has_gcn || !m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_train_moments_soilnet: model is not a SoilNet GCNClassifier");
  if (!features_dev || !moments_dev || n_rows <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_train_moments_soilnet: bad argument");
  return guarded([&] { m->dm.train_moments(features_dev, n_rows, moments_dev, static_cast<cudaStream_t>(stream)); });
}

int gnnqc_train_grads_soilnet(gnnqc_model* m, const float* features_dev, const int64_t* edges_dev, const int32_t* node_off_dev,
                              const int64_t* edge_off_dev, const int64_t* labels_dev, int64_t B, int32_t T, int64_t n_nodes_total,
                              int64_t n_edges_total, const double* moments_dev, float class_weight_0, float class_weight_1,
                              int64_t global_batch, float* rvec_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_train_grads_soilnet")) return rc;
  if (!m->dm.host.spec.has_gcn || !m->dm.host.spec.soilnet) return fail(GNNQC_ERR_STATE, "gnnqc_train_grads_soilnet: model is not a SoilNet GCNClassifier");
  if (B <= 0 || T <= 0 || n_nodes_total <= 0 || n_edges_total < 0 || global_batch <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_train_grads_soilnet: bad sizes");
  if (!features_dev || !edges_dev || !node_off_dev || !edge_off_dev || !labels_dev || !moments_dev || !rvec_dev)
    return fail(GNNQC_ERR_ARG, "gnnqc_train_grads_soilnet: null buffer");
  return guarded([&] {
    m->dm.train_grads(features_dev, nullptr, reinterpret_cast<const long long*>(edges_dev), node_off_dev,
                      reinterpret_cast<const long long*>(edge_off_dev), nullptr, reinterpret_cast<const long long*>(labels_dev), B, T,
                      n_nodes_total, moments_dev, class_weight_0, class_weight_1, global_batch, rvec_dev, static_cast<cudaStream_t>(stream),
                      n_edges_total);
  });
}

int gnnqc_train_grads_seq(gnnqc_model* m, const float* x_dev, const int64_t* labels_dev, int64_t n_seq, int32_t T, float class_weight_0,
                          float class_weight_1, int64_t global_batch, float* rvec_dev, void* stream) {
  if (int rc = need_device(m, "gnnqc_train_grads_seq")) return rc;
  if (m->dm.host.spec.has_gcn) return fail(GNNQC_ERR_STATE, "gnnqc_train_grads_seq: model is not a BaselineClassifier");
  if (n_seq <= 0 || T <= 0 || global_batch <= 0 || !x_dev || !labels_dev || !rvec_dev) return fail(GNNQC_ERR_ARG, "gnnqc_train_grads_seq: bad argument");
  return guarded([&] {
    m->dm.train_grads(nullptr, nullptr, nullptr, nullptr, nullptr, x_dev, reinterpret_cast<const long long*>(labels_dev), n_seq, T, 0, nullptr,
                      class_weight_0, class_weight_1, global_batch, rvec_dev, static_cast<cudaStream_t>(stream));
  });
}

int gnnqc_train_set_optimizer(gnnqc_model* m, int kind) {
  if (int rc = need_device(m, "gnnqc_train_set_optimizer")) return rc;
  if (kind < 0 || kind > 2) return fail(GNNQC_ERR_ARG, "gnnqc_train_set_optimizer: kind must be 0 (adam), 1 (sgd) or 2 (rmsprop)");
  return guarded([&] {
    GNNQC_CUDA(cudaSetDevice(m->dm.device));
    m->dm.set_optimizer(kind);
  });
}

int64_t gnnqc_train_iterations(const gnnqc_model* m) {
  if (!m) return 0;
  return m->dm.train_iterations();
}

int gnnqc_train_apply(gnnqc_model* m, float* rvec_dev, const double* moments_dev, float learning_rate, void* stream) {
  if (int rc = need_device(m, "gnnqc_train_apply")) return rc;
  if (!rvec_dev || !(learning_rate > 0.f)) return fail(GNNQC_ERR_ARG, "gnnqc_train_apply: bad argument");
  return guarded([&] { m->dm.train_apply(rvec_dev, moments_dev, learning_rate, static_cast<cudaStream_t>(stream)); });
}

int64_t gnnqc_launch_count(const gnnqc_model* m) { return m ? m->dm.launches : 0; }

int gnnqc_profile_enable(gnnqc_model* m, int on) {
  if (int rc = need_device(m, "gnnqc_profile_enable")) return rc;
  return guarded([&] {
    GNNQC_CUDA(cudaSetDevice(m->dm.device));
    m->dm.profile_reset();
    m->dm.profiling = on != 0;
  });
}

int gnnqc_profile_read(gnnqc_model* m, int max_entries, char (*names)[32], double* total_ms, int64_t* counts, int* n_out) {
  if (int rc = need_device(m, "gnnqc_profile_read")) return rc;
  if (!names || !total_ms || !counts || !n_out || max_entries <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_profile_read: null argument");
  return guarded([&] {
    GNNQC_CUDA(cudaSetDevice(m->dm.device));
    GNNQC_CUDA(cudaDeviceSynchronize());
    int n = 0;
    for (const gnnqc::LaunchRecord& r : m->dm.records) {
      float ms = 0.f;
      GNNQC_CUDA(cudaEventElapsedTime(&ms, r.start, r.stop));
      int slot = -1;
      for (int i = 0; i < n; ++i)
        if (std::strncmp(names[i], r.name, 31) == 0) slot = i;
      if (slot < 0) {
        if (n >= max_entries) continue;
        slot = n++;
        std::memset(names[slot], 0, 32);
        std::strncpy(names[slot], r.name, 31);
        total_ms[slot] = 0.0;
        counts[slot] = 0;
      }
      total_ms[slot] += ms;
      counts[slot] += 1;
    }
    *n_out = n;
    m->dm.profile_reset();
  });
}

int64_t gnnqc_score_scratch_bytes(const gnnqc_model* m, int64_t n_seq, int32_t T) {
  if (!m || n_seq < 0 || T <= 0) return fail(GNNQC_ERR_ARG, "gnnqc_score_scratch_bytes: bad argument");
  int64_t bytes = 0;
  int rc = guarded([&] {
    gnnqc::StackPlan plan(m->dm.host.spec, T);
    bytes = static_cast<int64_t>(plan.bytes_per_tile()) * ((n_seq + gnnqc::kTile - 1) / gnnqc::kTile);
  });
  return rc == GNNQC_OK ? bytes : rc;
}

}  // extern "C"
