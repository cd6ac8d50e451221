// gnnqc_score_cml_host: the CML GCN scoring call for callers whose windows live in HOST memory (the
// shape model.predict has in the reference: numpy batches from prepare_batch_cml,
// libs/preprocessing_functions.py:638-666).  The batch is cut into chunks; each chunk is copied to one
// of two device staging slots on a copy stream while the previous chunk is scored on a compute stream,
// and its scores are copied back behind the kernels.  Only what can reach the score crosses PCIe:
// when the windows of a chunk have the same node count, the feature and target-series copies are
// pitched 2-D copies of the live timesteps (162 of 181 for the shipped models, StackPlan).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/gnnqc.h"
#include "capi_internal.h"

namespace gnnqc {

struct HostStage {
  cudaStream_t copy = nullptr, compute = nullptr;
  cudaEvent_t ready[2] = {nullptr, nullptr}, done[2] = {nullptr, nullptr};
  char* dev[2] = {nullptr, nullptr};
  size_t cap[2] = {0, 0};
  char* pinned[2] = {nullptr, nullptr};   // relative node / edge offsets of the chunk
  size_t pinned_cap[2] = {0, 0};
};

void free_host_stage(HostStage* h) {
  if (!h) return;
  for (int i = 0; i < 2; ++i) {
    if (h->ready[i]) cudaEventDestroy(h->ready[i]);
    if (h->done[i]) cudaEventDestroy(h->done[i]);
    cudaFree(h->dev[i]);
    cudaFreeHost(h->pinned[i]);
  }
  if (h->copy) cudaStreamDestroy(h->copy);
  if (h->compute) cudaStreamDestroy(h->compute);
  delete h;
}

namespace {

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

HostStage* stage_of(DeviceModel& dm) {
  if (!dm.host_stage) {
    HostStage* h = new HostStage();
    dm.host_stage = h;
    GNNQC_CUDA(cudaStreamCreateWithFlags(&h->copy, cudaStreamNonBlocking));
    GNNQC_CUDA(cudaStreamCreateWithFlags(&h->compute, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      GNNQC_CUDA(cudaEventCreateWithFlags(&h->ready[i], cudaEventDisableTiming));
      GNNQC_CUDA(cudaEventCreateWithFlags(&h->done[i], cudaEventDisableTiming));
    }
  }
  return dm.host_stage;
}

void score_cml_host(DeviceModel& dm, const float* feat, const float* anom, const long long* edges, const int* node_off,
                    const long long* edge_off, long long B, int T, long long chunk, float* scores, long long* h2d_bytes) {
  GNNQC_CUDA(cudaSetDevice(dm.device));
  HostStage* h = stage_of(dm);
  const int t_run = StackPlan(dm.host.spec, T).t_run[0];
  long long moved = 0;
  int used[2] = {0, 0};
  for (long long w0 = 0, c = 0; w0 < B; w0 += chunk, ++c) {
    const int i = static_cast<int>(c & 1);
    const long long W = std::min(chunk, B - w0);
    const long long n0 = node_off[w0], nodes = node_off[w0 + W] - n0;
    const long long e0 = edge_off[w0], n_edges = edge_off[w0 + W] - e0;
    const size_t o_feat = 0, o_anom = up256(size_t(nodes) * T * 8), o_edges = o_anom + up256(size_t(W) * T * 8),
                 o_noff = o_edges + up256(size_t(n_edges) * 16), o_eoff = o_noff + up256(size_t(W + 1) * 4),
                 o_scores = o_eoff + up256(size_t(W + 1) * 8), need = o_scores + up256(size_t(W) * 4);
    // the slot is reused every other chunk: wait until its previous scores have been copied out
    if (used[i]) GNNQC_CUDA(cudaEventSynchronize(h->done[i]));
    if (need > h->cap[i]) {
      GNNQC_CUDA(cudaFree(h->dev[i]));
      h->dev[i] = nullptr;
      GNNQC_CUDA(cudaMalloc(&h->dev[i], need));
      h->cap[i] = need;
    }
    const size_t pin_need = size_t(W + 1) * 12 + 8;
    if (pin_need > h->pinned_cap[i]) {
      GNNQC_CUDA(cudaFreeHost(h->pinned[i]));
      h->pinned[i] = nullptr;
      GNNQC_CUDA(cudaMallocHost(&h->pinned[i], pin_need));
      h->pinned_cap[i] = pin_need;
    }
    int* rel_no = reinterpret_cast<int*>(h->pinned[i]);
    long long* rel_eo = reinterpret_cast<long long*>(h->pinned[i] + (size_t(W + 1) * 4 + 7) / 8 * 8);
    const int nn0 = node_off[w0 + 1] - node_off[w0];
    bool uniform = nn0 > 0;
    for (long long w = 0; w <= W; ++w) {
      rel_no[w] = static_cast<int>(node_off[w0 + w] - n0);
      rel_eo[w] = edge_off[w0 + w] - e0;
      if (w < W && node_off[w0 + w + 1] - node_off[w0 + w] != nn0) uniform = false;
    }
    char* d = h->dev[i];
    const float* fsrc = feat + size_t(n0) * T * 2;
    if (uniform && t_run < T) {
      const size_t pitch = size_t(T) * nn0 * 8, width = size_t(t_run) * nn0 * 8;
      GNNQC_CUDA(cudaMemcpy2DAsync(d + o_feat, pitch, fsrc, pitch, width, size_t(W), cudaMemcpyHostToDevice, h->copy));
      moved += static_cast<long long>(width) * W;
    } else {
      GNNQC_CUDA(cudaMemcpyAsync(d + o_feat, fsrc, size_t(nodes) * T * 8, cudaMemcpyHostToDevice, h->copy));
      moved += nodes * T * 8;
    }
    GNNQC_CUDA(cudaMemcpy2DAsync(d + o_anom, size_t(T) * 8, anom + size_t(w0) * T * 2, size_t(T) * 8, size_t(t_run) * 8, size_t(W),
                                 cudaMemcpyHostToDevice, h->copy));
    GNNQC_CUDA(cudaMemcpyAsync(d + o_edges, edges + size_t(e0) * 2, size_t(n_edges) * 16, cudaMemcpyHostToDevice, h->copy));
    GNNQC_CUDA(cudaMemcpyAsync(d + o_noff, rel_no, size_t(W + 1) * 4, cudaMemcpyHostToDevice, h->copy));
    GNNQC_CUDA(cudaMemcpyAsync(d + o_eoff, rel_eo, size_t(W + 1) * 8, cudaMemcpyHostToDevice, h->copy));
    moved += W * t_run * 8 + n_edges * 16 + (W + 1) * 12;
    GNNQC_CUDA(cudaEventRecord(h->ready[i], h->copy));
    GNNQC_CUDA(cudaStreamWaitEvent(h->compute, h->ready[i], 0));
    dm.score_cml(reinterpret_cast<const float*>(d + o_feat), reinterpret_cast<const float*>(d + o_anom),
                 reinterpret_cast<const long long*>(d + o_edges), reinterpret_cast<const int*>(d + o_noff),
                 reinterpret_cast<const long long*>(d + o_eoff), W, T, nodes, reinterpret_cast<float*>(d + o_scores), h->compute);
    GNNQC_CUDA(cudaMemcpyAsync(scores + w0, d + o_scores, size_t(W) * 4, cudaMemcpyDeviceToHost, h->compute));
    GNNQC_CUDA(cudaEventRecord(h->done[i], h->compute));
    used[i] = 1;
  }
  GNNQC_CUDA(cudaStreamSynchronize(h->compute));
  GNNQC_CUDA(cudaStreamSynchronize(h->copy));
  if (h2d_bytes) *h2d_bytes = moved;
}

}  // namespace
}  // namespace gnnqc

extern "C" int gnnqc_score_cml_host(gnnqc_model* m, const float* features_host, const float* anom_host, const int64_t* edges_host,
                                    const int32_t* node_off_host, const int64_t* edge_off_host, int64_t B, int32_t T,
                                    int64_t chunk_windows, float* scores_host, int64_t* h2d_bytes_out) {
  if (!m || m->dm.device < 0) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_cml_host: model has no device");
  if (!m->dm.host.spec.has_gcn || m->dm.host.spec.soilnet)
    return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_cml_host: model is not a CML GCNClassifier");
  if (B < 0 || T <= 0 || chunk_windows < 0) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_cml_host: bad B / T / chunk_windows");
  if (B > 0 && (!features_host || !anom_host || !edges_host || !node_off_host || !edge_off_host || !scores_host))
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_cml_host: null buffer");
  try {
    m->dm.refresh_tc();
    gnnqc::score_cml_host(m->dm, features_host, anom_host, reinterpret_cast<const long long*>(edges_host), node_off_host,
                          reinterpret_cast<const long long*>(edge_off_host), B, T, chunk_windows > 0 ? chunk_windows : 8192, scores_host,
                          reinterpret_cast<long long*>(h2d_bytes_out));
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
}
