// SoilNet windows built on the device from the resident raw network (SURVEY section 8(f) N1 for ds_type
// 'soilnet').  Replaces create_tfrecords_dataset's soilnet branch (libs/preprocessing_functions.py:418-480:
// one window per date over ALL sensors; a sensor is dropped from a window when its target at the window's
// date is NaN or any of its three features has a NaN inside the window; adjacency = the global
// distance/depth rule restricted to the remaining sensors), parse_soilnet_tfrecord_fn (:772-857,
// normalisation), prepare_batch_soilnet (:861-897) and GCNClassifier.call for 'soilnet'.
//
// Every (window, sensor) pair is scored as a sequence, whether the sensor is valid in that window or not:
// no compaction, no device-to-host read of a count.  Invalid sensors are left out of their neighbours'
// means and get NaN as score.  Neighbours are visited in ascending sensor order, which is the order
// np.where gives the reference's edge list, and the arithmetic is the materialised kernel's
// (soil_frontend.cuh), so the scores equal those of "build the windows on the host, then
// gnnqc_score_soilnet" bit for bit.
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/gnnqc.h"
#include "capi_internal.h"
#include "soil_frontend.cuh"

struct gnnqc_soil_series {
  int device = 0;
  long long S = 0, Tt = 0, n_adj = 0;
  float* raw = nullptr;        // [S][Tt][3] moisture, temp, battv (NaN = missing)
  float* target = nullptr;     // [S][Tt] anomaly flag as float (NaN = no label)
  int* nan_prefix = nullptr;   // [S][Tt+1] samples with a NaN in any feature before t (global scan, rows subtract)
  int* adj_off = nullptr;      // [S+1]
  int* adj_idx = nullptr;      // neighbours of each sensor in ascending order (the sensor itself included where the rule says so)
};

namespace gnnqc {
namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

void free_soil_series(gnnqc_soil_series* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaFree(s->raw); cudaFree(s->target); cudaFree(s->nan_prefix); cudaFree(s->adj_off); cudaFree(s->adj_idx);
  delete s;
}

__global__ void __launch_bounds__(256) soil_nan_flag_kernel(const float* __restrict__ raw, int* __restrict__ flags, long long S, long long Tt) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= S * (Tt + 1)) return;
  const long long s = i / (Tt + 1), t = i % (Tt + 1);
  int f = 0;
  if (t > 0) {
    const float* v = raw + (s * Tt + t - 1) * 3;
    f = (isnan(v[0]) || isnan(v[1]) || isnan(v[2])) ? 1 : 0;
  }
  flags[i] = f;
}

void rebuild_soil_nan_prefix(gnnqc_soil_series* s) {
  const long long n = s->S * (s->Tt + 1);
  soil_nan_flag_kernel<<<static_cast<unsigned>((n + 255) / 256), 256>>>(s->raw, s->nan_prefix, s->S, s->Tt);
  GNNQC_CUDA(cudaGetLastError());
  size_t tmp_bytes = 0;
  GNNQC_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, s->nan_prefix, s->nan_prefix, static_cast<int>(n)));
  void* tmp = nullptr;
  GNNQC_CUDA(cudaMalloc(&tmp, tmp_bytes));
  cudaError_t err = cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, s->nan_prefix, s->nan_prefix, static_cast<int>(n));
  cudaFree(tmp);
  GNNQC_CUDA(err);
  GNNQC_CUDA(cudaDeviceSynchronize());
}

struct SoilNorm {
  float sub[3], div[3];
};

// valid[w][s]: window w inside the series, sensor s labelled at the window's date and without a NaN in the window
__global__ void __launch_bounds__(256) soil_valid_kernel(const gnnqc_soil_series sr, const int* __restrict__ t_anom, int before, int after,
                                                         long long n_seq, unsigned char* __restrict__ valid) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_seq) return;
  const long long w = i / sr.S, s = i % sr.S;
  const long long ta = t_anom[w], t0 = ta - before, t1 = ta + after;
  bool ok = t0 >= 0 && t1 < sr.Tt;
  if (ok) {
    const int* pf = sr.nan_prefix + s * (sr.Tt + 1);
    ok = !isnan(sr.target[s * sr.Tt + ta]) && (pf[t1 + 1] - pf[t0] == 0);
  }
  valid[i] = ok ? 1 : 0;
}

// x1 tile [tile][t][19][64] of the sequences (window, sensor): channels 0..15 = mean over the sensor's valid
// neighbours of the projected features, 16..18 = its own normalised features.  One thread per (sequence, 8 steps).
__global__ void __launch_bounds__(256) soil_series_frontend_kernel(const gnnqc_soil_series sr, const int* __restrict__ t_anom, int before,
                                                                   const unsigned char* __restrict__ valid, const SoilNorm nm,
                                                                   const GcnFolded p, float* __restrict__ out, long long seq_base,
                                                                   long long n_seq_total, int t_run) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long seq = seq_base + tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
  constexpr int CH = kGcnChannelsDev + 3;
  bool live = seq < n_seq_total;
  long long w = 0, s = 0, base_t = 0;
  int a0 = 0, a1 = 0;
  if (live) {
    w = seq / sr.S;
    s = seq % sr.S;
    live = valid[seq] != 0;
    base_t = static_cast<long long>(t_anom[w]) - before;
    a0 = sr.adj_off[s];
    a1 = sr.adj_off[s + 1];
  }
  const unsigned char* vrow = valid + w * sr.S;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    float* o = out + ((tile * t_run + t) * CH) * kTile + b;
    if (!live) {   // a finite tile row; the score of this sequence is overwritten with NaN afterwards
#pragma unroll
      for (int c = 0; c < CH; ++c) o[c * kTile] = 0.f;
      continue;
    }
    SoilAcc acc;
    int d = 0;
    for (int e = a0; e < a1; ++e) {
      const int j = __ldg(sr.adj_idx + e);
      if (!vrow[j]) continue;
      const float* v = sr.raw + (static_cast<long long>(j) * sr.Tt + base_t + t) * 3;
      acc.add(p, (__ldg(v) - nm.sub[0]) / nm.div[0], (__ldg(v + 1) - nm.sub[1]) / nm.div[1], (__ldg(v + 2) - nm.sub[2]) / nm.div[2]);
      ++d;
    }
    const float inv = d > 0 ? 1.f / static_cast<float>(d) : 0.f, fd = static_cast<float>(d);
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) o[c * kTile] = acc.mean(p, c, fd, inv);
    const float* own = sr.raw + (s * sr.Tt + base_t + t) * 3;
    o[16 * kTile] = (__ldg(own) - nm.sub[0]) / nm.div[0];
    o[17 * kTile] = (__ldg(own + 1) - nm.sub[1]) / nm.div[1];
    o[18 * kTile] = (__ldg(own + 2) - nm.sub[2]) / nm.div[2];
  }
}

__global__ void __launch_bounds__(256) soil_mask_kernel(float* __restrict__ scores, const unsigned char* __restrict__ valid, long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n && !valid[i]) scores[i] = nanf("");
}


}  // namespace
}  // namespace gnnqc

extern "C" {

int gnnqc_soil_series_create(int device, const float* raw_host, const float* target_host, int64_t n_sensors, int64_t n_times,
                             const uint8_t* adjacency_host, gnnqc_soil_series** out) {
  if (!out) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_soil_series_create: null out");
  *out = nullptr;
  if (!raw_host || !target_host || !adjacency_host || n_sensors <= 0 || n_times <= 0)
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_soil_series_create: bad argument");
  if (n_sensors * (n_times + 1) >= (int64_t(1) << 31))
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_soil_series_create: more than 2^31 samples per handle; shard the time axis");
  gnnqc_soil_series* s = new gnnqc_soil_series();
  try {
    s->device = device;
    s->S = n_sensors;
    s->Tt = n_times;
    GNNQC_CUDA(cudaSetDevice(device));
    std::vector<int> off(size_t(n_sensors) + 1, 0), idx;
    for (int64_t i = 0; i < n_sensors; ++i) {
      for (int64_t j = 0; j < n_sensors; ++j)
        if (adjacency_host[i * n_sensors + j]) idx.push_back(static_cast<int>(j));
      off[size_t(i) + 1] = static_cast<int>(idx.size());
    }
    s->n_adj = static_cast<long long>(idx.size());
    const size_t n = size_t(n_sensors) * size_t(n_times);
    GNNQC_CUDA(cudaMalloc(&s->raw, n * 3 * sizeof(float)));
    GNNQC_CUDA(cudaMalloc(&s->target, n * sizeof(float)));
    GNNQC_CUDA(cudaMalloc(&s->nan_prefix, size_t(n_sensors) * size_t(n_times + 1) * sizeof(int)));
    GNNQC_CUDA(cudaMalloc(&s->adj_off, off.size() * sizeof(int)));
    GNNQC_CUDA(cudaMalloc(&s->adj_idx, std::max<size_t>(idx.size(), 1) * sizeof(int)));
    GNNQC_CUDA(cudaMemcpy(s->raw, raw_host, n * 3 * sizeof(float), cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->target, target_host, n * sizeof(float), cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->adj_off, off.data(), off.size() * sizeof(int), cudaMemcpyHostToDevice));
    if (!idx.empty()) GNNQC_CUDA(cudaMemcpy(s->adj_idx, idx.data(), idx.size() * sizeof(int), cudaMemcpyHostToDevice));
    gnnqc::rebuild_soil_nan_prefix(s);
  } catch (const gnnqc::CudaError& e) {
    gnnqc::free_soil_series(s);
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    gnnqc::free_soil_series(s);
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
  *out = s;
  return GNNQC_OK;
}

void gnnqc_soil_series_free(gnnqc_soil_series* s) { gnnqc::free_soil_series(s); }

int gnnqc_score_soilnet_series(gnnqc_model* m, const gnnqc_soil_series* series, const int32_t* t_anom_dev, int64_t B, int32_t steps_before,
                               int32_t steps_after, const float* norm_sub3_host, const float* norm_div3_host, float* scores_dev,
                               uint8_t* valid_dev, void* stream) {
  using namespace gnnqc;
  if (!m || m->dm.device < 0) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_soilnet_series: model has no device");
  if (!m->dm.host.spec.has_gcn || !m->dm.host.spec.soilnet)
    return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_soilnet_series: model is not a SoilNet GCNClassifier");
  if (!series || series->device != m->dm.device) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_soilnet_series: series is on another device");
  if (B < 0 || steps_before < 0 || steps_after < 0 || (B > 0 && (!t_anom_dev || !scores_dev)))
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_soilnet_series: bad argument");
  if (B == 0) return GNNQC_OK;
  try {
    DeviceModel& dm = m->dm;
    dm.refresh_tc();
    GNNQC_CUDA(cudaSetDevice(dm.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const ModelSpec& s = dm.host.spec;
    const int T = steps_before + steps_after + 1;
    StackPlan plan(s, T);
    const long long n_seq = B * series->S, n_tiles = (n_seq + kTile - 1) / kTile;
    const size_t sa = align_up(plan.bytes_a, 256), sb = align_up(plan.bytes_b, 256);
    const char* env = std::getenv("GNNQC_SCRATCH_MB");
    const size_t budget = std::max<size_t>(env ? std::strtoull(env, nullptr, 10) : 8192, 16) << 20;
    const long long chunk = balanced_chunk(n_tiles, static_cast<long long>(budget / (sa + sb)));
    const size_t vbytes = align_up(size_t(n_seq), 256);
    char* base = static_cast<char*>(dm.scratch(vbytes + size_t(chunk) * (sa + sb)));
    unsigned char* valid = reinterpret_cast<unsigned char*>(base);
    float* buf_a = reinterpret_cast<float*>(base + vbytes);
    float* buf_b = reinterpret_cast<float*>(base + vbytes + size_t(chunk) * sa);
    SoilNorm nm = {{0.f, -20.f, 2800.f}, {60.f, 60.f, 800.f}};   // scale_range (parse_soilnet_tfrecord_fn :825-832)
    if (norm_sub3_host && norm_div3_host)
      for (int c = 0; c < 3; ++c) { nm.sub[c] = norm_sub3_host[c]; nm.div[c] = norm_div3_host[c]; }
    dm.mark_begin("soil_series_valid", st);
    soil_valid_kernel<<<static_cast<unsigned>((n_seq + 255) / 256), 256, 0, st>>>(*series, t_anom_dev, steps_before, steps_after, n_seq, valid);
    GNNQC_CUDA(cudaGetLastError());
    dm.mark_end(st);
    const GcnFolded folded = dm.folded(st);
    for (long long t0 = 0; t0 < n_tiles; t0 += chunk) {
      const long long nt = std::min(chunk, n_tiles - t0);
      const long long seq0 = t0 * kTile, ns = std::min<long long>(n_seq - seq0, nt * kTile);
      dim3 grid(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
      dm.mark_begin("soil_series_frontend", st);
      soil_series_frontend_kernel<<<grid, 256, 0, st>>>(*series, t_anom_dev, steps_before, valid, nm, folded, buf_a, seq0, n_seq, plan.t_run[0]);
      GNNQC_CUDA(cudaGetLastError());
      dm.mark_end(st);
      dm.run_stack(plan, buf_a, buf_b, nt, ns, scores_dev + seq0, st);
    }
    soil_mask_kernel<<<static_cast<unsigned>((n_seq + 255) / 256), 256, 0, st>>>(scores_dev, valid, n_seq);
    GNNQC_CUDA(cudaGetLastError());
    if (valid_dev) GNNQC_CUDA(cudaMemcpyAsync(valid_dev, valid, size_t(n_seq), cudaMemcpyDeviceToDevice, st));
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
}

}  // extern "C"
