// On-device window builder for the CML GCN model (gnnqc_series_*, gnnqc_score_cml_series): scores
// sliding windows straight from the raw per-sensor series resident in HBM, so that a stride-1 sweep
// moves ~8 bytes per window over PCIe instead of the ~35 kB of a materialised window.
//
// Restates what the reference does on the host when it writes and parses its TFRecords:
//   libs/preprocessing_functions.py:388-412  window = [t - timestep_before, t + timestep_after];
//        skipped when it is not complete or the target sensor has a NaN in it; neighbours with a NaN
//        in the window are dropped; adjacency = pairwise distance < max_sample_distance among the
//        sensors that are left (here: the caller's per-target edge list restricted to them)
//   :123-140, :259-262, :620-622             normalisation 'rolling_median': every sensor's series
//        minus that sensor's rolling median AT THE WINDOW'S TARGET TIME (one scalar per sensor,
//        channel and window); the rolling medians are an input here, as they are precomputed per
//        dataset in the reference (calculate_statistics)
//   :638-666                                 rows in (window, timestep, node) order, target series too
// followed by the same GeneralConv + node pooling + LSTM stack + head as gnnqc_score_cml.
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>

#include "../../include/gnnqc.h"
#include "capi_internal.h"
#include "frontend.cuh"
#include "ig_frontend_bwd.cuh"

struct gnnqc_series {
  int device = 0;
  long long S = 0, Tt = 0, n_nbr = 0, n_edges = 0;
  int max_nbr = 0;
  float2* raw = nullptr;       // [S][Tt]
  float2* rmed = nullptr;      // [S][Tt]
  int* nan_prefix = nullptr;   // [S][Tt+1] number of samples with a NaN (either channel) before t
  int* nbr_off = nullptr;      // [S+1]
  int* nbr_idx = nullptr;      // [n_nbr] global sensor ids of each target's neighbourhood (target included)
  long long* edge_off = nullptr;  // [S+1]
  long long* edges = nullptr;     // [n_edges][2] indices into the target's neighbour list
  void* scan_tmp = nullptr;       // cub scratch of the prefix rebuild, kept so that updates allocate nothing
  size_t scan_tmp_bytes = 0;
};

namespace gnnqc {
namespace {

constexpr int kMaxNbr = 32;   // neighbourhood size the shared-memory coefficient table supports

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__global__ void __launch_bounds__(256) nan_flag_kernel(const float2* __restrict__ raw, int* __restrict__ flags, long long S, long long Tt) {
  // flags laid out [S][Tt+1]; flags[s][0] stays 0 so that an inclusive scan per row gives the prefix
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= S * (Tt + 1)) return;
  const long long s = i / (Tt + 1), t = i % (Tt + 1);
  int f = 0;
  if (t > 0) {
    const float2 v = raw[s * Tt + t - 1];
    f = (isnan(v.x) || isnan(v.y)) ? 1 : 0;
  }
  flags[i] = f;
}

// After a global inclusive scan over the flattened [S][Tt+1] flags, prefix[s][t] - prefix[s][0] is
// the per-sensor count; store the row bases so windows can subtract them (kept implicit: the
// difference of two entries of the same row is all that is ever used).

// One warp per window: validity of the window and of each neighbour, node coefficients (same dense
// table as cml_node_coef_kernel, restricted to valid nodes), per-node medians.
// Per-window outputs (stride kMaxNbr): sensor[n] (global id or -1), coef[n], med[n].
__global__ void __launch_bounds__(256) series_prepare_kernel(const gnnqc_series sr, const int* __restrict__ target,
                                                             const int* __restrict__ t_anom, int before, int after, long long n_windows,
                                                             int* __restrict__ w_sensor, float* __restrict__ w_coef, float2* __restrict__ w_med,
                                                             int* __restrict__ w_count, int* __restrict__ w_tgt_slot) {
  __shared__ int mult_s[8][kMaxNbr * kMaxNbr];
  __shared__ int slot_s[8][kMaxNbr];
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  if (w >= n_windows) return;
  const int s = target[w];
  const long long t0 = static_cast<long long>(t_anom[w]) - before, t1 = static_cast<long long>(t_anom[w]) + after;  // inclusive
  int* sensors = w_sensor + w * kMaxNbr;
  float* coef = w_coef + w * kMaxNbr;
  float2* med = w_med + w * kMaxNbr;
  sensors[lane] = -1;
  coef[lane] = 0.f;
  med[lane] = make_float2(0.f, 0.f);
  bool window_ok = s >= 0 && s < sr.S && t0 >= 0 && t1 < sr.Tt;
  const int n0 = window_ok ? sr.nbr_off[s] : 0;
  const int nn = window_ok ? sr.nbr_off[s + 1] - n0 : 0;
  // validity of my neighbour (lane = position in the neighbour list)
  int sensor = -1;
  bool valid = false;
  if (lane < nn) {
    sensor = sr.nbr_idx[n0 + lane];
    const int* pf = sr.nan_prefix + static_cast<long long>(sensor) * (sr.Tt + 1);
    valid = (pf[t1 + 1] - pf[t0]) == 0;
  }
  // the target must be in its own neighbourhood and NaN-free
  const unsigned tgt_mask = __ballot_sync(0xffffffffu, lane < nn && sensor == s);
  const unsigned valid_mask = __ballot_sync(0xffffffffu, valid);
  if (tgt_mask == 0 || !(valid_mask & (tgt_mask & -tgt_mask))) window_ok = false;
  if (!window_ok) {
    if (lane == 0) { w_count[w] = 0; w_tgt_slot[w] = -1; }
    return;
  }
  // compact slot of every valid neighbour (order preserved, like drop_isel in the reference)
  const int slot = valid ? __popc(valid_mask & ((1u << lane) - 1)) : -1;
  const int nv = __popc(valid_mask);
  slot_s[wib][lane] = slot;
  int* M = mult_s[wib];
  for (int i = lane; i < nv * nv; i += 32) M[i] = 0;
  __syncwarp();
  const long long e0 = sr.edge_off[s], ne = sr.edge_off[s + 1] - e0;
  for (long long e = lane; e < ne; e += 32) {
    const long long r = sr.edges[2 * (e0 + e)], c = sr.edges[2 * (e0 + e) + 1];
    if (r >= 0 && r < nn && c >= 0 && c < nn) {
      const int sr_ = slot_s[wib][r], sc = slot_s[wib][c];
      if (sr_ >= 0 && sc >= 0) atomicAdd(&M[sr_ * nv + sc], 1);
    }
  }
  __syncwarp();
  int my_deg = 0;
  if (lane < nv)
    for (int j = 0; j < nv; ++j) my_deg += M[lane * nv + j];
  float acc = 0.f;
  for (int i = 0; i < nv; ++i) {
    const int d = __shfl_sync(0xffffffffu, my_deg, i);
    if (lane < nv && d > 0) {
      const int m = M[i * nv + lane];
      if (m) acc += static_cast<float>(m) / static_cast<float>(d);
    }
  }
  if (lane < nv) coef[lane] = acc * (1.f / static_cast<float>(nv));   // same rounding as cml_node_coef_kernel
  if (valid) {
    sensors[slot] = sensor;
    med[slot] = sr.rmed[static_cast<long long>(sensor) * sr.Tt + t_anom[w]];
  }
  if (lane == 0) {
    w_count[w] = nv;
    const int tl = __ffs(tgt_mask) - 1;
    w_tgt_slot[w] = __popc(valid_mask & ((1u << tl) - 1));
  }
}

// x1 tile from the resident series: channels 0..1 target series minus its median, 2..17 pooled conv.
__global__ void __launch_bounds__(256) series_frontend_kernel(const gnnqc_series sr, const int* __restrict__ t_anom, int before,
                                                              const int* __restrict__ w_sensor, const float* __restrict__ w_coef,
                                                              const float2* __restrict__ w_med, const int* __restrict__ w_count,
                                                              const int* __restrict__ w_tgt_slot, const GcnFolded p, float* __restrict__ out,
                                                              long long n_windows, int t_run, float alpha = 1.f) {
  // The window's node list (sensor row, coefficient, rolling median) is the same for all its steps:
  // staged once per CTA, one padded row per window so lanes hit distinct banks.
  constexpr int STR = kMaxNbr + 1;
  __shared__ int sens_s[kTile * STR];
  __shared__ float coef_s[kTile * STR];
  __shared__ float2 med_s[kTile * STR];
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long win = tile * kTile + b;
  const int tt0 = blockIdx.y * 32 + tq * 8;
  const bool live = win < n_windows && w_count[win] > 0;
  const int nn = live ? w_count[win] : 0;
  const long long base_t = live ? static_cast<long long>(t_anom[win]) - before : 0;
  for (int n = tq; n < nn; n += 4) {
    sens_s[b * STR + n] = __ldg(w_sensor + win * kMaxNbr + n);
    coef_s[b * STR + n] = __ldg(w_coef + win * kMaxNbr + n);
    med_s[b * STR + n] = __ldg(w_med + win * kMaxNbr + n);
  }
  __syncthreads();
  const int tslot = live ? w_tgt_slot[win] : 0;
  constexpr int CH = 2 + kGcnChannelsDev;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = tt0 + dt;
    if (t >= t_run) break;
    PoolAcc acc;
    float2 tv = make_float2(0.f, 0.f);
#pragma unroll 2
    for (int n = 0; n < nn; ++n) {
      const float2 raw = __ldg(sr.raw + static_cast<long long>(sens_s[b * STR + n]) * sr.Tt + base_t + t);
      const float2 m = med_s[b * STR + n];
      float2 v = make_float2(raw.x - m.x, raw.y - m.y);
      v.x *= alpha; v.y *= alpha;   // integrated-gradients path scaling (as cml_frontend_body); exact no-op for alpha == 1
      if (n == tslot) tv = v;
      acc.add_packed(p, v, coef_s[b * STR + n]);
    }
    float* o = out + ((tile * t_run + t) * CH) * kTile + b;
    o[0] = tv.x;
    o[kTile] = tv.y;
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) o[(2 + c) * kTile] = acc.pooled(p, c);
  }
}

// Integrated gradients, backward through the front-end for windows that live in the resident series: the arithmetic of
// cml_frontend_bwd_multi_kernel (ig_frontend_bwd.cuh, a whole group of interpolation points per pass) with x = raw - median
// read from the network.  Attributions go to dense, zero-padded per-window blocks: attr_nodes [window][t][kMaxNbr][2]
// (node = position in the window's node list, which node_sensor names), attr_anom [window][t][2].
// Grid (ceil(windows / 8), T), 256 threads = 8 windows x kMaxNbr node slots.
__global__ void __launch_bounds__(256) series_frontend_bwd_kernel(const gnnqc_series sr, const int* __restrict__ t_anom, int before,
                                                                  const int* __restrict__ w_sensor, const float* __restrict__ w_coef,
                                                                  const float2* __restrict__ w_med, const int* __restrict__ w_count,
                                                                  const int* __restrict__ w_tgt_slot, const GcnFolded p, const IgGroup g,
                                                                  float2* __restrict__ attr_nodes, float2* __restrict__ attr_anom,
                                                                  long long n_windows, int T, int t_run) {
  extern __shared__ __align__(16) unsigned char ig_raw[];
  const IgSmem sm = ig_smem(ig_raw, g.nk);
  const long long win0 = static_cast<long long>(blockIdx.x) * kIgWin;
  const int t = blockIdx.y;
  if (t >= t_run && !g.first) return;
  ig_stage(g, win0, t, t_run, sm);
  __syncthreads();
  const int w_l = threadIdx.x / kMaxNbr, n = threadIdx.x % kMaxNbr;
  const long long win = win0 + w_l;
  if (win >= n_windows) return;
  const int nn = w_count[win];
  const long long base_t = static_cast<long long>(t_anom[win]) - before;
  float2* out = attr_nodes + (win * T + t) * kMaxNbr + n;
  if (t >= t_run || n >= nn) {                    // padding, skipped windows and steps that cannot reach the score: 0
    if (g.first) *out = make_float2(0.f, 0.f);
  } else {
    const float2 raw = __ldg(sr.raw + static_cast<long long>(w_sensor[win * kMaxNbr + n]) * sr.Tt + base_t + t);
    const float2 med = w_med[win * kMaxNbr + n];
    const float2 v = make_float2(raw.x - med.x, raw.y - med.y);
    const float cf = w_coef[win * kMaxNbr + n];
    const float2 s = ig_row(p, g, v.x, v.y, sm, w_l);
    float2 a = g.first ? make_float2(0.f, 0.f) : *out;
    a.x = fmaf(g.scale_by_input ? v.x * cf : cf, s.x, a.x);
    a.y = fmaf(g.scale_by_input ? v.y * cf : cf, s.y, a.y);
    *out = a;
  }
  if (n == 0) {
    float2 a = make_float2(0.f, 0.f);
    if (t < t_run && nn > 0) {
      const int tslot = w_tgt_slot[win];
      const float2 raw = __ldg(sr.raw + static_cast<long long>(w_sensor[win * kMaxNbr + tslot]) * sr.Tt + base_t + t);
      const float2 med = w_med[win * kMaxNbr + tslot];
      if (!g.first) a = attr_anom[win * T + t];
      const float2 ga = ig_anom(g, sm, w_l);
      a.x = fmaf(g.scale_by_input ? raw.x - med.x : 1.f, ga.x, a.x);
      a.y = fmaf(g.scale_by_input ? raw.y - med.y : 1.f, ga.y, a.y);
    }
    attr_anom[win * T + t] = a;
  }
}

__global__ void __launch_bounds__(256) series_mask_scores_kernel(float* __restrict__ scores, const int* __restrict__ w_count, int* __restrict__ n_nodes_out,
                                                                 long long n_windows) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_windows) return;
  const int c = w_count[i];
  if (c == 0) scores[i] = nanf("");
  if (n_nodes_out) n_nodes_out[i] = c;
}


void free_series(gnnqc_series* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaFree(s->raw); cudaFree(s->rmed); cudaFree(s->nan_prefix); cudaFree(s->nbr_off); cudaFree(s->nbr_idx);
  cudaFree(s->edge_off); cudaFree(s->edges); cudaFree(s->scan_tmp);
  delete s;
}

// Stream-ordered; the cub scratch is allocated once (at creation) and reused by every update.
void rebuild_nan_prefix(gnnqc_series* s, cudaStream_t st) {
  const long long n = s->S * (s->Tt + 1);
  nan_flag_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(s->raw, s->nan_prefix, s->S, s->Tt);
  GNNQC_CUDA(cudaGetLastError());
  size_t tmp_bytes = 0;
  GNNQC_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, s->nan_prefix, s->nan_prefix, static_cast<int>(n), st));
  if (tmp_bytes > s->scan_tmp_bytes) {
    GNNQC_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->scan_tmp);
    s->scan_tmp = nullptr;
    GNNQC_CUDA(cudaMalloc(&s->scan_tmp, tmp_bytes));
    s->scan_tmp_bytes = tmp_bytes;
  }
  GNNQC_CUDA(cub::DeviceScan::InclusiveSum(s->scan_tmp, tmp_bytes, s->nan_prefix, s->nan_prefix, static_cast<int>(n), st));
}

}  // namespace
}  // namespace gnnqc

extern "C" {

int gnnqc_series_create(int device, const float* raw_host, const float* rmed_host, int64_t n_sensors, int64_t n_times,
                        const int32_t* nbr_off_host, const int32_t* nbr_idx_host, const int64_t* edge_off_host, const int64_t* edges_host,
                        gnnqc_series** out) {
  if (!out) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_series_create: null out");
  *out = nullptr;
  if (!raw_host || !rmed_host || !nbr_off_host || !nbr_idx_host || !edge_off_host || !edges_host || n_sensors <= 0 || n_times <= 0)
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_series_create: bad argument");
  if (n_sensors * (n_times + 1) >= (int64_t(1) << 31))
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_series_create: more than 2^31 samples per handle; shard the sensors over GPUs");
  gnnqc_series* s = new gnnqc_series();
  try {
    s->device = device;
    s->S = n_sensors;
    s->Tt = n_times;
    s->n_nbr = nbr_off_host[n_sensors];
    s->n_edges = edge_off_host[n_sensors];
    for (int64_t i = 0; i < n_sensors; ++i) s->max_nbr = std::max(s->max_nbr, nbr_off_host[i + 1] - nbr_off_host[i]);
    if (s->max_nbr > gnnqc::kMaxNbr)
      throw std::invalid_argument("gnnqc_series_create: a neighbourhood has " + std::to_string(s->max_nbr) + " sensors; at most " +
                                  std::to_string(gnnqc::kMaxNbr) + " are supported by the window builder");
    GNNQC_CUDA(cudaSetDevice(device));
    const size_t nb = size_t(n_sensors) * n_times * sizeof(float2);
    GNNQC_CUDA(cudaMalloc(&s->raw, nb));
    GNNQC_CUDA(cudaMalloc(&s->rmed, nb));
    GNNQC_CUDA(cudaMalloc(&s->nan_prefix, size_t(n_sensors) * (n_times + 1) * sizeof(int)));
    GNNQC_CUDA(cudaMalloc(&s->nbr_off, size_t(n_sensors + 1) * sizeof(int)));
    GNNQC_CUDA(cudaMalloc(&s->nbr_idx, std::max<size_t>(s->n_nbr, 1) * sizeof(int)));
    GNNQC_CUDA(cudaMalloc(&s->edge_off, size_t(n_sensors + 1) * sizeof(long long)));
    GNNQC_CUDA(cudaMalloc(&s->edges, std::max<size_t>(s->n_edges, 1) * 2 * sizeof(long long)));
    GNNQC_CUDA(cudaMemcpy(s->raw, raw_host, nb, cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->rmed, rmed_host, nb, cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->nbr_off, nbr_off_host, size_t(n_sensors + 1) * sizeof(int), cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->nbr_idx, nbr_idx_host, size_t(s->n_nbr) * sizeof(int), cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->edge_off, edge_off_host, size_t(n_sensors + 1) * sizeof(long long), cudaMemcpyHostToDevice));
    GNNQC_CUDA(cudaMemcpy(s->edges, edges_host, size_t(s->n_edges) * 2 * sizeof(long long), cudaMemcpyHostToDevice));
    gnnqc::rebuild_nan_prefix(s, nullptr);
    GNNQC_CUDA(cudaDeviceSynchronize());
  } catch (const gnnqc::CudaError& e) {
    gnnqc::free_series(s);
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    gnnqc::free_series(s);
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
  *out = s;
  return GNNQC_OK;
}

int gnnqc_series_update_async(gnnqc_series* s, int64_t t0, int64_t n_t, const float* raw_host, const float* rmed_host, void* stream) {
  if (!s || !raw_host || !rmed_host || t0 < 0 || n_t <= 0 || t0 + n_t > s->Tt) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_series_update: bad range");
  try {
    GNNQC_CUDA(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // host slab is [S][n_t][2]; device rows have pitch Tt
    GNNQC_CUDA(cudaMemcpy2DAsync(s->raw + t0, size_t(s->Tt) * sizeof(float2), raw_host, size_t(n_t) * sizeof(float2), size_t(n_t) * sizeof(float2),
                                 size_t(s->S), cudaMemcpyHostToDevice, st));
    GNNQC_CUDA(cudaMemcpy2DAsync(s->rmed + t0, size_t(s->Tt) * sizeof(float2), rmed_host, size_t(n_t) * sizeof(float2), size_t(n_t) * sizeof(float2),
                                 size_t(s->S), cudaMemcpyHostToDevice, st));
    gnnqc::rebuild_nan_prefix(s, st);
  } catch (const std::exception& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  }
  return GNNQC_OK;
}

int gnnqc_series_update(gnnqc_series* s, int64_t t0, int64_t n_t, const float* raw_host, const float* rmed_host) {
  const int rc = gnnqc_series_update_async(s, t0, n_t, raw_host, rmed_host, nullptr);
  if (rc != GNNQC_OK) return rc;
  return cudaStreamSynchronize(nullptr) == cudaSuccess ? GNNQC_OK : gnnqc_internal_fail(GNNQC_ERR_CUDA, "gnnqc_series_update: copy failed");
}

void gnnqc_series_free(gnnqc_series* s) { gnnqc::free_series(s); }

int gnnqc_score_cml_series(gnnqc_model* m, const gnnqc_series* series, const int32_t* target_dev, const int32_t* t_anom_dev, int64_t B,
                           int32_t timestep_before, int32_t timestep_after, float* scores_dev, int32_t* n_nodes_dev, void* stream) {
  if (!m || !series) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_cml_series: null handle");
  gnnqc::DeviceModel& dm = m->dm;
  if (dm.device < 0) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_cml_series: handle was opened host-only; there is no CPU path");
  if (dm.device != series->device) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_cml_series: model and series live on different GPUs");
  if (!dm.host.spec.has_gcn || dm.host.spec.soilnet) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_cml_series: model is not a CML GCNClassifier");
  if (B < 0 || timestep_before < 0 || timestep_after < 0) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_cml_series: bad sizes");
  if (B == 0) return GNNQC_OK;
  if (!target_dev || !t_anom_dev || !scores_dev) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_cml_series: null buffer");
  try {
    using namespace gnnqc;
    dm.refresh_tc();
    GNNQC_CUDA(cudaSetDevice(dm.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int T = timestep_before + timestep_after + 1;
    StackPlan plan(dm.host.spec, T);
    const long long n_tiles = (B + kTile - 1) / kTile;
    const size_t sa = align_up(plan.bytes_a, 256), sb = align_up(plan.bytes_b, 256);
    const size_t per_win = align_up(size_t(kMaxNbr) * (4 + 4 + 8) + 8, 16);
    const char* env = std::getenv("GNNQC_SCRATCH_MB");
    const size_t budget = std::max<size_t>(env ? std::strtoull(env, nullptr, 10) : 8192, 16) << 20;
    const long long chunk = balanced_chunk(n_tiles, static_cast<long long>(budget / (sa + sb)));
    const size_t meta = align_up(size_t(B) * kMaxNbr * 4, 256) * 2 + align_up(size_t(B) * kMaxNbr * 8, 256) + 2 * align_up(size_t(B) * 4, 256);
    (void)per_win;
    char* base = static_cast<char*>(dm.scratch(meta + size_t(chunk) * (sa + sb)));
    char* p = base;
    int* w_sensor = reinterpret_cast<int*>(p); p += align_up(size_t(B) * kMaxNbr * 4, 256);
    float* w_coef = reinterpret_cast<float*>(p); p += align_up(size_t(B) * kMaxNbr * 4, 256);
    float2* w_med = reinterpret_cast<float2*>(p); p += align_up(size_t(B) * kMaxNbr * 8, 256);
    int* w_count = reinterpret_cast<int*>(p); p += align_up(size_t(B) * 4, 256);
    int* w_tgt = reinterpret_cast<int*>(p); p += align_up(size_t(B) * 4, 256);
    float* buf_a = reinterpret_cast<float*>(p);
    float* buf_b = reinterpret_cast<float*>(p + size_t(chunk) * sa);

    dm.mark_begin("series_prepare", st);
    series_prepare_kernel<<<static_cast<unsigned>((B * 32 + 255) / 256), 256, 0, st>>>(*series, target_dev, t_anom_dev, timestep_before, timestep_after, B,
                                                                                       w_sensor, w_coef, w_med, w_count, w_tgt);
    GNNQC_CUDA(cudaGetLastError());
    dm.mark_end(st);
    const GcnFolded folded = dm.folded(st);
    for (long long t0 = 0; t0 < n_tiles; t0 += chunk) {
      const long long nt = std::min(chunk, n_tiles - t0);
      const long long win0 = t0 * kTile, nw = std::min<long long>(B - win0, nt * kTile);
      dim3 grid(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
      dm.mark_begin("series_frontend", st);
      series_frontend_kernel<<<grid, 256, 0, st>>>(*series, t_anom_dev + win0, timestep_before, w_sensor + win0 * kMaxNbr, w_coef + win0 * kMaxNbr,
                                                   w_med + win0 * kMaxNbr, w_count + win0, w_tgt + win0, folded, buf_a, nw, plan.t_run[0]);
      GNNQC_CUDA(cudaGetLastError());
      dm.mark_end(st);
      dm.run_stack(plan, buf_a, buf_b, nt, nw, scores_dev + win0, st);
    }
    series_mask_scores_kernel<<<static_cast<unsigned>((B + 255) / 256), 256, 0, st>>>(scores_dev, w_count, n_nodes_dev, B);
    GNNQC_CUDA(cudaGetLastError());
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
}

int gnnqc_ig_cml_series(gnnqc_model* m, const gnnqc_series* series, const int32_t* target_dev, const int32_t* t_anom_dev, int64_t B,
                        int32_t timestep_before, int32_t timestep_after, int32_t m_steps, float* attr_nodes_dev, float* attr_anom_dev,
                        float* scores_dev, int32_t* node_sensor_dev, int32_t* n_nodes_dev, void* stream) {
  if (!m || !series) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_ig_cml_series: null handle");
  gnnqc::DeviceModel& dm = m->dm;
  if (dm.device < 0) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_ig_cml_series: host-only model handle - there is no CPU path");
  if (!dm.host.spec.has_gcn || dm.host.spec.soilnet) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_ig_cml_series: model is not a CML GCNClassifier");
  if (dm.device != series->device) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_ig_cml_series: model and series live on different devices");
  if (B < 0 || timestep_before < 0 || timestep_after < 0 || m_steps < 1) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_ig_cml_series: bad sizes");
  if (B == 0) return GNNQC_OK;
  if (!target_dev || !t_anom_dev || !attr_nodes_dev || !attr_anom_dev || !scores_dev) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_ig_cml_series: null buffer");
  try {
    using namespace gnnqc;
    GNNQC_CUDA(cudaSetDevice(dm.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int T = timestep_before + timestep_after + 1;
    const StackPlan plan(dm.host.spec, T);
    const size_t b_sens = align_up(size_t(B) * kMaxNbr * 4, 256), b_med = align_up(size_t(B) * kMaxNbr * 8, 256), b_one = align_up(size_t(B) * 4, 256);
    const GcnFolded folded = dm.folded(st);
    int *w_sensor = nullptr, *w_count = nullptr, *w_tgt = nullptr;
    float* w_coef = nullptr;
    float2* w_med = nullptr;
    dm.ig_core(B, T, m_steps, scores_dev, st, 2 * b_sens + b_med + 2 * b_one,
               [&](char* p) {
                 w_sensor = reinterpret_cast<int*>(p); p += b_sens;
                 w_coef = reinterpret_cast<float*>(p); p += b_sens;
                 w_med = reinterpret_cast<float2*>(p); p += b_med;
                 w_count = reinterpret_cast<int*>(p); p += b_one;
                 w_tgt = reinterpret_cast<int*>(p);
                 dm.mark_begin("series_prepare", st);
                 series_prepare_kernel<<<static_cast<unsigned>((B * 32 + 255) / 256), 256, 0, st>>>(*series, target_dev, t_anom_dev, timestep_before,
                                                                                                    timestep_after, B, w_sensor, w_coef, w_med, w_count, w_tgt);
                 GNNQC_CUDA(cudaGetLastError());
                 dm.mark_end(st);
               },
               [&](float alpha, long long win0, long long nw, long long nt, float* x1) {
                 dim3 grid(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
                 series_frontend_kernel<<<grid, 256, 0, st>>>(*series, t_anom_dev + win0, timestep_before, w_sensor + win0 * kMaxNbr, w_coef + win0 * kMaxNbr,
                                                              w_med + win0 * kMaxNbr, w_count + win0, w_tgt + win0, folded, x1, nw, plan.t_run[0], alpha);
               },
               [&](const IgGroup& g, long long win0, long long nw) {
                 static const bool configured = [] {
                   return cudaFuncSetAttribute(series_frontend_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ig_bwd_smem_bytes(kIgMaxGroup))) == cudaSuccess;
                 }();
                 (void)configured;
                 const dim3 grid(static_cast<unsigned>((nw + kIgWin - 1) / kIgWin), static_cast<unsigned>(T));
                 series_frontend_bwd_kernel<<<grid, 256, ig_bwd_smem_bytes(g.nk), st>>>(
                     *series, t_anom_dev + win0, timestep_before, w_sensor + win0 * kMaxNbr, w_coef + win0 * kMaxNbr, w_med + win0 * kMaxNbr, w_count + win0,
                     w_tgt + win0, folded, g, reinterpret_cast<float2*>(attr_nodes_dev) + win0 * T * kMaxNbr, reinterpret_cast<float2*>(attr_anom_dev) + win0 * T,
                     nw, T, plan.t_run[0]);
               });
    series_mask_scores_kernel<<<static_cast<unsigned>((B + 255) / 256), 256, 0, st>>>(scores_dev, w_count, n_nodes_dev, B);
    GNNQC_CUDA(cudaGetLastError());
    if (node_sensor_dev) GNNQC_CUDA(cudaMemcpyAsync(node_sensor_dev, w_sensor, size_t(B) * kMaxNbr * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
}

}  // extern "C"
