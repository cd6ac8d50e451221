// SoilNet GCN front-end and scoring entry (gnnqc_score_soilnet).
//
// Restates GCNClassifier.call for ds_type 'soilnet' (reference libs/create_model.py:224-231):
//   x = GeneralConv([features, adjacency])         per node i, per timestep: mean over the edges
//                                                  received by i of PReLU(BN(x_j W + b))
//   x = Concatenate([x, features])                 GCN output first, then the 3 raw channels
//   x = graph_reshape(x, sample_indices)           [sum N_b, T, 19], sequences in (window, node) order
// followed by the shared LSTM stack + head: one score per (window, sensor).
// The graph is static over a window's timesteps (libs/preprocessing_functions.py:861-892 tiles the
// same edge list T times), so a CSR by receiving node is built once per call and the per-timestep
// aggregation gathers the neighbours' features and recomputes their 16-channel projection.
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>

#include "device_model.cuh"
#include "frontend.cuh"
#include "soil_frontend.cuh"

namespace gnnqc {
namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// deg[node] = number of edges received; seq_win[node] = window of the node.
__global__ void __launch_bounds__(256) soil_count_kernel(const long long* __restrict__ edges, const int* __restrict__ node_off,
                                                         const long long* __restrict__ edge_off, int* __restrict__ deg,
                                                         int* __restrict__ seq_win, long long n_windows) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_windows) return;
  const int n0 = node_off[w], nn = node_off[w + 1] - n0;
  const long long e0 = edge_off[w], ne = edge_off[w + 1] - e0;
  for (int i = lane; i < nn; i += 32) seq_win[n0 + i] = static_cast<int>(w);
  for (long long e = lane; e < ne; e += 32) {
    const long long row = edges[2 * (e0 + e)], col = edges[2 * (e0 + e) + 1];
    if (row >= 0 && row < nn && col >= 0 && col < nn) atomicAdd(&deg[n0 + row], 1);
  }
}

// nbr[row_ptr[node] ...] = sending nodes (local index), filled through per-node cursors.
__global__ void __launch_bounds__(256) soil_fill_kernel(const long long* __restrict__ edges, const int* __restrict__ node_off,
                                                        const long long* __restrict__ edge_off, const int* __restrict__ row_ptr,
                                                        int* __restrict__ cursor, int* __restrict__ nbr, long long n_windows) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_windows) return;
  const int n0 = node_off[w], nn = node_off[w + 1] - n0;
  const long long e0 = edge_off[w], ne = edge_off[w + 1] - e0;
  for (long long e = lane; e < ne; e += 32) {
    const long long row = edges[2 * (e0 + e)], col = edges[2 * (e0 + e) + 1];
    if (row >= 0 && row < nn && col >= 0 && col < nn) {
      const int pos = atomicAdd(&cursor[n0 + row], 1);
      nbr[row_ptr[n0 + row] + pos] = static_cast<int>(col);
    }
  }
}

// Sorts every node's neighbour list so the floating-point sum below has a fixed order whatever
// order the caller's edge list (and the atomics above) produced.
__global__ void __launch_bounds__(256) soil_sort_kernel(const int* __restrict__ row_ptr, const int* __restrict__ deg,
                                                        int* __restrict__ nbr, long long n_nodes) {
  const long long s = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (s >= n_nodes) return;
  int* a = nbr + row_ptr[s];
  const int n = deg[s];
  for (int i = 1; i < n; ++i) {
    const int v = a[i];
    int j = i - 1;
    while (j >= 0 && a[j] > v) { a[j + 1] = a[j]; --j; }
    a[j + 1] = v;
  }
}

// x1 tile [tile][t][19][64]: channels 0..15 = neighbour mean of the projected features,
// 16..18 = the node's own raw features.  One thread per (sequence, 8 timesteps).
__global__ void __launch_bounds__(256) soil_frontend_kernel(const float* __restrict__ feat, const int* __restrict__ node_off,
                                                            const int* __restrict__ seq_win, const int* __restrict__ row_ptr,
                                                            const int* __restrict__ deg, const int* __restrict__ nbr,
                                                            const GcnFolded p, float* __restrict__ out, long long seq_base,
                                                            long long n_seq_total, int T, int t_run) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long s = seq_base + tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
  const bool live = s < n_seq_total;
  int n0 = 0, nn = 0, me = 0, d = 0, rp = 0;
  if (live) {
    const int w = seq_win[s];
    n0 = node_off[w];
    nn = node_off[w + 1] - n0;
    me = static_cast<int>(s - n0);
    d = deg[s];
    rp = row_ptr[s];
  }
  const float inv = d > 0 ? 1.f / static_cast<float>(d) : 0.f;
  constexpr int CH = kGcnChannelsDev + 3;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    SoilAcc acc;   // soil_frontend.cuh
    const float* row = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn) * 3;
    for (int e = 0; e < d; ++e) {
      const float* v = row + 3 * __ldg(nbr + rp + e);
      acc.add(p, __ldg(v), __ldg(v + 1), __ldg(v + 2));
    }
    float* o = out + ((tile * t_run + t) * CH) * kTile + b;
    const float fd = static_cast<float>(d);
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) o[c * kTile] = acc.mean(p, c, fd, inv);
    const float* own = row + 3 * me;
    o[16 * kTile] = live ? __ldg(own) : 0.f;
    o[17 * kTile] = live ? __ldg(own + 1) : 0.f;
    o[18 * kTile] = live ? __ldg(own + 2) : 0.f;
  }
}

GcnFolded fold(const HostModel& m) {
  const ModelSpec& s = m.spec;
  const float* p = m.params.data();
  GcnFolded f;
  std::memset(&f, 0, sizeof(f));
  for (int c = 0; c < kGcnChannels; ++c) {
    const float scale = p[s.off_bn_gamma + c] / std::sqrt(p[s.off_bn_var + c] + s.bn_eps);
    for (int k = 0; k < s.n_features; ++k) f.a[k][c] = p[s.off_gcn_kernel + k * kGcnChannels + c] * scale;
    f.c[c] = (p[s.off_gcn_bias + c] - p[s.off_bn_mean + c]) * scale + p[s.off_bn_beta + c];
    f.alpha[c] = p[s.off_prelu + c];
  }
  return f;
}

size_t budget_bytes() {
  const char* env = std::getenv("GNNQC_SCRATCH_MB");
  size_t mb = env ? std::strtoull(env, nullptr, 10) : 8192;
  return std::max<size_t>(mb, 16) << 20;
}

}  // namespace

void DeviceModel::score_soilnet(const float* feat, const long long* edges, const int* node_off, const long long* edge_off,
                                long long B, int T, long long n_nodes, long long n_edges, float* scores, cudaStream_t st) {
  if (B <= 0 || n_nodes <= 0) return;
  const ModelSpec& s = host.spec;
  GNNQC_CUDA(cudaSetDevice(device));
  StackPlan plan(s, T);
  const long long n_tiles = (n_nodes + kTile - 1) / kTile;
  const size_t sa = align_up(plan.bytes_a, 256), sb = align_up(plan.bytes_b, 256);
  const size_t nb = align_up(size_t(n_nodes + 1) * sizeof(int), 256);
  size_t scan_bytes = 0;
  GNNQC_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, static_cast<int*>(nullptr), static_cast<int*>(nullptr), static_cast<int>(n_nodes + 1), st));
  scan_bytes = align_up(scan_bytes, 256);
  const size_t nbr_bytes = align_up(size_t(std::max<long long>(n_edges, 1)) * sizeof(int), 256);
  const size_t graph_bytes = 4 * nb + scan_bytes + nbr_bytes;
  const long long chunk = balanced_chunk(n_tiles, static_cast<long long>(budget_bytes() / (sa + sb)));
  char* base = static_cast<char*>(scratch(graph_bytes + size_t(chunk) * (sa + sb)));
  int* deg = reinterpret_cast<int*>(base);
  int* cursor = reinterpret_cast<int*>(base + nb);
  int* row_ptr = reinterpret_cast<int*>(base + 2 * nb);
  int* seq_win = reinterpret_cast<int*>(base + 3 * nb);
  void* scan_tmp = base + 4 * nb;
  int* nbr = reinterpret_cast<int*>(base + 4 * nb + scan_bytes);
  float* buf_a = reinterpret_cast<float*>(base + graph_bytes);
  float* buf_b = reinterpret_cast<float*>(base + graph_bytes + size_t(chunk) * sa);

  GNNQC_CUDA(cudaMemsetAsync(deg, 0, 2 * nb, st));   // deg and cursor
  const unsigned wblocks = static_cast<unsigned>((B * 32 + 255) / 256);
  mark_begin("soil_csr_count", st);
  soil_count_kernel<<<wblocks, 256, 0, st>>>(edges, node_off, edge_off, deg, seq_win, B);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);
  GNNQC_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, deg, row_ptr, static_cast<int>(n_nodes + 1), st));
  mark_begin("soil_csr_fill", st);
  soil_fill_kernel<<<wblocks, 256, 0, st>>>(edges, node_off, edge_off, row_ptr, cursor, nbr, B);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);
  mark_begin("soil_csr_sort", st);
  soil_sort_kernel<<<static_cast<unsigned>((n_nodes + 255) / 256), 256, 0, st>>>(row_ptr, deg, nbr, n_nodes);
  GNNQC_CUDA(cudaGetLastError());
  mark_end(st);

  const GcnFolded folded = fold(host);
  for (long long t0 = 0; t0 < n_tiles; t0 += chunk) {
    const long long nt = std::min(chunk, n_tiles - t0);
    const long long seq0 = t0 * kTile, ns = std::min<long long>(n_nodes - seq0, nt * kTile);
    dim3 grid(static_cast<unsigned>(nt), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
    mark_begin("soil_frontend", st);
    soil_frontend_kernel<<<grid, 256, 0, st>>>(feat, node_off, seq_win, row_ptr, deg, nbr, folded, buf_a, seq0, n_nodes, T, plan.t_run[0]);
    GNNQC_CUDA(cudaGetLastError());
    mark_end(st);
    run_stack(plan, buf_a, buf_b, nt, ns, scores + seq0, st);
  }
}

}  // namespace gnnqc
