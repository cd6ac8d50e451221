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
#include "soil_kernels.cuh"

namespace gnnqc {
namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }


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

  const GcnFolded folded = this->folded(st);
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
