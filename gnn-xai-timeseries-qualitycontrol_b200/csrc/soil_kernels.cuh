// Kernels of the SoilNet graph front-end shared by scoring (soilnet.cu) and training (train.cu): CSR of the per-window
// adjacency by receiving node (or, transposed, by sending node), and the gather-and-project front-end itself
// (reference libs/create_model.py:224-231; spektral GeneralConv with mean aggregation).
#pragma once
#include "frontend.cuh"
#include "soil_frontend.cuh"

namespace gnnqc {

// deg[node] = number of edges received; seq_win[node] = window of the node.
static __global__ void __launch_bounds__(256) soil_count_kernel(const long long* __restrict__ edges, const int* __restrict__ node_off,
                                                                const long long* __restrict__ edge_off, int* __restrict__ deg,
                                                                int* __restrict__ seq_win, long long n_windows, int transpose = 0) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_windows) return;
  const int n0 = node_off[w], nn = node_off[w + 1] - n0;
  const long long e0 = edge_off[w], ne = edge_off[w + 1] - e0;
  for (int i = lane; i < nn; i += 32) seq_win[n0 + i] = static_cast<int>(w);
  for (long long e = lane; e < ne; e += 32) {
    long long row = edges[2 * (e0 + e)], col = edges[2 * (e0 + e) + 1];
    if (transpose) { const long long tmp = row; row = col; col = tmp; }     // lists of RECEIVERS per sending node (backward pass)
    if (row >= 0 && row < nn && col >= 0 && col < nn) atomicAdd(&deg[n0 + row], 1);
  }
}

// nbr[row_ptr[node] ...] = sending nodes (local index), filled through per-node cursors.
static __global__ void __launch_bounds__(256) soil_fill_kernel(const long long* __restrict__ edges, const int* __restrict__ node_off,
                                                               const long long* __restrict__ edge_off, const int* __restrict__ row_ptr,
                                                               int* __restrict__ cursor, int* __restrict__ nbr, long long n_windows, int transpose = 0) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_windows) return;
  const int n0 = node_off[w], nn = node_off[w + 1] - n0;
  const long long e0 = edge_off[w], ne = edge_off[w + 1] - e0;
  for (long long e = lane; e < ne; e += 32) {
    long long row = edges[2 * (e0 + e)], col = edges[2 * (e0 + e) + 1];
    if (transpose) { const long long tmp = row; row = col; col = tmp; }
    if (row >= 0 && row < nn && col >= 0 && col < nn) {
      const int pos = atomicAdd(&cursor[n0 + row], 1);
      nbr[row_ptr[n0 + row] + pos] = static_cast<int>(col);
    }
  }
}

// Sorts every node's neighbour list so the floating-point sum below has a fixed order whatever
// order the caller's edge list (and the atomics above) produced.
static __global__ void __launch_bounds__(256) soil_sort_kernel(const int* __restrict__ row_ptr, const int* __restrict__ deg,
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
__device__ __forceinline__ void soil_frontend_body(const float* __restrict__ feat, const int* __restrict__ node_off,
                                                   const int* __restrict__ seq_win, const int* __restrict__ row_ptr,
                                                   const int* __restrict__ deg, const int* __restrict__ nbr, const GcnFolded& p,
                                                   float* __restrict__ out, long long seq_base, long long n_seq_total, int T, int t_run) {
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

static __global__ void __launch_bounds__(256) soil_frontend_kernel(const float* __restrict__ feat, const int* __restrict__ node_off,
                                                                   const int* __restrict__ seq_win, const int* __restrict__ row_ptr,
                                                                   const int* __restrict__ deg, const int* __restrict__ nbr,
                                                                   const GcnFolded p, float* __restrict__ out, long long seq_base,
                                                                   long long n_seq_total, int T, int t_run) {
  soil_frontend_body(feat, node_off, seq_win, row_ptr, deg, nbr, p, out, seq_base, n_seq_total, T, t_run);
}

// Same with the folded projection read from device memory (training: it is derived from this step's batch statistics on the device).
static __global__ void __launch_bounds__(256) soil_frontend_devparam_kernel(const float* __restrict__ feat, const int* __restrict__ node_off,
                                                                            const int* __restrict__ seq_win, const int* __restrict__ row_ptr,
                                                                            const int* __restrict__ deg, const int* __restrict__ nbr,
                                                                            const GcnFolded* __restrict__ pp, float* __restrict__ out,
                                                                            long long seq_base, long long n_seq_total, int T, int t_run) {
  __shared__ GcnFolded p;
  for (int i = threadIdx.x; i < static_cast<int>(sizeof(GcnFolded) / 4); i += blockDim.x) reinterpret_cast<float*>(&p)[i] = reinterpret_cast<const float*>(pp)[i];
  __syncthreads();
  soil_frontend_body(feat, node_off, seq_win, row_ptr, deg, nbr, p, out, seq_base, n_seq_total, T, t_run);
}

}  // namespace gnnqc
