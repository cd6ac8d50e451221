// Kernels either side of the LSTM stack: layout conversion into sequence tiles, the CML graph
// front-end (GeneralConv + per-window node pooling + concat), and the dense head.
#pragma once
#include <cstdint>

#include "lstm_fwd.cuh"
#include "model_spec.h"

namespace gnnqc {

constexpr int kGcnChannelsDev = 16;  // GeneralConv channels (model_config.yml graph_convolution.units)

// ---- [n_seq, T, C] row-major  ->  sequence tiles [tile][t][c][64], t < t_run ---------------
// Input contract of the baseline models: libs/preprocessing_functions.py:756-760 (CML,
// [B,181,2]) and graph_reshape's output (libs/create_model.py:242-258, SoilNet [S,337,3]).
template <int C>
__global__ void __launch_bounds__(256) seq_to_tiles_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                           long long n_seq, int T, int t_run) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long seq = tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    float v[C];
#pragma unroll
    for (int c = 0; c < C; ++c) v[c] = (seq < n_seq) ? __ldg(x + (seq * T + t) * C + c) : 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) out[((tile * t_run + t) * C + c) * kTile + b] = v[c];
  }
}

// ---- CML node coefficients -----------------------------------------------------------------
// GeneralConv aggregates with a mean over each receiving node's edges (spektral 1.3.0
// MessagePassing: gather by a.indices[:,1], unsorted_segment_mean by a.indices[:,0]; call
// site libs/create_model.py:185-189,219) and timeseries_pooling then averages over the
// window's nodes (libs/create_model.py:23-27).  Both are linear, so
//   pooled[t] = sum_j coef_j * z_j[t],   coef_j = (1/N) * sum_{(i,j) in E} 1/deg_i,
// with deg_i the number of edges received by node i.  One warp per window; `deg` is integer
// scratch that must be zero on entry.
constexpr int kCoefDenseMax = 32;  // windows with at most this many nodes use the shared-memory path

static __global__ void __launch_bounds__(256) cml_node_coef_kernel(const long long* __restrict__ edges,
                                                                   const int* __restrict__ node_off,
                                                                   const long long* __restrict__ edge_off,
                                                                   int* __restrict__ deg, float* __restrict__ coef,
                                                                   long long n_windows) {
  __shared__ int mult_s[8][kCoefDenseMax * kCoefDenseMax];  // per warp: edge multiplicity M[row][col]
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  if (w >= n_windows) return;
  const int n0 = node_off[w], nn = node_off[w + 1] - n0;
  const long long e0 = edge_off[w], ne = edge_off[w + 1] - e0;
  const float inv_n = nn > 0 ? 1.f / static_cast<float>(nn) : 0.f;
  if (nn <= kCoefDenseMax) {
    // Small graph (the CML case, N_b <= 23): count edges into a dense N x N table with integer
    // shared-memory atomics (order-independent), then lane j sums column j in row order, so the
    // result is bit-reproducible whatever order the edge list is in.
    int* M = mult_s[wib];
    for (int i = lane; i < nn * nn; i += 32) M[i] = 0;
    __syncwarp();
    for (long long e = lane; e < ne; e += 32) {
      const longlong2 rc = *reinterpret_cast<const longlong2*>(edges + 2 * (e0 + e));
      if (rc.x >= 0 && rc.x < nn && rc.y >= 0 && rc.y < nn) atomicAdd(&M[static_cast<int>(rc.x) * nn + static_cast<int>(rc.y)], 1);
    }
    __syncwarp();
    // deg_i = row sum; every lane needs all of them: lane i computes deg_i, shuffle to share.
    int my_deg = 0;
    if (lane < nn)
      for (int j = 0; j < nn; ++j) my_deg += M[lane * nn + j];
    float s = 0.f;
    for (int i = 0; i < nn; ++i) {
      const int d = __shfl_sync(0xffffffffu, my_deg, i);
      if (lane < nn && d > 0) {
        const int m = M[i * nn + lane];
        if (m) s += static_cast<float>(m) / static_cast<float>(d);
      }
    }
    if (lane < nn) coef[n0 + lane] = s * inv_n;
    return;
  }
  // General path (any graph size): degree counts in global scratch `deg` (zero on entry).
  for (long long e = lane; e < ne; e += 32) {
    const long long row = edges[2 * (e0 + e)];
    if (row >= 0 && row < nn) atomicAdd(&deg[n0 + row], 1);
  }
  __syncwarp();
  __threadfence_block();
  for (int j = lane; j < nn; j += 32) {
    float s = 0.f;
    for (long long e = 0; e < ne; ++e) {
      const long long row = edges[2 * (e0 + e)], col = edges[2 * (e0 + e) + 1];
      if (col == j && row >= 0 && row < nn) s += 1.f / static_cast<float>(__ldcg(deg + n0 + row));
    }
    coef[n0 + j] = s * inv_n;
  }
}

// ---- graph_reshape (libs/create_model.py:242-258) for the SoilNet baseline ---------------------
// feat [sum_b N_b * T, C] rows in (window, t, node) order  ->  sequence tiles of the (window, node) sequences.
// The window of a sequence is found by bisection of node_off (B is a few hundred at most).
template <int C>
__global__ void __launch_bounds__(256) rows_to_tiles_kernel(const float* __restrict__ feat, const int* __restrict__ node_off, int n_windows,
                                                            float* __restrict__ out, long long seq_base, long long n_seq_total, int T, int t_run) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long s = seq_base + tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
  int n0 = 0, nn = 0;
  const bool live = s < n_seq_total;
  if (live) {
    int lo = 0, hi = n_windows;                       // node_off[lo] <= s < node_off[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(node_off + mid) <= s) lo = mid; else hi = mid;
    }
    n0 = __ldg(node_off + lo);
    nn = __ldg(node_off + lo + 1) - n0;
  }
  const int me = static_cast<int>(s - n0);
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    const float* x = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn + me) * C;
    float* o = out + ((tile * t_run + t) * C) * kTile + b;
#pragma unroll
    for (int c = 0; c < C; ++c) o[c * kTile] = live ? __ldg(x + c) : 0.f;
  }
}

// GcnFolded (BatchNorm folded into the GeneralConv projection) is declared in model_spec.h.
static_assert(kGcnChannelsDev == kGcnChannels, "GeneralConv width");

// PReLU pooled over nodes without a branch per element: prelu(z) = k1*z + k2*|z| with
// k1 = (1+alpha)/2, k2 = (1-alpha)/2, and z is affine in the two input channels, so
//   sum_n cf_n prelu(z_n) = k1 * (a0*S0 + a1*S1 + c*Sc) + k2 * sum_n cf_n |z_n|
// with S0 = sum cf_n x0_n, S1 = sum cf_n x1_n, Sc = sum cf_n.  Per (node, channel) that is two
// FFMA for z and one FFMA (|z| is an operand modifier) instead of six instructions.
struct PoolAcc {
  float s0 = 0.f, s1 = 0.f, sc = 0.f;
  float sabs[kGcnChannelsDev];
  __device__ __forceinline__ PoolAcc() {
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) sabs[c] = 0.f;
  }
  __device__ __forceinline__ void add(const GcnFolded& p, float2 v, float cf) {
    s0 = fmaf(cf, v.x, s0);
    s1 = fmaf(cf, v.y, s1);
    sc += cf;
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) {
      const float z = fmaf(v.x, p.a[0][c], fmaf(v.y, p.a[1][c], p.c[c]));
      sabs[c] = fmaf(cf, fabsf(z), sabs[c]);
    }
  }
  // Same update with z formed two channels at a time (fma.rn.f32x2: the scalar input is a broadcast
  // operand and the channel-pair constants stay in uniform registers when `p` is a kernel parameter).
  __device__ __forceinline__ void add_packed(const GcnFolded& p, float2 v, float cf) {
    s0 = fmaf(cf, v.x, s0);
    s1 = fmaf(cf, v.y, s1);
    sc += cf;
    unsigned long long vx, vy;
    asm("mov.b64 %0, {%1, %1};" : "=l"(vx) : "f"(v.x));
    asm("mov.b64 %0, {%1, %1};" : "=l"(vy) : "f"(v.y));
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; c += 2) {
      unsigned long long z, a0, a1;
      asm("mov.b64 %0, {%1, %2};" : "=l"(z) : "f"(p.c[c]), "f"(p.c[c + 1]));
      asm("mov.b64 %0, {%1, %2};" : "=l"(a0) : "f"(p.a[0][c]), "f"(p.a[0][c + 1]));
      asm("mov.b64 %0, {%1, %2};" : "=l"(a1) : "f"(p.a[1][c]), "f"(p.a[1][c + 1]));
      asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(z) : "l"(vy), "l"(a1));
      asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(z) : "l"(vx), "l"(a0));
      float z0, z1;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(z0), "=f"(z1) : "l"(z));
      sabs[c] = fmaf(cf, fabsf(z0), sabs[c]);
      sabs[c + 1] = fmaf(cf, fabsf(z1), sabs[c + 1]);
    }
  }
  __device__ __forceinline__ float pooled(const GcnFolded& p, int c) const {
    const float lin = fmaf(p.a[0][c], s0, fmaf(p.a[1][c], s1, p.c[c] * sc));
    return fmaf(0.5f * (1.f + p.alpha[c]), lin, 0.5f * (1.f - p.alpha[c]) * sabs[c]);
  }
};

// ---- CML front-end: features (b,t,n) + coef + target series -> x1 tiles [tile][t][18][64] --
// Restates GCNClassifier.call for 'cml' up to the Concatenate (libs/create_model.py:217-223):
// channels 0..1 = anomalous_sensor_features, 2..17 = pooled GeneralConv output.
__device__ __forceinline__ void cml_frontend_body(const float2* __restrict__ feat, const float2* __restrict__ anom,
                                                           const float* __restrict__ coef, const int* __restrict__ node_off,
                                                           const GcnFolded& p, float* __restrict__ out,
                                                           long long n_windows, int T, int t_run, float alpha) {
  const int b = threadIdx.x & 63, tq = threadIdx.x >> 6;
  const long long tile = blockIdx.x;
  const long long win = tile * kTile + b;
  const int t0 = blockIdx.y * 32 + tq * 8;
  const bool live = win < n_windows;
  const int n0 = live ? node_off[win] : 0;
  const int nn = live ? node_off[win + 1] - n0 : 0;
  constexpr int CH = 2 + kGcnChannelsDev;
#pragma unroll 1
  for (int dt = 0; dt < 8; ++dt) {
    const int t = t0 + dt;
    if (t >= t_run) break;
    PoolAcc acc;
    const float2* row = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn);
    for (int n = 0; n < nn; ++n) {
      float2 v = __ldg(row + n);
      v.x *= alpha; v.y *= alpha;   // integrated-gradients path scaling; exact no-op for alpha == 1
      acc.add(p, v, __ldg(coef + n0 + n));
    }
    const float2 av = live ? __ldg(anom + win * T + t) : make_float2(0.f, 0.f);
    float* o = out + ((tile * t_run + t) * CH) * kTile + b;
    o[0] = av.x * alpha;
    o[kTile] = av.y * alpha;
#pragma unroll
    for (int c = 0; c < kGcnChannelsDev; ++c) o[(2 + c) * kTile] = acc.pooled(p, c);
  }
}

static __global__ void __launch_bounds__(256) cml_frontend_kernel(const float2* __restrict__ feat, const float2* __restrict__ anom,
                                                                  const float* __restrict__ coef, const int* __restrict__ node_off,
                                                                  const GcnFolded p, float* __restrict__ out, long long n_windows,
                                                                  int T, int t_run, float alpha = 1.f) {
  cml_frontend_body(feat, anom, coef, node_off, p, out, n_windows, T, t_run, alpha);
}

// Same, with the folded projection read from device memory (training: it depends on the batch
// statistics computed on the device a moment earlier).
static __global__ void __launch_bounds__(256) cml_frontend_devparam_kernel(const float2* __restrict__ feat, const float2* __restrict__ anom,
                                                                           const float* __restrict__ coef, const int* __restrict__ node_off,
                                                                           const GcnFolded* __restrict__ pp, float* __restrict__ out,
                                                                           long long n_windows, int T, int t_run) {
  __shared__ GcnFolded p;
  if (threadIdx.x < sizeof(GcnFolded) / 4) reinterpret_cast<float*>(&p)[threadIdx.x] = reinterpret_cast<const float*>(pp)[threadIdx.x];
  __syncthreads();
  cml_frontend_body(feat, anom, coef, node_off, p, out, n_windows, T, t_run, 1.f);
}

// ---- CML front-end, persistent, TMA-staged and double-buffered --------------------------------
// Same arithmetic as cml_frontend_kernel.  A window's slab of 8 timesteps (a contiguous
// 8*N_b*8-byte run of the features array) is fetched with ONE cp.async.bulk (TMA) into its own
// shared-memory slot, the 16 windows of an item signalling one mbarrier; the threads then read the
// slab from shared memory instead of streaming 8-byte pieces of many different rows.  The features
// array only guarantees 8-byte alignment per window (N_b*T can be odd), so the copy starts at the
// enclosing 16-byte boundary and the reader skips the 0 or 8 lead bytes.  Windows with more than
// kFsCap nodes, or whose rounded-up copy would run past the end of the array, take direct loads.
// A CTA of 128 threads (lane = 16 windows x 8 steps) walks a contiguous range of (16-window block,
// step octet) items with two slab buffers, so the copy of the next item is always in flight under
// the arithmetic of the current one, and a buffer is refilled with item i+2 as soon as every
// thread is done with item i.  The block's node coefficients sit in shared memory for its ~20 items.
constexpr int kFsWin = 16;                                  // windows per item
constexpr int kFsSteps = 8;                                 // timesteps per item
constexpr int kFsThreads = kFsWin * kFsSteps;
// CAP = nodes per window staged through smem (windows with more take direct loads).  Two builds: 24 nodes
// (50 kB per CTA, four CTAs per SM) for batches that average at most 24 nodes per window, 32 otherwise
// (three CTAs per SM).
template <int CAP>
struct FsCfg {
  static constexpr int kSlot = kFsSteps * CAP * 8 + 16;     // bytes per window slot (16 * odd)
  static constexpr int kSlab = kFsWin * kSlot;
  static constexpr int kCoefStride = CAP + 1;               // per-window coefficient row in smem (odd: no bank conflicts)
  static constexpr size_t kSmem = 2 * size_t(kSlab) + 2 * kFsWin * sizeof(int) + 2 * 8 + 16 + kFsWin * kCoefStride * sizeof(float);
  static constexpr int kCtasPerSm = CAP <= 24 ? 4 : 3;
};

template <int CAP>
static __global__ void __launch_bounds__(kFsThreads) cml_frontend_stream_kernel(const float2* __restrict__ feat, long long feat_rows,
                                                                                const float2* __restrict__ anom, const float* __restrict__ coef,
                                                                                const int* __restrict__ node_off, const GcnFolded p,
                                                                                float* __restrict__ out, long long n_windows, int T, int t_run,
                                                                                int n_tiles, float alpha) {
  extern __shared__ __align__(128) unsigned char fs_smem[];
  constexpr int kFsCap = CAP, kFsSlot = FsCfg<CAP>::kSlot, kFsSlab = FsCfg<CAP>::kSlab, kFsCoefStride = FsCfg<CAP>::kCoefStride;
  int* lead_s = reinterpret_cast<int*>(fs_smem + 2 * kFsSlab);                         // [2][kFsWin]
  uint64_t* full = reinterpret_cast<uint64_t*>(fs_smem + 2 * kFsSlab + 2 * kFsWin * sizeof(int));   // [2]
  float* coef_s = reinterpret_cast<float*>(fs_smem + 2 * kFsSlab + 2 * kFsWin * sizeof(int) + 32);  // [kFsWin][kFsCoefStride]
  const int r = threadIdx.x, tq = r / kFsWin, bl = r % kFsWin;
  const int NTO = (t_run + kFsSteps - 1) / kFsSteps;
  const long long items = static_cast<long long>(n_tiles) * (kTile / kFsWin) * NTO;
  const long long i0 = items * blockIdx.x / gridDim.x, i1 = items * (blockIdx.x + 1) / gridDim.x;
  if (r == 0) {
    ptx::mbar_init(&full[0], kFsWin);
    ptx::mbar_init(&full[1], kFsWin);
    ptx::mbar_init_fence();
  }
  __syncthreads();

  // thread r < 16 requests window r's slab of an item into buffer s.
  // lead_s: >= 0 byte offset of the first row inside the slot (0 for a window without nodes), -1 direct loads.
  auto issue_slab = [&](long long item, int s) {
    const long long blk_n = item / NTO;
    const int t0n = static_cast<int>(item % NTO) * kFsSteps;
    const int tchn = min(kFsSteps, t_run - t0n);
    const long long winn = blk_n * kFsWin + r;
    int lead = 0;
    bool copied = false;
    if (winn < n_windows) {
      const int n0n = node_off[winn], nnn = node_off[winn + 1] - n0n;
      if (nnn > 0) {
        lead = -1;
        const long long first_row = static_cast<long long>(n0n) * T + static_cast<long long>(t0n) * nnn;   // float2 index
        const unsigned long long addr = reinterpret_cast<unsigned long long>(feat + first_row);
        const int ld = static_cast<int>(addr & 15ull);
        const unsigned bytes = (static_cast<unsigned>(ld) + static_cast<unsigned>(tchn * nnn) * 8u + 15u) & ~15u;
        const unsigned long long end = reinterpret_cast<unsigned long long>(feat + feat_rows);
        if (nnn <= kFsCap && addr - ld + bytes <= end) {
          lead = ld;
          copied = true;
          ptx::mbar_arrive_expect_tx(&full[s], bytes);
          ptx::bulk_load(fs_smem + size_t(s) * kFsSlab + size_t(r) * kFsSlot, reinterpret_cast<const void*>(addr - ld), bytes, &full[s]);
        }
      }
    }
    lead_s[s * kFsWin + r] = lead;
    if (!copied) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ptx::smem_addr(&full[s])) : "memory");
  };
  if (r < kFsWin) {
    if (i0 < i1) issue_slab(i0, 0);
    if (i0 + 1 < i1) issue_slab(i0 + 1, 1);
  }

  constexpr int CH = 2 + kGcnChannelsDev;
  long long cur_blk = -1;
  int n0 = 0, nn = 0;
  for (long long it = i0; it < i1; ++it) {
    const int s = static_cast<int>((it - i0) & 1);
    const long long blk = it / NTO;
    const int t = static_cast<int>(it % NTO) * kFsSteps + tq;
    const long long win = blk * kFsWin + bl;
    const bool live = win < n_windows;
    if (blk != cur_blk) {
      cur_blk = blk;
      n0 = live ? node_off[win] : 0;
      nn = live ? node_off[win + 1] - n0 : 0;
      // the block's node coefficients stay in shared memory for its ~20 items
      for (int n = tq; n < min(nn, kFsCap); n += kFsSteps) coef_s[bl * kFsCoefStride + n] = __ldg(coef + n0 + n);
      __syncthreads();
    }
    ptx::mbar_wait(&full[s], static_cast<uint32_t>(((it - i0) >> 1) & 1));
    if (t < t_run) {
      const int lead = lead_s[s * kFsWin + bl];
      const float2* srow = reinterpret_cast<const float2*>(fs_smem + size_t(s) * kFsSlab + size_t(bl) * kFsSlot + (lead < 0 ? 0 : lead)) + tq * nn;
      const float2* grow = feat + (static_cast<long long>(n0) * T + static_cast<long long>(t) * nn);
      const float* crow = coef + n0;
      const float* cs = coef_s + bl * kFsCoefStride;
      PoolAcc acc;
      if (lead >= 0) {   // implies nn <= kFsCap
#pragma unroll 4
        for (int n = 0; n < nn; ++n) {
          float2 v = srow[n];
          v.x *= alpha; v.y *= alpha;   // integrated-gradients path scaling; exact no-op for alpha == 1
          acc.add_packed(p, v, cs[n]);
        }
      } else {
        for (int n = 0; n < nn; ++n) {
          float2 v = __ldg(grow + n);
          v.x *= alpha; v.y *= alpha;
          acc.add_packed(p, v, __ldg(crow + n));
        }
      }
      const float2 av = live ? __ldg(anom + win * T + t) : make_float2(0.f, 0.f);
      float* o = out + (((win / kTile) * t_run + t) * CH) * kTile + (win % kTile);
      o[0] = av.x * alpha;
      o[kTile] = av.y * alpha;
#pragma unroll
      for (int c = 0; c < kGcnChannelsDev; ++c) o[(2 + c) * kTile] = acc.pooled(p, c);
    }
    __syncthreads();   // everyone is done with buffer s
    if (r < kFsWin && it + 2 < i1) issue_slab(it + 2, s);
  }
}

// Launch helper: grid sized to the SM count, dynamic shared memory opted in by configure_cml_frontend().
inline void configure_cml_frontend() {
  cudaFuncSetAttribute(cml_frontend_stream_kernel<24>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(FsCfg<24>::kSmem));
  cudaFuncSetAttribute(cml_frontend_stream_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(FsCfg<32>::kSmem));
}
template <int CAP>
inline void launch_cml_frontend_cap(const float2* feat, long long feat_rows, const float2* anom, const float* coef, const int* node_off,
                                    const GcnFolded& p, float* out, long long n_windows, int T, int t_run, long long n_tiles, float alpha,
                                    int sm_count, cudaStream_t st) {
  const long long items = n_tiles * (kTile / kFsWin) * ((t_run + kFsSteps - 1) / kFsSteps);
  const long long slots = static_cast<long long>(sm_count) * FsCfg<CAP>::kCtasPerSm;
  const long long grid = items < slots ? items : slots;
  cml_frontend_stream_kernel<CAP><<<static_cast<unsigned>(grid > 0 ? grid : 1), kFsThreads, FsCfg<CAP>::kSmem, st>>>(
      feat, feat_rows, anom, coef, node_off, p, out, n_windows, T, t_run, static_cast<int>(n_tiles), alpha);
}
// feat_rows = T * (total nodes): the average node count per window picks the build.
inline void launch_cml_frontend(const float2* feat, long long feat_rows, const float2* anom, const float* coef, const int* node_off,
                                const GcnFolded& p, float* out, long long n_windows, int T, int t_run, long long n_tiles, float alpha,
                                int sm_count, cudaStream_t st) {
  if (feat_rows <= 24LL * T * n_windows)
    launch_cml_frontend_cap<24>(feat, feat_rows, anom, coef, node_off, p, out, n_windows, T, t_run, n_tiles, alpha, sm_count, st);
  else
    launch_cml_frontend_cap<32>(feat, feat_rows, anom, coef, node_off, p, out, n_windows, T, t_run, n_tiles, alpha, sm_count, st);
}

// ---- dense head: [tile][128][64] -> sigmoid(Dense(LeakyReLU(Dense(LeakyReLU(Dense))))) ------
// libs/create_model.py:235-239 (GCN) and :372-376 (baseline).  One CTA per tile.
template <int HL, int D>
__global__ void __launch_bounds__(128) head_kernel(const float* __restrict__ y7, const float* __restrict__ w1,
                                                   const float* __restrict__ b1, const float* __restrict__ w2,
                                                   const float* __restrict__ b2, const float* __restrict__ wo,
                                                   const float* __restrict__ bo, float leaky, float* __restrict__ scores,
                                                   long long n_seq) {
  // Register tile: thread = 8 outputs x 4 windows (l, l+16, l+32, l+48), so one pair of 16-byte weight
  // loads feeds 32 FMAs; 128 threads cover the 64 outputs x 64 windows of a tile.
  static_assert(D == 64, "tile mapping assumes 64 dense units");
  __shared__ float d1_s[D * kTile];
  __shared__ float d2_s[D * kTile];
  const long long tile = blockIdx.x;
  const int l = threadIdx.x & 15, q = threadIdx.x >> 4;
  float acc[8][4];
  auto dense = [&](const float* __restrict__ w, const float* __restrict__ bias, int K, auto load_in, float* __restrict__ dst) {
#pragma unroll
    for (int o = 0; o < 8; ++o) {
      const float bv = __ldg(bias + q * 8 + o);
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[o][j] = bv;
    }
#pragma unroll 4
    for (int k = 0; k < K; ++k) {
      const float4 wa = __ldg(reinterpret_cast<const float4*>(w + k * D + q * 8));
      const float4 wb = __ldg(reinterpret_cast<const float4*>(w + k * D + q * 8 + 4));
      const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
      float h[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = load_in(k, l + 16 * j);
#pragma unroll
      for (int o = 0; o < 8; ++o)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[o][j] = fmaf(h[j], wv[o], acc[o][j]);
    }
#pragma unroll
    for (int o = 0; o < 8; ++o)
#pragma unroll
      for (int j = 0; j < 4; ++j) dst[(q * 8 + o) * kTile + l + 16 * j] = acc[o][j] > 0.f ? acc[o][j] : leaky * acc[o][j];
  };
  dense(w1, b1, HL, [&](int k, int b) { return __ldg(y7 + (tile * HL + k) * kTile + b); }, d1_s);
  __syncthreads();
  dense(w2, b2, D, [&](int k, int b) { return d1_s[k * kTile + b]; }, d2_s);
  __syncthreads();
  if (threadIdx.x < kTile) {
    const int b = threadIdx.x;
    float z = __ldg(bo);
#pragma unroll 8
    for (int k = 0; k < D; ++k) z = fmaf(d2_s[k * kTile + b], __ldg(wo + k), z);
    const long long seq = tile * kTile + b;
    if (seq < n_seq) scores[seq] = 1.f / (1.f + expf(-z));
  }
}

// (the backward through this front-end for integrated gradients lives in ig_frontend_bwd.cuh)

// ---- head forward + gradient of the score w.r.t. the last LSTM state ------------------------------
// dy7[k] = d p / d h7[k] through Dense-LeakyReLU-Dense-LeakyReLU-Dense(sigmoid)
// (libs/create_model.py:235-239).  Writes the scores too when `scores` is not null.
template <int HL, int D>
__global__ void __launch_bounds__(256) head_grad_kernel(const float* __restrict__ y7, const float* __restrict__ w1,
                                                        const float* __restrict__ b1, const float* __restrict__ w2,
                                                        const float* __restrict__ b2, const float* __restrict__ wo,
                                                        const float* __restrict__ bo, float leaky, float* __restrict__ dy7,
                                                        float* __restrict__ scores, long long n_seq) {
  __shared__ float d1_s[D * kTile];   // post-activation dense 1, later its gradient
  __shared__ float d2_s[D * kTile];   // post-activation dense 2, later its gradient
  __shared__ float m1_s[D * kTile];   // LeakyReLU slopes of dense 1
  const long long tile = blockIdx.x;
  const int b = threadIdx.x & 63, q = threadIdx.x >> 6;
  constexpr int DQ = D / 4;
  float acc[DQ];
#pragma unroll
  for (int o = 0; o < DQ; ++o) acc[o] = __ldg(b1 + q * DQ + o);
  // the weights are warp-uniform: 16-byte loads, four FMAs each (the kernel was bound by one 4-byte load per FMA)
#pragma unroll 4
  for (int k = 0; k < HL; ++k) {
    const float h = __ldg(y7 + (tile * HL + k) * kTile + b);
    const float4* wr = reinterpret_cast<const float4*>(w1 + k * D + q * DQ);
#pragma unroll
    for (int o = 0; o < DQ; o += 4) {
      const float4 w = __ldg(wr + o / 4);
      acc[o] = fmaf(h, w.x, acc[o]); acc[o + 1] = fmaf(h, w.y, acc[o + 1]); acc[o + 2] = fmaf(h, w.z, acc[o + 2]); acc[o + 3] = fmaf(h, w.w, acc[o + 3]);
    }
  }
#pragma unroll
  for (int o = 0; o < DQ; ++o) {
    d1_s[(q * DQ + o) * kTile + b] = acc[o] > 0.f ? acc[o] : leaky * acc[o];
    m1_s[(q * DQ + o) * kTile + b] = acc[o] > 0.f ? 1.f : leaky;
  }
  __syncthreads();
#pragma unroll
  for (int o = 0; o < DQ; ++o) acc[o] = __ldg(b2 + q * DQ + o);
#pragma unroll 4
  for (int k = 0; k < D; ++k) {
    const float h = d1_s[k * kTile + b];
    const float4* wr = reinterpret_cast<const float4*>(w2 + k * D + q * DQ);
#pragma unroll
    for (int o = 0; o < DQ; o += 4) {
      const float4 w = __ldg(wr + o / 4);
      acc[o] = fmaf(h, w.x, acc[o]); acc[o + 1] = fmaf(h, w.y, acc[o + 1]); acc[o + 2] = fmaf(h, w.z, acc[o + 2]); acc[o + 3] = fmaf(h, w.w, acc[o + 3]);
    }
  }
  float m2[DQ];
#pragma unroll
  for (int o = 0; o < DQ; ++o) {
    d2_s[(q * DQ + o) * kTile + b] = acc[o] > 0.f ? acc[o] : leaky * acc[o];
    m2[o] = acc[o] > 0.f ? 1.f : leaky;
  }
  __syncthreads();
  float z = __ldg(bo);
  for (int k = 0; k < D; ++k) z = fmaf(d2_s[k * kTile + b], __ldg(wo + k), z);
  const float pr = 1.f / (1.f + expf(-z));
  const long long seq = tile * kTile + b;
  if (q == 0 && scores != nullptr && seq < n_seq) scores[seq] = pr;
  const float dz3 = pr * (1.f - pr);
  __syncthreads();   // all reads of d2_s done
#pragma unroll
  for (int o = 0; o < DQ; ++o) d2_s[(q * DQ + o) * kTile + b] = dz3 * __ldg(wo + q * DQ + o) * m2[o];   // grad at dense-2 pre-activation
  __syncthreads();
  // grad at dense-1 pre-activation: sum_o dd2[o] * w2[k][o] * slope1[k]
#pragma unroll
  for (int kk = 0; kk < DQ; ++kk) acc[kk] = 0.f;
  for (int o = 0; o < D; o += 4) {   // four gradients from shared memory serve all 16 rows of this thread; same summation order as before
    const float g0 = d2_s[o * kTile + b], g1 = d2_s[(o + 1) * kTile + b], g2 = d2_s[(o + 2) * kTile + b], g3 = d2_s[(o + 3) * kTile + b];
#pragma unroll
    for (int kk = 0; kk < DQ; ++kk) {
      const float4 w = __ldg(reinterpret_cast<const float4*>(w2 + (q * DQ + kk) * D + o));
      acc[kk] = fmaf(g3, w.w, fmaf(g2, w.z, fmaf(g1, w.y, fmaf(g0, w.x, acc[kk]))));
    }
  }
#pragma unroll
  for (int kk = 0; kk < DQ; ++kk) acc[kk] *= m1_s[(q * DQ + kk) * kTile + b];
  __syncthreads();   // all reads of d1_s (none pending) - reuse it for the dense-1 gradients
#pragma unroll
  for (int kk = 0; kk < DQ; ++kk) d1_s[(q * DQ + kk) * kTile + b] = acc[kk];
  __syncthreads();
  float dy[HL / 4];
#pragma unroll
  for (int kk = 0; kk < HL / 4; ++kk) dy[kk] = 0.f;
  for (int o = 0; o < D; o += 4) {
    const float g0 = d1_s[o * kTile + b], g1 = d1_s[(o + 1) * kTile + b], g2 = d1_s[(o + 2) * kTile + b], g3 = d1_s[(o + 3) * kTile + b];
#pragma unroll
    for (int kk = 0; kk < HL / 4; ++kk) {
      const float4 w = __ldg(reinterpret_cast<const float4*>(w1 + (q * (HL / 4) + kk) * D + o));
      dy[kk] = fmaf(g3, w.w, fmaf(g2, w.z, fmaf(g1, w.y, fmaf(g0, w.x, dy[kk]))));
    }
  }
#pragma unroll
  for (int kk = 0; kk < HL / 4; ++kk) dy7[(tile * HL + q * (HL / 4) + kk) * kTile + b] = dy[kk];
}

}  // namespace gnnqc
