// Threshold selection and ROC-AUC on the device - the step right after scoring in the reference's
// evaluation (libs/test_model.py:9-17 select_threshold, :38-59 calculate_metrics -> sklearn
// roc_auc_score / matthews_corrcoef).  Scores stay where the scoring call left them.
//
// select_threshold: the candidates are the distinct values of round(score, 3) (float32, round half to
// even like numpy); `score > thr` is evaluated in float32 for every candidate and the one with the
// highest Matthews correlation wins, the smallest on ties.  All candidates lie on the grid k/1000, so one
// pass builds, per class, a histogram of m(s) = #{k : float32(k)/1000 < s}; the confusion matrix of
// candidate k is a suffix sum of the histograms, exactly as if the comparison had been made per score.
// roc_auc: scores sorted descending (cub radix sort), the ROC points are the ends of runs of equal
// scores, and twice the trapezoid area times P*N is accumulated as an exact integer.
#include <cub/cub.cuh>

#include <cmath>
#include <cstdint>

#include "../../include/gnnqc.h"
#include "capi_internal.h"

namespace gnnqc {
namespace {

constexpr int kGrid = 1001;   // thresholds 0.000 ... 1.000

__device__ __forceinline__ float grid_value(int k) { return static_cast<float>(k) / 1000.f; }

__global__ void __launch_bounds__(256) threshold_hist_kernel(const float* __restrict__ scores, const long long* __restrict__ labels, long long n,
                                                             unsigned long long* __restrict__ hist,   // [2][kGrid + 1]
                                                             int* __restrict__ present, int* __restrict__ bad) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float s = scores[i];
    if (!(s >= 0.f && s <= 1.f)) {   // NaN or outside the sigmoid range: not a score this routine handles
      atomicExch(bad, 1);
      continue;
    }
    const int r = static_cast<int>(rintf(s * 1000.f));   // numpy.round(s, 3) * 1000
    present[r] = 1;
    int m = static_cast<int>(floorf(s * 1000.f));        // then fix up against the float32 grid values
    m = m < 0 ? 0 : (m > kGrid - 1 ? kGrid - 1 : m);
    while (m < kGrid && grid_value(m) < s) ++m;          // m = first k with grid(k) >= s  (kGrid if none)
    while (m > 0 && !(grid_value(m - 1) < s)) --m;
    atomicAdd(&hist[(labels[i] != 0 ? 1 : 0) * (kGrid + 1) + m], 1ull);
  }
}

// One block: suffix sums, MCC per present candidate, arg-max with "smallest threshold on ties".
__global__ void __launch_bounds__(1024) threshold_pick_kernel(const unsigned long long* __restrict__ hist, const int* __restrict__ present,
                                                              float* __restrict__ thr_out, double* __restrict__ mcc_out,
                                                              long long* __restrict__ confusion_out) {
  __shared__ unsigned long long suf[2][kGrid + 2];
  __shared__ double best_s[1024];
  __shared__ int best_k[1024];
  const int k = threadIdx.x;
  if (k < 2) {   // serial suffix sums: 1002 adds per class
    unsigned long long acc = 0;
    suf[k][kGrid + 1] = 0;
    for (int m = kGrid; m >= 0; --m) {
      acc += hist[k * (kGrid + 1) + m];
      suf[k][m] = acc;
    }
  }
  __syncthreads();
  double mcc = -2.0;
  double tp = 0, fp = 0, tn = 0, fn = 0;
  if (k < kGrid && present[k]) {
    // predicted positive: grid(k) < s  <=>  m(s) >= k + 1
    tp = static_cast<double>(suf[1][k + 1]);
    fp = static_cast<double>(suf[0][k + 1]);
    fn = static_cast<double>(suf[1][0]) - tp;
    tn = static_cast<double>(suf[0][0]) - fp;
    const double den = sqrt((tp + fp) * (tp + fn) * (tn + fp) * (tn + fn));
    mcc = den > 0 ? (tp * tn - fp * fn) / den : 0.0;
  }
  best_s[k] = mcc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (k < o) best_s[k] = fmax(best_s[k], best_s[k + o]);
    __syncthreads();
  }
  const double top = best_s[0];
  __syncthreads();
  best_k[k] = (k < kGrid && present[k] && mcc >= top - 1e-12) ? k : kGrid;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (k < o) best_k[k] = min(best_k[k], best_k[k + o]);
    __syncthreads();
  }
  if (k == best_k[0]) {
    *thr_out = grid_value(k);
    *mcc_out = mcc;
    confusion_out[0] = static_cast<long long>(tp);
    confusion_out[1] = static_cast<long long>(fp);
    confusion_out[2] = static_cast<long long>(tn);
    confusion_out[3] = static_cast<long long>(fn);
  }
}

__global__ void __launch_bounds__(256) auc_label_kernel(const long long* __restrict__ labels, long long n, unsigned char* __restrict__ y) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) y[i] = labels[i] != 0 ? 1 : 0;
}
__global__ void __launch_bounds__(256) auc_run_end_kernel(const float* __restrict__ key, int n, int* __restrict__ run_end) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) run_end[i] = (i == n - 1 || key[i] != key[i + 1]) ? i : -1;
}
__global__ void __launch_bounds__(256) auc_area_kernel(const int* __restrict__ run_end, const int* __restrict__ prev_end, const int* __restrict__ tps,
                                                       int n, unsigned long long* __restrict__ twice_area) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || run_end[i] < 0) return;
  const int p = prev_end[i];
  const long long tp = tps[i], fp = static_cast<long long>(i) + 1 - tp;
  const long long tp0 = p >= 0 ? tps[p] : 0, fp0 = p >= 0 ? static_cast<long long>(p) + 1 - tp0 : 0;
  atomicAdd(twice_area, static_cast<unsigned long long>((fp - fp0) * (tp + tp0)));
}

struct MaxOp {
  __device__ __forceinline__ int operator()(int a, int b) const { return a > b ? a : b; }
};

struct Scratch {   // frees on scope exit
  void* p = nullptr;
  ~Scratch() { cudaFree(p); }
};

}  // namespace
}  // namespace gnnqc

using namespace gnnqc;

extern "C" int gnnqc_select_threshold(int device, const float* scores_dev, const int64_t* labels_dev, int64_t n, float* threshold_out,
                                      double* mcc_out, int64_t* confusion_out, void* stream) {
  if (n <= 0 || !scores_dev || !labels_dev || !threshold_out) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_select_threshold: bad arguments");
  try {
    GNNQC_CUDA(cudaSetDevice(device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Scratch s;
    const size_t hist_bytes = 2 * (kGrid + 1) * sizeof(unsigned long long), flag_bytes = (kGrid + 1) * sizeof(int);
    const size_t total = hist_bytes + flag_bytes + 64;
    GNNQC_CUDA(cudaMalloc(&s.p, total));
    GNNQC_CUDA(cudaMemsetAsync(s.p, 0, total, st));
    char* base = static_cast<char*>(s.p);
    unsigned long long* hist = reinterpret_cast<unsigned long long*>(base);
    int* present = reinterpret_cast<int*>(base + hist_bytes);
    int* bad = present + kGrid;
    char* outp = base + hist_bytes + flag_bytes;   // thr (4) | pad | mcc (8) | confusion (32)
    float* thr_d = reinterpret_cast<float*>(outp);
    double* mcc_d = reinterpret_cast<double*>(outp + 8);
    long long* conf_d = reinterpret_cast<long long*>(outp + 16);
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((n + 255) / 256, 148 * 8));
    threshold_hist_kernel<<<blocks, 256, 0, st>>>(scores_dev, reinterpret_cast<const long long*>(labels_dev), n, hist, present, bad);
    threshold_pick_kernel<<<1, 1024, 0, st>>>(hist, present, thr_d, mcc_d, conf_d);
    GNNQC_CUDA(cudaGetLastError());
    struct { float thr; float pad; double mcc; long long conf[4]; } h;
    int bad_h = 0;
    GNNQC_CUDA(cudaMemcpyAsync(&h, outp, sizeof(h), cudaMemcpyDeviceToHost, st));
    GNNQC_CUDA(cudaMemcpyAsync(&bad_h, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    GNNQC_CUDA(cudaStreamSynchronize(st));
    if (bad_h) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_select_threshold: scores must be finite and within [0, 1]");
    *threshold_out = h.thr;
    if (mcc_out) *mcc_out = h.mcc;
    if (confusion_out)
      for (int i = 0; i < 4; ++i) confusion_out[i] = h.conf[i];
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  }
}

extern "C" int gnnqc_roc_auc(int device, const float* scores_dev, const int64_t* labels_dev, int64_t n, double* auc_out, void* stream) {
  if (n <= 0 || n >= (int64_t(1) << 31) - 1 || !scores_dev || !labels_dev || !auc_out)
    return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_roc_auc: bad arguments (0 < n < 2^31)");
  try {
    GNNQC_CUDA(cudaSetDevice(device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int ni = static_cast<int>(n);
    size_t sort_bytes = 0, scan_bytes = 0, scan2_bytes = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, sort_bytes, static_cast<const float*>(nullptr), static_cast<float*>(nullptr),
                                              static_cast<const unsigned char*>(nullptr), static_cast<unsigned char*>(nullptr), ni, 0, 32, st);
    cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, static_cast<const unsigned char*>(nullptr), static_cast<int*>(nullptr), ni, st);
    cub::DeviceScan::ExclusiveScan(nullptr, scan2_bytes, static_cast<const int*>(nullptr), static_cast<int*>(nullptr), MaxOp(), -1, ni, st);
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t o_key = 0, o_y = o_key + up(size_t(n) * 4), o_ys = o_y + up(size_t(n)), o_tps = o_ys + up(size_t(n)),
                 o_re = o_tps + up(size_t(n) * 4), o_pe = o_re + up(size_t(n) * 4), o_acc = o_pe + up(size_t(n) * 4),
                 o_tmp = o_acc + 256, total = o_tmp + up(std::max(sort_bytes, std::max(scan_bytes, scan2_bytes)));
    Scratch s;
    GNNQC_CUDA(cudaMalloc(&s.p, total));
    char* base = static_cast<char*>(s.p);
    float* key = reinterpret_cast<float*>(base + o_key);
    unsigned char* y = reinterpret_cast<unsigned char*>(base + o_y);
    unsigned char* ys = reinterpret_cast<unsigned char*>(base + o_ys);
    int* tps = reinterpret_cast<int*>(base + o_tps);
    int* run_end = reinterpret_cast<int*>(base + o_re);
    int* prev_end = reinterpret_cast<int*>(base + o_pe);
    unsigned long long* acc = reinterpret_cast<unsigned long long*>(base + o_acc);
    void* tmp = base + o_tmp;
    const unsigned blocks = static_cast<unsigned>((n + 255) / 256);
    GNNQC_CUDA(cudaMemsetAsync(acc, 0, 8, st));
    auc_label_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const long long*>(labels_dev), n, y);
    size_t b = sort_bytes;
    GNNQC_CUDA(cub::DeviceRadixSort::SortPairsDescending(tmp, b, scores_dev, key, y, ys, ni, 0, 32, st));
    b = scan_bytes;
    GNNQC_CUDA(cub::DeviceScan::InclusiveSum(tmp, b, ys, tps, ni, st));
    auc_run_end_kernel<<<blocks, 256, 0, st>>>(key, ni, run_end);
    b = scan2_bytes;
    GNNQC_CUDA(cub::DeviceScan::ExclusiveScan(tmp, b, run_end, prev_end, MaxOp(), -1, ni, st));
    auc_area_kernel<<<blocks, 256, 0, st>>>(run_end, prev_end, tps, ni, acc);
    GNNQC_CUDA(cudaGetLastError());
    unsigned long long twice = 0;
    int pos = 0;
    GNNQC_CUDA(cudaMemcpyAsync(&twice, acc, 8, cudaMemcpyDeviceToHost, st));
    GNNQC_CUDA(cudaMemcpyAsync(&pos, tps + (ni - 1), 4, cudaMemcpyDeviceToHost, st));
    GNNQC_CUDA(cudaStreamSynchronize(st));
    const double P = pos, N = static_cast<double>(n) - pos;
    *auc_out = (P > 0 && N > 0) ? static_cast<double>(twice) / (2.0 * P * N) : 0.0;
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  }
}
