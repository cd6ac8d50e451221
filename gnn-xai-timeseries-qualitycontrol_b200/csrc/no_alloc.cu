// gnnqc_score_seq_noalloc: the sequence-scoring pipeline for callers that must not allocate.
//
// Who this is for: code that runs where cudaMalloc is off limits - a real-time acquisition loop
// with a fixed memory budget, a host that has handed all device memory to another framework's
// pool, a callback on a thread that may not take the allocator's locks.  gnnqc_score_seq keeps
// its inter-layer tiles in an arena that it grows with cudaMalloc whenever a bigger batch shows
// up; this variant keeps them in the fixed array below instead, so a call performs no
// cudaMalloc / cudaFree and uses no allocating library routine.  Inputs and outputs are still
// the caller's buffers.
//
// Cost and limits: the array is part of the library image, so every GPU that loads
// libgnnqc.so carries kNoAllocScratchBytes (32 MiB) until the process exits, used or not.  A
// call is rejected when its tiles do not fit (about 1,600 windows of 181 steps, about 800 of
// 337 steps).  All calls on a device share the one array and it keeps whatever the last call
// left in it: calls that overlap on the GPU (different streams, handles or threads) race and
// give garbage, so serialise them.  Nothing in the array has to start at zero - every tile is
// written in full before it is read - so no clearing pass is needed.
#include <cuda_runtime.h>

#include <algorithm>
#include <mutex>
#include <string>

#include "../../include/gnnqc.h"
#include "capi_internal.h"
#include "frontend.cuh"

namespace {

constexpr size_t kNoAllocScratchBytes = size_t(32) << 20;
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for the runtime; the device address is looked up
// once per device below and that pointer is what kernels receive.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

std::mutex g_setup_mutex;
unsigned char* g_scratch_ptr[kMaxDevices] = {nullptr};

// First call on `device`: resolve the array's address on that device (with it current).
// Later calls return the cached pointer.
cudaError_t scratch_for_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_scratch_ptr[device]) {
    cudaError_t err = cudaSetDevice(device);
    if (err != cudaSuccess) return err;
    void* p = nullptr;
    err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (err != cudaSuccess) return err;
    g_scratch_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_ptr[device];
  return cudaSuccess;
}

inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" int gnnqc_score_seq_noalloc(gnnqc_model* m, const float* x_dev, int64_t n_seq, int32_t T, float* scores_dev,
                                       void* stream) {
  if (!m) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_seq_noalloc: null model handle");
  gnnqc::DeviceModel& dm = m->dm;
  if (dm.device < 0) return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_seq_noalloc: handle was opened host-only; there is no CPU path");
  if (dm.device >= kMaxDevices) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_seq_noalloc: device index too large");
  if (n_seq < 0 || T <= 0) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_seq_noalloc: bad n_seq / T");
  if (n_seq == 0) return GNNQC_OK;
  if (!x_dev || !scores_dev) return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_seq_noalloc: null buffer");
  if (dm.profiling)
    return gnnqc_internal_fail(GNNQC_ERR_STATE, "gnnqc_score_seq_noalloc: per-launch profiling creates CUDA events; switch it off first");
  try {
    const gnnqc::ModelSpec& s = dm.host.spec;
    gnnqc::StackPlan plan(s, T);
    const long long n_tiles = (n_seq + gnnqc::kTile - 1) / gnnqc::kTile;
    const size_t sa = round_up(plan.bytes_a, 256), sb = round_up(plan.bytes_b, 256);
    const long long cap_tiles = static_cast<long long>(kNoAllocScratchBytes / (sa + sb));
    if (n_tiles > cap_tiles)
      return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_seq_noalloc: " + std::to_string(n_seq) + " sequences of " + std::to_string(T) +
                                                    " steps need " + std::to_string(size_t(n_tiles) * (sa + sb)) +
                                                    " bytes of scratch; the built-in array holds " + std::to_string(kNoAllocScratchBytes) +
                                                    " (at most " + std::to_string(cap_tiles * gnnqc::kTile) +
                                                    " sequences of this length per call)");
    unsigned char* base = nullptr;
    cudaError_t err = scratch_for_device(dm.device, &base);
    if (err != cudaSuccess)
      return gnnqc_internal_fail(GNNQC_ERR_CUDA, std::string("gnnqc_score_seq_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(err));
    GNNQC_CUDA(cudaSetDevice(dm.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* buf_a = reinterpret_cast<float*>(base);
    float* buf_b = reinterpret_cast<float*>(base + size_t(n_tiles) * sa);
    const int c_in = s.lstm_in[0];
    dim3 grid(static_cast<unsigned>(n_tiles), static_cast<unsigned>((plan.t_run[0] + 31) / 32));
    dm.mark_begin("seq_to_tiles", st);
    if (c_in == 2) gnnqc::seq_to_tiles_kernel<2><<<grid, 256, 0, st>>>(x_dev, buf_a, n_seq, T, plan.t_run[0]);
    else if (c_in == 3) gnnqc::seq_to_tiles_kernel<3><<<grid, 256, 0, st>>>(x_dev, buf_a, n_seq, T, plan.t_run[0]);
    else if (c_in == 18) gnnqc::seq_to_tiles_kernel<18><<<grid, 256, 0, st>>>(x_dev, buf_a, n_seq, T, plan.t_run[0]);
    else if (c_in == 19) gnnqc::seq_to_tiles_kernel<19><<<grid, 256, 0, st>>>(x_dev, buf_a, n_seq, T, plan.t_run[0]);
    else return gnnqc_internal_fail(GNNQC_ERR_ARG, "gnnqc_score_seq_noalloc: unsupported sequence width");
    GNNQC_CUDA(cudaGetLastError());
    dm.mark_end(st);
    dm.run_stack(plan, buf_a, buf_b, n_tiles, n_seq, scores_dev, st);
    return GNNQC_OK;
  } catch (const gnnqc::CudaError& e) {
    return gnnqc_internal_fail(GNNQC_ERR_CUDA, e.what());
  } catch (const std::exception& e) {
    return gnnqc_internal_fail(GNNQC_ERR_ARG, e.what());
  }
}
