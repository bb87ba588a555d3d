// FP32 FMA throughput probe: the measured denominator of the FP32 roofline (MEASURED_PEAKS.json has
// HBM and BF16 only).  Register-resident, 16 independent FMA chains per thread, 8 CTAs of 256 threads
// per SM; timed with CUDA events on the default stream.
#include <algorithm>

#include "qnn_internal.h"

namespace {
__global__ void __launch_bounds__(256) k_fma_probe(float* out, int iters, float a, float b) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = (float)(threadIdx.x + i) * 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(v[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
  if (s == 123.456f) out[0] = s;  // never true in practice; keeps the chains live
}
}  // namespace

extern "C" int qnn_fp32_peak_probe(int32_t iters, double* tflops, double* sm_clock_mhz_unused) {
  if (!tflops || iters < 1) { qnn_set_error("qnn_fp32_peak_probe: bad arguments"); return QNN_ERR_INVALID; }
  if (sm_clock_mhz_unused) *sm_clock_mhz_unused = 0.0;
  int dev = 0, sms = 0;
  QNN_CUDA_CHECK(cudaGetDevice(&dev));
  QNN_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  float* d = nullptr;
  QNN_CUDA_CHECK(cudaMalloc(&d, 16));
  cudaEvent_t e0, e1;
  QNN_CUDA_CHECK(cudaEventCreate(&e0));
  QNN_CUDA_CHECK(cudaEventCreate(&e1));
  const int grid = sms * 8, threads = 256;
  k_fma_probe<<<grid, threads>>>(d, iters, 0.999f, 1e-4f);  // warm-up
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    QNN_CUDA_CHECK(cudaEventRecord(e0));
    k_fma_probe<<<grid, threads>>>(d, iters, 0.999f, 1e-4f);
    QNN_CUDA_CHECK(cudaEventRecord(e1));
    QNN_CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0.f;
    QNN_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = (double)grid * threads * (double)iters * 8.0 * 16.0 * 2.0;
    best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  *tflops = best;
  return QNN_OK;
}
