// Fused optimiser tail over the flat gradient bucket.
//
// The reference's step is: on-chip-QNN gradient pruning (Estimators_QuantumNAT_onchipQNN.py:205-228:
// zero every gradient entry with |g| <= threshold, over ALL parameters, with a `.item()` host
// synchronisation per parameter tensor) followed by torch.optim.AdamW(lr, weight_decay)
// (Runner_P128_QuantumNAT_onchipQNN.py:320,369).  Once the QNN itself takes microseconds those
// per-parameter loops dominate a step; here both run as ONE elementwise kernel over the flat
// parameter / gradient / moment buffers (HBM-bound: 5 reads + 4 writes of 4 bytes per parameter), the
// pruned count is accumulated on the device and never forces a synchronisation.
//
// Arithmetic follows torch.optim.AdamW (amsgrad=False, maximize=False) operation for operation:
//   p *= 1 - lr*wd;  m = lerp(m, g, 1-b1);  v = b2*v + (1-b2)*g*g;
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
#include <algorithm>
#include <cmath>

#include "qnn_internal.h"

namespace {
__global__ void __launch_bounds__(256) k_pruned_adamw(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                                      float* __restrict__ v, long long n, float lr, float b1, float b2,
                                                      float eps, float wd, float step_size, float bc2_sqrt,
                                                      float threshold, unsigned long long* __restrict__ pruned,
                                                      const long long* __restrict__ steps_done) {
  if (steps_done != nullptr) {   // step count kept on the device: the bias corrections are computed here
    const double t = (double)(*steps_done + 1);
    step_size = (float)((double)lr / (1.0 - pow((double)b1, t)));
    bc2_sqrt = (float)sqrt(1.0 - pow((double)b2, t));
  }
  unsigned int local = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gi = g[i];
    if (threshold >= 0.f && !(fabsf(gi) > threshold)) { gi = 0.f; g[i] = 0.f; ++local; }
    float pi = p[i] * (1.f - lr * wd);
    const float mi = m[i] + (gi - m[i]) * (1.f - b1);            // lerp
    const float vi = fmaf(gi * gi, 1.f - b2, v[i] * b2);
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    pi -= step_size * (mi / denom);
    p[i] = pi; m[i] = mi; v[i] = vi;
  }
  if (pruned != nullptr && threshold >= 0.f) {
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(pruned, (unsigned long long)local);   // integer: order-free
  }
}
__global__ void k_count_step(long long* steps_done) { *steps_done += 1; }
}  // namespace

extern "C" int qnn_pruned_adamw_step_counted(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                             float lr, float beta1, float beta2, float eps, float weight_decay,
                                             long long* steps_done, float prune_threshold,
                                             unsigned long long* pruned_count, void* cuda_stream) {
  if (!params || !grads || !exp_avg || !exp_avg_sq || !steps_done) { qnn_set_error("qnn_pruned_adamw_step_counted: NULL buffer"); return QNN_ERR_INVALID; }
  if (n <= 0) { qnn_set_error("qnn_pruned_adamw_step_counted: n=%lld must be positive", (long long)n); return QNN_ERR_INVALID; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)qnn_sm_count() * 8));
  k_pruned_adamw<<<grid, 256, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, 0.f, 1.f,
                                       prune_threshold, pruned_count, steps_done);
  QNN_CUDA_CHECK(cudaGetLastError());
  k_count_step<<<1, 1, 0, st>>>(steps_done);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

extern "C" int qnn_pruned_adamw_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                     float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                                     float prune_threshold, unsigned long long* pruned_count, void* cuda_stream) {
  if (!params || !grads || !exp_avg || !exp_avg_sq) { qnn_set_error("qnn_pruned_adamw_step: NULL buffer"); return QNN_ERR_INVALID; }
  if (n <= 0 || step < 1) { qnn_set_error("qnn_pruned_adamw_step: n=%lld, step=%lld must be positive", (long long)n, (long long)step); return QNN_ERR_INVALID; }
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)qnn_sm_count() * 8));
  k_pruned_adamw<<<grid, 256, 0, static_cast<cudaStream_t>(cuda_stream)>>>(params, grads, exp_avg, exp_avg_sq, n, lr, beta1,
                                                                           beta2, eps, weight_decay, step_size, bc2_sqrt,
                                                                           prune_threshold, pruned_count, nullptr);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}
