// QuantumNAT post-measurement pipeline (include/qnn_b200.h: qnat_*): per-qubit batch statistics,
// normalise, inject caller-supplied noise, quantise; and the matching backward (straight-through
// estimator + batch-norm backward).  HBM-bound elementwise work on [B, n] float32, n <= 16.
//
// Batch sums are two-stage and order-fixed (per-CTA partials in double, then one CTA adds them in
// index order) so a given (B, n) always reduces the same way.  The two stages are separate entry
// points so that data-parallel ranks can all-reduce the 2n sums in between.
#include <algorithm>

#include "qnn_internal.h"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxCtasPerSm = 4;

struct Stats {
  float mean, rstd;
};

__device__ __forceinline__ Stats stats_from_sums(const double* sums, int q, int n, long long Bt, float eps) {
  const double m = sums[q] / (double)Bt;
  double var = (sums[n + q] - (double)Bt * m * m) / (double)max(Bt - 1, 1LL);
  var = var < 0.0 ? 0.0 : var;
  Stats s;
  s.mean = (float)m;
  s.rstd = (float)(1.0 / sqrt(var + (double)eps));
  return s;
}

// MODE 0: (y, y^2).  MODE 1: (g, g*yh) with g the STE-masked upstream gradient.
template <int MODE>
__global__ void __launch_bounds__(kThreads) k_col_sums(const float* __restrict__ y, const float* __restrict__ gout,
                                                       const float* __restrict__ yn, const float* __restrict__ stats,
                                                       long long B, int n, qnat_params p, double* __restrict__ partial) {
  __shared__ double s1[kThreads], s2[kThreads];
  const int rows_per_pass = kThreads / n;  // n <= 16 -> >= 16
  const int q = threadIdx.x % n, rl = threadIdx.x / n;
  const bool active = rl < rows_per_pass;
  float mean = 0.f, rstd = 1.f;
  if (MODE == 1 && p.normalise) { mean = stats[q]; rstd = stats[n + q]; }
  double a1 = 0.0, a2 = 0.0;
  if (active) {
    for (long long r = (long long)blockIdx.x * rows_per_pass + rl; r < B; r += (long long)gridDim.x * rows_per_pass) {
      const size_t i = (size_t)r * n + q;
      if (MODE == 0) {
        const float v = y[i];
        a1 += (double)v;
        a2 += (double)v * (double)v;
      } else {
        float g = gout[i];
        if (p.quantise) { const float t = yn[i]; g = (t >= p.lo && t <= p.hi) ? g : 0.f; }
        const float yh = (y[i] - mean) * rstd;
        a1 += (double)g;
        a2 += (double)g * (double)yh;
      }
    }
  }
  s1[threadIdx.x] = a1; s2[threadIdx.x] = a2;
  __syncthreads();
  if (threadIdx.x < n) {
    double t1 = 0.0, t2 = 0.0;
    for (int j = 0; j < rows_per_pass; ++j) { t1 += s1[j * n + threadIdx.x]; t2 += s2[j * n + threadIdx.x]; }
    partial[(size_t)blockIdx.x * 2 * n + threadIdx.x] = t1;
    partial[(size_t)blockIdx.x * 2 * n + n + threadIdx.x] = t2;
  }
}

// One warp per column; lanes stride over the CTA partials, then a shuffle tree: fixed order.
__global__ void k_sum_partials(const double* __restrict__ partial, int n_ctas, int n2, double* __restrict__ sums) {
  const int j = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (j >= n2) return;
  double t = 0.0;
  for (int c = lane; c < n_ctas; c += 32) t += partial[(size_t)c * n2 + j];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if (lane == 0) sums[j] = t;
}

__global__ void __launch_bounds__(kThreads) k_apply(const float* __restrict__ y, const float* __restrict__ noise,
                                                    const double* __restrict__ sums, long long Bt, float* __restrict__ out,
                                                    float* __restrict__ yn_out, float* __restrict__ stats, long long total,
                                                    int n, qnat_params p) {
  __shared__ float smean[QNN_MAX_QUBITS], srstd[QNN_MAX_QUBITS];
  if (Bt <= 0 && p.normalise) Bt = llrint(sums[2 * n]);   // row count kept (and all-reduced) next to the sums
  if (threadIdx.x < n) {
    Stats s = {0.f, 1.f};
    if (p.normalise) s = stats_from_sums(sums, threadIdx.x, n, Bt, p.eps);
    smean[threadIdx.x] = s.mean; srstd[threadIdx.x] = s.rstd;
    if (blockIdx.x == 0 && stats) { stats[threadIdx.x] = s.mean; stats[n + threadIdx.x] = s.rstd; }
  }
  __syncthreads();
  const float step = (p.hi - p.lo) / (float)(p.levels - 1), inv_step = (float)(p.levels - 1) / (p.hi - p.lo);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % n);
    float v = (y[i] - smean[q]) * srstd[q];
    if (noise) v = fmaf(p.sigma, noise[i], v);
    if (yn_out) yn_out[i] = v;
    if (p.quantise) {
      const float cl = fminf(fmaxf(v, p.lo), p.hi);
      v = fmaf(step, rintf((cl - p.lo) * inv_step), p.lo);
    }
    out[i] = v;
  }
}

__global__ void __launch_bounds__(kThreads) k_bwd_apply(const float* __restrict__ gout, const float* __restrict__ y,
                                                        const float* __restrict__ yn, const float* __restrict__ stats,
                                                        const double* __restrict__ bsums, long long Bt, float* __restrict__ gy,
                                                        long long total, int n, qnat_params p) {
  __shared__ float smean[QNN_MAX_QUBITS], srstd[QNN_MAX_QUBITS], sa[QNN_MAX_QUBITS], sb[QNN_MAX_QUBITS];
  if (Bt <= 0 && p.normalise) Bt = llrint(bsums[2 * n]);
  if (threadIdx.x < n && p.normalise) {
    smean[threadIdx.x] = stats[threadIdx.x];
    srstd[threadIdx.x] = stats[n + threadIdx.x];
    sa[threadIdx.x] = (float)(bsums[threadIdx.x] / (double)Bt);
    sb[threadIdx.x] = (float)(bsums[n + threadIdx.x] / (double)max(Bt - 1, 1LL));
  }
  __syncthreads();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % n);
    float g = gout[i];
    if (p.quantise) { const float t = yn[i]; g = (t >= p.lo && t <= p.hi) ? g : 0.f; }
    if (p.normalise) {
      const float yh = (y[i] - smean[q]) * srstd[q];
      g = srstd[q] * (g - sa[q] - yh * sb[q]);
    }
    gy[i] = g;
  }
}

int n_ctas_for(long long B, int n) {
  const int rows_per_pass = kThreads / n;
  long long c = (B + rows_per_pass - 1) / rows_per_pass;
  return (int)std::max<long long>(1, std::min<long long>(c, (long long)qnn_sm_count() * kMaxCtasPerSm));
}

int check(const char* fn, int64_t B, int32_t n, const qnat_params* p) {
  if (B <= 0 || n < 1 || n > QNN_MAX_QUBITS) { qnn_set_error("%s: bad shape B=%lld n=%d", fn, (long long)B, n); return QNN_ERR_INVALID; }
  if (p && p->quantise && (p->levels < 2 || !(p->hi > p->lo))) { qnn_set_error("%s: quantiser needs levels >= 2 and hi > lo", fn); return QNN_ERR_INVALID; }
  return QNN_OK;
}

}  // namespace

extern "C" size_t qnat_workspace_bytes(int64_t B, int32_t n) {
  if (B <= 0 || n < 1 || n > QNN_MAX_QUBITS) return 0;
  return (size_t)n_ctas_for(B, n) * 2 * n * sizeof(double);
}

extern "C" int qnat_batch_sums(const float* y, double* sums, int64_t B, int32_t n, void* workspace,
                               size_t workspace_bytes, void* cuda_stream) {
  int rc = check("qnat_batch_sums", B, n, nullptr);
  if (rc) return rc;
  if (!y || !sums) { qnn_set_error("qnat_batch_sums: NULL pointer"); return QNN_ERR_INVALID; }
  if (!workspace || workspace_bytes < qnat_workspace_bytes(B, n)) { qnn_set_error("qnat_batch_sums: workspace too small"); return QNN_ERR_WORKSPACE; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int g = n_ctas_for(B, n);
  qnat_params p{};
  k_col_sums<0><<<g, kThreads, 0, st>>>(y, nullptr, nullptr, nullptr, B, n, p, static_cast<double*>(workspace));
  QNN_CUDA_CHECK(cudaGetLastError());
  k_sum_partials<<<1, 32 * 2 * n, 0, st>>>(static_cast<double*>(workspace), g, 2 * n, sums);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

extern "C" int qnat_apply(const float* y, const float* noise, const double* sums, int64_t B_total, float* out,
                          float* yn, float* stats, int64_t B, int32_t n, const qnat_params* p, void* cuda_stream) {
  int rc = check("qnat_apply", B, n, p);
  if (rc) return rc;
  if (!y || !out || !p || (p->normalise && !sums)) { qnn_set_error("qnat_apply: NULL pointer"); return QNN_ERR_INVALID; }
  if (B_total < 0) { qnn_set_error("qnat_apply: B_total=%lld is negative", (long long)B_total); return QNN_ERR_INVALID; }
  const long long total = (long long)B * n;
  const int g = (int)std::max<long long>(1, std::min<long long>((total + kThreads - 1) / kThreads, (long long)qnn_sm_count() * 8));
  k_apply<<<g, kThreads, 0, static_cast<cudaStream_t>(cuda_stream)>>>(y, p->sigma != 0.f ? noise : nullptr, sums, B_total, out, yn, stats, total, n, *p);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

extern "C" int qnat_backward_sums(const float* gout, const float* y, const float* yn, const float* stats,
                                  double* bsums, int64_t B, int32_t n, const qnat_params* p, void* workspace,
                                  size_t workspace_bytes, void* cuda_stream) {
  int rc = check("qnat_backward_sums", B, n, p);
  if (rc) return rc;
  if (!gout || !y || !bsums || !p || (p->quantise && !yn) || (p->normalise && !stats)) { qnn_set_error("qnat_backward_sums: NULL pointer"); return QNN_ERR_INVALID; }
  if (!workspace || workspace_bytes < qnat_workspace_bytes(B, n)) { qnn_set_error("qnat_backward_sums: workspace too small"); return QNN_ERR_WORKSPACE; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int g = n_ctas_for(B, n);
  k_col_sums<1><<<g, kThreads, 0, st>>>(y, gout, yn, stats, B, n, *p, static_cast<double*>(workspace));
  QNN_CUDA_CHECK(cudaGetLastError());
  k_sum_partials<<<1, 32 * 2 * n, 0, st>>>(static_cast<double*>(workspace), g, 2 * n, bsums);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

extern "C" int qnat_backward_apply(const float* gout, const float* y, const float* yn, const float* stats,
                                   const double* bsums, int64_t B_total, float* gy, int64_t B, int32_t n,
                                   const qnat_params* p, void* cuda_stream) {
  int rc = check("qnat_backward_apply", B, n, p);
  if (rc) return rc;
  if (!gout || !gy || !p || (p->quantise && !yn) || (p->normalise && (!stats || !bsums || !y))) { qnn_set_error("qnat_backward_apply: NULL pointer"); return QNN_ERR_INVALID; }
  const long long total = (long long)B * n;
  const int g = (int)std::max<long long>(1, std::min<long long>((total + kThreads - 1) / kThreads, (long long)qnn_sm_count() * 8));
  k_bwd_apply<<<g, kThreads, 0, static_cast<cudaStream_t>(cuda_stream)>>>(gout, y, yn, stats, bsums, B_total, gy, total, n, *p);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}
