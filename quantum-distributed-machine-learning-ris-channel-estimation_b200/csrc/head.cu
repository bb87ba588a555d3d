// Angle head of the scenario classifier: the last two stages of the reference's
// `preprocess` stack, nn.Linear(32*(h/4)*(w/4), n_qubits) followed by nn.Tanh
// (Estimators_QuantumNAT_onchipQNN.py:161-162), which produce the encoder angles x of the circuit.
//
//   forward:   x[b, j] = tanh(bias[j] + sum_k f[b, k] W[j, k])
//   backward:  gz = gx * (1 - x^2);  gf = gz W;  gW = gz^T f;  gbias = sum_b gz
//
// Stock PyTorch runs this as addmm + tanh forward and tanh_backward + two GEMMs + a column sum backward,
// each a separate pass over [B, K] or [B, n].  Here it is one pass each way: HBM-bound on the feature
// matrix (4 K bytes per sample read forward; 8 K read+written backward when gf is wanted).
//
// Forward: one warp per sample; W^T staged in shared memory; each lane owns the columns k = lane (mod 32)
// and keeps n partial sums in registers; butterfly reduction over the warp.
// Backward: one thread per feature column k, a CTA walks chunks of 32 samples; gz of the chunk is staged in
// shared memory (broadcast reads), the thread keeps W[:, k] and its n gW accumulators in registers, writes
// gf[b, k] coalesced, and leaves per-CTA partial sums of gW and gbias that a second kernel adds in CTA
// order (deterministic, no atomics).
#include <algorithm>

#include "qnn_internal.h"

namespace {

constexpr int kHeadMaxN = 16;           // n_qubits <= 16 everywhere in this library
constexpr int kHeadMaxK = 4096;         // feature columns
constexpr int kHeadBwdCtasPerSm = 2;    // persistent backward grid (two CTAs per SM): partial sums per CTA
constexpr int kChunk = 32;              // samples per backward chunk

__global__ void __launch_bounds__(256) k_head_fwd(const float* __restrict__ f, const float* __restrict__ W,
                                                  const float* __restrict__ bias, float* __restrict__ x,
                                                  long long B, int K, int n) {
  extern __shared__ float sW[];          // [n][K]
  for (int i = threadIdx.x; i < n * K; i += blockDim.x) sW[i] = W[i];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
  for (long long b = (long long)blockIdx.x * wpc + warp; b < B; b += (long long)gridDim.x * wpc) {
    float acc[kHeadMaxN];
#pragma unroll
    for (int j = 0; j < kHeadMaxN; ++j) acc[j] = 0.f;
    const float* fr = f + (size_t)b * K;
    for (int k = lane; k < K; k += 32) {
      const float fv = fr[k];
#pragma unroll
      for (int j = 0; j < kHeadMaxN; ++j)
        if (j < n) acc[j] = fmaf(fv, sW[j * K + k], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < kHeadMaxN; ++j) {
      if (j < n) {
        float v = acc[j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        acc[j] = v;
      }
    }
    if (lane < n) {
      float v = 0.f;
#pragma unroll
      for (int j = 0; j < kHeadMaxN; ++j) if (j == lane) v = acc[j];
      x[(size_t)b * n + lane] = tanhf(v + bias[lane]);
    }
  }
}

// partial: [gridDim.x][n * K + n]  (gW row-major, then gbias)
__global__ void __launch_bounds__(1024) k_head_bwd(const float* __restrict__ f, const float* __restrict__ W,
                                                   const float* __restrict__ x, const float* __restrict__ gx,
                                                   float* __restrict__ gf, float* __restrict__ partial,
                                                   long long B, int K, int n) {
  __shared__ float sgz[kChunk][kHeadMaxN];
  const int tid = threadIdx.x;
  float* const mine = partial + (size_t)blockIdx.x * ((size_t)n * K + n);
  const long long n_chunks = (B + kChunk - 1) / kChunk;
  float gb = 0.f;                                   // thread j < n: bias gradient of output j
  for (int k0 = 0; k0 < K; k0 += blockDim.x) {      // column tiles (one tile when K <= blockDim.x)
    const int k = k0 + tid;
    float wk[kHeadMaxN], acc[kHeadMaxN];
#pragma unroll
    for (int j = 0; j < kHeadMaxN; ++j) { wk[j] = (j < n && k < K) ? W[(size_t)j * K + k] : 0.f; acc[j] = 0.f; }
    for (long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
      const long long b0 = c * kChunk;
      const int rows = (int)min((long long)kChunk, B - b0);
      __syncthreads();                              // previous chunk's sgz no longer read
      for (int i = tid; i < rows * n; i += blockDim.x) {
        const float xv = x[(size_t)b0 * n + i];
        sgz[i / n][i % n] = gx[(size_t)b0 * n + i] * (1.f - xv * xv);
      }
      __syncthreads();
      if (k0 == 0 && tid < n)
        for (int r = 0; r < rows; ++r) gb += sgz[r][tid];
      if (k < K) {
        for (int r = 0; r < rows; ++r) {
          const float fv = f[(size_t)(b0 + r) * K + k];
          float g = 0.f;
#pragma unroll
          for (int j = 0; j < kHeadMaxN; ++j)
            if (j < n) { const float z = sgz[r][j]; g = fmaf(z, wk[j], g); acc[j] = fmaf(z, fv, acc[j]); }
          if (gf != nullptr) gf[(size_t)(b0 + r) * K + k] = g;
        }
      }
    }
    if (k < K) {
#pragma unroll
      for (int j = 0; j < kHeadMaxN; ++j) if (j < n) mine[(size_t)j * K + k] = acc[j];
    }
  }
  if (tid < n) mine[(size_t)n * K + tid] = gb;
}

__global__ void __launch_bounds__(256) k_head_reduce(const float* __restrict__ partial, int ctas, int count,
                                                     int nK, float* __restrict__ gW, float* __restrict__ gb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float s = 0.f;
  for (int c = 0; c < ctas; ++c) s += partial[(size_t)c * count + i];   // fixed order
  if (i < nK) gW[i] = s; else gb[i - nK] = s;
}

int check_shape(const char* fn, int64_t B, int K, int n) {
  if (B <= 0 || K <= 0 || n <= 0) { qnn_set_error("%s: B=%lld, K=%d, n=%d must be positive", fn, (long long)B, K, n); return QNN_ERR_INVALID; }
  if (n > kHeadMaxN) { qnn_set_error("%s: n=%d exceeds %d outputs", fn, n, kHeadMaxN); return QNN_ERR_INVALID; }
  if (K > kHeadMaxK) { qnn_set_error("%s: K=%d exceeds %d feature columns", fn, K, kHeadMaxK); return QNN_ERR_INVALID; }
  if ((size_t)n * K > 49152) { qnn_set_error("%s: n*K=%lld exceeds 49152 (the weight matrix is staged in shared memory)", fn, (long long)n * K); return QNN_ERR_INVALID; }
  return QNN_OK;
}

int bwd_grid(int64_t B) { return (int)std::max<int64_t>(1, std::min<int64_t>((B + kChunk - 1) / kChunk, (int64_t)qnn_sm_count() * kHeadBwdCtasPerSm)); }

}  // namespace

extern "C" size_t qnn_head_workspace_bytes(int64_t B, int32_t K, int32_t n) {
  if (B <= 0 || K <= 0 || n <= 0 || n > kHeadMaxN || K > kHeadMaxK || (size_t)n * K > 49152) return 0;
  return (size_t)bwd_grid(B) * ((size_t)n * K + n) * sizeof(float);
}

extern "C" int qnn_head_forward(const float* f, const float* W, const float* bias, float* x, int64_t B, int32_t K,
                                int32_t n, void* cuda_stream) {
  if (int rc = check_shape("qnn_head_forward", B, K, n)) return rc;
  if (!f || !W || !bias || !x) { qnn_set_error("qnn_head_forward: NULL buffer"); return QNN_ERR_INVALID; }
  const size_t smem = (size_t)n * K * sizeof(float);
  if (int rc = qnn_allow_smem(reinterpret_cast<const void*>(k_head_fwd), smem)) return rc;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((B + 7) / 8, (int64_t)qnn_sm_count() * 4));
  k_head_fwd<<<grid, 256, smem, static_cast<cudaStream_t>(cuda_stream)>>>(f, W, bias, x, B, K, n);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

extern "C" int qnn_head_backward(const float* f, const float* W, const float* x, const float* gx, float* gf, float* gW,
                                 float* gbias, int64_t B, int32_t K, int32_t n, void* workspace, size_t workspace_bytes,
                                 void* cuda_stream) {
  if (int rc = check_shape("qnn_head_backward", B, K, n)) return rc;
  if (!f || !W || !x || !gx || !gW || !gbias) { qnn_set_error("qnn_head_backward: NULL buffer (only gf may be NULL)"); return QNN_ERR_INVALID; }
  const size_t need = qnn_head_workspace_bytes(B, K, n);
  if (!workspace || workspace_bytes < need) {
    qnn_set_error("qnn_head_backward: workspace of %zu bytes needed, %zu given", need, workspace_bytes);
    return QNN_ERR_WORKSPACE;
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int grid = bwd_grid(B);
  const int threads = std::min(1024, ((K + 31) / 32) * 32);
  k_head_bwd<<<grid, threads, 0, st>>>(f, W, x, gx, gf, static_cast<float*>(workspace), B, K, n);
  QNN_CUDA_CHECK(cudaGetLastError());
  const int count = n * K + n;
  k_head_reduce<<<(count + 255) / 256, 256, 0, st>>>(static_cast<const float*>(workspace), grid, count, n * K, gW, gbias);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}
