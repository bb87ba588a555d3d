// Generic statevector kernels: any gate list, any n <= 16.
//
// The state of one sample lives in shared memory (n <= 13/14) or, for the largest n, in a slice of
// the caller's workspace; a warp (n <= 10) or a whole CTA owns one sample at a time and walks the
// lowered gate list.  Backward is the adjoint method: recompute psi, seed lam = (sum_q gy_q Z_q) psi,
// then walk the gates in reverse applying U^dagger to both vectors while accumulating the 2x2
// transition matrix T of each parametrised gate (qnn_gates.cuh).  Nothing is stored per gate.
//
// These kernels are the coverage path (every circuit, every size).  The register-resident kernels
// in qnn_vm.cu take over for the shapes they support.
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "qnn_gates.cuh"

namespace {

constexpr int kMaxRed = 16;  // values reduced at once by a CTA-wide reduction

__device__ __forceinline__ unsigned insert_zero(unsigned i, int p) {
  return ((i >> p) << (p + 1)) | (i & ((1u << p) - 1u));
}

template <bool WARP>
struct Group {
  int tid, size;
  float* red;  // CTA mode: kMaxRed * 32 floats of shared scratch
  __device__ __forceinline__ void sync() const {
    if (WARP) __syncwarp(); else __syncthreads();
  }
  // Sum K values over the group; every thread receives the totals.  Fixed order: deterministic.
  template <int K>
  __device__ __forceinline__ void reduce(float* v) const {
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if (!WARP) {
      const int warp = tid >> 5, lane = tid & 31, nw = size >> 5;
      __syncthreads();
      if (lane == 0)
#pragma unroll
        for (int k = 0; k < K; ++k) red[k * 32 + warp] = v[k];
      __syncthreads();
#pragma unroll
      for (int k = 0; k < K; ++k) {
        float s = 0.f;
        for (int wv = 0; wv < nw; ++wv) s += red[k * 32 + wv];
        v[k] = s;
      }
    }
  }
};

__device__ __forceinline__ float2 cfma(float2 u, float2 a, float2 acc) {  // acc + u*a
  acc.x = fmaf(u.x, a.x, acc.x);
  acc.x = fmaf(-u.y, a.y, acc.x);
  acc.y = fmaf(u.x, a.y, acc.y);
  acc.y = fmaf(u.y, a.x, acc.y);
  return acc;
}
__device__ __forceinline__ float2 cfma_conj(float2 u, float2 a, float2 acc) {  // acc + conj(u)*a
  acc.x = fmaf(u.x, a.x, acc.x);
  acc.x = fmaf(u.y, a.y, acc.x);
  acc.y = fmaf(u.x, a.y, acc.y);
  acc.y = fmaf(-u.y, a.x, acc.y);
  return acc;
}

// ---- coefficient table for batch-invariant gates (double-precision trig, once per call) ----
__global__ void k_prep_coef(const DevOp* __restrict__ ops, int n_ops, const float* __restrict__ w,
                            float* __restrict__ coef) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_ops) return;
  DevOp op = ops[i];
  if (op.mode != 1) return;
  double a[3] = {0, 0, 0};
  for (int k = 0; k < op.nang; ++k)
    a[k] = (op.src[k] == QNN_SRC_CONST) ? (double)op.value[k] : (double)w[op.src[k]];
  make_coef<double>(op.kind, a, coef + (size_t)op.coef * QNN_COEF_STRIDE);
}

// Per-sample gate: thread 0 of the group builds U (and G_k) into shared memory.
template <bool WARP>
__device__ __forceinline__ const float* sample_coef(const Group<WARP>& g, const DevOp& op,
                                                    const float* __restrict__ xrow,
                                                    const float* __restrict__ w, float* scoef) {
  g.sync();
  if (g.tid == 0) {
    float a[3] = {0, 0, 0};
    for (int k = 0; k < op.nang; ++k) {
      int s = op.src[k];
      a[k] = (s == QNN_SRC_CONST) ? op.value[k] : (s >= 0 ? w[s] : xrow[-1 - s]);
    }
    make_coef<float>(op.kind, a, scoef);
  }
  g.sync();
  return scoef;
}

template <bool WARP>
__device__ __forceinline__ void init_state(const Group<WARP>& g, float2* psi, int n, int encoder,
                                           const float* __restrict__ xrow, int n_inputs,
                                           float* nrm_out) {
  const unsigned dim = 1u << n;
  if (encoder == QNN_ENC_AMPLITUDE) {
    const unsigned kk = min(dim, (unsigned)n_inputs);
    float s[1] = {0.f};
    for (unsigned k = g.tid; k < kk; k += g.size) {
      float v = xrow[k];
      s[0] = fmaf(v, v, s[0]);
    }
    g.template reduce<1>(s);
    const float nrm = sqrtf(s[0]);
    const float inv = nrm > 0.f ? 1.f / nrm : 0.f;   // an all-zero row stays the zero state (y = 0, zero gradients)
    for (unsigned k = g.tid; k < dim; k += g.size)
      psi[k] = make_float2(k < kk ? xrow[k] * inv : 0.f, 0.f);
    *nrm_out = nrm;
  } else {
    for (unsigned k = g.tid; k < dim; k += g.size) psi[k] = make_float2(k == 0 ? 1.f : 0.f, 0.f);
  }
  g.sync();
}

// Apply one gate (forward direction) to one vector.
template <bool WARP>
__device__ __forceinline__ void apply_fwd(const Group<WARP>& g, float2* st, int n, const DevOp& op,
                                          const float* __restrict__ U) {
  const unsigned tb = 1u << op.tpos;
  if (op.kind == QNN_CNOT || op.kind == QNN_CZ || op.kind == QNN_CU3) {
    const unsigned cb = 1u << op.cpos;
    const int lo = min(op.tpos, op.cpos), hi = max(op.tpos, op.cpos);
    const unsigned cnt = 1u << (n - 2);
    float2 u00, u01, u10, u11;
    if (op.kind == QNN_CU3) {
      u00 = make_float2(U[0], U[1]); u01 = make_float2(U[2], U[3]);
      u10 = make_float2(U[4], U[5]); u11 = make_float2(U[6], U[7]);
    }
    for (unsigned i = g.tid; i < cnt; i += g.size) {
      unsigned i0 = insert_zero(insert_zero(i, lo), hi) | cb;
      unsigned i1 = i0 | tb;
      float2 a0 = st[i0], a1 = st[i1];
      if (op.kind == QNN_CNOT) {
        st[i0] = a1; st[i1] = a0;
      } else if (op.kind == QNN_CZ) {
        st[i1] = make_float2(-a1.x, -a1.y);
      } else {
        st[i0] = cfma(u01, a1, cfma(u00, a0, make_float2(0.f, 0.f)));
        st[i1] = cfma(u11, a1, cfma(u10, a0, make_float2(0.f, 0.f)));
      }
    }
  } else {
    const unsigned cnt = 1u << (n - 1);
    const float2 u00 = make_float2(U[0], U[1]), u01 = make_float2(U[2], U[3]);
    const float2 u10 = make_float2(U[4], U[5]), u11 = make_float2(U[6], U[7]);
    for (unsigned i = g.tid; i < cnt; i += g.size) {
      unsigned i0 = insert_zero(i, op.tpos), i1 = i0 | tb;
      float2 a0 = st[i0], a1 = st[i1];
      st[i0] = cfma(u01, a1, cfma(u00, a0, make_float2(0.f, 0.f)));
      st[i1] = cfma(u11, a1, cfma(u10, a0, make_float2(0.f, 0.f)));
    }
  }
  g.sync();
}

// Reverse step for a parametrised gate: accumulate T from the AFTER vectors, then apply U^dagger
// to psi and lam.  T layout: (T00, T01, T10, T11) as (re, im) pairs.
template <bool WARP>
__device__ __forceinline__ void apply_bwd(const Group<WARP>& g, float2* psi, float2* lam, int n,
                                          const DevOp& op, const float* __restrict__ U, float* T) {
  const unsigned tb = 1u << op.tpos;
  const bool ctl = op.cpos >= 0;
  const unsigned cb = ctl ? (1u << op.cpos) : 0u;
  const int lo = ctl ? min(op.tpos, op.cpos) : 0, hi = ctl ? max(op.tpos, op.cpos) : 0;
  const unsigned cnt = 1u << (ctl ? n - 2 : n - 1);
  const float2 u00 = make_float2(U[0], U[1]), u01 = make_float2(U[2], U[3]);
  const float2 u10 = make_float2(U[4], U[5]), u11 = make_float2(U[6], U[7]);
  float2 t00 = {0, 0}, t01 = {0, 0}, t10 = {0, 0}, t11 = {0, 0};
  const float2 z = make_float2(0.f, 0.f);
  for (unsigned i = g.tid; i < cnt; i += g.size) {
    unsigned i0 = ctl ? (insert_zero(insert_zero(i, lo), hi) | cb) : insert_zero(i, op.tpos);
    unsigned i1 = i0 | tb;
    float2 a0 = psi[i0], a1 = psi[i1], l0 = lam[i0], l1 = lam[i1];
    t00 = cfma_conj(l0, a0, t00); t01 = cfma_conj(l0, a1, t01);
    t10 = cfma_conj(l1, a0, t10); t11 = cfma_conj(l1, a1, t11);
    psi[i0] = cfma_conj(u10, a1, cfma_conj(u00, a0, z));
    psi[i1] = cfma_conj(u11, a1, cfma_conj(u01, a0, z));
    lam[i0] = cfma_conj(u10, l1, cfma_conj(u00, l0, z));
    lam[i1] = cfma_conj(u11, l1, cfma_conj(u01, l0, z));
  }
  T[0] = t00.x; T[1] = t00.y; T[2] = t01.x; T[3] = t01.y;
  T[4] = t10.x; T[5] = t10.y; T[6] = t11.x; T[7] = t11.y;
  g.template reduce<8>(T);
  g.sync();
}

struct KArgs {
  const DevOp* ops;
  int n_ops, n, n_inputs, encoder, n_slots;
  const float* x;
  const float* w;
  const float* coef;
  float* y;           // forward
  const float* gy;    // backward
  float* gx;
  float* partial;     // [n_groups][n_slots]
  float2* gstate;     // mode 2: global state, per CTA
  long long B;
};

template <bool WARP, bool GLOBAL_STATE, bool BACKWARD>
__global__ void __launch_bounds__(WARP ? 256 : 1024) k_generic(KArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n = a.n;
  const unsigned dim = 1u << n;
  const int groups_per_cta = WARP ? (blockDim.x >> 5) : 1;
  const int gl = WARP ? (threadIdx.x >> 5) : 0;  // group index within the CTA
  Group<WARP> g;
  g.tid = WARP ? (threadIdx.x & 31) : threadIdx.x;
  g.size = WARP ? 32 : blockDim.x;

  // shared-memory carve-up: [red | per-group: psi, lam, gacc, gxacc, scoef]
  float* red = reinterpret_cast<float*>(smem_raw);
  g.red = red;
  size_t off = WARP ? 0 : kMaxRed * 32 * sizeof(float);
  const size_t state_bytes = GLOBAL_STATE ? 0 : (size_t)dim * sizeof(float2) * (BACKWARD ? 2 : 1);
  const int n_gx = (a.encoder == QNN_ENC_AMPLITUDE) ? 0 : a.n_inputs;
  const size_t aux_floats = (BACKWARD ? (a.n_slots + n_gx) : 0) + QNN_COEF_STRIDE;
  const size_t per_group = state_bytes + ((aux_floats * sizeof(float) + 15) & ~size_t(15));
  unsigned char* base = smem_raw + off + per_group * gl;
  float2* psi;
  float2* lam = nullptr;
  if (GLOBAL_STATE) {
    psi = a.gstate + (size_t)blockIdx.x * dim * (BACKWARD ? 2 : 1);
    if (BACKWARD) lam = psi + dim;
  } else {
    psi = reinterpret_cast<float2*>(base);
    if (BACKWARD) lam = psi + dim;
  }
  float* aux = reinterpret_cast<float*>(base + state_bytes);
  float* scoef = aux;
  float* gacc = aux + QNN_COEF_STRIDE;       // [n_slots]   (backward)
  float* gxacc = gacc + a.n_slots;           // [n_gx]      (backward)

  const long long n_groups = (long long)gridDim.x * groups_per_cta;
  const long long gid = (long long)blockIdx.x * groups_per_cta + gl;

  if (BACKWARD) {
    for (int s = g.tid; s < a.n_slots; s += g.size) gacc[s] = 0.f;
    g.sync();
  }

  for (long long b = gid; b < a.B; b += n_groups) {
    const float* xrow = a.x + (size_t)b * a.n_inputs;
    float nrm = 1.f;
    init_state<WARP>(g, psi, n, a.encoder, xrow, a.n_inputs, &nrm);

    for (int i = 0; i < a.n_ops; ++i) {
      const DevOp op = a.ops[i];
      const float* U = nullptr;
      if (op.mode == 1) U = a.coef + (size_t)op.coef * QNN_COEF_STRIDE;
      else if (op.mode == 2) U = sample_coef<WARP>(g, op, xrow, a.w, scoef);
      apply_fwd<WARP>(g, psi, n, op, U);
    }

    if constexpr (!BACKWARD) {
      float acc[kMaxRed];
#pragma unroll
      for (int q = 0; q < kMaxRed; ++q) acc[q] = 0.f;
      for (unsigned k = g.tid; k < dim; k += g.size) {
        float2 v = psi[k];
        float p = fmaf(v.x, v.x, v.y * v.y);
#pragma unroll
        for (int q = 0; q < kMaxRed; ++q)
          if (q < n) acc[q] += ((k >> (n - 1 - q)) & 1u) ? -p : p;
      }
      g.template reduce<kMaxRed>(acc);
#pragma unroll
      for (int q = 0; q < kMaxRed; ++q)
        if (q < n && g.tid == q) a.y[(size_t)b * n + q] = acc[q];
      g.sync();
    } else {
    // ---- backward ----
    {
      float gyr[kMaxRed];
#pragma unroll
      for (int q = 0; q < kMaxRed; ++q) gyr[q] = (q < n) ? a.gy[(size_t)b * n + q] : 0.f;
      for (unsigned k = g.tid; k < dim; k += g.size) {
        float c = 0.f;
#pragma unroll
        for (int q = 0; q < kMaxRed; ++q)
          if (q < n) c += ((k >> (n - 1 - q)) & 1u) ? -gyr[q] : gyr[q];
        float2 v = psi[k];
        lam[k] = make_float2(c * v.x, c * v.y);
      }
      for (int j = g.tid; j < n_gx; j += g.size) gxacc[j] = 0.f;
      g.sync();
    }
    for (int i = a.n_ops - 1; i >= 0; --i) {
      const DevOp op = a.ops[i];
      if (op.mode == 0) {  // CNOT / CZ: involutions
        apply_fwd<WARP>(g, psi, n, op, nullptr);
        apply_fwd<WARP>(g, lam, n, op, nullptr);
        continue;
      }
      const float* C = (op.mode == 1) ? a.coef + (size_t)op.coef * QNN_COEF_STRIDE
                                      : sample_coef<WARP>(g, op, xrow, a.w, scoef);
      float T[8];
      apply_bwd<WARP>(g, psi, lam, n, op, C, T);
      if (g.tid == 0) {  // one thread: two angles may share a source
        for (int k = 0; k < op.nang; ++k) {
          const int s = op.src[k];
          if (s == QNN_SRC_CONST) continue;
          const float* G = C + 8 + 8 * k;
          float gr = 0.f;
#pragma unroll
          for (int e = 0; e < 4; ++e) gr += G[2 * e] * T[2 * e] - G[2 * e + 1] * T[2 * e + 1];
          gr *= 2.f;
          if (s >= 0) gacc[op.slot + k] += gr;
          else gxacc[-1 - s] += gr;
        }
      }
      g.sync();
    }
    if (a.gx != nullptr) {
      float* gxrow = a.gx + (size_t)b * a.n_inputs;
      if (a.encoder == QNN_ENC_AMPLITUDE) {
        const unsigned kk = min(dim, (unsigned)a.n_inputs);
        const float inv = nrm > 0.f ? 1.f / nrm : 0.f;
        float d[1] = {0.f};
        for (unsigned k = g.tid; k < kk; k += g.size) d[0] = fmaf(xrow[k] * inv, 2.f * lam[k].x, d[0]);
        g.template reduce<1>(d);
        for (unsigned k = g.tid; k < (unsigned)a.n_inputs; k += g.size)
          gxrow[k] = (k < kk) ? (2.f * lam[k].x - xrow[k] * inv * d[0]) * inv : 0.f;
      } else {
        for (int j = g.tid; j < n_gx; j += g.size) gxrow[j] = gxacc[j];
      }
    }
    g.sync();
    }  // BACKWARD
  }

  if (BACKWARD) {
    g.sync();
    for (int s = g.tid; s < a.n_slots; s += g.size) a.partial[(size_t)gid * a.n_slots + s] = gacc[s];
  }
}

// gw[i] = sum over groups and over the slots that map to weight i, in a fixed order.
__global__ void k_reduce_gw(const float* __restrict__ partial, int n_groups, int n_slots,
                            const int32_t* __restrict__ slot_ptr, const int32_t* __restrict__ slot_idx,
                            float* __restrict__ gw) {
  __shared__ double sred[128];
  const int i = blockIdx.x;
  double acc = 0.0;
  for (int e = slot_ptr[i]; e < slot_ptr[i + 1]; ++e) {
    const int s = slot_idx[e];
    for (int gidx = threadIdx.x; gidx < n_groups; gidx += blockDim.x)
      acc += (double)partial[(size_t)gidx * n_slots + s];
  }
  sred[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if (threadIdx.x < o) sred[threadIdx.x] += sred[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) gw[i] = (float)sred[0];
}

size_t smem_needed(const qnn_circuit* c, bool warp, bool global_state, bool backward, int gpc) {
  const size_t dim = (size_t)1 << c->n_qubits;
  const size_t state = global_state ? 0 : dim * sizeof(float2) * (backward ? 2 : 1);
  const int n_gx = (c->encoder == QNN_ENC_AMPLITUDE) ? 0 : c->n_inputs;
  const size_t auxf = (backward ? (size_t)(c->n_slots + n_gx) : 0) + QNN_COEF_STRIDE;
  const size_t per_group = state + ((auxf * sizeof(float) + 15) & ~size_t(15));
  return (warp ? 0 : kMaxRed * 32 * sizeof(float)) + per_group * gpc;
}

constexpr size_t kSmemBudget = 200 * 1024;  // per CTA; leaves room for static shared + carve-out slack

struct SubPlan {
  int mode, threads, gpc, grid;
  size_t smem, ws_state;
};

SubPlan sub_plan(const qnn_circuit* c, int64_t B, bool backward) {
  SubPlan s{};
  const int n = c->n_qubits;
  const size_t dim = (size_t)1 << n;
  const int num_sm = c->num_sm;
  if (n <= 10 && smem_needed(c, true, false, backward, 1) <= kSmemBudget) {
    s.mode = 0;
    int gpc = 8;
    while (gpc > 1 && smem_needed(c, true, false, backward, gpc) > kSmemBudget / 2) gpc >>= 1;
    while (gpc > 1 && smem_needed(c, true, false, backward, gpc) > kSmemBudget) gpc >>= 1;
    s.gpc = gpc;
    s.threads = 32 * gpc;
    s.smem = smem_needed(c, true, false, backward, gpc);
    int64_t ctas = (B + gpc - 1) / gpc;
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (220 * 1024) / std::max<size_t>(s.smem, 1)));
    s.grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, (int64_t)num_sm * per_sm));
    return s;
  }
  s.gpc = 1;
  if (smem_needed(c, false, false, backward, 1) <= kSmemBudget) {
    s.mode = 1;
    s.threads = (int)std::min<size_t>(1024, std::max<size_t>(64, dim / 8));
    s.smem = smem_needed(c, false, false, backward, 1);
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(4, (220 * 1024) / s.smem));
    per_sm = std::min(per_sm, 2048 / s.threads);
    s.grid = (int)std::max<int64_t>(1, std::min<int64_t>(B, (int64_t)num_sm * per_sm));
    return s;
  }
  s.mode = 2;
  s.threads = 1024;
  s.smem = smem_needed(c, false, true, backward, 1);
  s.grid = (int)std::max<int64_t>(1, std::min<int64_t>(B, (int64_t)num_sm * 2));
  s.ws_state = (size_t)s.grid * dim * sizeof(float2) * (backward ? 2 : 1);
  return s;
}

size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

template <bool BACKWARD>
int launch(const SubPlan& s, const KArgs& a, cudaStream_t st) {
  auto go = [&](auto kern) -> int {
    if (int rc = qnn_allow_smem(reinterpret_cast<const void*>(kern), s.smem)) return rc;
    kern<<<s.grid, s.threads, s.smem, st>>>(a);
    QNN_CUDA_CHECK(cudaGetLastError());
    return QNN_OK;
  };
  if (s.mode == 0) return go(k_generic<true, false, BACKWARD>);
  if (s.mode == 1) return go(k_generic<false, false, BACKWARD>);
  return go(k_generic<false, true, BACKWARD>);
}

}  // namespace

int generic_plan(const qnn_circuit* c, int64_t B, GenericPlan* p) {
  SubPlan f = sub_plan(c, B, false), b = sub_plan(c, B, true);
  p->mode = b.mode;
  p->threads = b.threads;
  p->groups_per_cta = b.gpc;
  p->grid = b.grid;
  p->smem_fwd = f.smem;
  p->smem_bwd = b.smem;
  p->ws_coef = align256((size_t)std::max(c->n_coef, 1) * QNN_COEF_STRIDE * sizeof(float));
  p->ws_state_fwd = align256(f.ws_state);
  p->ws_state_bwd = align256(b.ws_state);
  p->ws_partial = align256((size_t)b.grid * b.gpc * std::max(c->n_slots, 1) * sizeof(float));
  static const char* names[3] = {"generic-warp", "generic-cta", "generic-cta-global"};
  p->name = names[b.mode];
  return QNN_OK;
}

int generic_forward(const qnn_circuit* c, const GenericPlan& p, const float* x, const float* w,
                    float* y, int64_t B, void* ws, cudaStream_t st) {
  SubPlan s = sub_plan(c, B, false);
  unsigned char* wsb = static_cast<unsigned char*>(ws);
  float* coef = reinterpret_cast<float*>(wsb);
  if (c->n_coef > 0) {
    k_prep_coef<<<(c->dev_ops_h.size() + 127) / 128, 128, 0, st>>>(c->d_ops, (int)c->dev_ops_h.size(), w, coef);
    QNN_CUDA_CHECK(cudaGetLastError());
  }
  KArgs a{};
  a.ops = c->d_ops; a.n_ops = (int)c->dev_ops_h.size(); a.n = c->n_qubits; a.n_inputs = c->n_inputs;
  a.encoder = c->encoder; a.n_slots = c->n_slots; a.x = x; a.w = w; a.coef = coef; a.y = y; a.B = B;
  a.gstate = reinterpret_cast<float2*>(wsb + p.ws_coef);
  return launch<false>(s, a, st);
}

int generic_backward(const qnn_circuit* c, const GenericPlan& p, const float* x, const float* w,
                     const float* gy, float* gx, float* gw, int64_t B, void* ws, cudaStream_t st) {
  SubPlan s = sub_plan(c, B, true);
  unsigned char* wsb = static_cast<unsigned char*>(ws);
  float* coef = reinterpret_cast<float*>(wsb);
  if (c->n_coef > 0) {
    k_prep_coef<<<(c->dev_ops_h.size() + 127) / 128, 128, 0, st>>>(c->d_ops, (int)c->dev_ops_h.size(), w, coef);
    QNN_CUDA_CHECK(cudaGetLastError());
  }
  KArgs a{};
  a.ops = c->d_ops; a.n_ops = (int)c->dev_ops_h.size(); a.n = c->n_qubits; a.n_inputs = c->n_inputs;
  a.encoder = c->encoder; a.n_slots = c->n_slots; a.x = x; a.w = w; a.coef = coef; a.gy = gy; a.gx = gx;
  a.B = B;
  a.gstate = reinterpret_cast<float2*>(wsb + p.ws_coef);
  a.partial = reinterpret_cast<float*>(wsb + p.ws_coef + p.ws_state_bwd);
  int rc = launch<true>(s, a, st);
  if (rc != QNN_OK) return rc;
  if (c->n_weights > 0) {
    k_reduce_gw<<<c->n_weights, 128, 0, st>>>(a.partial, s.grid * s.gpc, c->n_slots, c->d_slot_ptr,
                                              c->d_slot_idx, gw);
    QNN_CUDA_CHECK(cudaGetLastError());
  }
  return QNN_OK;
}

// ------------------------------------------------------------------------------------------------
// qnn_backward_static: the adjoint backward for callers that cannot hand over scratch memory.
//
// Who that is: a per-slot channel-estimation loop in a base-station PHY, or any real-time service,
// sizes all of its device memory at start-up and then forbids allocation, because cudaMalloc may
// synchronise the whole device and takes an unbounded time; a process whose HBM has been handed in
// full to another owner's pool (an inference server, a framework arena) would simply get
// out-of-memory from cudaMalloc; and a plain C or FFI caller has no caching allocator to borrow a
// workspace from.  For them the scratch that qnn_backward takes from its `workspace` argument (gate
// coefficient table, per-group gradient partials, and for n >= 14 the state vectors) lives here, in
// one array that is part of the library image: every GPU that loads libqnn_b200.so carries
// QNN_STATIC_SCRATCH_BYTES of it until the process exits.
//
// In host code `g_static_scratch` is only a handle for the symbol, not a device address, so the
// address is looked up once per device with cudaGetSymbolAddress (that device current) and cached.
// The lookup happens on the first call on each device: make that call before starting any stream
// capture.  All calls share the one array and it keeps whatever the previous call left, so two calls
// whose kernels can overlap on the GPU (different streams, or threads, without ordering) race; issue
// them on one stream or order them yourself.  Nothing in the array has to start at zero: the kernels
// write every coefficient row, every partial and every amplitude they later read.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[QNN_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_static_ptr[kMaxDevices] = {nullptr};
std::mutex g_static_mu;

int static_scratch_for_current_device(void** out) {
  int dev = -1;
  QNN_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    qnn_set_error("qnn_backward_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    return QNN_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    QNN_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return QNN_OK;
}
}  // namespace

extern "C" size_t qnn_static_scratch_bytes(void) { return QNN_STATIC_SCRATCH_BYTES; }

extern "C" int qnn_backward_static(const qnn_circuit* c, const float* x, const float* w, const float* gy,
                                   float* gx, float* gw, int64_t B, void* cuda_stream) {
  if (int rc = qnn_check_common("qnn_backward_static", c, x, w, B)) return rc;
  if (!gy || (!gw && c->n_weights > 0)) {
    qnn_set_error("qnn_backward_static: NULL gy or gw");
    return QNN_ERR_INVALID;
  }
  GenericPlan p;
  generic_plan(c, B, &p);
  const size_t need = p.ws_coef + p.ws_state_bwd + p.ws_partial;
  if (need > QNN_STATIC_SCRATCH_BYTES) {
    qnn_set_error("qnn_backward_static: this circuit and batch need %zu bytes of scratch but the built-in "
                  "array holds %zu (QNN_STATIC_SCRATCH_BYTES); use qnn_backward with a caller workspace, "
                  "or a smaller circuit", need, (size_t)QNN_STATIC_SCRATCH_BYTES);
    return QNN_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc != QNN_OK) return rc;
  return generic_backward(c, p, x, w, gy, gx, gw, B, scratch, static_cast<cudaStream_t>(cuda_stream));
}
