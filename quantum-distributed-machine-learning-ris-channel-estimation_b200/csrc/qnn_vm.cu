// Register-resident statevector VM kernels (program format: qnn_vm.h, compiler: qnn_vm_compile.cpp).
//
// One warp carries 1024 amplitudes (32 / 2^tb samples of n = 5 + tb qubits); every lane keeps 32 complex
// amplitudes in registers.  The kernel is a persistent interpreter: per micro-op one warp-uniform switch
// selects a fully unrolled body specialised on the register bit(s) it touches, so all state indexing is
// static and the state never leaves the register file except in EXCHANGE, which re-lays it out through a
// warp-private 8 KB shared buffer with XOR (GF(2)-linear) addressing that is bank-conflict free on both
// sides by construction.  FP32 pipe bound; HBM traffic is x, y (and gy, gx) only.
//
// Backward = adjoint method: the forward program is re-run to obtain psi, lam = (sum_q gy_q Z_q) psi is
// seeded, and the program is walked in reverse applying each micro-op's inverse to both vectors while
// gradients are taken from (lam, psi) pairs:
//   real rotation    dL/dtheta = Re sum_pairs conj(lam1) psi0 - conj(lam0) psi1
//   diagonal term    dL/dphi   = sum_p z_bit(p) Im(conj(lam_p) psi_p)
//   general 2x2      T_ab = sum conj(lam_a) psi_b, finished as 2 Re sum G_ab T_ab in k_vm_finalize.
// Weight-gradient accumulators are reduced over the warp by shuffles, kept per warp in shared memory,
// summed per CTA in a fixed order and finished by k_vm_finalize: deterministic.
#include <algorithm>
#include <cstdio>

#include "qnn_gates.cuh"
#include "qnn_vm.h"

namespace vm {
namespace {

constexpr int NA = kAmps;  // 32

struct VmArgs {
  const MicroOp* ops; int n_ops;            // including the END marker
  const Exchange* exchanges; int n_exchanges;
  const DiagDesc* diags; int n_diags;
  const float* coef; int n_coef;            // [n_coef][QNN_COEF_STRIDE] (workspace)
  const float2* big; int n_big;             // [n_big][1024]
  const float2* small; int n_small;         // [n_small][32]
  int n, tb, n_inputs, n_acc;
  int8_t wire_of_bit[16];                   // final (measurement) layout
  const float* x; float* y;                 // forward
  const float* gy; float* gx; float* partial;  // backward; partial [grid][n_acc]
  long long B;
};

// ---- shared-memory access by 32-bit shared address (XOR addressing needs integer addresses) --------
__device__ __forceinline__ void sts64(uint32_t addr, float a, float b) {
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void lds64(uint32_t addr, float& a, float& b) {
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "r"(addr) : "memory");
}

template <int RB>
__device__ __forceinline__ constexpr int pair_lo(int i) { return ((i >> RB) << (RB + 1)) | (i & ((1 << RB) - 1)); }

// index of the i-th register index (i < 8) with bit RT = 0 and bit RC = 1
template <int RT, int RC>
__device__ __forceinline__ constexpr int cpair_lo(int i) {
  constexpr int lo = RT < RC ? RT : RC, hi = RT < RC ? RC : RT;
  int v = ((i >> lo) << (lo + 1)) | (i & ((1 << lo) - 1));
  v = ((v >> hi) << (hi + 1)) | (v & ((1 << hi) - 1));
  return v | (1 << RC);
}

// ---- gate bodies ---------------------------------------------------------------------------------
template <int RB>
__device__ __forceinline__ void g1_real(float (&re)[NA], float (&im)[NA], float c, float s) {
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    const float a0 = re[j0], a1 = re[j1], b0 = im[j0], b1 = im[j1];
    re[j0] = fmaf(-s, a1, c * a0); re[j1] = fmaf(s, a0, c * a1);
    im[j0] = fmaf(-s, b1, c * b0); im[j1] = fmaf(s, b0, c * b1);
  }
}

struct U2 { float r00, i00, r01, i01, r10, i10, r11, i11; };

__device__ __forceinline__ U2 load_u(const float* c) {
  const float4 a = *reinterpret_cast<const float4*>(c), b = *reinterpret_cast<const float4*>(c + 4);
  return U2{a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
}
__device__ __forceinline__ U2 dagger(const U2& u) { return U2{u.r00, -u.i00, u.r10, -u.i10, u.r01, -u.i01, u.r11, -u.i11}; }

__device__ __forceinline__ void apply_pair(const U2& u, float& r0, float& i0, float& r1, float& i1) {
  const float a0 = r0, b0 = i0, a1 = r1, b1 = i1;
  r0 = fmaf(-u.i01, b1, fmaf(u.r01, a1, fmaf(-u.i00, b0, u.r00 * a0)));
  i0 = fmaf(u.i01, a1, fmaf(u.r01, b1, fmaf(u.i00, a0, u.r00 * b0)));
  r1 = fmaf(-u.i11, b1, fmaf(u.r11, a1, fmaf(-u.i10, b0, u.r10 * a0)));
  i1 = fmaf(u.i11, a1, fmaf(u.r11, b1, fmaf(u.i10, a0, u.r10 * b0)));
}

template <int RB>
__device__ __forceinline__ void g1_gen(float (&re)[NA], float (&im)[NA], const U2& u) {
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    apply_pair(u, re[j0], im[j0], re[j1], im[j1]);
  }
}
template <int RT, int RC>
__device__ __forceinline__ void cg1_loc(float (&re)[NA], float (&im)[NA], const U2& u) {
#pragma unroll
  for (int i = 0; i < NA / 4; ++i) {
    const int j0 = cpair_lo<RT, RC>(i), j1 = j0 | (1 << RT);
    apply_pair(u, re[j0], im[j0], re[j1], im[j1]);
  }
}

// T += conj(l_a) p_b over the pairs of register bit RB (T as T00,T01,T10,T11, (re, im))
__device__ __forceinline__ void t_acc(float (&T)[8], float p0r, float p0i, float p1r, float p1i, float l0r, float l0i,
                                      float l1r, float l1i) {
  T[0] = fmaf(l0i, p0i, fmaf(l0r, p0r, T[0])); T[1] = fmaf(-l0i, p0r, fmaf(l0r, p0i, T[1]));
  T[2] = fmaf(l0i, p1i, fmaf(l0r, p1r, T[2])); T[3] = fmaf(-l0i, p1r, fmaf(l0r, p1i, T[3]));
  T[4] = fmaf(l1i, p0i, fmaf(l1r, p0r, T[4])); T[5] = fmaf(-l1i, p0r, fmaf(l1r, p0i, T[5]));
  T[6] = fmaf(l1i, p1i, fmaf(l1r, p1r, T[6])); T[7] = fmaf(-l1i, p1r, fmaf(l1r, p1i, T[7]));
}
template <int RB>
__device__ __forceinline__ void g1_gen_T(const float (&pr)[NA], const float (&pi)[NA], const float (&lr)[NA],
                                         const float (&li)[NA], float (&T)[8]) {
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    t_acc(T, pr[j0], pi[j0], pr[j1], pi[j1], lr[j0], li[j0], lr[j1], li[j1]);
  }
}
template <int RT, int RC>
__device__ __forceinline__ void cg1_loc_T(const float (&pr)[NA], const float (&pi)[NA], const float (&lr)[NA],
                                          const float (&li)[NA], float (&T)[8]) {
#pragma unroll
  for (int i = 0; i < NA / 4; ++i) {
    const int j0 = cpair_lo<RT, RC>(i), j1 = j0 | (1 << RT);
    t_acc(T, pr[j0], pi[j0], pr[j1], pi[j1], lr[j0], li[j0], lr[j1], li[j1]);
  }
}
// Re sum conj(l1) p0 - conj(l0) p1  (gradient of a real rotation angle, from the AFTER vectors)
template <int RB>
__device__ __forceinline__ float g1_real_grad(const float (&pr)[NA], const float (&pi)[NA], const float (&lr)[NA],
                                              const float (&li)[NA]) {
  float v0 = 0.f, v1 = 0.f;
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    v0 = fmaf(lr[j1], pr[j0], v0); v0 = fmaf(li[j1], pi[j0], v0);
    v1 = fmaf(lr[j0], pr[j1], v1); v1 = fmaf(li[j0], pi[j1], v1);
  }
  return v0 - v1;
}

#define RB_SWITCH(rb, CALL)          \
  switch (rb) {                      \
    case 0: { CALL(0); } break;      \
    case 1: { CALL(1); } break;      \
    case 2: { CALL(2); } break;      \
    case 3: { CALL(3); } break;      \
    default: { CALL(4); } break;     \
  }
// (target, control) pairs, control != target
#define RC_CASES(RT, CALL)                                 \
  switch (rc) {                                            \
    case 0: if constexpr (RT != 0) { CALL(RT, 0); } break; \
    case 1: if constexpr (RT != 1) { CALL(RT, 1); } break; \
    case 2: if constexpr (RT != 2) { CALL(RT, 2); } break; \
    case 3: if constexpr (RT != 3) { CALL(RT, 3); } break; \
    default: if constexpr (RT != 4) { CALL(RT, 4); } break; \
  }
#define RTC_SWITCH(rt, CALL)               \
  switch (rt) {                            \
    case 0: RC_CASES(0, CALL) break;       \
    case 1: RC_CASES(1, CALL) break;       \
    case 2: RC_CASES(2, CALL) break;       \
    case 3: RC_CASES(3, CALL) break;       \
    default: RC_CASES(4, CALL) break;      \
  }

// ---- sums over the 32 register amplitudes: total and, per register bit r, the signed sum -------------
// S[r] = sum_j (-1)^{j_r} v_j, computed with a shared tree (61 adds).
template <typename F>
__device__ __forceinline__ void signed_sums(F f, float& total, float (&S)[kR]) {
  float a[16], o0 = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { const float e = f(2 * i), o = f(2 * i + 1); a[i] = e + o; o0 += o; }
  float b[8], o1 = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { b[i] = a[2 * i] + a[2 * i + 1]; o1 += a[2 * i + 1]; }
  float c[4], o2 = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) { c[i] = b[2 * i] + b[2 * i + 1]; o2 += b[2 * i + 1]; }
  const float d0 = c[0] + c[1], d1 = c[2] + c[3];
  const float o3 = c[1] + c[3], o4 = d1;
  total = d0 + d1;
  S[0] = fmaf(-2.f, o0, total); S[1] = fmaf(-2.f, o1, total); S[2] = fmaf(-2.f, o2, total);
  S[3] = fmaf(-2.f, o3, total); S[4] = fmaf(-2.f, o4, total);
}

__device__ __forceinline__ float warp_sum(float v) {   // over all 32 lanes
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float sample_sum(float v, int tb) {   // over the 2^tb lanes of one sample
  for (int b = 0; b < tb; ++b) v += __shfl_xor_sync(0xffffffffu, v, 1 << b);
  return v;
}
// 8 values summed over the warp with the halving butterfly: afterwards lane l holds component
// (l >> 2) & 7 ... precisely: comp = 4*bit4 + 2*bit3 + bit2 of the lane, fully reduced.
__device__ __forceinline__ float warp_sum8(const float (&T)[8], int lane) {
  float k4[4];
  {
    const bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float keep = up ? T[4 + i] : T[i], send = up ? T[i] : T[4 + i];
      k4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  float k2[2];
  {
    const bool up = lane & 8;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float keep = up ? k4[2 + i] : k4[i], send = up ? k4[i] : k4[2 + i];
      k2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  float k1;
  {
    const bool up = lane & 4;
    const float keep = up ? k2[1] : k2[0], send = up ? k2[0] : k2[1];
    k1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  k1 += __shfl_xor_sync(0xffffffffu, k1, 2);
  k1 += __shfl_xor_sync(0xffffffffu, k1, 1);
  return k1;
}

// ---- exchange ---------------------------------------------------------------------------------------
__device__ __forceinline__ void exchange(float (&re)[NA], float (&im)[NA], uint32_t buf, const Exchange* X, int lane,
                                         bool reverse) {
  const uint32_t* wbase = reverse ? X->rbase : X->wbase;
  const uint32_t* woff = reverse ? X->roff : X->woff;
  const uint32_t* rbase = reverse ? X->wbase : X->rbase;
  const uint32_t* roff = reverse ? X->woff : X->roff;
  const uint32_t wa = buf | wbase[lane], ra = buf | rbase[lane];
  __syncwarp();
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(woff + 4 * q);
    sts64(wa ^ o.x, re[4 * q + 0], im[4 * q + 0]);
    sts64(wa ^ o.y, re[4 * q + 1], im[4 * q + 1]);
    sts64(wa ^ o.z, re[4 * q + 2], im[4 * q + 2]);
    sts64(wa ^ o.w, re[4 * q + 3], im[4 * q + 3]);
  }
  __syncwarp();
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(roff + 4 * q);
    lds64(ra ^ o.x, re[4 * q + 0], im[4 * q + 0]);
    lds64(ra ^ o.y, re[4 * q + 1], im[4 * q + 1]);
    lds64(ra ^ o.z, re[4 * q + 2], im[4 * q + 2]);
    lds64(ra ^ o.w, re[4 * q + 3], im[4 * q + 3]);
  }
}

// multiply by a diagonal (optionally its conjugate): big table [j][lane] or register-only table [j]
template <bool CONJ>
__device__ __forceinline__ void diag_mul(float (&re)[NA], float (&im)[NA], const float2* tbl, int stride, int lane_off) {
#pragma unroll
  for (int j = 0; j < NA; ++j) {
    const float2 d = tbl[j * stride + lane_off];
    const float dr = d.x, di = CONJ ? -d.y : d.y;
    const float a = re[j], b = im[j];
    re[j] = fmaf(-di, b, dr * a);
    im[j] = fmaf(di, a, dr * b);
  }
}

struct Smem {
  MicroOp* ops;
  Exchange* exchanges;
  DiagDesc* diags;
  float* coef;        // [n_coef][8]
  float2* big;        // [n_big][1024]
  float2* small;      // [n_small][32]
  float* wacc;        // [warps][n_acc]        (backward)
  float* gxacc;       // [warps][32 * n_inputs / ... ]  = [warps][spw * n_inputs]
  uint32_t xbuf;      // shared address of this warp's 8 KB exchange buffer
};

size_t smem_bytes(const Program& p, int n_inputs, int warps, bool bwd) {
  size_t b = 8192;  // alignment slack for the exchange buffers
  b += (size_t)warps * 8192;
  b += p.ops.size() * sizeof(MicroOp);
  b += p.exchanges.size() * sizeof(Exchange);
  b += p.diags.size() * sizeof(DiagDesc);
  b += p.coef_ops.size() * 8 * sizeof(float);
  b += (size_t)p.n_big_tables * 1024 * sizeof(float2);
  b += (size_t)p.n_small_tables * 32 * sizeof(float2);
  if (bwd) {
    b += (size_t)warps * std::max(p.n_acc, 1) * sizeof(float);
    b += (size_t)warps * (32 >> p.tb) * std::max(n_inputs, 1) * sizeof(float);
  }
  return b + 256;  // per-array 16-byte alignment slack
}

template <bool BWD, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_vm(const VmArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int WARPS = THREADS / 32;
  const int tb = a.tb, T = 1 << tb, spw = 32 >> tb;
  const int lis = lane & (T - 1), siw = lane >> tb;

  // ---- carve shared memory: exchange buffers first (8 KB aligned), then the tables -------------
  Smem S;
  {
    const uint32_t s0 = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t xb = (s0 + 8191u) & ~8191u;
    S.xbuf = xb + (uint32_t)warp * 8192u;
    unsigned char* p = smem_raw + (xb - s0) + (size_t)WARPS * 8192;
    auto take = [&](size_t bytes) { unsigned char* r = p; p += (bytes + 15) & ~size_t(15); return r; };
    S.ops = reinterpret_cast<MicroOp*>(take((size_t)a.n_ops * sizeof(MicroOp)));
    S.exchanges = reinterpret_cast<Exchange*>(take((size_t)a.n_exchanges * sizeof(Exchange)));
    S.diags = reinterpret_cast<DiagDesc*>(take((size_t)a.n_diags * sizeof(DiagDesc)));
    S.coef = reinterpret_cast<float*>(take((size_t)a.n_coef * 8 * sizeof(float)));
    S.big = reinterpret_cast<float2*>(take((size_t)a.n_big * 1024 * sizeof(float2)));
    S.small = reinterpret_cast<float2*>(take((size_t)a.n_small * 32 * sizeof(float2)));
    if (BWD) {
      S.wacc = reinterpret_cast<float*>(take((size_t)WARPS * max(a.n_acc, 1) * sizeof(float)));
      S.gxacc = reinterpret_cast<float*>(take((size_t)WARPS * spw * max(a.n_inputs, 1) * sizeof(float)));
    }
  }
  {
    const int tid = threadIdx.x;
    const uint32_t* src; uint32_t* dst; int words;
    src = reinterpret_cast<const uint32_t*>(a.ops); dst = reinterpret_cast<uint32_t*>(S.ops); words = a.n_ops * (int)(sizeof(MicroOp) / 4);
    for (int i = tid; i < words; i += THREADS) dst[i] = src[i];
    src = reinterpret_cast<const uint32_t*>(a.exchanges); dst = reinterpret_cast<uint32_t*>(S.exchanges); words = a.n_exchanges * (int)(sizeof(Exchange) / 4);
    for (int i = tid; i < words; i += THREADS) dst[i] = src[i];
    src = reinterpret_cast<const uint32_t*>(a.diags); dst = reinterpret_cast<uint32_t*>(S.diags); words = a.n_diags * (int)(sizeof(DiagDesc) / 4);
    for (int i = tid; i < words; i += THREADS) dst[i] = src[i];
    for (int i = tid; i < a.n_coef * 8; i += THREADS) S.coef[i] = a.coef[(size_t)(i >> 3) * QNN_COEF_STRIDE + (i & 7)];
    for (int i = tid; i < a.n_big * 1024; i += THREADS) S.big[i] = a.big[i];
    for (int i = tid; i < a.n_small * 32; i += THREADS) S.small[i] = a.small[i];
    if (BWD) {
      for (int i = tid; i < WARPS * max(a.n_acc, 1); i += THREADS) S.wacc[i] = 0.f;
      for (int i = tid; i < WARPS * spw * max(a.n_inputs, 1); i += THREADS) S.gxacc[i] = 0.f;
    }
  }
  __syncthreads();
  float* const wacc = BWD ? S.wacc + (size_t)warp * max(a.n_acc, 1) : nullptr;
  float* const gxacc = BWD ? S.gxacc + (size_t)warp * spw * max(a.n_inputs, 1) : nullptr;

  float pr[NA], pi[NA];
  const long long n_batches = (a.B + spw - 1) / spw;
  for (long long wb = (long long)blockIdx.x * WARPS + warp; wb < n_batches; wb += (long long)gridDim.x * WARPS) {
    const long long bs = wb * spw + siw;
    const bool active = bs < a.B;
    const float* xrow = a.x + (size_t)(active ? bs : a.B - 1) * a.n_inputs;

#pragma unroll
    for (int j = 0; j < NA; ++j) { pr[j] = 0.f; pi[j] = 0.f; }
    pr[0] = lis == 0 ? 1.f : 0.f;

    // ================= forward program =================
    for (int pc = 0;; ++pc) {
      const MicroOp op = S.ops[pc];
      if (op.kind == MOP_END) break;
      switch (op.kind) {
        case MOP_G1_REAL: {
          const float c = S.coef[op.idx * 8 + 0], s = S.coef[op.idx * 8 + 4];
#define CALL(RB) g1_real<RB>(pr, pi, c, s)
          RB_SWITCH(op.a, CALL)
#undef CALL
        } break;
        case MOP_G1_REAL_X: {
          float s, c;
          sincosf(0.5f * xrow[op.idx], &s, &c);
#define CALL(RB) g1_real<RB>(pr, pi, c, s)
          RB_SWITCH(op.a, CALL)
#undef CALL
        } break;
        case MOP_G1_GEN: {
          const U2 u = load_u(S.coef + op.idx * 8);
#define CALL(RB) g1_gen<RB>(pr, pi, u)
          RB_SWITCH(op.a, CALL)
#undef CALL
        } break;
        case MOP_CG1_L: {
          const U2 u = load_u(S.coef + op.idx * 8);
          const int rc = op.b;
#define CALL(RT, RC) cg1_loc<RT, RC>(pr, pi, u)
          RTC_SWITCH(op.a, CALL)
#undef CALL
        } break;
        case MOP_CG1_P: {
          if (__popc(lis & op.aux) & 1) {
            const U2 u = load_u(S.coef + op.idx * 8);
#define CALL(RB) g1_gen<RB>(pr, pi, u)
            RB_SWITCH(op.a, CALL)
#undef CALL
          }
        } break;
        case MOP_DIAG: diag_mul<false>(pr, pi, S.big + op.idx * 1024, 32, lane); break;
        case MOP_DIAG_REG: diag_mul<false>(pr, pi, S.small + op.idx * 32, 1, 0); break;
        case MOP_XCHG: exchange(pr, pi, S.xbuf, S.exchanges + op.idx, lane, false); break;
        default: break;
      }
    }

    if constexpr (!BWD) {
      // ================= measurement: <Z> per physical bit =================
      float tot, Sr[kR];
      signed_sums([&](int j) { return fmaf(pr[j], pr[j], pi[j] * pi[j]); }, tot, Sr);
      float val[10];
#pragma unroll
      for (int r = 0; r < kR; ++r) val[r] = sample_sum(Sr[r], tb);
#pragma unroll
      for (int b = 0; b < 5; ++b) val[kR + b] = (b < tb) ? sample_sum(((lis >> b) & 1) ? -tot : tot, tb) : 0.f;
      if (active && lis == 0) {
        float* yrow = a.y + (size_t)bs * a.n;
#pragma unroll
        for (int pb = 0; pb < 10; ++pb)
          if (pb < a.n) yrow[a.wire_of_bit[pb]] = val[pb];
      }
    } else {
      // ================= seed lam = (sum_q gy_q Z_q) psi =================
      float lr[NA], li[NA];
      {
        float g[10];
        const float* gyrow = a.gy + (size_t)(active ? bs : 0) * a.n;
#pragma unroll
        for (int pb = 0; pb < 10; ++pb) g[pb] = (active && pb < a.n) ? gyrow[a.wire_of_bit[pb]] : 0.f;
        float cl = 0.f;
#pragma unroll
        for (int b = 0; b < 5; ++b) cl += ((lis >> b) & 1) ? -g[kR + b] : g[kR + b];
        float cj[NA];
        cj[0] = cl + g[0] + g[1] + g[2] + g[3] + g[4];
#pragma unroll
        for (int j = 1; j < NA; ++j) {
          const int low = j & (-j);                      // lowest set bit
          const int r = low == 1 ? 0 : low == 2 ? 1 : low == 4 ? 2 : low == 8 ? 3 : 4;
          cj[j] = fmaf(-2.f, g[r], cj[j ^ low]);
        }
#pragma unroll
        for (int j = 0; j < NA; ++j) { lr[j] = cj[j] * pr[j]; li[j] = cj[j] * pi[j]; }
      }
      // ================= reverse program =================
      for (int pc = a.n_ops - 2; pc >= 0; --pc) {
        const MicroOp op = S.ops[pc];
        switch (op.kind) {
          case MOP_G1_REAL:
          case MOP_G1_REAL_X: {
            float s, c;
            if (op.kind == MOP_G1_REAL) { c = S.coef[op.idx * 8 + 0]; s = S.coef[op.idx * 8 + 4]; }
            else sincosf(0.5f * xrow[op.idx], &s, &c);
            float v = 0.f;
#define CALL(RB) { v = g1_real_grad<RB>(pr, pi, lr, li); g1_real<RB>(pr, pi, c, -s); g1_real<RB>(lr, li, c, -s); }
            RB_SWITCH(op.a, CALL)
#undef CALL
            if (op.kind == MOP_G1_REAL) {
              if (op.acc >= 0) { v = warp_sum(v); if (lane == 0) wacc[op.acc] += v; }
            } else {
              v = sample_sum(v, tb);
              if (lis == 0) gxacc[siw * a.n_inputs + op.idx] += v;
            }
          } break;
          case MOP_G1_GEN: {
            const U2 u = load_u(S.coef + op.idx * 8);
            const U2 ud = dagger(u);
            float Tm[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#define CALL(RB) { if (op.acc >= 0) g1_gen_T<RB>(pr, pi, lr, li, Tm); g1_gen<RB>(pr, pi, ud); g1_gen<RB>(lr, li, ud); }
            RB_SWITCH(op.a, CALL)
#undef CALL
            if (op.acc >= 0) { const float r = warp_sum8(Tm, lane); if ((lane & 3) == 0) wacc[op.acc + (lane >> 2)] += r; }
          } break;
          case MOP_CG1_L: {
            const U2 u = load_u(S.coef + op.idx * 8);
            const U2 ud = dagger(u);
            const int rc = op.b;
            float Tm[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#define CALL(RT, RC) { if (op.acc >= 0) cg1_loc_T<RT, RC>(pr, pi, lr, li, Tm); cg1_loc<RT, RC>(pr, pi, ud); cg1_loc<RT, RC>(lr, li, ud); }
            RTC_SWITCH(op.a, CALL)
#undef CALL
            if (op.acc >= 0) { const float r = warp_sum8(Tm, lane); if ((lane & 3) == 0) wacc[op.acc + (lane >> 2)] += r; }
          } break;
          case MOP_CG1_P: {
            float Tm[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (__popc(lis & op.aux) & 1) {
              const U2 u = load_u(S.coef + op.idx * 8);
              const U2 ud = dagger(u);
#define CALL(RB) { if (op.acc >= 0) g1_gen_T<RB>(pr, pi, lr, li, Tm); g1_gen<RB>(pr, pi, ud); g1_gen<RB>(lr, li, ud); }
              RB_SWITCH(op.a, CALL)
#undef CALL
            }
            if (op.acc >= 0) { const float r = warp_sum8(Tm, lane); if ((lane & 3) == 0) wacc[op.acc + (lane >> 2)] += r; }
          } break;
          case MOP_DIAG:
          case MOP_DIAG_REG: {
            const DiagDesc* D = S.diags + op.aux;
            float tot, Sr[kR];
            signed_sums([&](int j) { return fmaf(lr[j], pi[j], -li[j] * pr[j]); }, tot, Sr);   // Im(conj(lam) psi)
#pragma unroll
            for (int r = 0; r < kR; ++r) {
              const int ai = D->acc_reg[r];
              if (ai >= 0) { const float v = warp_sum(Sr[r]); if (lane == 0) wacc[ai] += v; }
            }
            if (op.kind == MOP_DIAG) {
#pragma unroll
              for (int b = 0; b < 5; ++b) {
                const int ai = D->acc_lane[b];
                if (ai >= 0) { const float v = warp_sum(((lis >> b) & 1) ? -tot : tot); if (lane == 0) wacc[ai] += v; }
              }
              diag_mul<true>(pr, pi, S.big + op.idx * 1024, 32, lane);
              diag_mul<true>(lr, li, S.big + op.idx * 1024, 32, lane);
            } else {
              diag_mul<true>(pr, pi, S.small + op.idx * 32, 1, 0);
              diag_mul<true>(lr, li, S.small + op.idx * 32, 1, 0);
            }
          } break;
          case MOP_XCHG:
            exchange(pr, pi, S.xbuf, S.exchanges + op.idx, lane, true);
            exchange(lr, li, S.xbuf, S.exchanges + op.idx, lane, true);
            break;
          default: break;
        }
      }
      // ================= input gradients of this warp's samples =================
      __syncwarp();
      if (a.gx != nullptr) {
        for (int i = lane; i < spw * a.n_inputs; i += 32) {
          const long long bb = wb * spw + i / a.n_inputs;
          if (bb < a.B) a.gx[(size_t)bb * a.n_inputs + (i % a.n_inputs)] = gxacc[i];
        }
      }
      for (int i = lane; i < spw * a.n_inputs; i += 32) gxacc[i] = 0.f;
      __syncwarp();
    }
  }

  if constexpr (BWD) {
    __syncthreads();
    for (int i = threadIdx.x; i < a.n_acc; i += THREADS) {
      float s = 0.f;
#pragma unroll 1
      for (int w = 0; w < WARPS; ++w) s += S.wacc[(size_t)w * a.n_acc + i];
      a.partial[(size_t)blockIdx.x * a.n_acc + i] = s;
    }
  }
}

// ---- per-call tables: gate coefficients (double-precision trig) and merged diagonals -----------------
__global__ void k_vm_prep(const DevOp* __restrict__ coef_ops, int n_coef, const DiagDesc* __restrict__ diags, int n_diags,
                          const DiagTerm* __restrict__ terms, const float* __restrict__ w, int tb, float* __restrict__ coef,
                          float2* __restrict__ big, float2* __restrict__ small, int coef_blocks) {
  if ((int)blockIdx.x < coef_blocks) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_coef) return;
    const DevOp op = coef_ops[i];
    double a[3] = {0, 0, 0};
    for (int k = 0; k < op.nang; ++k) a[k] = (op.src[k] == QNN_SRC_CONST) ? (double)op.value[k] : (double)w[op.src[k]];
    make_coef<double>(op.kind, a, coef + (size_t)i * QNN_COEF_STRIDE);
    return;
  }
  const int d = blockIdx.x - coef_blocks;
  if (d >= n_diags) return;
  const DiagDesc D = diags[d];
  const int entries = D.is_reg ? 32 : 1024;
  const int T = 1 << tb;
  for (int e = threadIdx.x; e < entries; e += blockDim.x) {
    const int j = D.is_reg ? e : (e >> 5), lane = D.is_reg ? 0 : (e & 31);
    const unsigned p = (unsigned)j | ((unsigned)(lane & (T - 1)) << kR);
    double ang = 0.0;
    for (int t = 0; t < D.n_terms; ++t) {
      const DiagTerm tm = terms[D.first_term + t];
      if (tm.kind == 0) {
        const double phi = (tm.src == QNN_SRC_CONST) ? (double)tm.value : (double)w[tm.src];
        ang += ((p >> tm.bit0) & 1u) ? 0.5 * phi : -0.5 * phi;
      } else if (((p >> tm.bit0) & 1u) && ((p >> tm.bit1) & 1u)) {
        ang += 3.14159265358979323846;
      }
    }
    double s, c;
    sincos(ang, &s, &c);
    const float2 v = make_float2((float)c, (float)s);
    if (D.is_reg) small[(size_t)D.table * 32 + e] = v; else big[(size_t)D.table * 1024 + e] = v;
  }
}

// gw[i] = sum over the gradient slots of weight i, each finished from the CTA partials (fixed order).
__global__ void k_vm_finalize(const float* __restrict__ partial, int n_ctas, int n_acc, const GradSlot* __restrict__ slots,
                              const int32_t* __restrict__ slot_ptr, const int32_t* __restrict__ slot_idx,
                              const float* __restrict__ coef, float* __restrict__ gw) {
  __shared__ double sred[128];
  const int i = blockIdx.x;
  double acc = 0.0;
  for (int e = slot_ptr[i]; e < slot_ptr[i + 1]; ++e) {
    const GradSlot s = slots[slot_idx[e]];
    if (s.kind == 0) {
      for (int c = threadIdx.x; c < n_ctas; c += blockDim.x) acc += (double)partial[(size_t)c * n_acc + s.acc];
    } else {
      const float* G = coef + (size_t)s.coef * QNN_COEF_STRIDE + 8 + 8 * s.angle;
      for (int c = threadIdx.x; c < n_ctas; c += blockDim.x) {
        const float* T = partial + (size_t)c * n_acc + s.acc;
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 4; ++k) v += (double)G[2 * k] * (double)T[2 * k] - (double)G[2 * k + 1] * (double)T[2 * k + 1];
        acc += 2.0 * v;
      }
    }
  }
  sred[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if (threadIdx.x < o) sred[threadIdx.x] += sred[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) gw[i] = (float)sred[0];
}

size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

constexpr int kFwdThreads = 256, kFwdMinB = 2;
constexpr int kBwdThreads = 256, kBwdMinB = 1;
constexpr size_t kSmemMax = 227 * 1024;

template <typename K>
int set_smem(K kern, size_t bytes) {
  if (bytes > 48 * 1024) QNN_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return QNN_OK;
}

void fill_args(const qnn_circuit* c, const Program& p, const Plan& pl, void* ws, VmArgs* a) {
  unsigned char* wsb = static_cast<unsigned char*>(ws);
  a->ops = p.d_ops; a->n_ops = (int)p.ops.size();
  a->exchanges = p.d_exchanges; a->n_exchanges = (int)p.exchanges.size();
  a->diags = p.d_diags; a->n_diags = (int)p.diags.size();
  a->coef = reinterpret_cast<const float*>(wsb); a->n_coef = (int)p.coef_ops.size();
  a->big = reinterpret_cast<const float2*>(wsb + pl.ws_coef); a->n_big = p.n_big_tables;
  a->small = a->big + (size_t)p.n_big_tables * 1024; a->n_small = p.n_small_tables;
  a->n = p.n; a->tb = p.tb; a->n_inputs = c->n_inputs; a->n_acc = p.n_acc;
  for (int b = 0; b < 16; ++b) a->wire_of_bit[b] = 0;
  for (int w = 0; w < p.n; ++w) a->wire_of_bit[p.final_bit_of_wire[w]] = (int8_t)w;
  a->partial = reinterpret_cast<float*>(wsb + pl.ws_coef + pl.ws_diag);
}

int launch_prep(const Program& p, const float* w, const VmArgs& a, cudaStream_t st) {
  const int coef_blocks = ((int)p.coef_ops.size() + 127) / 128;
  const int blocks = coef_blocks + (int)p.diags.size();
  if (blocks == 0) return QNN_OK;
  k_vm_prep<<<blocks, 128, 0, st>>>(p.d_coef_ops, (int)p.coef_ops.size(), p.d_diags, (int)p.diags.size(), p.d_terms, w, p.tb,
                                    const_cast<float*>(a.coef), const_cast<float2*>(a.big), const_cast<float2*>(a.small),
                                    coef_blocks);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

}  // namespace

bool plan(const qnn_circuit* c, const Program& p, int64_t B, Plan* out) {
  if (!p.ok) return false;
  const size_t sf = smem_bytes(p, c->n_inputs, kFwdThreads / 32, false);
  const size_t sb = smem_bytes(p, c->n_inputs, kBwdThreads / 32, true);
  if (sf > kSmemMax || sb > kSmemMax) return false;
  out->threads = kBwdThreads;
  out->warps_per_cta = kBwdThreads / 32;
  out->smem_fwd = sf;
  out->smem_bwd = sb;
  const int64_t batches = (B + (32 >> p.tb) - 1) / (32 >> p.tb);
  const int64_t ctas = (batches + out->warps_per_cta - 1) / out->warps_per_cta;
  out->grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, 148));   // backward: one CTA per SM
  out->ws_coef = align256(std::max<size_t>(p.coef_ops.size(), 1) * QNN_COEF_STRIDE * sizeof(float));
  out->ws_diag = align256(((size_t)p.n_big_tables * 1024 + (size_t)p.n_small_tables * 32 + 1) * sizeof(float2));
  out->ws_partial = align256((size_t)out->grid * std::max(p.n_acc, 1) * sizeof(float));
  return true;
}

static int upload_csr(Program* p, int n_weights);

int upload(Program* p, int n_weights) {
  auto up = [&](void** d, const void* src, size_t bytes) -> int {
    QNN_CUDA_CHECK(cudaMalloc(d, bytes ? bytes : 16));
    if (bytes) QNN_CUDA_CHECK(cudaMemcpy(*d, src, bytes, cudaMemcpyHostToDevice));
    return QNN_OK;
  };
  int rc;
  if ((rc = up((void**)&p->d_ops, p->ops.data(), p->ops.size() * sizeof(MicroOp)))) return rc;
  if ((rc = up((void**)&p->d_coef_ops, p->coef_ops.data(), p->coef_ops.size() * sizeof(DevOp)))) return rc;
  if ((rc = up((void**)&p->d_terms, p->terms.data(), p->terms.size() * sizeof(DiagTerm)))) return rc;
  if ((rc = up((void**)&p->d_diags, p->diags.data(), p->diags.size() * sizeof(DiagDesc)))) return rc;
  if ((rc = up((void**)&p->d_exchanges, p->exchanges.data(), p->exchanges.size() * sizeof(Exchange)))) return rc;
  if ((rc = up((void**)&p->d_slots, p->slots.data(), p->slots.size() * sizeof(GradSlot)))) return rc;
  return upload_csr(p, n_weights);
}

static int upload_csr(Program* p, int n_weights) {
  std::vector<int32_t> ptr(n_weights + 1, 0), idx;
  for (const GradSlot& s : p->slots) ptr[s.weight + 1]++;
  for (int i = 0; i < n_weights; ++i) ptr[i + 1] += ptr[i];
  idx.resize(p->slots.size());
  std::vector<int32_t> fill(ptr.begin(), ptr.end() - 1);
  for (size_t s = 0; s < p->slots.size(); ++s) idx[fill[p->slots[s].weight]++] = (int32_t)s;
  QNN_CUDA_CHECK(cudaMalloc((void**)&p->d_slot_ptr, ptr.size() * sizeof(int32_t)));
  QNN_CUDA_CHECK(cudaMemcpy(p->d_slot_ptr, ptr.data(), ptr.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  QNN_CUDA_CHECK(cudaMalloc((void**)&p->d_slot_idx, std::max<size_t>(idx.size(), 1) * sizeof(int32_t)));
  if (!idx.empty()) QNN_CUDA_CHECK(cudaMemcpy(p->d_slot_idx, idx.data(), idx.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  return QNN_OK;
}

void release(Program* p) {
  for (void* d : {(void*)p->d_ops, (void*)p->d_coef_ops, (void*)p->d_terms, (void*)p->d_diags, (void*)p->d_exchanges,
                  (void*)p->d_slots, (void*)p->d_slot_ptr, (void*)p->d_slot_idx})
    if (d) cudaFree(d);
  p->d_ops = nullptr; p->d_coef_ops = nullptr; p->d_terms = nullptr; p->d_diags = nullptr;
  p->d_exchanges = nullptr; p->d_slots = nullptr; p->d_slot_ptr = nullptr; p->d_slot_idx = nullptr;
}

int forward(const qnn_circuit* c, const Program& p, const Plan& pl, const float* x, const float* w, float* y,
            int64_t B, void* ws, cudaStream_t st) {
  VmArgs a{};
  fill_args(c, p, pl, ws, &a);
  a.x = x; a.y = y; a.B = B;
  int rc = launch_prep(p, w, a, st);
  if (rc) return rc;
  auto kern = k_vm<false, kFwdThreads, kFwdMinB>;
  if ((rc = set_smem(kern, pl.smem_fwd))) return rc;
  const int spw = 32 >> p.tb, wpc = kFwdThreads / 32;
  const int64_t ctas = ((B + spw - 1) / spw + wpc - 1) / wpc;
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(kFwdMinB, kSmemMax / pl.smem_fwd));
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, 148 * per_sm));
  kern<<<grid, kFwdThreads, pl.smem_fwd, st>>>(a);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

int backward(const qnn_circuit* c, const Program& p, const Plan& pl, const float* x, const float* w,
             const float* gy, float* gx, float* gw, int64_t B, void* ws, cudaStream_t st) {
  VmArgs a{};
  fill_args(c, p, pl, ws, &a);
  a.x = x; a.gy = gy; a.gx = gx; a.B = B;
  int rc = launch_prep(p, w, a, st);
  if (rc) return rc;
  auto kern = k_vm<true, kBwdThreads, kBwdMinB>;
  if ((rc = set_smem(kern, pl.smem_bwd))) return rc;
  kern<<<pl.grid, kBwdThreads, pl.smem_bwd, st>>>(a);
  QNN_CUDA_CHECK(cudaGetLastError());
  if (c->n_weights > 0) {
    k_vm_finalize<<<c->n_weights, 128, 0, st>>>(a.partial, pl.grid, p.n_acc, p.d_slots, p.d_slot_ptr, p.d_slot_idx, a.coef, gw);
    QNN_CUDA_CHECK(cudaGetLastError());
  }
  return QNN_OK;
}

}  // namespace vm
