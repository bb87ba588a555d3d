// Register-resident statevector VM kernels (program format: qnn_vm.h, compiler: qnn_vm_compile.cpp).
//
// One warp carries 1024 amplitudes (32 / 2^tb samples of n = 5 + tb qubits); every lane keeps 32 complex
// amplitudes in registers (the narrow variants: 8 or 16 per lane, template parameter KR).  The kernel is a persistent interpreter: per micro-op one warp-uniform switch
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
//   phase on |1>     dL/dphi   = -2 sum_pairs Im(conj(lam1) psi1)        (the two phases of a CU3)
// Weight-gradient accumulators are reduced over the warp by shuffles, kept per warp in shared memory,
// summed per CTA in a fixed order and finished by k_vm_finalize: deterministic.
#pragma once
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "qnn_vm.h"

namespace vm {


struct VmArgs {
  const MicroOp* ops; int n_ops;            // including the END marker
  const Exchange* exchanges; int n_exchanges;
  const DiagDesc* diags; int n_diags;
  const float* coef; int n_coef;            // [n_coef][QNN_COEF_STRIDE] (workspace)
  const float2* big; int n_big;             // [n_big][1024]
  const float2* small; int n_small;         // [n_small][32]
  int n, tb, wb, n_inputs, n_acc;           // tb = n - 5 thread bits; wb = warp bits (a group = 2^wb warps)
  const uint32_t* xbase; int n_xbase;       // per-thread exchange bases
  unsigned table_bytes;                     // shared-memory bytes of all tables (the exchange buffers follow them)
  struct { unsigned ops, exch, xbase, diags, coef, big, small, xcoef, red, wacc, gxacc; } off;   // table offsets (bytes)
  int cluster;                              // CTAs per sample (cluster geometry), else 0
  int amplitude;                            // amplitude encoder (GEOM 0 only): x row -> initial state
  int n_angle;                              // input columns used as angles (0 with the amplitude encoder)
  int8_t init_bit_of_wire[16];              // initial layout (amplitude encoder)
  unsigned char* cbuf;                      // cluster geometry: [clusters][2^n * 8] exchange buffers (workspace)
  float* cred;                              // cluster geometry: [clusters][group warps][48] cross-CTA partials
  int8_t wire_of_bit[16];                   // final (measurement) layout
  const float* x; float* y;                 // forward
  const float* gy; float* gx; float* partial;  // backward; partial [grid][n_acc]
  float4* state;                            // optional: final psi of every warp-batch (forward writes, backward reads)
  long long B;
  // narrow variants: the per-call tables (coefficient rows, diagonals) are made in the kernel's prologue by every
  // CTA instead of by a k_vm_prep launch (small programs, small batches: the launch costs more than the work)
  int inline_prep;
  const DevOp* coef_ops; const DiagTerm* terms; const float* w; int n_terms, n_w;
};

// ---- per-call tables: gate coefficients and merged diagonals (double-precision trig) -------------------------
// row: [al, be, m0, 0, cos l, sin l, cos p, sin p]  (RY rows use al, be only)
__device__ __forceinline__ void make_coef_row(const DevOp& op, const float* __restrict__ w, float* row) {
  double a[3] = {0, 0, 0};
  for (int k = 0; k < op.nang; ++k) a[k] = (op.src[k] == QNN_SRC_CONST) ? (double)op.value[k] : (double)w[op.src[k]];
  const double pi = 3.14159265358979323846;
  double phi = 0.5 * a[0];
  const double k = rint(phi / pi);
  phi -= k * pi;
  row[0] = (float)(-tan(0.5 * phi));
  row[1] = (float)sin(phi);
  row[2] = (((long long)k) & 1LL) ? -1.f : 1.f;
  row[3] = 0.f;
  if (op.nang < 3) { row[4] = 1.f; row[5] = 0.f; row[6] = 1.f; row[7] = 0.f; return; }   // plain rotation: no phases
  double sl, cl, sp, cp;
  sincos(a[2], &sl, &cl);
  sincos(a[1], &sp, &cp);
  row[4] = (float)cl; row[5] = (float)sl; row[6] = (float)cp; row[7] = (float)sp;
}
// phase angle of a merged diagonal at physical index p (register bits below, the sample's thread bits above)
__device__ __forceinline__ double diag_angle(const DiagDesc& D, const DiagTerm* __restrict__ terms, const float* __restrict__ w, unsigned p) {
  double ang = 0.0;
  for (int t = 0; t < D.n_terms; ++t) {
    const DiagTerm tm = terms[D.first_term + t];
    const double v = (tm.src == QNN_SRC_CONST) ? (double)tm.value : (double)w[tm.src];
    if (tm.kind == 0) ang += ((p >> tm.bit0) & 1u) ? 0.5 * v : -0.5 * v;
    else if (((p >> tm.bit0) & 1u) && ((p >> tm.bit1) & 1u)) ang += tm.kind == 1 ? 3.14159265358979323846 : v;
  }
  return ang;
}

// ---- optional self-checks (make checked: -DQNN_VM_CHECKS -> libqnn_b200_checked.so) ----------------------------
// compute-sanitizer is not available on the GPU pool this was developed on; this build checks every exchange
// address against the bounds of the buffer it must fall into (alignment, window, destination CTA) on the device and
// prints violations (`QNN_VM_CHECKS: violation ...`; profiles/experiments/sanitize_driver.py looks for that line).  Not
// compiled into the product library.
#ifdef QNN_VM_CHECKS
static __device__ unsigned long long g_vm_check_failures = 0ULL;   // (one per translation unit: only limits the printing)
__device__ __forceinline__ void vm_check(bool ok, int code, uint32_t v) {
  if (!ok && atomicAdd(&g_vm_check_failures, 1ULL) < 8ULL)
    printf("QNN_VM_CHECKS: violation %d (value 0x%x) block %d thread %d\n", code, v, (int)blockIdx.x, (int)threadIdx.x);
}
// all NA addresses base ^ off[j] (byte offsets relative to `lo`) are 8-byte aligned and below `size`
#define VM_CHECK_OFFSETS(CODE, BASE, OFFS, SIZE)                                                     \
  for (int j_ = 0; j_ < NA; ++j_) { const uint32_t v_ = (uint32_t)(BASE) ^ (OFFS)[j_]; vm_check((v_ & 7u) == 0u && v_ < (uint32_t)(SIZE), CODE, v_); }
#else
#define VM_CHECK_OFFSETS(CODE, BASE, OFFS, SIZE)
#endif

// ---- shared-memory access by 32-bit shared address (XOR addressing needs integer addresses) --------
__device__ __forceinline__ void sts64(uint32_t addr, float a, float b) {
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void lds64(uint32_t addr, float& a, float& b) {
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "r"(addr) : "memory");
}

__device__ __forceinline__ void stg64(unsigned char* base, uint32_t off, float a, float b) {
  __stcg(reinterpret_cast<float2*>(base + off), make_float2(a, b));
}
// Plain (L1-allocating) load: only ever issued after the cluster barrier's acquire, which makes the other
// CTAs' stores visible whatever the cache policy; a reader's neighbouring registers share 32-byte sectors.
__device__ __forceinline__ void ldg64(const unsigned char* base, uint32_t off, float& a, float& b) {
  asm volatile("ld.global.v2.f32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "l"(base + off) : "memory");
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
// The state is kept as (re, im) register pairs and all arithmetic uses Blackwell's packed FP32x2
// instructions (FFMA2 / FMUL2 / FADD2, sm_100+): one issue slot per two FMAs.  Scalar three-register
// FFMAs issue at only ~2/3 rate on this part (register-bank read ports; profiles/experiments/fma_rate.cu),
// so the packed form is what lets the FMA pipe, not the issue port, be the limit.
__device__ __forceinline__ float2 mul2(float s, float2 a) { return __fmul2_rn(make_float2(s, s), a); }
__device__ __forceinline__ float2 fma2(float s, float2 a, float2 c) { return __ffma2_rn(make_float2(s, s), a, c); }
__device__ __forceinline__ float2 swp(float2 a) { return make_float2(a.y, a.x); }

// Real rotation [[c,-s],[s,c]] (c,s = cos,sin(phi)) as three in-place shears
//   a0 += al*a1;  a1 += be*a0;  a0 += al*a1       al = -tan(phi/2), be = sin(phi)
// 3 packed FMAs per amplitude pair, no temporaries; the inverse uses (-al, -be).  phi is range-reduced to
// [-pi/2, pi/2] when the coefficients are made, so |al| = |tan(phi/2)| <= 1.
template <int RB, int NA>
__device__ __forceinline__ void rot(float2 (&st)[NA], float al, float be) {
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    st[j0] = fma2(al, st[j1], st[j0]);
    st[j1] = fma2(be, st[j0], st[j1]);
    st[j0] = fma2(al, st[j1], st[j0]);
  }
}
// Re sum conj(l1) p0 - conj(l0) p1  (gradient of a rotation angle, from the AFTER vectors)
__device__ __forceinline__ void rot_grad_acc(float2& v0, float2& v1, float2 p0, float2 p1, float2 l0, float2 l1) {
  v0 = __ffma2_rn(l1, p0, v0);
  v1 = __ffma2_rn(l0, p1, v1);
}
template <int RB, int NA>
__device__ __forceinline__ float rot_grad(const float2 (&ps)[NA], const float2 (&ls)[NA]) {
  const float2 z = make_float2(0.f, 0.f);
  float2 v0[4] = {z, z, z, z}, v1[4] = {z, z, z, z};   // four independent chains: the FMA latency is exposed otherwise
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    rot_grad_acc(v0[i & 3], v1[i & 3], ps[j0], ps[j1], ls[j0], ls[j1]);
  }
  const float2 a = __fadd2_rn(__fadd2_rn(v0[0], v0[1]), __fadd2_rn(v0[2], v0[3]));
  const float2 b = __fadd2_rn(__fadd2_rn(v1[0], v1[1]), __fadd2_rn(v1[2], v1[3]));
  return (a.x + a.y) - (b.x + b.y);
}

// a <- e * a for a unit complex e = (er, ei), in place: er*a + (-ei, ei) (.) swap(a)
__device__ __forceinline__ float2 cmul(float er, float2 ei2, float2 a) { return __ffma2_rn(ei2, swp(a), mul2(er, a)); }

// Controlled U3 on one control=1 pair: diag(1, e^{i phi}) * m0 R(phi') * diag(1, e^{i lambda}); the sign
// m0 = +-1 of the range reduction is folded into the first phase (a1) and one multiply (a0).
struct Cu3 {
  float al, be, m0, elr, epr;
  float2 eli, epi;     // (-im, +im) of m0*e^{i lambda} and e^{i phi}
};
__device__ __forceinline__ Cu3 make_cu3(float4 a, float4 b, bool inverse) {
  Cu3 g;
  g.m0 = a.z;
  if (!inverse) {
    g.al = a.x; g.be = a.y; g.elr = a.z * b.x; g.eli = make_float2(-a.z * b.y, a.z * b.y);
    g.epr = b.z; g.epi = make_float2(-b.w, b.w);
  } else {   // conjugate phases, opposite shears
    g.al = -a.x; g.be = -a.y; g.elr = a.z * b.x; g.eli = make_float2(a.z * b.y, -a.z * b.y);
    g.epr = b.z; g.epi = make_float2(b.w, -b.w);
  }
  return g;
}
__device__ __forceinline__ Cu3 load_cu3(const float* c, bool inverse) {
  return make_cu3(*reinterpret_cast<const float4*>(c), *reinterpret_cast<const float4*>(c + 4), inverse);
}
// The e^{i phi} post-phase is either applied here (`post`, a separate short loop) or was deferred by the
// compiler into a later merged diagonal (it commutes with everything in between), which removes one
// complex multiply per pair from the forward pass and two from the backward pass.
__device__ __forceinline__ void cu3_pair(const Cu3& g, float2& a0, float2& a1) {
  a1 = cmul(g.elr, g.eli, a1);
  a0 = mul2(g.m0, a0);
  a0 = fma2(g.al, a1, a0);
  a1 = fma2(g.be, a0, a1);
  a0 = fma2(g.al, a1, a0);
}
// Reverse step on one pair, for psi (p) and lam (l) together, g = load_cu3(.., inverse=true).
// Accumulates (as packed pairs, finished by the caller): q_phi, v_theta0/1, q_lambda.  The phi gradient
// -2 sum Im(conj(l1) p1) is invariant under diagonal gates, so it is taken here whether or not the phase
// itself was deferred.
struct Cu3Acc { float2 qp, v0, v1, ql; };
__device__ __forceinline__ void cu3_pair_bwd(const Cu3& g, float2& p0, float2& p1, float2& l0, float2& l1, Cu3Acc& A) {
  A.qp = __ffma2_rn(l1, swp(p1), A.qp);           // (l.re p.im, l.im p.re): Im(conj(l1) p1) = x - y
  rot_grad_acc(A.v0, A.v1, p0, p1, l0, l1);
  p0 = fma2(g.al, p1, p0); p1 = fma2(g.be, p0, p1); p0 = fma2(g.al, p1, p0);
  l0 = fma2(g.al, l1, l0); l1 = fma2(g.be, l0, l1); l0 = fma2(g.al, l1, l0);
  p0 = mul2(g.m0, p0); l0 = mul2(g.m0, l0);
  A.ql = __ffma2_rn(l1, swp(p1), A.ql);
  p1 = cmul(g.elr, g.eli, p1); l1 = cmul(g.elr, g.eli, l1);
}
template <int RT, int RC, int NA>
__device__ __forceinline__ void cu3_loc(float2 (&st)[NA], const Cu3& g, bool post) {
#pragma unroll
  for (int i = 0; i < NA / 4; ++i) {
    const int j0 = cpair_lo<RT, RC>(i), j1 = j0 | (1 << RT);
    cu3_pair(g, st[j0], st[j1]);
  }
  if (post) {
#pragma unroll
    for (int i = 0; i < NA / 4; ++i) { const int j1 = cpair_lo<RT, RC>(i) | (1 << RT); st[j1] = cmul(g.epr, g.epi, st[j1]); }
  }
}
template <int RB, int NA>
__device__ __forceinline__ void cu3_all(float2 (&st)[NA], const Cu3& g, bool post) {
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    cu3_pair(g, st[j0], st[j1]);
  }
  if (post) {
#pragma unroll
    for (int i = 0; i < NA / 2; ++i) { const int j1 = pair_lo<RB>(i) | (1 << RB); st[j1] = cmul(g.epr, g.epi, st[j1]); }
  }
}
template <int RT, int RC, int NA>
__device__ __forceinline__ void cu3_loc_bwd(float2 (&ps)[NA], float2 (&ls)[NA], const Cu3& g, Cu3Acc& A, bool post) {
  if (post) {   // undo the inline post-phase first (g holds the conjugate phases)
#pragma unroll
    for (int i = 0; i < NA / 4; ++i) {
      const int j1 = cpair_lo<RT, RC>(i) | (1 << RT);
      ps[j1] = cmul(g.epr, g.epi, ps[j1]); ls[j1] = cmul(g.epr, g.epi, ls[j1]);
    }
  }
#pragma unroll
  for (int i = 0; i < NA / 4; ++i) {
    const int j0 = cpair_lo<RT, RC>(i), j1 = j0 | (1 << RT);
    cu3_pair_bwd(g, ps[j0], ps[j1], ls[j0], ls[j1], A);
  }
}
template <int RB, int NA>
__device__ __forceinline__ void cu3_all_bwd(float2 (&ps)[NA], float2 (&ls)[NA], const Cu3& g, Cu3Acc& A, bool post) {
  if (post) {
#pragma unroll
    for (int i = 0; i < NA / 2; ++i) {
      const int j1 = pair_lo<RB>(i) | (1 << RB);
      ps[j1] = cmul(g.epr, g.epi, ps[j1]); ls[j1] = cmul(g.epr, g.epi, ls[j1]);
    }
  }
#pragma unroll
  for (int i = 0; i < NA / 2; ++i) {
    const int j0 = pair_lo<RB>(i), j1 = j0 | (1 << RB);
    cu3_pair_bwd(g, ps[j0], ps[j1], ls[j0], ls[j1], A);
  }
}

// ---- sums over the 32 register amplitudes: total and, per register bit r, the signed sum -------------
// S[r] = sum_j (-1)^{j_r} v_j with a shared tree (61 packed adds); v_j are pairs, reduced per half.
// KR < 5 (fewer amplitudes per thread): the same tree, written as a loop over the levels; S[r >= KR] stay unset.
template <int KR, typename F>
__device__ __forceinline__ void signed_sums2_small(F f, float2& total, float2 (&S)[kR]) {
  constexpr int N = 1 << KR;
  const float2 z = make_float2(0.f, 0.f);
  float2 a[N], odd[KR];
#pragma unroll
  for (int i = 0; i < N; ++i) a[i] = f(i);
#pragma unroll
  for (int r = 0; r < KR; ++r) {
    float2 o = z;
#pragma unroll
    for (int i = 0; i < (N >> (r + 1)); ++i) { o = __fadd2_rn(o, a[2 * i + 1]); a[i] = __fadd2_rn(a[2 * i], a[2 * i + 1]); }
    odd[r] = o;
  }
  total = a[0];
#pragma unroll
  for (int r = 0; r < KR; ++r) S[r] = fma2(-2.f, odd[r], total);
}
template <typename F>
__device__ __forceinline__ void signed_sums2(F f, float2& total, float2 (&S)[kR]) {
  const float2 z = make_float2(0.f, 0.f);
  float2 a[16], o0 = z;
#pragma unroll
  for (int i = 0; i < 16; ++i) { const float2 e = f(2 * i), o = f(2 * i + 1); a[i] = __fadd2_rn(e, o); o0 = __fadd2_rn(o0, o); }
  float2 b[8], o1 = z;
#pragma unroll
  for (int i = 0; i < 8; ++i) { b[i] = __fadd2_rn(a[2 * i], a[2 * i + 1]); o1 = __fadd2_rn(o1, a[2 * i + 1]); }
  float2 c[4], o2 = z;
#pragma unroll
  for (int i = 0; i < 4; ++i) { c[i] = __fadd2_rn(b[2 * i], b[2 * i + 1]); o2 = __fadd2_rn(o2, b[2 * i + 1]); }
  const float2 d0 = __fadd2_rn(c[0], c[1]), d1 = __fadd2_rn(c[2], c[3]);
  const float2 o3 = __fadd2_rn(c[1], c[3]), o4 = d1;
  total = __fadd2_rn(d0, d1);
  S[0] = fma2(-2.f, o0, total); S[1] = fma2(-2.f, o1, total); S[2] = fma2(-2.f, o2, total);
  S[3] = fma2(-2.f, o3, total); S[4] = fma2(-2.f, o4, total);
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
// 16 values summed over the warp with the halving butterfly (8+4+2+1+1 = 16 shuffles, depth 5):
// afterwards lane l holds component l >> 1, fully reduced.
__device__ __forceinline__ float warp_sum16(const float (&t)[16], int lane) {
  float k8[8], k4[4], k2[2];
  {
    const bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) k8[i] = (up ? t[8 + i] : t[i]) + __shfl_xor_sync(0xffffffffu, up ? t[i] : t[8 + i], 16);
  }
  {
    const bool up = lane & 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) k4[i] = (up ? k8[4 + i] : k8[i]) + __shfl_xor_sync(0xffffffffu, up ? k8[i] : k8[4 + i], 8);
  }
  {
    const bool up = lane & 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) k2[i] = (up ? k4[2 + i] : k4[i]) + __shfl_xor_sync(0xffffffffu, up ? k4[i] : k4[2 + i], 4);
  }
  float k;
  {
    const bool up = lane & 2;
    k = (up ? k2[1] : k2[0]) + __shfl_xor_sync(0xffffffffu, up ? k2[0] : k2[1], 2);
  }
  k += __shfl_xor_sync(0xffffffffu, k, 1);
  return k;
}

// 4 values summed over the warp with the halving butterfly: afterwards lane l holds component
// 2*bit4 + bit3 of l, fully reduced (9 shuffles instead of 20).
__device__ __forceinline__ float warp_sum4(float t0, float t1, float t2, float t3, int lane) {
  float k0, k1;
  {
    const bool up = lane & 16;
    k0 = (up ? t2 : t0) + __shfl_xor_sync(0xffffffffu, up ? t0 : t2, 16);
    k1 = (up ? t3 : t1) + __shfl_xor_sync(0xffffffffu, up ? t1 : t3, 16);
  }
  float k;
  {
    const bool up = lane & 8;
    k = (up ? k1 : k0) + __shfl_xor_sync(0xffffffffu, up ? k0 : k1, 8);
  }
  k += __shfl_xor_sync(0xffffffffu, k, 4);
  k += __shfl_xor_sync(0xffffffffu, k, 2);
  k += __shfl_xor_sync(0xffffffffu, k, 1);
  return k;
}

// ---- exchange ---------------------------------------------------------------------------------------
// wa / ra: this thread's write / read base (buffer address | per-thread offset); `sync` is the barrier
// of the group that shares the buffer (a warp for n <= 10, 2^wb warps above).
// buf: 0 when wa / ra already carry the (size-aligned) buffer address; else the buffer's shared address, added
// after the XOR (cluster geometry: its two 64 KB buffers are not aligned to their size).
template <int NA>
__device__ __forceinline__ void smem_store_all(const float2 (&st)[NA], uint32_t buf, uint32_t wa, const uint32_t* woff) {
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(woff + 4 * q);
    sts64(buf + (wa ^ o.x), st[4 * q + 0].x, st[4 * q + 0].y);
    sts64(buf + (wa ^ o.y), st[4 * q + 1].x, st[4 * q + 1].y);
    sts64(buf + (wa ^ o.z), st[4 * q + 2].x, st[4 * q + 2].y);
    sts64(buf + (wa ^ o.w), st[4 * q + 3].x, st[4 * q + 3].y);
  }
}
template <int NA>
__device__ __forceinline__ void smem_load_all(float2 (&st)[NA], uint32_t buf, uint32_t ra, const uint32_t* roff) {
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(roff + 4 * q);
    lds64(buf + (ra ^ o.x), st[4 * q + 0].x, st[4 * q + 0].y);
    lds64(buf + (ra ^ o.y), st[4 * q + 1].x, st[4 * q + 1].y);
    lds64(buf + (ra ^ o.z), st[4 * q + 2].x, st[4 * q + 2].y);
    lds64(buf + (ra ^ o.w), st[4 * q + 3].x, st[4 * q + 3].y);
  }
}
template <typename Sync, int NA>
__device__ __forceinline__ void exchange(float2 (&st)[NA], uint32_t buf, uint32_t wa, uint32_t ra, const uint32_t* woff,
                                         const uint32_t* roff, Sync sync) {
  sync();
  smem_store_all(st, buf, wa, woff);
  sync();
  smem_load_all(st, buf, ra, roff);
}

// Cluster geometry, exchange across CTAs through distributed shared memory: every amplitude is stored straight
// into the buffer of the CTA that will hold it.  A slot's byte offset carries the destination CTA's rank in bits
// 16.. and the offset inside that CTA's 64 KB buffer below; `buf` is the buffer's shared address (the same in
// every CTA of the cluster).  TMA is not involved: these are plain st.shared::cluster stores.
__device__ __forceinline__ void st_cluster64(uint32_t local_addr, uint32_t rank, float a, float b) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(remote), "f"(a), "f"(b) : "memory");
}
template <int NA>
__device__ __forceinline__ void dsmem_store_all(const float2 (&st)[NA], uint32_t buf, uint32_t wa, const uint32_t* woff) {
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(woff + 4 * q);
    uint32_t s;
    s = wa ^ o.x; st_cluster64(buf + (s & 0xFFFFu), s >> 16, st[4 * q + 0].x, st[4 * q + 0].y);
    s = wa ^ o.y; st_cluster64(buf + (s & 0xFFFFu), s >> 16, st[4 * q + 1].x, st[4 * q + 1].y);
    s = wa ^ o.z; st_cluster64(buf + (s & 0xFFFFu), s >> 16, st[4 * q + 2].x, st[4 * q + 2].y);
    s = wa ^ o.w; st_cluster64(buf + (s & 0xFFFFu), s >> 16, st[4 * q + 3].x, st[4 * q + 3].y);
  }
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// Cluster geometry (n >= 14): the buffers are slices of the workspace (L2), accessed around L1.  Buffer
// sets alternate from one exchange to the next, so ONE cluster barrier per exchange suffices: a set is
// written again only after every thread has passed the barrier of the exchange in between, i.e. after
// all reads of it have completed.  TWO: psi and lam travel together (backward), sharing that barrier.
template <bool TWO, typename Sync, int NA>
__device__ __forceinline__ void exchange_global(float2 (&st)[NA], float2 (&st2)[NA], unsigned char* buf, unsigned char* buf2,
                                                uint32_t wa, uint32_t ra, const uint32_t* woff, const uint32_t* roff, Sync sync) {
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(woff + 4 * q);
    stg64(buf, wa ^ o.x, st[4 * q + 0].x, st[4 * q + 0].y);
    stg64(buf, wa ^ o.y, st[4 * q + 1].x, st[4 * q + 1].y);
    stg64(buf, wa ^ o.z, st[4 * q + 2].x, st[4 * q + 2].y);
    stg64(buf, wa ^ o.w, st[4 * q + 3].x, st[4 * q + 3].y);
    if (TWO) {
      stg64(buf2, wa ^ o.x, st2[4 * q + 0].x, st2[4 * q + 0].y);
      stg64(buf2, wa ^ o.y, st2[4 * q + 1].x, st2[4 * q + 1].y);
      stg64(buf2, wa ^ o.z, st2[4 * q + 2].x, st2[4 * q + 2].y);
      stg64(buf2, wa ^ o.w, st2[4 * q + 3].x, st2[4 * q + 3].y);
    }
  }
  sync();
#pragma unroll
  for (int q = 0; q < NA / 4; ++q) {
    const uint4 o = *reinterpret_cast<const uint4*>(roff + 4 * q);
    ldg64(buf, ra ^ o.x, st[4 * q + 0].x, st[4 * q + 0].y);
    ldg64(buf, ra ^ o.y, st[4 * q + 1].x, st[4 * q + 1].y);
    ldg64(buf, ra ^ o.z, st[4 * q + 2].x, st[4 * q + 2].y);
    ldg64(buf, ra ^ o.w, st[4 * q + 3].x, st[4 * q + 3].y);
    if (TWO) {
      ldg64(buf2, ra ^ o.x, st2[4 * q + 0].x, st2[4 * q + 0].y);
      ldg64(buf2, ra ^ o.y, st2[4 * q + 1].x, st2[4 * q + 1].y);
      ldg64(buf2, ra ^ o.z, st2[4 * q + 2].x, st2[4 * q + 2].y);
      ldg64(buf2, ra ^ o.w, st2[4 * q + 3].x, st2[4 * q + 3].y);
    }
  }
}

// multiply by a diagonal (optionally its conjugate): big table [j][lane] or register-only table [j].
// d*a = d.re * a + (-d.im, d.im) (.) swap(a); one table load serves psi and lam in the backward pass.
template <bool CONJ, bool TWO, int NA>
__device__ __forceinline__ void diag_mul(float2 (&st)[NA], float2 (&st2)[NA], const float2* tbl, int stride, int lane_off) {
#pragma unroll
  for (int j = 0; j < NA; ++j) {
    const float2 d = tbl[j * stride + lane_off];
    const float2 di = CONJ ? make_float2(d.y, -d.y) : make_float2(-d.y, d.y);
    st[j] = __ffma2_rn(di, swp(st[j]), mul2(d.x, st[j]));
    if (TWO) st2[j] = __ffma2_rn(di, swp(st2[j]), mul2(d.x, st2[j]));
  }
}

// Shear coefficients of RY(x) = R(x/2) for a per-sample angle: (al, be) = (-tan(phi'/2), sin phi') with
// phi' = x/2 reduced to [-pi/2, pi/2]; the dropped sign (-1)^k is a global phase of the sample's state.
__device__ __forceinline__ float2 rot_coef_x(float xv) {
  float phi = 0.5f * xv;
  phi = fmaf(-3.14159265358979323846f, rintf(phi * 0.318309886183790672f), phi);
  float sh, ch;
  sincosf(0.5f * phi, &sh, &ch);
  return make_float2(-sh / ch, 2.f * sh * ch);
}

// Shared-memory form of a micro-op, built once per CTA: what the handlers read, plus the first half of the
// op's coefficient row, so that one dispatch is two consecutive 16-byte loads (no dependent second load).
struct __align__(16) SOp {
  int h;        // handler id | kPostBit when a CU3 applies its post-phase inline
  int idx, acc, aux;
  float4 cf;    // coefficient row [0..3] of the op (zero row if it has none)
};
static_assert(sizeof(SOp) == sizeof(MicroOp), "the table layout reserves sizeof(MicroOp) per op");
constexpr int kPostBit = 256;

struct Smem {
  SOp* ops;
  Exchange* exchanges;
  uint32_t* xbase;
  DiagDesc* diags;
  float* coef;        // [n_coef][8]
  const float2* big;  // [n_big][32 * G]: shared memory for n <= 10, the workspace (L2) above
  float2* small;      // [n_small][32]
  float* wacc;        // [warps][n_acc]        (backward)
  float* gxacc;       // [warps][spw * n_inputs]
  float2* xcoef;      // [warps][spw * n_inputs]: shear coefficients of RY(x[b, col]), made once per batch
  float* red;         // [groups][group warps][16]: cross-warp measurement partials (n > 10)
  uint32_t xbuf;      // shared address of this group's exchange buffer (8 KB per warp, aligned to its size)
};


// GEOM 0: a warp carries whole samples (n <= 10): constant strides, __syncwarp barriers.
// GEOM 1: a sample spans 2^wb warps of one CTA (n = 11..13): shared exchange buffer, named barriers.
// GEOM 2: a sample spans a thread-block cluster (n = 14..16): exchanges through the CTA's own or, across CTAs,
//         distributed shared memory (L2 buffers as the fallback and for lam in the backward pass), cluster barriers,
//         per-thread bases and cross-CTA partial sums in global memory.
// AMP: amplitude encoder (GEOM 0 only); a separate instantiation keeps its gather/scatter code and live
// values out of the angle-encoder kernels.
// KR: register bits (2^KR amplitudes per thread).  5 is the throughput geometry; 3 and 4 spread a batch over four / two
// times as many warps (a warp then carries 256 / 512 amplitudes): lower latency for small batches, less state per thread.
template <bool BWD, bool STATE, int GEOM, bool AMP, int KR = 5>
__global__ void __launch_bounds__(BWD ? 256 : 512, (BWD && KR < 5) ? 2 : 1) k_vm(const VmArgs a) {
  constexpr int NA = 1 << KR;
  // backward: let the finalize kernel behind this one (launched with programmatic stream serialization) be set up
  // now; it waits for this grid's completion before it reads the partial sums
  if constexpr (BWD) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // GEOM 0 always runs with the full CTA (compile-time constants); the others may be launched with fewer warps
  const int WARPS = GEOM == 0 ? ((BWD || KR < 5) ? 8 : 16) : (int)(blockDim.x >> 5);   // narrow variants: 8-warp CTAs both ways
  constexpr bool MW = GEOM >= 1, CL = GEOM == 2;
  const int tb = a.tb, wbits = MW ? a.wb : 0;
  const int gwarps = 1 << wbits, G = 32 << wbits;          // warps / threads per group (one sample for n >= 10)
  unsigned crank = 0, cid = blockIdx.x, ncl = gridDim.x;   // cluster rank / id / count (GEOM 2)
  if constexpr (CL) {
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    cid = blockIdx.x / (unsigned)a.cluster; ncl = gridDim.x / (unsigned)a.cluster;
  }
  const int gw = CL ? (int)crank * WARPS + warp : (warp & (gwarps - 1));   // warp in group
  const int grp = CL ? 0 : (warp >> wbits), groups = CL ? 1 : (WARPS >> wbits);
  const int spw = wbits ? 1 : (32 >> tb);                  // samples per warp (n <= 10)
  const int tig = (gw << 5) | lane;                        // thread in group
  const int tis = wbits ? tig : (lane & ((1 << tb) - 1));  // thread in sample: the sample's thread bits
  const int siw = wbits ? 0 : (lane >> tb);                // sample in warp
  const int lbits = tb < 5 ? tb : 5;                       // thread bits that are lanes of this warp

  // ---- carve shared memory: tables first, at host-computed offsets from the start of dynamic shared memory
  //      (a constant address: nothing about them has to stay in registers), then the group exchange buffers
  //      at the next multiple of their size ------------------------------------------------------------
  Smem S;
  {
    const uint32_t s0 = (uint32_t)__cvta_generic_to_shared(smem_raw);
    // cluster geometry: two CTA-wide 64 KB buffers right behind the tables (addressed as base + offset, so they
    // need no alignment to their size); otherwise one buffer per group at the next multiple of its size
    const uint32_t gbuf = CL ? 65536u : ((256u * NA) << wbits);
    const uint32_t xb = CL ? ((s0 + a.table_bytes + 15u) & ~15u) : ((s0 + a.table_bytes + gbuf - 1u) & ~(gbuf - 1u));
    S.xbuf = xb + (uint32_t)grp * gbuf;
    unsigned char* const p = smem_raw;
    // table offsets come from the host (constant bank): the pointers are then cheap to rematerialise
    S.ops = reinterpret_cast<SOp*>(p + a.off.ops);
    S.exchanges = reinterpret_cast<Exchange*>(p + a.off.exch);
    // n > 10: the per-thread bases (2 * group threads words per exchange, the same for every sample) stay in
    // global memory - two L1-resident loads per exchange instead of up to 27 KB of shared memory
    S.xbase = MW ? const_cast<uint32_t*>(a.xbase) : reinterpret_cast<uint32_t*>(p + a.off.xbase);
    S.diags = reinterpret_cast<DiagDesc*>(p + a.off.diags);
    S.coef = reinterpret_cast<float*>(p + a.off.coef);
    float2* sbig = wbits == 0 ? reinterpret_cast<float2*>(p + a.off.big) : nullptr;
    S.big = wbits == 0 ? sbig : a.big;
    S.small = reinterpret_cast<float2*>(p + a.off.small);
    S.xcoef = reinterpret_cast<float2*>(p + a.off.xcoef);
    S.red = reinterpret_cast<float*>(p + a.off.red);
    if (BWD) {
      S.wacc = reinterpret_cast<float*>(p + a.off.wacc);
      S.gxacc = reinterpret_cast<float*>(p + a.off.gxacc);
    }
    const int tid = threadIdx.x;
    const uint32_t* src; uint32_t* dst; int words;
    bool inl = false;
    if constexpr (KR < 5) inl = a.inline_prep != 0;
    if constexpr (KR < 5) {
      // Small batches: the prologue below reads half a dozen small arrays one after the other, each a cold miss (the
      // caller's other kernels have flushed them out of L2 since the last call), some of them dependent (terms -> w).
      // One round of prefetches brings all of them, and this CTA's first input rows, towards L2 at once.
      auto prefetch = [&](const void* base, size_t bytes) {
        for (size_t o = (size_t)tid * 128; o < bytes; o += (size_t)blockDim.x * 128)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(base) + o));
      };
      prefetch(a.ops, (size_t)a.n_ops * sizeof(MicroOp));
      prefetch(a.exchanges, (size_t)a.n_exchanges * sizeof(Exchange));
      prefetch(a.xbase, (size_t)a.n_xbase * 4);
      prefetch(a.diags, (size_t)a.n_diags * sizeof(DiagDesc));
      if (inl) {
        prefetch(a.terms, (size_t)a.n_terms * sizeof(DiagTerm));
        prefetch(a.coef_ops, (size_t)a.n_coef * sizeof(DevOp));
        prefetch(a.w, (size_t)a.n_w * 4);
      }
      const size_t row0 = (size_t)blockIdx.x * WARPS * spw;     // first sample of this CTA
      if ((long long)row0 < a.B) {
        const size_t rows = min((size_t)(a.B - (long long)row0), (size_t)WARPS * spw);
        prefetch(a.x + row0 * a.n_inputs, rows * a.n_inputs * 4);
        if (BWD) prefetch(a.gy + row0 * a.n, rows * a.n * 4);
      }
    }
    if (!inl) {
      for (int i = tid; i < a.n_ops; i += blockDim.x) {
        const MicroOp m = a.ops[i];
        SOp o;
        o.h = m.pad0 | ((m.b & kCu3InlinePost) ? kPostBit : 0); o.idx = m.idx; o.acc = m.acc; o.aux = m.aux;
        o.cf = *reinterpret_cast<const float4*>(a.coef + (size_t)m.pad1 * QNN_COEF_STRIDE);
        S.ops[i] = o;
      }
    }
    src = reinterpret_cast<const uint32_t*>(a.exchanges); dst = reinterpret_cast<uint32_t*>(S.exchanges); words = a.n_exchanges * (int)(sizeof(Exchange) / 4);
    for (int i = tid; i < words; i += blockDim.x) dst[i] = src[i];
    if (!MW) for (int i = tid; i < a.n_xbase; i += blockDim.x) S.xbase[i] = a.xbase[i];
    src = reinterpret_cast<const uint32_t*>(a.diags); dst = reinterpret_cast<uint32_t*>(S.diags); words = a.n_diags * (int)(sizeof(DiagDesc) / 4);
    for (int i = tid; i < words; i += blockDim.x) dst[i] = src[i];
    if constexpr (KR < 5) {
      if (inl) {   // (host: only for wbits == 0, where every table lives in shared memory)
        // every term's angle and bit masks once, staged in the (not yet used) exchange-buffer area: the table
        // entries below then read shared memory instead of chasing terms[] and w[] through L2 per entry
        struct STerm { double phi; unsigned m0, m1; };
        STerm* const st = reinterpret_cast<STerm*>(smem_raw + (((size_t)a.table_bytes + 15) & ~(size_t)15));
        const int nst = min(a.n_terms, 128);     // 2 KB: fits the smallest buffer area (16 warps x 2 KB)
        for (int i = tid; i < nst; i += blockDim.x) {
          const DiagTerm tm = a.terms[i];
          const double v = (tm.src == QNN_SRC_CONST) ? (double)tm.value : (double)a.w[tm.src];
          st[i].phi = tm.kind == 0 ? 0.5 * v : (tm.kind == 1 ? 3.14159265358979323846 : v);
          st[i].m0 = 1u << tm.bit0;
          st[i].m1 = tm.kind == 0 ? 0u : (1u << tm.bit1);
        }
        for (int i = tid; i < a.n_coef; i += blockDim.x) make_coef_row(a.coef_ops[i], a.w, S.coef + i * 8);
        __syncthreads();
        const unsigned tmask = (1u << tb) - 1u;
        for (int d = 0; d < a.n_diags; ++d) {
          const DiagDesc D = a.diags[d];
          const int entries = D.is_reg ? NA : NA * G;
          for (int e = tid; e < entries; e += blockDim.x) {
            const int j = D.is_reg ? e : (e / G), t = D.is_reg ? 0 : (e % G);
            const unsigned p = (unsigned)j | (((unsigned)t & tmask) << KR);
            double ang = 0.0;
            if (D.first_term + D.n_terms <= nst) {
              for (int k = D.first_term; k < D.first_term + D.n_terms; ++k) {
                if (st[k].m1 == 0u) ang += (p & st[k].m0) ? st[k].phi : -st[k].phi;
                else if ((p & st[k].m0) && (p & st[k].m1)) ang += st[k].phi;
              }
            } else {
              ang = diag_angle(D, a.terms, a.w, p);
            }
            double sn, cs;
            sincos(ang, &sn, &cs);
            const float2 v = make_float2((float)cs, (float)sn);
            if (D.is_reg) S.small[(size_t)D.table * NA + e] = v; else sbig[(size_t)D.table * NA * G + e] = v;
          }
        }
        __syncthreads();
        for (int i = tid; i < a.n_ops; i += blockDim.x) {
          const MicroOp m = a.ops[i];
          SOp o;
          o.h = m.pad0 | ((m.b & kCu3InlinePost) ? kPostBit : 0); o.idx = m.idx; o.acc = m.acc; o.aux = m.aux;
          o.cf = *reinterpret_cast<const float4*>(S.coef + (size_t)m.pad1 * 8);
          S.ops[i] = o;
        }
      }
    }
    if (!inl) {
      for (int i = tid; i < a.n_coef * 8; i += blockDim.x) S.coef[i] = a.coef[(size_t)(i >> 3) * QNN_COEF_STRIDE + (i & 7)];
      if (wbits == 0) for (int i = tid; i < a.n_big * NA * G; i += blockDim.x) sbig[i] = a.big[i];
      for (int i = tid; i < a.n_small * NA; i += blockDim.x) S.small[i] = a.small[i];
    }
    if (BWD) {
      for (int i = tid; i < WARPS * max(a.n_acc, 1); i += blockDim.x) S.wacc[i] = 0.f;
      for (int i = tid; i < WARPS * spw * max(a.n_angle, 1); i += blockDim.x) S.gxacc[i] = 0.f;
    }
  }
  __syncthreads();
  float* const wacc = BWD ? S.wacc + (size_t)warp * max(a.n_acc, 1) : nullptr;
  float* const gxacc = BWD ? S.gxacc + (size_t)warp * spw * max(a.n_angle, 1) : nullptr;
  float* const red = S.red + (size_t)warp * 16;
  // barrier over the threads that share one exchange buffer
  auto gsync = [&]() {
    if constexpr (!MW) __syncwarp();
    else if constexpr (!CL) asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G) : "memory");
    else {   // every thread of every CTA of the cluster; release/acquire orders the global exchange traffic
      asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
  };
  auto csync = [&]() { asm volatile("bar.sync 0;" ::: "memory"); };   // this CTA only (local exchanges of a cluster)
  // cluster geometry: four exchange buffers per cluster (two alternating sets of {psi, lam})
  unsigned char* const cbuf = CL ? a.cbuf + ((size_t)cid << (a.n + 5)) : nullptr;
  const size_t cbuf_one = (size_t)1 << (a.n + 3);
  unsigned xpar = 0, xl2 = 0;   // buffer parities: shared-memory pair (forward program), L2 pair
  float* const cred = CL ? a.cred + (size_t)cid * gwarps * 48 : nullptr;

  float2* const xcoef = S.xcoef + (size_t)warp * spw * max(a.n_angle, 1);
  const int n_ang = a.n_angle;
  // amplitude encoder: logical index (wire 0 = most significant bit) of this thread's register j is klog_t | klog_j(j)
  auto klog = [&](unsigned p) { unsigned k = 0; for (int w = 0; w < a.n; ++w) k |= ((p >> a.init_bit_of_wire[w]) & 1u) << (a.n - 1 - w); return k; };
  const unsigned klog_t = (AMP) ? klog((unsigned)tis << KR) : 0u;
  float2 ps[NA];
  const long long n_batches = (a.B + spw - 1) / spw;
  for (long long wb = CL ? (long long)cid : (long long)blockIdx.x * groups + grp; wb < n_batches;
       wb += CL ? (long long)ncl : (long long)gridDim.x * groups) {
    const long long bs = wb * spw + siw;
    const bool active = bs < a.B;
    gsync();   // nothing of the previous batch (scratch, exchange buffer) is still being read

    // shear coefficients of the per-sample rotations RY(x[b, col]): each (sample, column) once per batch
    for (int i = lane; i < spw * n_ang; i += 32) {
      const long long bb = wb * spw + i / n_ang;
      xcoef[i] = rot_coef_x(bb < a.B ? a.x[(size_t)bb * a.n_inputs + (i % n_ang)] : 0.f);
    }
    __syncwarp();
    const float2* const xc = xcoef + siw * n_ang;

    // Saved final state: [batch][j / 2][lane] float4 = amplitudes (2 * (j/2), 2 * (j/2) + 1) of that lane;
    // a warp reads or writes 512 contiguous bytes per instruction.
    float4* const st_rows = STATE ? a.state + (size_t)wb * (NA / 2) * G + tig : nullptr;
    if constexpr (BWD && STATE) {
#pragma unroll
      for (int q = 0; q < NA / 2; ++q) {
        const float4 v = __ldcs(st_rows + (size_t)q * G);
        ps[2 * q] = make_float2(v.x, v.y); ps[2 * q + 1] = make_float2(v.z, v.w);
      }
    } else {
#pragma unroll
    for (int j = 0; j < NA; ++j) ps[j] = make_float2(0.f, 0.f);
    if (AMP) {
      // |psi> = x / |x| (first 2^n entries of the row, zero padded), gathered into the initial layout
      const float* xr = a.x + (size_t)(active ? bs : 0) * a.n_inputs;
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < NA; ++j) {
        const unsigned k = klog_t | klog((unsigned)j);
        const float v = (active && (int)k < a.n_inputs) ? xr[k] : 0.f;
        ps[j].x = v;
        ss = fmaf(v, v, ss);
      }
      // padding samples of a ragged last batch (and all-zero rows) carry the zero state: rsqrt(0) would put NaNs
      // into the warp-wide weight-gradient sums
      const float ssum = sample_sum(ss, lbits);
      const float inv = (active && ssum > 0.f) ? rsqrtf(ssum) : 0.f;
#pragma unroll
      for (int j = 0; j < NA; ++j) ps[j].x *= inv;
    } else {
      ps[0].x = tis == 0 ? 1.f : 0.f;
    }

    // ================= forward program =================
    // One jump per micro-op (handler id = kind and register bits); the next op and its coefficient row
    // are fetched while the current body runs, so the dispatch latency overlaps with arithmetic.
    {
      SOp op = S.ops[0];
      for (int pc = 0; (op.h & 255) != H_END; ++pc) {
        const SOp next = S.ops[pc + 1];
        const float4 cfa = op.cf;
        // Handlers with a lane-dependent branch (parity-controlled CU3) may leave the warp split; without this
        // the two halves run the following handlers one after the other until the next exchange (measured: the
        // merged diagonal after a controlled U3 executed twice per warp).
        __syncwarp();
        switch (op.h & 255) {
#define H_CASE_ROT(RB) case H_ROT + RB: if constexpr (RB < KR) rot<RB>(ps, cfa.x, cfa.y); break;
          H_CASE_ROT(0) H_CASE_ROT(1) H_CASE_ROT(2) H_CASE_ROT(3) H_CASE_ROT(4)
#undef H_CASE_ROT
#define H_CASE_ROTX(RB) case H_ROTX + RB: if constexpr (RB < KR) { const float2 ab = xc[op.idx]; rot<RB>(ps, ab.x, ab.y); } break;
          H_CASE_ROTX(0) H_CASE_ROTX(1) H_CASE_ROTX(2) H_CASE_ROTX(3) H_CASE_ROTX(4)
#undef H_CASE_ROTX
#define H_CASE_CU3L(RT, RC) case H_CU3L + RT * 4 + (RC < RT ? RC : RC - 1): if constexpr (RT < KR && RC < KR) { const Cu3 g = make_cu3(cfa, *reinterpret_cast<const float4*>(S.coef + op.idx * 8 + 4), false); cu3_loc<RT, RC>(ps, g, op.h & kPostBit); } break;
          H_CASE_CU3L(0, 1) H_CASE_CU3L(0, 2) H_CASE_CU3L(0, 3) H_CASE_CU3L(0, 4)
          H_CASE_CU3L(1, 0) H_CASE_CU3L(1, 2) H_CASE_CU3L(1, 3) H_CASE_CU3L(1, 4)
          H_CASE_CU3L(2, 0) H_CASE_CU3L(2, 1) H_CASE_CU3L(2, 3) H_CASE_CU3L(2, 4)
          H_CASE_CU3L(3, 0) H_CASE_CU3L(3, 1) H_CASE_CU3L(3, 2) H_CASE_CU3L(3, 4)
          H_CASE_CU3L(4, 0) H_CASE_CU3L(4, 1) H_CASE_CU3L(4, 2) H_CASE_CU3L(4, 3)
#undef H_CASE_CU3L
#define H_CASE_CU3P(RB) case H_CU3P + RB: if constexpr (RB < KR) { if (__popc(tis & op.aux) & 1) { const Cu3 g = make_cu3(cfa, *reinterpret_cast<const float4*>(S.coef + op.idx * 8 + 4), false); cu3_all<RB>(ps, g, op.h & kPostBit); } } break;
          H_CASE_CU3P(0) H_CASE_CU3P(1) H_CASE_CU3P(2) H_CASE_CU3P(3) H_CASE_CU3P(4)
#undef H_CASE_CU3P
#define ROT_BIT(K) if constexpr (K < KR) rot<K>(ps, ab[K].x, ab[K].y);
          case H_ROT5: {   // rotations on all KR register bits
            float2 ab[5];
#pragma unroll
            for (int k = 0; k < KR; ++k) ab[k] = *reinterpret_cast<const float2*>(S.coef + (op.idx + ((op.aux >> (3 * k)) & 7)) * 8);
            ROT_BIT(0) ROT_BIT(1) ROT_BIT(2) ROT_BIT(3) ROT_BIT(4)
          } break;
          case H_ROTX5: {
            float2 ab[5];
#pragma unroll
            for (int k = 0; k < KR; ++k) ab[k] = xc[(op.idx >> (5 * k)) & 31];
            ROT_BIT(0) ROT_BIT(1) ROT_BIT(2) ROT_BIT(3) ROT_BIT(4)
          } break;
#undef ROT_BIT
          case H_CHAIN: {   // controlled U3 (target, control) = (1,0), (2,1), ... up to the last register bit
            { const Cu3 g = make_cu3(cfa, *reinterpret_cast<const float4*>(S.coef + op.idx * 8 + 4), false); cu3_loc<1, 0>(ps, g, op.aux & 1); }
            { const Cu3 g = load_cu3(S.coef + (op.idx + 1) * 8, false); cu3_loc<2, 1>(ps, g, op.aux & 2); }
            if constexpr (KR > 3) { const Cu3 g = load_cu3(S.coef + (op.idx + 2) * 8, false); cu3_loc<3, 2>(ps, g, op.aux & 4); }
            if constexpr (KR > 4) { const Cu3 g = load_cu3(S.coef + (op.idx + 3) * 8, false); cu3_loc<4, 3>(ps, g, op.aux & 8); }
          } break;
          case H_DIAG: diag_mul<false, false>(ps, ps, S.big + (size_t)op.idx * NA * G, G, tig); break;
          case H_DIAGR: diag_mul<false, false>(ps, ps, S.small + op.idx * NA, 1, 0); break;
          case H_XCHG: {
            const Exchange* X = S.exchanges + op.idx;
            if constexpr (CL) {
              if (op.aux == 1) {   // local: the CTA bits stay put, the exchange runs through this CTA's shared memory
                VM_CHECK_OFFSETS(1, S.xbase[X->base + tig], X->woff, 65536u) VM_CHECK_OFFSETS(2, S.xbase[X->base + G + tig], X->roff, 65536u)
                exchange(ps, S.xbuf + (xpar << 16), S.xbase[X->base + tig], S.xbase[X->base + G + tig], X->woff, X->roff, csync);
              } else if (op.aux == 2) {
                // Across CTAs through distributed shared memory.  The two buffers alternate from one such exchange
                // to the next and the local exchanges in between use the one received into last, so ONE cluster
                // barrier per exchange suffices: a buffer is written remotely only by CTAs that have passed the
                // barrier of the exchange in between, which this CTA reaches after its last use of that buffer.
                // The forward kernel keeps ONE buffer (so that two of its CTAs, of different clusters, fit on an SM and
                // hide each other's barrier and exchange latency): a second cluster barrier in front of the stores
                // then makes sure every CTA of the cluster is done with its previous use of the buffer.
                if constexpr (!BWD) gsync(); else xpar ^= 1u;
                const uint32_t buf = S.xbuf + (xpar << 16);
                VM_CHECK_OFFSETS(3, S.xbase[X->base + tig], X->woff, (uint32_t)a.cluster << 16)       // destination rank < cluster size
                VM_CHECK_OFFSETS(4, S.xbase[X->base + G + tig] ^ (crank << 16), X->roff, 65536u)    // loads stay in this CTA
                dsmem_store_all(ps, buf, S.xbase[X->base + tig], X->woff);
                gsync();
                smem_load_all(ps, buf, S.xbase[X->base + G + tig] & 0xFFFFu, X->roff);
              } else {
                xl2 ^= 1u;
                unsigned char* b0 = cbuf + (size_t)(2 * xl2) * cbuf_one;
                VM_CHECK_OFFSETS(5, S.xbase[X->base + tig], X->woff, cbuf_one) VM_CHECK_OFFSETS(6, S.xbase[X->base + G + tig], X->roff, cbuf_one)
                exchange_global<false>(ps, ps, b0, b0, S.xbase[X->base + tig], S.xbase[X->base + G + tig], X->woff, X->roff, gsync);
              }
            } else {
              VM_CHECK_OFFSETS(7, S.xbase[X->base + tig], X->woff, (256u * NA) << wbits) VM_CHECK_OFFSETS(8, S.xbase[X->base + G + tig], X->roff, (256u * NA) << wbits)
              exchange(ps, 0u, S.xbuf | S.xbase[X->base + tig], S.xbuf | S.xbase[X->base + G + tig], X->woff, X->roff, gsync);
            }
          } break;
          default: break;
        }
        op = next;
      }
    }
    if constexpr (CL && BWD && !STATE) { gsync(); xpar = 0; }   // the reverse sweep starts with both shared buffers free
    }  // forward program (skipped when the backward pass starts from a saved state)

    if constexpr (!BWD && STATE) {
#pragma unroll
      for (int q = 0; q < NA / 2; ++q)
        __stcs(st_rows + (size_t)q * G, make_float4(ps[2 * q].x, ps[2 * q].y, ps[2 * q + 1].x, ps[2 * q + 1].y));
    }

    if constexpr (!BWD) {
      // ================= measurement: <Z> per physical bit =================
      float2 tot2, S2[kR];
      if constexpr (KR == 5) signed_sums2([&](int j) { return __fmul2_rn(ps[j], ps[j]); }, tot2, S2);
      else signed_sums2_small<KR>([&](int j) { return __fmul2_rn(ps[j], ps[j]); }, tot2, S2);
      const float tot = tot2.x + tot2.y;
      // per-warp partials, indexed by physical bit: [0, KR) register bits, [KR, KR + 5) lane bits, [10] plain total
      // (for the warp bits)
      constexpr int PW = KR + 5;   // first warp bit
      float val[11];
#pragma unroll
      for (int r = 0; r < KR; ++r) val[r] = sample_sum(S2[r].x + S2[r].y, lbits);
#pragma unroll
      for (int b = 0; b < 5; ++b) val[KR + b] = (b < lbits) ? sample_sum(((lane >> b) & 1) ? -tot : tot, lbits) : 0.f;
#pragma unroll
      for (int c = PW; c < 10; ++c) val[c] = 0.f;
      val[10] = sample_sum(tot, lbits);
      if (wbits == 0) {
        if (active && tis == 0) {
          float* yrow = a.y + (size_t)bs * a.n;
#pragma unroll
          for (int pb = 0; pb < PW; ++pb)
            if (pb < a.n) yrow[a.wire_of_bit[pb]] = val[pb];
        }
      } else {
        // a sample spans 2^wb warps: combine their partials through shared (or, for a cluster, global)
        // memory in a fixed order
        float* const mine = CL ? cred + (size_t)gw * 48 : red;
        if (lane == 0) {
#pragma unroll
          for (int c = 0; c < 11; ++c) { if (CL) __stcg(mine + c, val[c]); else mine[c] = val[c]; }
        }
        gsync();
        if (gw == 0 && lane < a.n && active) {
          const float* r0 = CL ? cred : S.red + (size_t)(grp << wbits) * 16;
          const int stride = CL ? 48 : 16;
          float t = 0.f;
          for (int w = 0; w < gwarps; ++w) {
            const float v = CL ? __ldcg(r0 + w * stride + (lane < PW ? lane : 10)) : r0[w * stride + (lane < PW ? lane : 10)];
            t += (lane >= PW && ((w >> (lane - PW)) & 1)) ? -v : v;
          }
          a.y[(size_t)bs * a.n + a.wire_of_bit[lane]] = t;
        }
      }
    } else {
      // ================= seed lam = (sum_q gy_q Z_q) psi =================
      float2 ls[NA];
      {
        float g[KR + kMaxThreadBits];
        const float* gyrow = a.gy + (size_t)(active ? bs : 0) * a.n;
#pragma unroll
        for (int pb = 0; pb < KR + kMaxThreadBits; ++pb) g[pb] = (active && pb < a.n) ? gyrow[a.wire_of_bit[pb]] : 0.f;
        float cl = 0.f;
#pragma unroll
        for (int b = 0; b < kMaxThreadBits; ++b) cl += ((tis >> b) & 1) ? -g[KR + b] : g[KR + b];
        float cj[NA];
        cj[0] = cl;
#pragma unroll
        for (int r = 0; r < KR; ++r) cj[0] += g[r];
#pragma unroll
        for (int j = 1; j < NA; ++j) {
          const int low = j & (-j);                      // lowest set bit
          const int r = low == 1 ? 0 : low == 2 ? 1 : low == 4 ? 2 : low == 8 ? 3 : 4;
          cj[j] = fmaf(-2.f, g[r], cj[j ^ low]);
        }
#pragma unroll
        for (int j = 0; j < NA; ++j) ls[j] = mul2(cj[j], ps[j]);
      }
      // ================= reverse program =================
      for (int pc = a.n_ops - 2; pc >= 0; --pc) {
        // no prefetch of the next op here: at 255 registers it would travel through local memory anyway
        // (measured: cfg3 backward 3.37 -> 3.22 ms without it)
        const SOp op = S.ops[pc];
        const float4 cfa = op.cf;
        // Handlers with a lane-dependent branch (parity-controlled CU3) may leave the warp split; without this
        // the two halves run the following handlers one after the other until the next exchange (measured: the
        // merged diagonal after a controlled U3 executed twice per warp).
        __syncwarp();
        switch (op.h & 255) {
#define H_CASE_ROT(RB) case H_ROT + RB: if constexpr (RB < KR) { float v = rot_grad<RB>(ps, ls); rot<RB>(ps, -cfa.x, -cfa.y); rot<RB>(ls, -cfa.x, -cfa.y); \
            if (op.acc >= 0) { v = warp_sum(v); if (lane == 0) wacc[op.acc] += v; } } break;
          H_CASE_ROT(0) H_CASE_ROT(1) H_CASE_ROT(2) H_CASE_ROT(3) H_CASE_ROT(4)
#undef H_CASE_ROT
#define H_CASE_ROTX(RB) case H_ROTX + RB: if constexpr (RB < KR) { const float2 ab = xc[op.idx]; float v = rot_grad<RB>(ps, ls); rot<RB>(ps, -ab.x, -ab.y); rot<RB>(ls, -ab.x, -ab.y); \
            v = sample_sum(v, lbits); if ((lane & ((1 << lbits) - 1)) == 0) gxacc[siw * n_ang + op.idx] += v; } break;
          H_CASE_ROTX(0) H_CASE_ROTX(1) H_CASE_ROTX(2) H_CASE_ROTX(3) H_CASE_ROTX(4)
#undef H_CASE_ROTX
#define CU3_REDUCE(A, ACC)                                                                                  \
            if ((ACC) >= 0) {                                                                               \
              const float r = warp_sum4((A.v0.x + A.v0.y) - (A.v1.x + A.v1.y), -2.f * (A.qp.x - A.qp.y),    \
                                        -2.f * (A.ql.x - A.ql.y), 0.f, lane);                               \
              if ((lane & 7) == 0 && (lane >> 3) < 3) wacc[(ACC) + (lane >> 3)] += r;                       \
            }
#define H_CASE_CU3L(RT, RC) case H_CU3L + RT * 4 + (RC < RT ? RC : RC - 1): if constexpr (RT < KR && RC < KR) { const float2 z = make_float2(0.f, 0.f); Cu3Acc A{z, z, z, z}; \
            const Cu3 g = make_cu3(cfa, *reinterpret_cast<const float4*>(S.coef + op.idx * 8 + 4), true); cu3_loc_bwd<RT, RC>(ps, ls, g, A, op.h & kPostBit); CU3_REDUCE(A, op.acc) } break;
          H_CASE_CU3L(0, 1) H_CASE_CU3L(0, 2) H_CASE_CU3L(0, 3) H_CASE_CU3L(0, 4)
          H_CASE_CU3L(1, 0) H_CASE_CU3L(1, 2) H_CASE_CU3L(1, 3) H_CASE_CU3L(1, 4)
          H_CASE_CU3L(2, 0) H_CASE_CU3L(2, 1) H_CASE_CU3L(2, 3) H_CASE_CU3L(2, 4)
          H_CASE_CU3L(3, 0) H_CASE_CU3L(3, 1) H_CASE_CU3L(3, 2) H_CASE_CU3L(3, 4)
          H_CASE_CU3L(4, 0) H_CASE_CU3L(4, 1) H_CASE_CU3L(4, 2) H_CASE_CU3L(4, 3)
#undef H_CASE_CU3L
#define H_CASE_CU3P(RB) case H_CU3P + RB: if constexpr (RB < KR) { const float2 z = make_float2(0.f, 0.f); Cu3Acc A{z, z, z, z}; \
            if (__popc(tis & op.aux) & 1) { const Cu3 g = make_cu3(cfa, *reinterpret_cast<const float4*>(S.coef + op.idx * 8 + 4), true); cu3_all_bwd<RB>(ps, ls, g, A, op.h & kPostBit); } \
            CU3_REDUCE(A, op.acc) } break;
          H_CASE_CU3P(0) H_CASE_CU3P(1) H_CASE_CU3P(2) H_CASE_CU3P(3) H_CASE_CU3P(4)
#undef H_CASE_CU3P
          case H_ROT5: {
            float v[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#define ROT5_STEP(K)                                                                                       \
            if constexpr (K < KR) {                                                                        \
              const int pos = (op.aux >> (3 * K)) & 7;                                                     \
              const float2 ab = *reinterpret_cast<const float2*>(S.coef + (op.idx + pos) * 8);             \
              v[K] = rot_grad<K>(ps, ls);                                                                  \
              rot<K>(ps, -ab.x, -ab.y); rot<K>(ls, -ab.x, -ab.y);                                          \
            }
            ROT5_STEP(4) ROT5_STEP(3) ROT5_STEP(2) ROT5_STEP(1) ROT5_STEP(0)
#undef ROT5_STEP
            // five sums in one halving butterfly: lane 8k' ... holds component (lane >> 2)
            const float t8[16] = {v[0], v[1], v[2], v[3], v[4], 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            const float r = warp_sum16(t8, lane);
            if ((lane & 1) == 0 && (lane >> 1) < KR) wacc[op.acc + ((op.aux >> (3 * (lane >> 1))) & 7)] += r;
          } break;
          case H_ROTX5: {
#define ROTX5_STEP(K)                                                                                      \
            if constexpr (K < KR) {                                                                        \
              const int col = (op.idx >> (5 * K)) & 31;                                                    \
              const float2 ab = xc[col];                                                                   \
              float v = rot_grad<K>(ps, ls);                                                               \
              rot<K>(ps, -ab.x, -ab.y); rot<K>(ls, -ab.x, -ab.y);                                          \
              v = sample_sum(v, lbits);                                                                    \
              if ((lane & ((1 << lbits) - 1)) == 0) gxacc[siw * n_ang + col] += v;                          \
            }
            ROTX5_STEP(4) ROTX5_STEP(3) ROTX5_STEP(2) ROTX5_STEP(1) ROTX5_STEP(0)
#undef ROTX5_STEP
          } break;
          case H_CHAIN: {
#define CHAIN_STEP(K, RT, RC)                                                                              \
            if constexpr (RT < KR) {                                                                       \
              const float2 z = make_float2(0.f, 0.f);                                                      \
              Cu3Acc A{z, z, z, z};                                                                        \
              const Cu3 g = load_cu3(S.coef + (op.idx + K) * 8, true);                                     \
              cu3_loc_bwd<RT, RC>(ps, ls, g, A, (op.aux >> K) & 1);                                        \
              CU3_REDUCE(A, op.acc + 3 * K)                                                                \
            }
            CHAIN_STEP(3, 4, 3)
            CHAIN_STEP(2, 3, 2)
            CHAIN_STEP(1, 2, 1)
            CHAIN_STEP(0, 1, 0)
#undef CHAIN_STEP
          } break;
#undef CU3_REDUCE
          case H_DIAG:
          case H_DIAGR: {
            const DiagDesc* D = S.diags + op.aux;
            float2 tot2, S2[kR];
            // Im(conj(lam) psi) = lam.re psi.im - lam.im psi.re, kept as the pair (lam.re psi.im, lam.im psi.re)
            if constexpr (KR == 5) signed_sums2([&](int j) { return __fmul2_rn(ls[j], swp(ps[j])); }, tot2, S2);
            else signed_sums2_small<KR>([&](int j) { return __fmul2_rn(ls[j], swp(ps[j])); }, tot2, S2);
            const float tot = tot2.x - tot2.y;
            // all (up to ten) per-bit sums in one halving butterfly; lane l ends up with component l >> 1
            float t16[16];
#pragma unroll
            for (int r = 0; r < KR; ++r) t16[r] = S2[r].x - S2[r].y;
#pragma unroll
            for (int b = 0; b < kMaxThreadBits; ++b) t16[KR + b] = ((tis >> b) & 1) ? -tot : tot;
#pragma unroll
            for (int k = KR + kMaxThreadBits; k < 16; ++k) t16[k] = 0.f;
            const float r = warp_sum16(t16, lane);
            if ((lane & 1) == 0) {
              const int comp = lane >> 1;
              const int ai = comp < KR ? D->acc_reg[comp]
                                       : (comp < KR + kMaxThreadBits && (op.h & 255) == H_DIAG ? D->acc_thr[comp - KR] : -1);
              if (ai >= 0) wacc[ai] += r;
            }
            __syncwarp();   // reconverge before the table multiply (ptxas otherwise runs it once per lane subset)
            if ((op.h & 255) == H_DIAG) diag_mul<true, true>(ps, ls, S.big + (size_t)op.idx * NA * G, G, tig);
            else diag_mul<true, true>(ps, ls, S.small + op.idx * NA, 1, 0);
          } break;
          case H_XCHG: {   // inverse: roles of the two address sets swapped
            const Exchange* X = S.exchanges + op.idx;
            if constexpr (CL) {
              if (op.aux == 2) {
                // The inverse permutation has its own maps (entry pad[1]) so that here, too, stores are the remote
                // side and loads stay in this CTA.  psi travels through distributed shared memory and lam, at the
                // same time, through an L2 buffer of the workspace (same slot map: coalesced 256-byte stores,
                // sector-sized loads): the cluster fabric and the L2 path each carry half of the traffic, behind ONE
                // cluster barrier.  Both buffer pairs alternate from one such exchange to the next, and the local
                // exchanges in between use the shared buffer received into last (as in the forward program).
                const Exchange* Xb = S.exchanges + X->pad[1];
                const uint32_t wa = S.xbase[Xb->base + tig], ra = S.xbase[Xb->base + G + tig];
                xpar ^= 1u; xl2 ^= 1u;
                const uint32_t buf = S.xbuf + (xpar << 16);
                unsigned char* const lbuf = cbuf + (size_t)(2 * xl2 + 1) * cbuf_one;
                VM_CHECK_OFFSETS(9, wa, Xb->woff, (uint32_t)a.cluster << 16) VM_CHECK_OFFSETS(10, ra ^ (crank << 16), Xb->roff, 65536u)
                VM_CHECK_OFFSETS(11, wa, Xb->woff, cbuf_one) VM_CHECK_OFFSETS(12, ra, Xb->roff, cbuf_one)
                dsmem_store_all(ps, buf, wa, Xb->woff);
#pragma unroll
                for (int q = 0; q < NA / 4; ++q) {
                  const uint4 o = *reinterpret_cast<const uint4*>(Xb->woff + 4 * q);
                  stg64(lbuf, wa ^ o.x, ls[4 * q + 0].x, ls[4 * q + 0].y);
                  stg64(lbuf, wa ^ o.y, ls[4 * q + 1].x, ls[4 * q + 1].y);
                  stg64(lbuf, wa ^ o.z, ls[4 * q + 2].x, ls[4 * q + 2].y);
                  stg64(lbuf, wa ^ o.w, ls[4 * q + 3].x, ls[4 * q + 3].y);
                }
                gsync();
#pragma unroll
                for (int q = 0; q < NA / 4; ++q) {     // the L2 loads first: their latency overlaps the shared-memory loads
                  const uint4 o = *reinterpret_cast<const uint4*>(Xb->roff + 4 * q);
                  ldg64(lbuf, ra ^ o.x, ls[4 * q + 0].x, ls[4 * q + 0].y);
                  ldg64(lbuf, ra ^ o.y, ls[4 * q + 1].x, ls[4 * q + 1].y);
                  ldg64(lbuf, ra ^ o.z, ls[4 * q + 2].x, ls[4 * q + 2].y);
                  ldg64(lbuf, ra ^ o.w, ls[4 * q + 3].x, ls[4 * q + 3].y);
                }
                smem_load_all(ps, buf, ra & 0xFFFFu, Xb->roff);
              } else {
                const uint32_t wa = S.xbase[X->base + G + tig], ra = S.xbase[X->base + tig];
                if (op.aux == 1) {
                  VM_CHECK_OFFSETS(13, wa, X->roff, 65536u) VM_CHECK_OFFSETS(14, ra, X->woff, 65536u)
                  exchange(ps, S.xbuf + (xpar << 16), wa, ra, X->roff, X->woff, csync);
                  exchange(ls, S.xbuf + (xpar << 16), wa, ra, X->roff, X->woff, csync);
                } else {
                  xl2 ^= 1u;
                  unsigned char* b0 = cbuf + (size_t)(2 * xl2) * cbuf_one;
                  VM_CHECK_OFFSETS(15, wa, X->roff, cbuf_one) VM_CHECK_OFFSETS(16, ra, X->woff, cbuf_one)
                  exchange_global<true>(ps, ls, b0, b0 + cbuf_one, wa, ra, X->roff, X->woff, gsync);
                }
              }
            } else {
              const uint32_t wa = S.xbuf | S.xbase[X->base + G + tig], ra = S.xbuf | S.xbase[X->base + tig];
              VM_CHECK_OFFSETS(17, wa ^ S.xbuf, X->roff, (256u * NA) << wbits) VM_CHECK_OFFSETS(18, ra ^ S.xbuf, X->woff, (256u * NA) << wbits)
              exchange(ps, 0u, wa, ra, X->roff, X->woff, gsync);
              exchange(ls, 0u, wa, ra, X->roff, X->woff, gsync);
            }
          } break;
          default: break;
        }
      }
      // ================= input gradients of this group's samples =================
      if (AMP) {
        // psi0 = x / |x|:  dL/dx = (g - psi0 (psi0 . g)) / |x|  with g = 2 Re(lam0), in the initial layout
        if (a.gx != nullptr) {
          const float* xr = a.x + (size_t)(active ? bs : 0) * a.n_inputs;
          float* gxr = a.gx + (size_t)(active ? bs : 0) * a.n_inputs;
          float xv[NA], ss = 0.f, dot = 0.f;
#pragma unroll
          for (int j = 0; j < NA; ++j) {
            const unsigned k = klog_t | klog((unsigned)j);
            xv[j] = (active && (int)k < a.n_inputs) ? xr[k] : 0.f;
            ss = fmaf(xv[j], xv[j], ss);
            dot = fmaf(xv[j], 2.f * ls[j].x, dot);
          }
          const float ssum = sample_sum(ss, lbits);
          const float inv = (active && ssum > 0.f) ? rsqrtf(ssum) : 0.f;
          dot = sample_sum(dot, lbits) * inv;
#pragma unroll
          for (int j = 0; j < NA; ++j) {
            const unsigned k = klog_t | klog((unsigned)j);
            if (active && (int)k < a.n_inputs) gxr[k] = (2.f * ls[j].x - xv[j] * inv * dot) * inv;
          }
          if (active && a.n_inputs > (1 << a.n))          // columns beyond 2^n never enter the state
            for (int k = (1 << a.n) + tis; k < a.n_inputs; k += (1 << tb)) gxr[k] = 0.f;
        }
        continue;
      }
      if constexpr (CL) {   // publish this warp's partial sums to the cluster scratch
        for (int i = lane; i < n_ang; i += 32) __stcg(cred + (size_t)gw * 48 + 16 + i, gxacc[i]);
      }
      gsync();
      if (a.gx != nullptr && gw == 0) {
        for (int i = lane; i < spw * n_ang; i += 32) {
          const long long bb = wb * spw + i / n_ang;
          float t = 0.f;
          for (int w = 0; w < gwarps; ++w)                                   // the group's warps, fixed order
            t += CL ? __ldcg(cred + (size_t)w * 48 + 16 + i) : gxacc[(size_t)w * spw * max(n_ang, 1) + i];
          if (bb < a.B) a.gx[(size_t)bb * a.n_inputs + (i % n_ang)] = t;
        }
      }
      gsync();
      for (int i = lane; i < spw * n_ang; i += 32) gxacc[i] = 0.f;
    }
  }

  if constexpr (BWD) {
    __syncthreads();
    for (int i = threadIdx.x; i < a.n_acc; i += blockDim.x) {
      float s = 0.f;
#pragma unroll 1
      for (int w = 0; w < WARPS; ++w) s += S.wacc[(size_t)w * a.n_acc + i];
      a.partial[(size_t)blockIdx.x * a.n_acc + i] = s;
    }
  }
}


// Plain launch, or a cluster launch (cudaLaunchKernelEx) when a sample spans several CTAs.
struct LaunchCfg { int grid, threads, cluster; size_t smem; cudaStream_t stream; };

template <typename K>
int launch_kernel(K kern, const LaunchCfg& lc, const VmArgs& a) {
  if (int rc = qnn_allow_smem(reinterpret_cast<const void*>(kern), lc.smem)) return rc;
  if (lc.cluster == 0) {
    kern<<<lc.grid, lc.threads, lc.smem, lc.stream>>>(a);
  } else {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(lc.grid); cfg.blockDim = dim3(lc.threads); cfg.dynamicSmemBytes = lc.smem; cfg.stream = lc.stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = lc.cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr; cfg.numAttrs = 1;
    QNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, a));
  }
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

template <int GEOM, bool AMP, int KR = 5>
int launch_geom(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a) {
  if (bwd) return state ? launch_kernel(k_vm<true, true, GEOM, AMP, KR>, lc, a) : launch_kernel(k_vm<true, false, GEOM, AMP, KR>, lc, a);
  return state ? launch_kernel(k_vm<false, true, GEOM, AMP, KR>, lc, a) : launch_kernel(k_vm<false, false, GEOM, AMP, KR>, lc, a);
}

// one translation unit per geometry (compiled in parallel): qnn_vm_g0.cu, qnn_vm_g0a.cu, qnn_vm_g1.cu, qnn_vm_g2.cu
int launch_g0(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);    // n <= 10, angle inputs
int launch_g0_k3(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);  // n <= 8, 8 amplitudes per thread
int launch_g0_k4(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);  // n <= 9, 16 amplitudes per thread
int launch_g1_k4(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);  // n = 10..12, 16 amplitudes per thread
int launch_g0a(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);   // n <= 10, amplitude encoder
int launch_g1(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);    // n = 11..13
int launch_g2(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a);    // n = 14..16 (clusters)
int max_active_clusters_g2(int cluster, size_t smem, int device, int num_sm, bool bwd);

}  // namespace vm
