// Gate matrices and their generators, shared by every kernel family.
//
// For a gate U(a_0..a_{k-1}) the adjoint backward needs, per angle, G_k = (dU/da_k) U^dagger: with
// T_ab = sum over amplitude pairs of conj(lam_a) psi_b (psi, lam AFTER the gate) the gradient is
// 2 Re sum_ab G_k[a][b] T[a][b].  Conventions follow include/qnn_b200.h (PennyLane's).
#pragma once
#include <cuda_runtime.h>

#include "qnn_internal.h"

template <typename T>
struct Cx {
  T re, im;
};
template <typename T>
__host__ __device__ inline Cx<T> cmul(Cx<T> a, Cx<T> b) {
  return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re};
}
template <typename T>
__host__ __device__ inline Cx<T> cconj(Cx<T> a) {
  return {a.re, -a.im};
}
template <typename T>
__host__ __device__ inline Cx<T> cadd(Cx<T> a, Cx<T> b) {
  return {a.re + b.re, a.im + b.im};
}

__device__ inline void qsincos(float a, float* s, float* c) { sincosf(a, s, c); }
__device__ inline void qsincos(double a, double* s, double* c) { sincos(a, s, c); }

// U[4] row-major (U00,U01,U10,U11); dU[k][4].  Returns the number of angles.
template <typename T>
__device__ inline int make_gate(int kind, const T* a, Cx<T>* U, Cx<T> (*dU)[4]) {
  const T h = T(0.5);
  if (kind == QNN_RX || kind == QNN_RY || kind == QNN_RZ) {
    T s, c;
    qsincos(a[0] * h, &s, &c);
    if (kind == QNN_RY) {
      U[0] = {c, 0}; U[1] = {-s, 0}; U[2] = {s, 0}; U[3] = {c, 0};
      dU[0][0] = {-h * s, 0}; dU[0][1] = {-h * c, 0}; dU[0][2] = {h * c, 0}; dU[0][3] = {-h * s, 0};
    } else if (kind == QNN_RX) {
      U[0] = {c, 0}; U[1] = {0, -s}; U[2] = {0, -s}; U[3] = {c, 0};
      dU[0][0] = {-h * s, 0}; dU[0][1] = {0, -h * c}; dU[0][2] = {0, -h * c}; dU[0][3] = {-h * s, 0};
    } else {
      U[0] = {c, -s}; U[1] = {0, 0}; U[2] = {0, 0}; U[3] = {c, s};
      dU[0][0] = {-h * s, -h * c}; dU[0][1] = {0, 0}; dU[0][2] = {0, 0}; dU[0][3] = {-h * s, h * c};
    }
    return 1;
  }
  // U3 / CU3: (theta, phi, lambda)
  T s, c, sp, cp, sl, cl;
  qsincos(a[0] * h, &s, &c);
  qsincos(a[1], &sp, &cp);
  qsincos(a[2], &sl, &cl);
  Cx<T> ep = {cp, sp}, el = {cl, sl}, epl = cmul(ep, el);
  Cx<T> iep = {-sp, cp}, iel = {-sl, cl}, iepl = {-epl.im, epl.re};
  U[0] = {c, 0};
  U[1] = {-el.re * s, -el.im * s};
  U[2] = {ep.re * s, ep.im * s};
  U[3] = {epl.re * c, epl.im * c};
  // d/dtheta
  dU[0][0] = {-h * s, 0};
  dU[0][1] = {-h * el.re * c, -h * el.im * c};
  dU[0][2] = {h * ep.re * c, h * ep.im * c};
  dU[0][3] = {-h * epl.re * s, -h * epl.im * s};
  // d/dphi
  dU[1][0] = {0, 0}; dU[1][1] = {0, 0};
  dU[1][2] = {iep.re * s, iep.im * s};
  dU[1][3] = {iepl.re * c, iepl.im * c};
  // d/dlambda
  dU[2][0] = {0, 0};
  dU[2][1] = {-iel.re * s, -iel.im * s};
  dU[2][2] = {0, 0};
  dU[2][3] = {iepl.re * c, iepl.im * c};
  return 3;
}

// coef[0..7] = U, coef[8 + 8k ..] = G_k = dU_k * U^dagger, as float.
template <typename T>
__device__ inline void make_coef(int kind, const T* a, float* coef) {
  Cx<T> U[4], dU[3][4];
  int na = make_gate<T>(kind, a, U, dU);
  for (int i = 0; i < 4; ++i) {
    coef[2 * i] = (float)U[i].re;
    coef[2 * i + 1] = (float)U[i].im;
  }
  for (int k = 0; k < na; ++k) {
    for (int r = 0; r < 2; ++r)
      for (int cidx = 0; cidx < 2; ++cidx) {
        // G[r][cidx] = sum_m dU[r][m] * conj(U[cidx][m])
        Cx<T> g = cadd(cmul(dU[k][r * 2 + 0], cconj(U[cidx * 2 + 0])),
                       cmul(dU[k][r * 2 + 1], cconj(U[cidx * 2 + 1])));
        coef[8 + 8 * k + 2 * (r * 2 + cidx)] = (float)g.re;
        coef[8 + 8 * k + 2 * (r * 2 + cidx) + 1] = (float)g.im;
      }
  }
  for (int i = 8 + 8 * na; i < QNN_COEF_STRIDE; ++i) coef[i] = 0.f;
}
