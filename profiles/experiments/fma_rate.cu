// Experiment: FP32 issue rate of register-resident rotation / complex-multiply patterns vs the plain
// FMA-chain probe.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a fma_rate.cu -o fma_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE, int THREADS>
__global__ void __launch_bounds__(THREADS) k(float* out, const float* in, int iters, float pa, float pb) {
  float re[32], im[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) { re[i] = in[(threadIdx.x + i) & 1023]; im[i] = in[(threadIdx.x + 2 * i + 7) & 1023]; }
  float c = in[5], s = in[6];
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {  // probe: constant operands
#pragma unroll
      for (int i = 0; i < 32; ++i) { re[i] = fmaf(re[i], pa, pb); im[i] = fmaf(im[i], pa, pb); re[i] = fmaf(re[i], pa, pb); im[i] = fmaf(im[i], pa, pb); }
    } else if (MODE == 1) {  // register operands, c and s reused
#pragma unroll
      for (int i = 0; i < 32; ++i) { re[i] = fmaf(re[i], c, s); im[i] = fmaf(im[i], c, s); re[i] = fmaf(re[i], c, s); im[i] = fmaf(im[i], c, s); }
    } else if (MODE == 2) {  // real rotation on register bit 0 (as g1_real<0>): 4 instr per amplitude
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int j0 = 2 * i, j1 = j0 + 1;
        float a0 = re[j0], a1 = re[j1], b0 = im[j0], b1 = im[j1];
        re[j0] = fmaf(-s, a1, c * a0); re[j1] = fmaf(s, a0, c * a1);
        im[j0] = fmaf(-s, b1, c * b0); im[j1] = fmaf(s, b0, c * b1);
      }
    } else if (MODE == 3) {  // rotation on register bit 4
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int j0 = i, j1 = i + 16;
        float a0 = re[j0], a1 = re[j1], b0 = im[j0], b1 = im[j1];
        re[j0] = fmaf(-s, a1, c * a0); re[j1] = fmaf(s, a0, c * a1);
        im[j0] = fmaf(-s, b1, c * b0); im[j1] = fmaf(s, b0, c * b1);
      }
    } else if (MODE == 4) {  // complex multiply by a per-amplitude factor (diag): 4 instr per amplitude
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float a = re[i], b = im[i];
        re[i] = fmaf(-s, b, c * a); im[i] = fmaf(s, a, c * b);
        a = re[i]; b = im[i];
        re[i] = fmaf(-s, b, c * a); im[i] = fmaf(s, a, c * b);
      }
    }
  }
  if (MODE >= 5) {
    float2 a[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = make_float2(re[i], im[i]);
    const float2 c2 = make_float2(c, c), s2 = make_float2(s, s), ns2 = make_float2(-s, -s);
    const float2 ur = make_float2(in[9], in[9]), ui = make_float2(-in[10], in[10]);
    const float2 vr = make_float2(in[11], in[11]), vi = make_float2(-in[12], in[12]);
    for (int it = 0; it < iters; ++it) {
      if (MODE == 5) {   // packed chain probe
#pragma unroll
        for (int i = 0; i < 32; ++i) { a[i] = __ffma2_rn(a[i], c2, s2); a[i] = __ffma2_rn(a[i], c2, s2); }
      } else if (MODE == 6) {  // packed real rotation, register bit 0: 2 packed instr per amplitude
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float2 a0 = a[2 * i], a1 = a[2 * i + 1];
          a[2 * i] = __ffma2_rn(ns2, a1, __fmul2_rn(c2, a0));
          a[2 * i + 1] = __ffma2_rn(s2, a0, __fmul2_rn(c2, a1));
        }
      } else if (MODE == 8) {  // in-place 3-shear rotation on bit 0: 3 packed per pair
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          a[2 * i] = __ffma2_rn(c2, a[2 * i + 1], a[2 * i]);
          a[2 * i + 1] = __ffma2_rn(s2, a[2 * i], a[2 * i + 1]);
          a[2 * i] = __ffma2_rn(c2, a[2 * i + 1], a[2 * i]);
        }
      } else if (MODE == 9) {  // in-place 3-shear rotation on bit 4
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          a[i] = __ffma2_rn(c2, a[i + 16], a[i]);
          a[i + 16] = __ffma2_rn(s2, a[i], a[i + 16]);
          a[i] = __ffma2_rn(c2, a[i + 16], a[i]);
        }
      } else if (MODE == 10) {  // rotation-gradient contraction: acc += l[j1] (.) p[j0] (three pair operands), 4 chains
        float2 v0[4], v1[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { v0[k] = make_float2(0.f, 0.f); v1[k] = make_float2(0.f, 0.f); }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          v0[i & 3] = __ffma2_rn(a[2 * i + 17 > 31 ? 31 : 2 * i + 17], a[2 * i], v0[i & 3]);
          v1[i & 3] = __ffma2_rn(a[2 * i + 16], a[2 * i + 1], v1[i & 3]);
        }
        a[0] = __fadd2_rn(a[0], __fadd2_rn(__fadd2_rn(v0[0], v0[1]), __fadd2_rn(v0[2], v0[3])));
        a[1] = __fadd2_rn(a[1], __fadd2_rn(__fadd2_rn(v1[0], v1[1]), __fadd2_rn(v1[2], v1[3])));
      } else if (MODE == 11) {  // unit-phase multiply in place: a = er*a + (-ei,ei) (.) swap(a)   (DIAG / CU3 phases)
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float2 sw = make_float2(a[i].y, a[i].x);
          a[i] = __ffma2_rn(ui, sw, __fmul2_rn(ur, a[i]));
        }
      } else if (MODE == 7) {  // packed general complex 2x2 on bit 0 (first row only + second row): 8 packed + 4 swaps / pair
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float2 a0 = a[2 * i], a1 = a[2 * i + 1];
          const float2 a0s = make_float2(a0.y, a0.x), a1s = make_float2(a1.y, a1.x);
          float2 o0 = __fmul2_rn(ur, a0); o0 = __ffma2_rn(ui, a0s, o0); o0 = __ffma2_rn(vr, a1, o0); o0 = __ffma2_rn(vi, a1s, o0);
          float2 o1 = __fmul2_rn(vr, a0); o1 = __ffma2_rn(vi, a0s, o1); o1 = __ffma2_rn(ur, a1, o1); o1 = __ffma2_rn(ui, a1s, o1);
          a[2 * i] = o0; a[2 * i + 1] = o1;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) { re[i] = a[i].x; im[i] = a[i].y; }
  }
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) acc += re[i] + im[i];
  if (acc == 123.456f) out[0] = acc;
}

template <int MODE, int THREADS>
void run(const char* name, float* d_out, float* d_in, int ctas_per_sm) {
  const int iters = 2000, grid = 148 * ctas_per_sm;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE, THREADS><<<grid, THREADS>>>(d_out, d_in, 10, 0.999f, 1e-4f);
  cudaEventRecord(e0);
  k<MODE, THREADS><<<grid, THREADS>>>(d_out, d_in, iters, 0.999f, 1e-4f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double per_iter = (MODE == 8 || MODE == 9) ? 96.0 : (MODE == 10 ? 48.0 : 128.0);   // scalar-equivalent FMA-pipe instructions
  const double instr = (double)grid * (THREADS / 32) * iters * per_iter;
  const double per_clk_smsp = instr / (ms * 1e-3 * 1.965e9) / (148 * 4);
  printf("%-34s threads=%4d ctas/sm=%d: %.3f ms, %.3f FMA-pipe instr/clk/SMSP\n", name, THREADS, ctas_per_sm, ms, per_clk_smsp);
}

int main() {
  float *d_out, *d_in;
  cudaMalloc(&d_out, 16); cudaMalloc(&d_in, 4096);
  float h[1024]; for (int i = 0; i < 1024; ++i) h[i] = 0.001f * (i % 97) + 0.1f;
  cudaMemcpy(d_in, h, 4096, cudaMemcpyHostToDevice);
  run<0, 256>("probe const operands", d_out, d_in, 2);
  run<1, 256>("fma reg operands (reuse)", d_out, d_in, 2);
  run<2, 256>("rotation bit0", d_out, d_in, 1);
  run<2, 256>("rotation bit0", d_out, d_in, 2);
  run<2, 512>("rotation bit0", d_out, d_in, 1);
  run<3, 256>("rotation bit4", d_out, d_in, 2);
  run<3, 128>("rotation bit4", d_out, d_in, 1);
  run<5, 256>("packed chain (x2 counted as 2)", d_out, d_in, 2);
  run<6, 256>("packed rotation bit0", d_out, d_in, 2);
  run<6, 512>("packed rotation bit0", d_out, d_in, 1);
  run<6, 128>("packed rotation bit0", d_out, d_in, 1);
  run<7, 256>("packed general 2x2 (256 scalar-eq)", d_out, d_in, 2);
  run<8, 256>("packed shear rot bit0", d_out, d_in, 2);
  run<8, 512>("packed shear rot bit0", d_out, d_in, 1);
  run<8, 256>("packed shear rot bit0", d_out, d_in, 1);
  run<9, 256>("packed shear rot bit4", d_out, d_in, 2);
  run<9, 128>("packed shear rot bit4", d_out, d_in, 1);
  run<10, 256>("packed grad contraction (3 pairs)", d_out, d_in, 2);
  run<10, 256>("packed grad contraction (3 pairs)", d_out, d_in, 1);
  run<11, 256>("packed phase multiply (swap+2)", d_out, d_in, 2);
  run<11, 256>("packed phase multiply (swap+2)", d_out, d_in, 1);
  return 0;
}
