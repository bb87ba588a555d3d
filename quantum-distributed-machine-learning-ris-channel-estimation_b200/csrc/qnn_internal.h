// Internal declarations shared by the translation units of libqnn_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/qnn_b200.h"

#define QNN_COEF_STRIDE 32  // floats per gate: U (8) + up to three G_k = dU_k * U^dagger (24)

// One lowered gate as the generic kernels read it (device memory, read through the RO path).
struct DevOp {
  int32_t kind;      // enum qnn_gate
  int32_t tpos;      // bit position (from LSB) of the target
  int32_t cpos;      // bit position of the control, -1 if none
  int32_t mode;      // 0: no angles (CNOT/CZ); 1: batch-invariant (weights/constants); 2: per-sample
  int32_t src[3];    // as qnn_op.src
  float value[3];
  int32_t nang;      // number of angles
  int32_t coef;      // row in the coefficient table (mode 1), -1 otherwise
  int32_t slot;      // first gradient slot of this gate's weight angles (mode 1/2), -1 if none
  int32_t pad;
};

namespace vm { struct Program; }

struct qnn_circuit {
  int32_t n_qubits = 0, n_inputs = 0, n_weights = 0, encoder = 0;
  std::vector<qnn_op> ops;        // as given
  std::vector<DevOp> dev_ops_h;   // lowered
  int32_t n_coef = 0;             // rows of the coefficient table (mode-1 gates)
  int32_t n_slots = 0;            // gradient slots = weight-sourced angles, one per (gate, angle)
  std::vector<int32_t> slot_w;    // slot -> weight index
  // device copies (allocated once in qnn_circuit_create, on the device current at that time)
  int device = -1;
  DevOp* d_ops = nullptr;
  int32_t* d_slot_ptr = nullptr;  // CSR weight -> slots
  int32_t* d_slot_idx = nullptr;
  int num_sm = 148;               // SM count of `device` (cudaDevAttrMultiProcessorCount), read at creation
  vm::Program* vm = nullptr;      // register-resident program (qnn_vm.h); vm->ok says whether it applies
  vm::Program* vm_narrow = nullptr;  // the same circuit with 8 amplitudes per thread (n <= 8): small batches spread over
                                  //   four times as many warps; chosen per call by the batch size
  bool use_vm = false;            // the circuit was not created with QNN_CREATE_GENERIC_ONLY
  uint32_t flags = 0;             // creation flags
};

// SM count of the current device, cached per device ordinal (148 when it cannot be read).
int qnn_sm_count();
// Opt a kernel in to `smem` bytes of dynamic shared memory (cudaFuncAttributeMaxDynamicSharedMemorySize):
// issued once per (kernel, device) and again only if a later launch needs more, not on every launch.
int qnn_allow_smem(const void* kernel, size_t smem);
// Argument checks shared by every compute entry point (NULL circuit / pointers, B, device the circuit lives on).
int qnn_check_common(const char* fn, const qnn_circuit* c, const float* x, const float* w, int64_t B);

void qnn_set_error(const char* fmt, ...);
#define QNN_CUDA_CHECK(expr)                                                            \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      qnn_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return QNN_ERR_CUDA;                                                              \
    }                                                                                   \
  } while (0)

// ---- generic (interpreter) kernels: qnn_generic.cu --------------------------------------
struct GenericPlan {
  int mode;            // 0: one warp per sample, state in shared memory
                       // 1: one CTA per sample, state in shared memory
                       // 2: one CTA per sample, state in global workspace
  int threads;         // per CTA
  int groups_per_cta;  // samples in flight per CTA
  int grid;            // CTAs
  size_t smem_fwd, smem_bwd;
  size_t ws_coef;      // bytes of the coefficient table
  size_t ws_state_fwd, ws_state_bwd;  // bytes of global state (mode 2)
  size_t ws_partial;   // bytes of per-group gradient partials (backward)
  const char* name;
};
int generic_plan(const qnn_circuit* c, int64_t B, GenericPlan* p);
int generic_forward(const qnn_circuit* c, const GenericPlan& p, const float* x, const float* w,
                    float* y, int64_t B, void* ws, cudaStream_t st);
int generic_backward(const qnn_circuit* c, const GenericPlan& p, const float* x, const float* w,
                     const float* gy, float* gx, float* gw, int64_t B, void* ws, cudaStream_t st);
