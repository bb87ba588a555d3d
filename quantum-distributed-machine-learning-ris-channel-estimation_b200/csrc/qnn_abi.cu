// C ABI of libqnn_b200.so (include/qnn_b200.h): circuit lowering and kernel dispatch.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <map>
#include <mutex>
#include <new>

#include "qnn_internal.h"
#include "qnn_vm.h"

static thread_local char g_err[512] = "";

void qnn_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* qnn_last_error(void) { return g_err; }
extern "C" int32_t qnn_abi_version(void) { return QNN_ABI_VERSION; }

int qnn_sm_count() {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return 148; }
  if (dev >= 0 && dev < 64) { const int v = cache[dev].load(std::memory_order_relaxed); if (v > 0) return v; }
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) { (void)cudaGetLastError(); return 148; }
  if (dev >= 0 && dev < 64) cache[dev].store(sms, std::memory_order_relaxed);
  return sms;
}

int qnn_allow_smem(const void* kernel, size_t smem) {
  if (smem <= 48 * 1024) return QNN_OK;
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, size_t> allowed;
  int dev = 0;
  QNN_CUDA_CHECK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  size_t& have = allowed[{kernel, dev}];
  if (have >= smem) return QNN_OK;
  QNN_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  have = smem;
  return QNN_OK;
}

static int n_angles(int kind) {
  switch (kind) {
    case QNN_RX: case QNN_RY: case QNN_RZ: return 1;
    case QNN_U3: case QNN_CU3: return 3;
    case QNN_CNOT: case QNN_CZ: return 0;
    default: return -1;
  }
}
static bool two_qubit(int kind) { return kind == QNN_CNOT || kind == QNN_CZ || kind == QNN_CU3; }

extern "C" int qnn_circuit_create(const qnn_op* ops, int32_t n_ops, int32_t n_qubits, int32_t n_inputs,
                                  int32_t n_weights, int32_t encoder, qnn_circuit** out) {
  return qnn_circuit_create_ex(ops, n_ops, n_qubits, n_inputs, n_weights, encoder, 0u, out);
}

extern "C" int qnn_circuit_create_ex(const qnn_op* ops, int32_t n_ops, int32_t n_qubits, int32_t n_inputs,
                                     int32_t n_weights, int32_t encoder, uint32_t flags, qnn_circuit** out) {
  if (!out) { qnn_set_error("qnn_circuit_create: out is NULL"); return QNN_ERR_INVALID; }
  *out = nullptr;
  if (flags & ~(uint32_t)(QNN_CREATE_GENERIC_ONLY | QNN_CREATE_WIDE_ONLY | QNN_CREATE_NARROW_ONLY | QNN_CREATE_16_PER_THREAD)) { qnn_set_error("qnn_circuit_create_ex: unknown flags 0x%x", flags); return QNN_ERR_INVALID; }
  if (n_qubits < 1 || n_qubits > QNN_MAX_QUBITS) {
    qnn_set_error("n_qubits=%d outside [1, %d]", n_qubits, QNN_MAX_QUBITS);
    return QNN_ERR_INVALID;
  }
  if (n_ops < 0 || (n_ops > 0 && !ops) || n_inputs < 0 || n_weights < 0) {
    qnn_set_error("bad n_ops/n_inputs/n_weights (%d/%d/%d) or NULL ops", n_ops, n_inputs, n_weights);
    return QNN_ERR_INVALID;
  }
  if (encoder != QNN_ENC_NONE && encoder != QNN_ENC_AMPLITUDE) {
    qnn_set_error("unknown encoder %d", encoder);
    return QNN_ERR_INVALID;
  }
  if (encoder == QNN_ENC_AMPLITUDE && n_inputs < 1) {
    qnn_set_error("amplitude encoder needs n_inputs >= 1");
    return QNN_ERR_INVALID;
  }
  if (encoder == QNN_ENC_NONE && n_inputs > 4096) {
    qnn_set_error("n_inputs=%d > 4096 angle inputs", n_inputs);
    return QNN_ERR_INVALID;
  }
  qnn_circuit* c = new (std::nothrow) qnn_circuit();
  if (!c) { qnn_set_error("out of host memory"); return QNN_ERR_INVALID; }
  c->n_qubits = n_qubits; c->n_inputs = n_inputs; c->n_weights = n_weights; c->encoder = encoder;
  c->ops.assign(ops, ops + n_ops);
  for (int i = 0; i < n_ops; ++i) {
    const qnn_op& o = ops[i];
    const int na = n_angles(o.kind);
    if (na < 0) { qnn_set_error("op %d: unknown gate kind %d", i, o.kind); delete c; return QNN_ERR_INVALID; }
    const bool two = two_qubit(o.kind);
    if (o.wire0 < 0 || o.wire0 >= n_qubits || (two && (o.wire1 < 0 || o.wire1 >= n_qubits || o.wire1 == o.wire0))) {
      qnn_set_error("op %d: wires (%d, %d) invalid for %d qubits", i, o.wire0, o.wire1, n_qubits);
      delete c; return QNN_ERR_INVALID;
    }
    DevOp d{};
    d.kind = o.kind;
    const int target = two ? o.wire1 : o.wire0;
    d.tpos = n_qubits - 1 - target;
    d.cpos = two ? n_qubits - 1 - o.wire0 : -1;
    d.nang = na;
    d.coef = -1; d.slot = -1;
    bool any_x = false, any_w = false;
    for (int k = 0; k < 3; ++k) { d.src[k] = QNN_SRC_CONST; d.value[k] = 0.f; }
    for (int k = 0; k < na; ++k) {
      const int s = o.src[k];
      d.src[k] = s; d.value[k] = o.value[k];
      if (s == QNN_SRC_CONST) continue;
      if (s >= 0) {
        if (s >= n_weights) { qnn_set_error("op %d: weight index %d >= n_weights %d", i, s, n_weights); delete c; return QNN_ERR_INVALID; }
        any_w = true;
      } else {
        const int j = -1 - s;
        if (encoder == QNN_ENC_AMPLITUDE) { qnn_set_error("op %d: input-sourced angles cannot be combined with the amplitude encoder", i); delete c; return QNN_ERR_INVALID; }
        if (j >= n_inputs) { qnn_set_error("op %d: input column %d >= n_inputs %d", i, j, n_inputs); delete c; return QNN_ERR_INVALID; }
        any_x = true;
      }
    }
    d.mode = (na == 0) ? 0 : (any_x ? 2 : 1);
    if (d.mode == 1) d.coef = c->n_coef++;
    if (any_w) {
      d.slot = c->n_slots;
      for (int k = 0; k < na; ++k) c->slot_w.push_back(d.src[k] >= 0 ? d.src[k] : -1);
      c->n_slots += na;
    }
    c->dev_ops_h.push_back(d);
  }
  // CSR weight -> slots
  std::vector<int32_t> ptr(n_weights + 1, 0), idx;
  for (int s = 0; s < c->n_slots; ++s) if (c->slot_w[s] >= 0) ptr[c->slot_w[s] + 1]++;
  for (int i = 0; i < n_weights; ++i) ptr[i + 1] += ptr[i];
  idx.resize(ptr[n_weights]);
  { std::vector<int32_t> fill(ptr.begin(), ptr.end() - 1);
    for (int s = 0; s < c->n_slots; ++s) if (c->slot_w[s] >= 0) idx[fill[c->slot_w[s]]++] = s; }

  c->flags = flags;
  c->vm = new (std::nothrow) vm::Program();
  if (c->vm) vm::compile(c, c->vm, (flags & QNN_CREATE_16_PER_THREAD) ? 4 : vm::kR);
  if (n_qubits <= 8 && encoder == QNN_ENC_NONE && !(flags & (QNN_CREATE_WIDE_ONLY | QNN_CREATE_16_PER_THREAD))) {
    c->vm_narrow = new (std::nothrow) vm::Program();
    if (c->vm_narrow) vm::compile(c, c->vm_narrow, 3);
  }

  cudaError_t e = cudaGetDevice(&c->device);
  if (e != cudaSuccess) {
    // No CUDA device: keep the lowered circuit for introspection (qnn_circuit_describe); every compute
    // call on it reports QNN_ERR_CUDA.  There is no CPU execution path.
    (void)cudaGetLastError();
    c->device = -1;
    *out = c;
    return QNN_OK;
  }
  c->num_sm = qnn_sm_count();
  auto upload = [&](void** dptr, const void* src, size_t bytes) -> cudaError_t {
    cudaError_t err = cudaMalloc(dptr, bytes ? bytes : 16);
    if (err != cudaSuccess) return err;
    return bytes ? cudaMemcpy(*dptr, src, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
  };
  if ((e = upload((void**)&c->d_ops, c->dev_ops_h.data(), c->dev_ops_h.size() * sizeof(DevOp))) != cudaSuccess ||
      (e = upload((void**)&c->d_slot_ptr, ptr.data(), ptr.size() * sizeof(int32_t))) != cudaSuccess ||
      (e = upload((void**)&c->d_slot_idx, idx.data(), idx.size() * sizeof(int32_t))) != cudaSuccess) {
    qnn_set_error("uploading the lowered circuit failed: %s", cudaGetErrorString(e));
    qnn_circuit_destroy(c);
    return QNN_ERR_CUDA;
  }
  for (vm::Program* prog : {c->vm, c->vm_narrow}) {
    if (prog && prog->ok) {
      int rc = vm::upload(prog, n_weights);
      if (rc != QNN_OK) { qnn_circuit_destroy(c); return rc; }
    }
  }
  c->use_vm = !(flags & QNN_CREATE_GENERIC_ONLY);
  *out = c;
  return QNN_OK;
}

extern "C" void qnn_circuit_destroy(qnn_circuit* c) {
  if (!c) return;
  if (c->d_ops) cudaFree(c->d_ops);
  if (c->d_slot_ptr) cudaFree(c->d_slot_ptr);
  if (c->d_slot_idx) cudaFree(c->d_slot_idx);
  if (c->vm) { vm::release(c->vm); delete c->vm; }
  if (c->vm_narrow) { vm::release(c->vm_narrow); delete c->vm_narrow; }
  delete c;
}

extern "C" int32_t qnn_circuit_n_qubits(const qnn_circuit* c) { return c ? c->n_qubits : -1; }
extern "C" int32_t qnn_circuit_n_inputs(const qnn_circuit* c) { return c ? c->n_inputs : -1; }
extern "C" int32_t qnn_circuit_n_weights(const qnn_circuit* c) { return c ? c->n_weights : -1; }

extern "C" size_t qnn_circuit_describe(const qnn_circuit* c, char* buf, size_t len) {
  if (!c || !c->vm) return 0;
  // the program a large batch runs; circuits that only the narrow variant takes (n = 3, 4) describe that one
  const bool narrow_ok = c->vm_narrow && c->vm_narrow->ok;
  const vm::Program* prog = (narrow_ok && (!c->vm->ok || (c->flags & QNN_CREATE_NARROW_ONLY))) ? c->vm_narrow : c->vm;
  const std::string s = vm::describe(*prog);
  if (buf && len > 0) {
    const size_t k = std::min(len - 1, s.size());
    memcpy(buf, s.data(), k);
    buf[k] = 0;
  }
  return s.size() + 1;
}

// Which kernel family serves (circuit, B): the register-resident VM when the circuit lowered to it
// and its tables fit in shared memory, otherwise the generic interpreter.
// The narrow variant (8 amplitudes per thread, a warp carries 256) serves batches that would leave most warp slots
// of the wide variant (32 per thread, 1024 per warp) empty: fewer than a quarter of the backward pass's warp slots
// (measured cross-over, profiles/experiments/variants.py: n = 6 at B = 16,384 and n = 8 at B = 4,096 run alike).
static const vm::Program* vm_plan(const qnn_circuit* c, int64_t B, vm::Plan* pl) {
  if (!c->use_vm) return nullptr;
  const bool wide_ok = c->vm && c->vm->ok, narrow_ok = c->vm_narrow && c->vm_narrow->ok;
  bool narrow = narrow_ok && !wide_ok;
  if (narrow_ok && wide_ok) {
    const int64_t spw = c->vm->wb ? 1 : (32 >> c->vm->tb);
    narrow = (c->flags & QNN_CREATE_NARROW_ONLY) || (B + spw - 1) / spw <= (int64_t)c->num_sm * 2;
  }
  const vm::Program* first = narrow ? c->vm_narrow : c->vm;
  const vm::Program* second = narrow ? c->vm : c->vm_narrow;
  if (first && first->ok && vm::plan(c, *first, B, pl)) return first;
  if ((c->flags & (QNN_CREATE_NARROW_ONLY | QNN_CREATE_WIDE_ONLY)) == 0 && second && second->ok && vm::plan(c, *second, B, pl)) return second;
  return nullptr;
}

extern "C" const char* qnn_circuit_plan(const qnn_circuit* c, int64_t B) {
  if (!c) return "";
  vm::Plan vp;
  if (const vm::Program* prog = vm_plan(c, B, &vp)) return prog->kr == 3 ? "vm-register-narrow" : prog->kr == 4 ? "vm-register-16" : "vm-register";
  GenericPlan p;
  generic_plan(c, B, &p);
  return p.name;   // a string literal
}

extern "C" int32_t qnn_forward_launches(const qnn_circuit* c, int64_t B) {
  if (!c) return 0;
  vm::Plan vp;
  if (const vm::Program* prog = vm_plan(c, B, &vp)) return vm::inline_prep(*prog) ? 1 : 2;
  return 1 + (c->n_coef > 0 ? 1 : 0);
}
extern "C" int32_t qnn_backward_launches(const qnn_circuit* c, int64_t B) {
  if (!c) return 0;
  vm::Plan vp;
  if (const vm::Program* prog = vm_plan(c, B, &vp)) return (vm::inline_prep(*prog) ? 1 : 2) + (c->n_weights > 0 ? 1 : 0);
  return 1 + (c->n_coef > 0 ? 1 : 0) + (c->n_weights > 0 ? 1 : 0);
}

extern "C" size_t qnn_forward_workspace_bytes(const qnn_circuit* c, int64_t B) {
  if (!c || B <= 0) return 0;
  vm::Plan vp;
  if (vm_plan(c, B, &vp)) return vp.ws_coef + vp.ws_diag + vp.ws_partial + vp.ws_cluster;
  GenericPlan p;
  generic_plan(c, B, &p);
  return p.ws_coef + p.ws_state_fwd;
}
extern "C" size_t qnn_backward_workspace_bytes(const qnn_circuit* c, int64_t B) {
  if (!c || B <= 0) return 0;
  vm::Plan vp;
  if (vm_plan(c, B, &vp)) return vp.ws_coef + vp.ws_diag + vp.ws_partial + vp.ws_cluster;
  GenericPlan p;
  generic_plan(c, B, &p);
  return p.ws_coef + p.ws_state_bwd + p.ws_partial;
}

int qnn_check_common(const char* fn, const qnn_circuit* c, const float* x, const float* w, int64_t B) {
  if (!c) { qnn_set_error("%s: circuit is NULL", fn); return QNN_ERR_INVALID; }
  if (B <= 0) { qnn_set_error("%s: B=%lld must be positive", fn, (long long)B); return QNN_ERR_INVALID; }
  if (c->device < 0) { qnn_set_error("%s: the circuit was created without a CUDA device; there is no CPU path", fn); return QNN_ERR_CUDA; }
  int dev = -1;
  QNN_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev != c->device) {   // the lowered program and its tables live on the creating device
    qnn_set_error("%s: the circuit was created on device %d but device %d is current", fn, c->device, dev);
    return QNN_ERR_INVALID;
  }
  if (!x && c->n_inputs > 0) { qnn_set_error("%s: x is NULL", fn); return QNN_ERR_INVALID; }
  if (!w && c->n_weights > 0) { qnn_set_error("%s: w is NULL", fn); return QNN_ERR_INVALID; }
  return QNN_OK;
}

extern "C" size_t qnn_state_bytes(const qnn_circuit* c, int64_t B) {
  if (!c || B <= 0) return 0;
  vm::Plan vp;
  const vm::Program* prog = vm_plan(c, B, &vp);
  return prog ? vm::state_bytes(*prog, B) : 0;
}

static int forward_impl(const char* fn, const qnn_circuit* c, const float* x, const float* w, float* y, float* state,
                        int64_t B, void* workspace, size_t workspace_bytes, void* cuda_stream) {
  int rc = qnn_check_common(fn, c, x, w, B);
  if (rc) return rc;
  if (!y) { qnn_set_error("%s: y is NULL", fn); return QNN_ERR_INVALID; }
  const size_t need = qnn_forward_workspace_bytes(c, B);
  if (need > 0 && (!workspace || workspace_bytes < need)) {
    qnn_set_error("%s: workspace of %zu bytes needed, %zu given", fn, need, workspace ? workspace_bytes : (size_t)0);
    return QNN_ERR_WORKSPACE;
  }
  vm::Plan vp;
  if (const vm::Program* prog = vm_plan(c, B, &vp)) return vm::forward(c, *prog, vp, x, w, y, B, workspace, state, static_cast<cudaStream_t>(cuda_stream));
  if (state) { qnn_set_error("%s: this circuit runs on the generic kernels, which keep no state (qnn_state_bytes() is 0)", fn); return QNN_ERR_UNSUPPORTED; }
  GenericPlan p;
  generic_plan(c, B, &p);
  return generic_forward(c, p, x, w, y, B, workspace, static_cast<cudaStream_t>(cuda_stream));
}

static int backward_impl(const char* fn, const qnn_circuit* c, const float* x, const float* w, const float* gy,
                         const float* state, float* gx, float* gw, int64_t B, void* workspace, size_t workspace_bytes,
                         void* cuda_stream) {
  int rc = qnn_check_common(fn, c, x, w, B);
  if (rc) return rc;
  if (!gy) { qnn_set_error("%s: gy is NULL", fn); return QNN_ERR_INVALID; }
  if (!gw && c->n_weights > 0) { qnn_set_error("%s: gw is NULL", fn); return QNN_ERR_INVALID; }
  const size_t need = qnn_backward_workspace_bytes(c, B);
  if (!workspace || workspace_bytes < need) {
    qnn_set_error("%s: workspace of %zu bytes needed, %zu given", fn, need, workspace ? workspace_bytes : (size_t)0);
    return QNN_ERR_WORKSPACE;
  }
  vm::Plan vp;
  if (const vm::Program* prog = vm_plan(c, B, &vp)) return vm::backward(c, *prog, vp, x, w, gy, gx, gw, B, workspace, state, static_cast<cudaStream_t>(cuda_stream));
  if (state) { qnn_set_error("%s: this circuit runs on the generic kernels, which keep no state (qnn_state_bytes() is 0)", fn); return QNN_ERR_UNSUPPORTED; }
  GenericPlan p;
  generic_plan(c, B, &p);
  return generic_backward(c, p, x, w, gy, gx, gw, B, workspace, static_cast<cudaStream_t>(cuda_stream));
}

extern "C" int qnn_forward(const qnn_circuit* c, const float* x, const float* w, float* y, int64_t B,
                           void* workspace, size_t workspace_bytes, void* cuda_stream) {
  return forward_impl("qnn_forward", c, x, w, y, nullptr, B, workspace, workspace_bytes, cuda_stream);
}

extern "C" int qnn_forward_save(const qnn_circuit* c, const float* x, const float* w, float* y, float* state,
                                int64_t B, void* workspace, size_t workspace_bytes, void* cuda_stream) {
  if (!state) { qnn_set_error("qnn_forward_save: state is NULL"); return QNN_ERR_INVALID; }
  return forward_impl("qnn_forward_save", c, x, w, y, state, B, workspace, workspace_bytes, cuda_stream);
}

extern "C" int qnn_backward(const qnn_circuit* c, const float* x, const float* w, const float* gy,
                            float* gx, float* gw, int64_t B, void* workspace, size_t workspace_bytes,
                            void* cuda_stream) {
  return backward_impl("qnn_backward", c, x, w, gy, nullptr, gx, gw, B, workspace, workspace_bytes, cuda_stream);
}

extern "C" int qnn_backward_saved(const qnn_circuit* c, const float* x, const float* w, const float* gy,
                                  const float* state, float* gx, float* gw, int64_t B, void* workspace,
                                  size_t workspace_bytes, void* cuda_stream) {
  if (!state) { qnn_set_error("qnn_backward_saved: state is NULL"); return QNN_ERR_INVALID; }
  return backward_impl("qnn_backward_saved", c, x, w, gy, state, gx, gw, B, workspace, workspace_bytes, cuda_stream);
}
