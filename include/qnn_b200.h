/* qnn_b200.h -- C ABI of libqnn_b200.so: batched statevector QNN forward / adjoint backward and
 * the QuantumNAT post-measurement pipeline, as fused sm_100a CUDA kernels.
 *
 * This is the boundary a maintainer of the reference binds instead of PennyLane's
 * default.qubit simulator.  "Est" = Estimators_QuantumNAT_onchipQNN.py,
 * "Run" = Runner_P128_QuantumNAT_onchipQNN.py in the reference repository.
 *
 * Conventions
 *  - every function returns 0 (QNN_OK) or a negative QNN_ERR_* code and never throws;
 *    qnn_last_error() returns a thread-local, human-readable reason for the last failure;
 *  - x, w, y, gy, gx, gw, noise and workspace pointers are DEVICE pointers owned by the caller,
 *    contiguous row-major float32; the library allocates nothing per call and keeps no mutable
 *    global state between calls (several forwards may be outstanding before one backward,
 *    Run:341-364);
 *  - all work is enqueued on `cuda_stream` (a cudaStream_t); no call synchronises the device;
 *  - the device is the one current on the calling thread (the caller's tensors live there); it must be
 *    the device that was current when the circuit was created (its tables live there), else QNN_ERR_INVALID;
 *  - wire 0 is the most significant bit of the basis index (PennyLane's convention);
 *  - B may be any positive value.
 */
#ifndef QNN_B200_H
#define QNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define QNN_API __attribute__((visibility("default")))
#else
#define QNN_API
#endif

#define QNN_ABI_VERSION 1
#define QNN_MAX_QUBITS 16

enum qnn_status {
  QNN_OK = 0,
  QNN_ERR_INVALID = -1,      /* bad argument (null pointer, wire out of range, ...) */
  QNN_ERR_UNSUPPORTED = -2,  /* valid request this build has no kernel for */
  QNN_ERR_CUDA = -3,         /* a CUDA runtime call failed; see qnn_last_error() */
  QNN_ERR_WORKSPACE = -4     /* workspace missing or smaller than qnn_workspace_bytes() */
};

/* Gate kinds.  RY/RZ/CNOT are the reference's gate set (Est:133,134,138); U3/CU3 are
 * BASELINE.json's north_star; RX/CZ come for free.  Matrices:
 *   RX(t)=[[c,-is],[-is,c]]  RY(t)=[[c,-s],[s,c]]  RZ(p)=diag(e^{-ip/2},e^{+ip/2}),  c,s=cos,sin(t/2)
 *   U3(t,p,l)=[[c,-e^{il}s],[e^{ip}s,e^{i(p+l)}c]];  controlled gates: wire0 = control, wire1 = target. */
enum qnn_gate { QNN_RX = 0, QNN_RY = 1, QNN_RZ = 2, QNN_U3 = 3, QNN_CNOT = 4, QNN_CZ = 5, QNN_CU3 = 6 };

/* Initial state.  NONE: |0...0> (the reference's AngleEmbedding, Est:127, is expressed as RY
 * gates whose angle source is the input row).  AMPLITUDE: the input row (first 2^n entries,
 * zero padded), L2-normalised, is the initial state (north_star; not in the reference). */
enum qnn_encoder { QNN_ENC_NONE = 0, QNN_ENC_AMPLITUDE = 1 };

/* src[k] of an angle: >= 0 -> w[src]; QNN_SRC_X(j) -> x[b, j]; QNN_SRC_CONST -> value[k]. */
#define QNN_SRC_CONST INT32_MIN
#define QNN_SRC_X(j) (-1 - (int32_t)(j))

typedef struct qnn_op {
  int32_t kind;     /* enum qnn_gate */
  int32_t wire0;    /* target of a one-qubit gate; control of a two-qubit gate */
  int32_t wire1;    /* target of a two-qubit gate; -1 otherwise */
  int32_t src[3];   /* angle sources; RX/RY/RZ use src[0]; U3/CU3 use (theta, phi, lambda) */
  float value[3];   /* constants for QNN_SRC_CONST entries */
} qnn_op;

typedef struct qnn_circuit qnn_circuit; /* immutable after creation; shareable across streams/threads */

/* Replaces qml.device("default.qubit", wires=n) + qml.QNode(quantum_circuit, dev, interface="torch")
 * (Est:122,125-142,148): the gate list is what tracing `quantum_circuit` yields.  Allocates a few KB
 * of device memory for the lowered program (once, here -- never per call).
 * n_inputs = row length of x; n_weights = length of w. */
QNN_API int qnn_circuit_create(const qnn_op* ops, int32_t n_ops, int32_t n_qubits, int32_t n_inputs,
                       int32_t n_weights, int32_t encoder, qnn_circuit** out);
/* Same, with creation flags.  QNN_CREATE_GENERIC_ONLY keeps the circuit on the generic (shared-memory
 * interpreter) kernel family even where the register-resident family applies; the parity tests use it to
 * check one family against the other.  No counterpart in the reference. */
#define QNN_CREATE_GENERIC_ONLY 1u
/* Kernel-variant pins for measurements and tests.  By default a circuit of up to 8 qubits is lowered twice: to the wide
 * variant (32 amplitudes per thread, a warp carries 1024: the throughput geometry) and to the narrow one (8 per thread,
 * a warp carries 256), and each call picks by its batch size -- the reference's own sizes (256 to 4,096 samples of a
 * 6-qubit circuit, Run:341-352) would occupy 1 to 16 CTAs of the wide variant.  WIDE_ONLY / NARROW_ONLY pin one of them;
 * 16_PER_THREAD lowers the circuit (n <= 12) with 16 amplitudes per thread instead (an occupancy experiment). */
#define QNN_CREATE_WIDE_ONLY 2u
#define QNN_CREATE_NARROW_ONLY 4u
#define QNN_CREATE_16_PER_THREAD 8u
QNN_API int qnn_circuit_create_ex(const qnn_op* ops, int32_t n_ops, int32_t n_qubits, int32_t n_inputs,
                                  int32_t n_weights, int32_t encoder, uint32_t flags, qnn_circuit** out);
QNN_API void qnn_circuit_destroy(qnn_circuit* c);

/* Introspection. */
QNN_API int32_t qnn_circuit_n_qubits(const qnn_circuit* c);
QNN_API int32_t qnn_circuit_n_inputs(const qnn_circuit* c);
QNN_API int32_t qnn_circuit_n_weights(const qnn_circuit* c);
/* JSON description of the register-resident program the circuit was lowered to (micro-ops, merged
 * diagonals, exchanges, gradient slots), or {"ok": false, "why_not": ...} when only the generic kernels
 * apply.  Returns the buffer size needed (including the terminator); works without a CUDA device. */
QNN_API size_t qnn_circuit_describe(const qnn_circuit* c, char* buf, size_t len);
/* Name of the kernel family a call with batch B would use ("generic-warp", "generic-cta", ...). */
QNN_API const char* qnn_circuit_plan(const qnn_circuit* c, int64_t B);
/* Number of kernel launches one qnn_forward / qnn_backward call enqueues (for launch accounting). */
QNN_API int32_t qnn_forward_launches(const qnn_circuit* c, int64_t B);
QNN_API int32_t qnn_backward_launches(const qnn_circuit* c, int64_t B);

/* Scratch sizes in bytes for a batch of B samples (may be 0).  The caller allocates
 * (e.g. torch.empty) and may reuse one buffer across calls that do not overlap on the GPU. */
QNN_API size_t qnn_forward_workspace_bytes(const qnn_circuit* c, int64_t B);
QNN_API size_t qnn_backward_workspace_bytes(const qnn_circuit* c, int64_t B);

/* Replaces self.qlayer(x) (Est:191,199): y[b, q] = <Z_q> after running the circuit on row b of x
 * with weights w.   x [B, n_inputs], w [n_weights], y [B, n_qubits]. */
QNN_API int qnn_forward(const qnn_circuit* c, const float* x, const float* w, float* y, int64_t B,
                void* workspace, size_t workspace_bytes, void* cuda_stream);

/* Replaces torch autograd through default.qubit (loss.backward(), Run:364) with an adjoint-method
 * VJP that recomputes the state: given gy = dLoss/dy [B, n_qubits], writes gx = dLoss/dx
 * [B, n_inputs] (may be NULL) and gw = dLoss/dw [n_weights] (summed over the batch, overwritten,
 * deterministic reduction order).  `w` must be the weights the matching forward used (the caller
 * keeps a snapshot: the reference restores noisy weights before backward runs, Est:194-196). */
QNN_API int qnn_backward(const qnn_circuit* c, const float* x, const float* w, const float* gy, float* gx,
                 float* gw, int64_t B, void* workspace, size_t workspace_bytes, void* cuda_stream);

/* Forward that also keeps the final statevector of every sample, and the backward that starts from it
 * instead of re-running the forward program (the adjoint method needs psi_final; nothing per gate is
 * ever stored).  This is the pair autograd uses: what the reference's backprop keeps alive between
 * `self.qlayer(x)` (Est:191,199) and `loss.backward()` (Run:364) is ~60 intermediate states per sample,
 * here it is one.  `state` is an opaque device buffer of qnn_state_bytes(c, B) bytes (2^n * 8 per
 * sample, rounded up to whole warps); 0 means the circuit runs on the generic kernels, which recompute.
 * x, w and B given to qnn_backward_saved must be those of the qnn_forward_save that wrote `state`. */
QNN_API size_t qnn_state_bytes(const qnn_circuit* c, int64_t B);
QNN_API int qnn_forward_save(const qnn_circuit* c, const float* x, const float* w, float* y, float* state,
                             int64_t B, void* workspace, size_t workspace_bytes, void* cuda_stream);
QNN_API int qnn_backward_saved(const qnn_circuit* c, const float* x, const float* w, const float* gy,
                               const float* state, float* gx, float* gw, int64_t B, void* workspace,
                               size_t workspace_bytes, void* cuda_stream);

/* No counterpart in the reference.  Same gradients as qnn_backward (always computed by the generic kernel
 * family, so for n = 3..16 they agree with qnn_backward's register-resident kernels to rounding, not bit for
 * bit, and take several times longer), for callers that cannot supply a
 * workspace because they may not allocate device memory once running: real-time loops (a per-slot
 * channel-estimation service) where cudaMalloc's device-wide synchronisation and unbounded latency
 * are not acceptable, processes whose HBM is owned in full by another pool so that cudaMalloc
 * would fail, and C / FFI hosts with no caching allocator to borrow from.  The scratch is a fixed
 * array of QNN_STATIC_SCRATCH_BYTES inside the library image (held by every GPU that loads the
 * library, until exit); requests that need more are refused with QNN_ERR_WORKSPACE and a message.
 * The call allocates nothing.  Its first call on each device looks up the array's device address,
 * so make that first call before any stream capture begins.  All calls share the one array, which
 * keeps the previous call's contents: calls that can overlap on the GPU race -- keep them on one
 * stream or order them.  (Nothing in it needs clearing between calls.) */
#define QNN_STATIC_SCRATCH_BYTES (4u << 20)
QNN_API size_t qnn_static_scratch_bytes(void);
QNN_API int qnn_backward_static(const qnn_circuit* c, const float* x, const float* w, const float* gy,
                                float* gx, float* gw, int64_t B, void* cuda_stream);

/* QuantumNAT post-measurement pipeline (north_star; absent from the reference -- SURVEY.md 5.9):
 *   yh = (y - mean_b y) * rsqrt(var_b y + eps)   (unbiased variance; skipped if !normalise)
 *   yn = yh + sigma * noise                      (noise may be NULL -> no noise)
 *   out = lo + step * rint((clamp(yn, lo, hi) - lo) / step), step = (hi-lo)/(levels-1)  (if quantise)
 * stats [2*n] receives mean then rstd.  If sums_in is non-NULL it holds GLOBAL [sum y, sum y^2]
 * (2*n doubles) and B_global the global batch (data-parallel ranks all-reduce the sums first). */
typedef struct qnat_params {
  float sigma, lo, hi, eps;
  int32_t levels, normalise, quantise, reserved;
} qnat_params;

QNN_API size_t qnat_workspace_bytes(int64_t B, int32_t n);
/* stage 1: per-qubit [sum y, sum y^2] over this rank's B rows -> sums [2*n] doubles (device). */
QNN_API int qnat_batch_sums(const float* y, double* sums, int64_t B, int32_t n, void* workspace,
                    size_t workspace_bytes, void* cuda_stream);
/* stage 2: apply; sums as produced by stage 1 (possibly all-reduced), B_total the row count they cover.
 * B_total = 0: the count is read on the device from sums[2*n] (a double the caller keeps next to the sums and
 * all-reduces with them), so ranks with unequal shard sizes need no host round trip to learn the global count. */
QNN_API int qnat_apply(const float* y, const float* noise, const double* sums, int64_t B_total, float* out,
               float* yn, float* stats, int64_t B, int32_t n, const qnat_params* p, void* cuda_stream);
/* backward stage 1: per-qubit [sum g, sum g*yh] with g = STE-masked gout -> bsums [2*n] doubles. */
QNN_API int qnat_backward_sums(const float* gout, const float* y, const float* yn, const float* stats,
                       double* bsums, int64_t B, int32_t n, const qnat_params* p, void* workspace,
                       size_t workspace_bytes, void* cuda_stream);
/* backward stage 2: gy = rstd * (g - s1/B_total - yh * s2/(B_total-1)).  B_total = 0: read from bsums[2*n]. */
QNN_API int qnat_backward_apply(const float* gout, const float* y, const float* yn, const float* stats,
                        const double* bsums, int64_t B_total, float* gy, int64_t B, int32_t n,
                        const qnat_params* p, void* cuda_stream);

/* Fused optimiser tail over flat buffers: the reference's gradient pruning
 * (zero g where |g| <= threshold, Est:205-228; pass a negative threshold to skip it) followed by one
 * torch.optim.AdamW step (Run:320,369; amsgrad off), as one elementwise kernel.  `step` is the 1-based
 * step count.  pruned_count (device, may be NULL) is incremented by the number of zeroed entries; grads
 * are overwritten with their pruned values. */
QNN_API int qnn_pruned_adamw_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                  float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                                  float prune_threshold, unsigned long long* pruned_count, void* cuda_stream);

/* The same step with the step count kept on the device: *steps_done (int64, device) is the number of steps
 * taken so far; the kernel derives AdamW's bias corrections from it and a second one-thread kernel
 * increments it.  No argument changes from one step to the next, so a training step that ends with this
 * call can be recorded once and replayed (what runner.GraphedQSCStep does with a CUDA graph). */
QNN_API int qnn_pruned_adamw_step_counted(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                          float lr, float beta1, float beta2, float eps, float weight_decay,
                                          long long* steps_done, float prune_threshold,
                                          unsigned long long* pruned_count, void* cuda_stream);

/* Angle head of the classifier: the last two stages of the reference's `preprocess`
 * stack, nn.Linear(32*(h/4)*(w/4), n_qubits) + nn.Tanh (Est:161-162), which produce the encoder angles:
 *   x[b, j] = tanh(bias[j] + sum_k f[b, k] * W[j, k])        f [B, K], W [n, K] (torch's Linear layout), x [B, n]
 * n <= 16, K <= 4096, n*K <= 49152 (the weight matrix is staged in shared memory). */
QNN_API int qnn_head_forward(const float* f, const float* W, const float* bias, float* x, int64_t B, int32_t K,
                             int32_t n, void* cuda_stream);
/* Its backward from gx = dL/dx and the saved output x: gf [B, K] (may be NULL when the features need no
 * gradient), gW [n, K], gbias [n]; sums over the batch are two-stage and order-fixed.  The workspace holds the
 * per-CTA partial sums (qnn_head_workspace_bytes; 0 for shapes outside the limits). */
QNN_API size_t qnn_head_workspace_bytes(int64_t B, int32_t K, int32_t n);
QNN_API int qnn_head_backward(const float* f, const float* W, const float* x, const float* gx, float* gf, float* gW,
                              float* gbias, int64_t B, int32_t K, int32_t n, void* workspace, size_t workspace_bytes,
                              void* cuda_stream);

/* Measured FP32 FMA throughput of the current device in TFLOP/s (register-resident FMA chains,
 * timed with CUDA events; synchronises).  Used only by bench.py for the roofline denominator. */
QNN_API int qnn_fp32_peak_probe(int32_t iters, double* tflops, double* sm_clock_mhz_unused);

QNN_API const char* qnn_last_error(void);
QNN_API int32_t qnn_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* QNN_B200_H */
