// Register-resident statevector VM: program format shared by the host compiler (qnn_vm_compile.cpp)
// and the kernels (qnn_vm.cu).
//
// Geometry: every thread keeps 2^R = 32 complex amplitudes of one sample in registers (R = 5 "register
// bits"); 2^TB threads share a sample (TB = n - 5 "thread bits", n = 5..13): for n <= 10 these are lanes of
// one warp, which then carries 32 / 2^TB samples and always 1024 amplitudes; for n = 11..13 a sample spans
// 2^(n-10) warps of one CTA (a "group"), which exchange through a group-wide shared buffer and a named barrier;
// for n = 14..16 the group is a thread-block cluster of 2..8 CTAs (a sample needs the registers of several
// SMs), exchanging through (distributed) shared memory with cluster barriers.  A *layout* says which logical wire sits on which
// physical bit; gates act only on register bits, and an EXCHANGE micro-op re-lays the state out
// through a warp-private shared-memory buffer with GF(2)-linear, bank-conflict-free addressing.
// CNOTs cost nothing: they update the layout matrix and are materialised by the next exchange.
// RZ / CZ (and the RZ factors of U3 = RZ RY RZ) are merged into DIAG micro-ops.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "qnn_internal.h"

namespace vm {

constexpr int kR = 5;                  // register bits
constexpr int kAmps = 1 << kR;         // amplitudes per thread
constexpr int kWarpAmps = 1024;        // amplitudes per warp (any TB)
constexpr int kMinQubits = 5, kMaxQubits = 16;
constexpr int kMaxThreadBits = kMaxQubits - kR;   // 11

enum MopKind : int32_t {
  MOP_END = 0,
  MOP_ROT = 1,        // real rotation RY on register bit a, applied in place as three shears; coef row idx; acc (1)
  MOP_ROT_X = 2,      // same, angle = x[b, idx] (per sample); gradient goes to gx
  MOP_CU3_L = 4,      // controlled U3 = diag(1,e^{i phi}) RY(theta) diag(1,e^{i lambda}) on the control=1 pairs:
                      //   target register bit a, control register bit b & 7; coef row idx; acc (3: theta, phi, lambda);
                      //   b & kCu3InlinePost: the e^{i phi} phase is applied here (else it was deferred into a DIAG)
  MOP_CU3_P = 5,      // same with control = parity(lane_in_sample & aux)
  MOP_DIAG = 6,       // multiply by diagonal table idx ([j][lane], 1024 entries); desc aux
  MOP_DIAG_REG = 7,   // diagonal that depends on register bits only (32-entry table idx); desc aux
  MOP_XCHG = 8,       // exchange idx; cluster geometry: aux = 1 local (own CTA's shared memory), aux = 2 across CTAs
                      //   through distributed shared memory, aux = 0 through L2 buffers in the workspace
  // macro-ops: one dispatch for a whole group (the per-op dispatch latency is otherwise comparable to a body)
  MOP_ROT5 = 9,       // rotations on all five register bits; bit k uses coef row idx + pos_k and accumulator
                      //   acc + pos_k, pos_k = (aux >> 3k) & 7
  MOP_ROTX5 = 10,     // per-sample rotations on all five register bits; bit k uses x column (idx >> 5k) & 31
  MOP_CU3_CHAIN = 11, // CU3_L (t,c) = (1,0), (2,1), (3,2), (4,3) in that order; rows idx..idx+3, accs acc + 3k;
                      //   aux bit k: gate k applies its post-phase inline
};

constexpr int32_t kCu3InlinePost = 8;

struct MicroOp {       // 32 bytes
  int32_t kind, a, b, idx, acc, aux;
  int32_t pad0;        // handler id (enum Handler): kind and register bits folded into one dense jump index
  int32_t pad1;        // coefficient row the kernel prefetches for this op (0 if none)
};

enum Handler : int32_t {
  H_END = 0, H_ROT = 1, H_ROTX = 6, H_CU3L = 11, H_CU3P = 31, H_DIAG = 36, H_DIAGR = 37, H_XCHG = 38,
  H_ROT5 = 39, H_ROTX5 = 40, H_CHAIN = 41,
};

struct DiagTerm {      // one factor of a diagonal: RZ(angle) on a physical bit, or CZ on two bits
  int32_t kind;        // 0: RZ, 1: CZ, 2: controlled phase e^{i angle} where both bits are 1 (a deferred CU3 post-phase)
  int32_t src;         // RZ angle source (qnn_op convention), QNN_SRC_CONST -> value
  float value;
  int32_t bit0, bit1;  // physical bit(s), in the layout current when the diagonal is applied
};

struct DiagDesc {      // per DIAG / DIAG_REG micro-op
  int32_t first_term, n_terms;
  int32_t table;       // DIAG: offset (in float2) into the big-table area; DIAG_REG: into the small one
  int32_t is_reg;
  int32_t acc_reg[kR];   // accumulator index receiving sum_p z_bit(p) Im(conj(lam_p) psi_p), or -1
  int32_t acc_thr[kMaxThreadBits];   // same for thread bits (lane bits first, then warp bits)
};

struct Exchange {      // byte offsets into the group's buffer (256 * 2^kr bytes per warp); address = base[thread] ^ off[j]
  uint32_t woff[32], roff[32];
  uint32_t base;       // first of 2 * G entries in Program::xbase (G = threads per group): write bases, then read bases
  uint32_t pad[3];     // pad[0]: 0 plain / L2, 1 local exchange of the cluster geometry, 2 distributed shared memory
                       //   (forward-direction maps; pad[1] = entry holding the backward-direction maps), 3 = such a
                       //   backward-direction entry (pad[1] = its forward entry)
};

struct GradSlot {      // gw[weight] += accumulator[acc]  (every gradient is a direct, already-reduced sum)
  int32_t weight, acc;
};

struct Program {
  bool ok = false;
  std::string why_not;           // when !ok: why the VM does not take this circuit (generic path used)
  bool amplitude = false;        // amplitude encoder: the (normalised) input row is the initial state (n <= 10)
  int kr = kR;                   // register bits: every thread keeps 2^kr amplitudes (5; 3 or 4 in the narrow variants)
  int n = 0, tb = 0;             // tb = n - kr thread bits
  int wb = 0;                    // warp bits: max(0, n - kr - 5); a group is 2^wb warps
  std::vector<MicroOp> ops;      // forward order, MOP_END-terminated
  std::vector<DevOp> coef_ops;   // recipes (gate kind + angle sources), one per 8-float coefficient row
  std::vector<DiagTerm> terms;
  std::vector<DiagDesc> diags;
  std::vector<Exchange> exchanges;
  std::vector<uint32_t> xbase;   // per-thread base offsets of all exchanges
  std::vector<GradSlot> slots;
  int n_acc = 0;                 // floats of gradient accumulators
  int n_big_tables = 0, n_small_tables = 0;
  int final_bit_of_wire[16];     // physical bit holding each wire at measurement (clean layout)
  int initial_bit_of_wire[16];
  // device copies
  MicroOp* d_ops = nullptr;
  DevOp* d_coef_ops = nullptr;
  DiagTerm* d_terms = nullptr;
  DiagDesc* d_diags = nullptr;
  Exchange* d_exchanges = nullptr;
  uint32_t* d_xbase = nullptr;
  GradSlot* d_slots = nullptr;
  int32_t* d_slot_ptr = nullptr;  // CSR weight -> slots
  int32_t* d_slot_idx = nullptr;
};

// Lower a validated circuit.  Never fails hard: sets ok=false + why_not when out of the VM's scope.
void compile(const qnn_circuit* c, Program* out, int kr = kR);
std::string describe(const Program& p);   // JSON, for tests and introspection

struct Plan {
  int threads, grid, warps_per_cta, fwd_threads;
  int cluster, n_clusters;               // CTAs per sample for n >= 14 (0: no cluster), clusters in the backward grid
  int n_clusters_fwd;                    // clusters in the forward grid (two forward CTAs fit on an SM)
  size_t ws_cluster;                     // L2 exchange buffers (ws_cbuf, if the program has L2 exchanges) + cross-CTA scratch
  size_t ws_cbuf;
  size_t smem_fwd, smem_bwd;
  size_t ws_coef, ws_diag, ws_partial;   // bytes (each 256-aligned)
};
bool plan(const qnn_circuit* c, const Program& p, int64_t B, Plan* out);
bool inline_prep(const Program& p);      // the kernel makes its per-call tables itself: no k_vm_prep launch
int upload(Program* p, int n_weights);
void release(Program* p);
size_t state_bytes(const Program& p, int64_t B);
// state: optional device buffer of state_bytes(): forward stores the final statevector of every sample
// there, backward then starts from it instead of re-running the forward program.
int forward(const qnn_circuit* c, const Program& p, const Plan& pl, const float* x, const float* w, float* y,
            int64_t B, void* ws, float* state, cudaStream_t st);
int backward(const qnn_circuit* c, const Program& p, const Plan& pl, const float* x, const float* w,
             const float* gy, float* gx, float* gw, int64_t B, void* ws, const float* state, cudaStream_t st);

}  // namespace vm
