// Host side of the register-resident VM: per-call tables (k_vm_prep), gradient finish (k_vm_finalize),
// planning and launches.  The kernels themselves live in qnn_vm_kernel.cuh / qnn_vm_g*.cu.
#include "qnn_vm_kernel.cuh"

namespace vm {
// Narrow variants whose tables all live in shared memory make them in the kernel's prologue (VmArgs::inline_prep).
bool inline_prep(const Program& p) { return p.kr < kR && p.wb == 0; }

namespace {

// Layout of the shared-memory tables (byte offsets from their base); `total` is what they need.
struct TableLayout { unsigned ops, exch, xbase, diags, coef, big, small, xcoef, red, wacc, gxacc, total; };
TableLayout table_layout(const Program& p, int n_inputs, int warps, bool bwd, bool cluster) {
  const int G = 32 << p.wb, spw = p.wb ? 1 : (32 >> p.tb);
  unsigned at = 0;
  auto take = [&](size_t bytes) { const unsigned r = at; at += (unsigned)((bytes + 15) & ~size_t(15)); return r; };
  // diagonal tables are read as one 256-byte row per warp instruction: a row that straddles 128-byte lines
  // costs an extra shared-memory wavefront per read (dynamic shared memory starts 1 KB-aligned)
  auto align256 = [&]() { at = (at + 255u) & ~255u; };
  TableLayout t{};
  t.ops = take(p.ops.size() * sizeof(MicroOp));
  t.exch = take(p.exchanges.size() * sizeof(Exchange));
  t.xbase = (cluster || p.wb > 0) ? 0u : take(p.xbase.size() * sizeof(uint32_t));   // n > 10: read from global memory
  t.diags = take(p.diags.size() * sizeof(DiagDesc));
  t.coef = take(p.coef_ops.size() * 8 * sizeof(float));
  align256();
  const size_t NA = (size_t)1 << p.kr;                      // amplitudes per thread
  t.big = p.wb == 0 ? take((size_t)p.n_big_tables * NA * G * sizeof(float2)) : 0u;
  t.small = take((size_t)p.n_small_tables * NA * sizeof(float2));
  t.xcoef = take((size_t)warps * spw * std::max(n_inputs, 1) * sizeof(float2));
  t.red = take((size_t)warps * 16 * sizeof(float));
  if (bwd) {
    t.wacc = take((size_t)warps * std::max(p.n_acc, 1) * sizeof(float));
    t.gxacc = take((size_t)warps * spw * std::max(n_inputs, 1) * sizeof(float));
  }
  t.total = at;
  return t;
}
size_t table_bytes(const Program& p, int n_inputs, int warps, bool bwd, bool cluster) {
  return table_layout(p, n_inputs, warps, bwd, cluster).total;
}
// exchange buffers (8 KB per warp, each group's aligned to its size) + worst-case alignment gap + tables;
// the cluster geometry keeps its exchange buffer in the workspace
size_t smem_bytes(const Program& p, int n_inputs, int warps, bool bwd, bool cluster) {
  if (cluster) {   // + the two 64 KB buffers of the shared-memory exchanges (local and across CTAs), when the program has any
    bool has_shared = false;
    for (const Exchange& X : p.exchanges) has_shared = has_shared || X.pad[0] != 0;
    // (two alternating buffers in the backward kernel; the forward kernel keeps one, so that two of its CTAs fit on an SM)
    return table_bytes(p, n_inputs, warps, bwd, true) + (has_shared ? (bwd ? 2 : 1) * 65536 : 0) + 64;
  }
  const size_t warp_buf = (size_t)256 << p.kr;               // 32 lanes x 2^kr amplitudes x 8 bytes
  const size_t group_buf = warp_buf << p.wb;
  return group_buf + (size_t)warps * warp_buf + table_bytes(p, n_inputs, warps, bwd, false) + 64;
}

// ---- per-call tables: gate coefficients (double-precision trig) and merged diagonals -----------------
// Grid: coef_blocks CTAs for the coefficient rows, then `chunks` CTAs per diagonal, each filling kPrepChunk
// entries of its table (a 16-qubit table has 65,536 entries: one CTA per table took ~0.9 ms per call; two entries per
// thread keep the fill at a few microseconds, which matters when a rank's share of the batch is small).
constexpr int kPrepThreads = 128, kPrepChunk = 256, kPrepTermTile = 64;
__global__ void __launch_bounds__(kPrepThreads)
k_vm_prep(const DevOp* __restrict__ coef_ops, int n_coef, const DiagDesc* __restrict__ diags, int n_diags,
          const DiagTerm* __restrict__ terms, const float* __restrict__ w, int tb, int wbits, int kr, float* __restrict__ coef,
          float2* __restrict__ big, float2* __restrict__ small, int coef_blocks, int chunks) {
  if ((int)blockIdx.x < coef_blocks) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_coef) make_coef_row(coef_ops[i], w, coef + (size_t)i * QNN_COEF_STRIDE);
    return;
  }
  const int d = ((int)blockIdx.x - coef_blocks) / chunks, chunk = ((int)blockIdx.x - coef_blocks) % chunks;
  if (d >= n_diags) return;
  const DiagDesc D = diags[d];
  const int G = 32 << wbits;
  const int NA = 1 << kr;
  const int entries = D.is_reg ? NA : NA * G;
  if (chunk * kPrepChunk >= entries) return;
  const unsigned tmask = wbits ? (unsigned)(G - 1) : ((1u << tb) - 1u);   // thread bits of one sample
  // the terms of this diagonal, staged once per CTA: angle (the weight read happens here) and bit masks
  __shared__ double s_phi[kPrepTermTile];
  __shared__ unsigned s_m0[kPrepTermTile], s_m1[kPrepTermTile];
  __shared__ int s_kind[kPrepTermTile];
  constexpr int kPer = kPrepChunk / kPrepThreads;
  double ang[kPer];
#pragma unroll
  for (int q = 0; q < kPer; ++q) ang[q] = 0.0;
  for (int t0 = 0; t0 < D.n_terms; t0 += kPrepTermTile) {
    const int nt = min(kPrepTermTile, D.n_terms - t0);
    __syncthreads();
    if ((int)threadIdx.x < nt) {
      const DiagTerm tm = terms[D.first_term + t0 + threadIdx.x];
      const double v = (tm.src == QNN_SRC_CONST) ? (double)tm.value : (double)w[tm.src];
      s_kind[threadIdx.x] = tm.kind;
      s_phi[threadIdx.x] = tm.kind == 0 ? 0.5 * v : (tm.kind == 1 ? 3.14159265358979323846 : v);
      s_m0[threadIdx.x] = 1u << tm.bit0;
      s_m1[threadIdx.x] = tm.kind == 0 ? 0u : (1u << tm.bit1);
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < kPer; ++q) {
      const int e = chunk * kPrepChunk + q * kPrepThreads + threadIdx.x;
      const int j = D.is_reg ? e : (e / G), t = D.is_reg ? 0 : (e % G);
      const unsigned p = (unsigned)j | (((unsigned)t & tmask) << kr);
      double acc = ang[q];
      for (int k = 0; k < nt; ++k) {
        if (s_kind[k] == 0) acc += (p & s_m0[k]) ? s_phi[k] : -s_phi[k];
        else if ((p & s_m0[k]) && (p & s_m1[k])) acc += s_phi[k];
      }
      ang[q] = acc;
    }
  }
#pragma unroll
  for (int q = 0; q < kPer; ++q) {
    const int e = chunk * kPrepChunk + q * kPrepThreads + threadIdx.x;
    if (e >= entries) continue;
    double sn, cs;
    sincos(ang[q], &sn, &cs);
    const float2 v = make_float2((float)cs, (float)sn);
    if (D.is_reg) small[(size_t)D.table * NA + e] = v; else big[(size_t)D.table * NA * G + e] = v;
  }
}

// gw[i] = sum over the gradient slots of weight i of the CTA partials, in a fixed order.
__global__ void k_vm_finalize(const float* __restrict__ partial, int n_ctas, int n_acc, const GradSlot* __restrict__ slots,
                              const int32_t* __restrict__ slot_ptr, const int32_t* __restrict__ slot_idx,
                              float* __restrict__ gw) {
  __shared__ double sred[128];
  // launched with programmatic stream serialization right behind the backward kernel: its CTAs may be resident
  // early; they wait here until that kernel has completed and its partial sums are visible (a no-op otherwise)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int i = blockIdx.x;
  double acc = 0.0;
  for (int e = slot_ptr[i]; e < slot_ptr[i + 1]; ++e) {
    const int ai = slots[slot_idx[e]].acc;
    for (int c = threadIdx.x; c < n_ctas; c += blockDim.x) acc += (double)partial[(size_t)c * n_acc + ai];
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

constexpr size_t kSmemMax = 227 * 1024;
constexpr int kFwdThreads = 512, kBwdThreads = 256;   // 16 / 8 warps per CTA, one CTA per SM (124 / 231 registers)


void fill_args(const qnn_circuit* c, const Program& p, const Plan& pl, void* ws, bool bwd, VmArgs* a) {
  unsigned char* wsb = static_cast<unsigned char*>(ws);
  const size_t G = (size_t)32 << p.wb;
  a->ops = p.d_ops; a->n_ops = (int)p.ops.size();
  a->exchanges = p.d_exchanges; a->n_exchanges = (int)p.exchanges.size();
  a->xbase = p.d_xbase; a->n_xbase = (int)p.xbase.size();
  a->diags = p.d_diags; a->n_diags = (int)p.diags.size();
  a->coef = reinterpret_cast<const float*>(wsb); a->n_coef = (int)p.coef_ops.size();
  a->big = reinterpret_cast<const float2*>(wsb + pl.ws_coef); a->n_big = p.n_big_tables;
  a->small = a->big + ((size_t)p.n_big_tables << p.kr) * G; a->n_small = p.n_small_tables;
  a->n = p.n; a->tb = p.tb; a->wb = p.wb; a->n_inputs = c->n_inputs; a->n_acc = p.n_acc;
  a->cluster = pl.cluster;
  a->inline_prep = inline_prep(p) ? 1 : 0;
  a->coef_ops = p.d_coef_ops; a->terms = p.d_terms; a->n_terms = (int)p.terms.size(); a->n_w = c->n_weights;
  a->amplitude = p.amplitude ? 1 : 0;
  a->n_angle = p.amplitude ? 0 : c->n_inputs;
  for (int w = 0; w < 16; ++w) a->init_bit_of_wire[w] = w < p.n ? (int8_t)p.initial_bit_of_wire[w] : 0;
  const int threads = bwd ? kBwdThreads : pl.fwd_threads;
  const TableLayout tl = table_layout(p, p.amplitude ? 0 : c->n_inputs, threads / 32, bwd, pl.cluster > 0);
  a->table_bytes = tl.total;
  a->off.ops = tl.ops; a->off.exch = tl.exch; a->off.xbase = tl.xbase; a->off.diags = tl.diags; a->off.coef = tl.coef;
  a->off.big = tl.big; a->off.small = tl.small; a->off.xcoef = tl.xcoef; a->off.red = tl.red; a->off.wacc = tl.wacc;
  a->off.gxacc = tl.gxacc;
  for (int b = 0; b < 16; ++b) a->wire_of_bit[b] = 0;
  for (int w = 0; w < p.n; ++w) a->wire_of_bit[p.final_bit_of_wire[w]] = (int8_t)w;
  a->partial = reinterpret_cast<float*>(wsb + pl.ws_coef + pl.ws_diag);
  if (pl.cluster > 0) {
    unsigned char* cb = wsb + pl.ws_coef + pl.ws_diag + pl.ws_partial;
    a->cbuf = cb;
    a->cred = reinterpret_cast<float*>(cb + pl.ws_cbuf);
  }
}

int launch_prep(const Program& p, const float* w, const VmArgs& a, cudaStream_t st) {
  if (inline_prep(p)) return QNN_OK;
  const int coef_blocks = ((int)p.coef_ops.size() + kPrepThreads - 1) / kPrepThreads;
  const int chunks = p.n_big_tables > 0 ? std::max(1, ((32 << p.kr) << p.wb) / kPrepChunk) : 1;   // CTAs per diagonal table
  const int blocks = coef_blocks + (int)p.diags.size() * chunks;
  if (blocks == 0) return QNN_OK;
  k_vm_prep<<<blocks, kPrepThreads, 0, st>>>(p.d_coef_ops, (int)p.coef_ops.size(), p.d_diags, (int)p.diags.size(), p.d_terms, w, p.tb,
                                             p.wb, p.kr, const_cast<float*>(a.coef), const_cast<float2*>(a.big),
                                             const_cast<float2*>(a.small), coef_blocks, chunks);
  QNN_CUDA_CHECK(cudaGetLastError());
  return QNN_OK;
}

}  // namespace

static int64_t n_batches(const Program& p, int64_t B) {
  const int64_t spw = p.wb ? 1 : (32 >> p.tb);
  return (B + spw - 1) / spw;
}

size_t state_bytes(const Program& p, int64_t B) {
  return (size_t)n_batches(p, B) * (((size_t)32 << p.kr) << p.wb) * sizeof(float2);
}

bool plan(const qnn_circuit* c, const Program& p, int64_t B, Plan* out) {
  if (!p.ok) return false;
  const int bwd_warps = kBwdThreads / 32;
  const int n_ang_in = p.amplitude ? 0 : c->n_inputs;        // input columns that need per-warp scratch
  out->cluster = 0; out->n_clusters = 0; out->n_clusters_fwd = 0; out->ws_cluster = 0; out->ws_cbuf = 0;
  out->threads = kBwdThreads;
  out->warps_per_cta = bwd_warps;
  if ((1 << p.wb) > bwd_warps) {
    // a sample needs the registers of several SMs: thread-block cluster of 2^wb / 8 CTAs (2, 4, 8)
    out->cluster = (1 << p.wb) / bwd_warps;
    out->fwd_threads = kBwdThreads;
    out->smem_fwd = smem_bytes(p, n_ang_in, bwd_warps, false, true);
    out->smem_bwd = smem_bytes(p, n_ang_in, bwd_warps, true, true);
    if (out->smem_fwd > kSmemMax || out->smem_bwd > kSmemMax) return false;
    out->n_clusters = (int)std::max<int64_t>(1, std::min<int64_t>(B, max_active_clusters_g2(out->cluster, out->smem_bwd, c->device, c->num_sm, true)));
    out->n_clusters_fwd = (int)std::max<int64_t>(1, std::min<int64_t>(B, max_active_clusters_g2(out->cluster, out->smem_fwd, c->device, c->num_sm, false)));
    out->grid = out->n_clusters * out->cluster;
    // L2 exchange buffers (four state-sized slices per cluster) for programs with exchanges across CTAs,
    // then the cross-CTA partial sums
    bool has_l2 = false;
    for (const MicroOp& m : p.ops) has_l2 = has_l2 || (m.kind == MOP_XCHG && (m.aux == 0 || m.aux == 2));   // (2: lam of the backward pass)
    const size_t ncl = (size_t)std::max(out->n_clusters, out->n_clusters_fwd);
    out->ws_cbuf = has_l2 ? (ncl << (p.n + 5)) : 0;
    out->ws_cluster = align256(out->ws_cbuf + ncl * (1 << p.wb) * 48 * sizeof(float));
  } else {
    // narrow variants serve small batches: 8-warp CTAs in the forward pass too, so that a batch spreads over twice as many SMs
    out->fwd_threads = p.kr < kR ? kBwdThreads : kFwdThreads;
    size_t sf = smem_bytes(p, n_ang_in, out->fwd_threads / 32, false, false);
    if (sf > kSmemMax && p.wb > 0) {                               // large samples: fewer groups per CTA
      out->fwd_threads = kBwdThreads;
      sf = smem_bytes(p, n_ang_in, bwd_warps, false, false);
    }
    const size_t sb = smem_bytes(p, n_ang_in, bwd_warps, true, false);
    if (sf > kSmemMax || sb > kSmemMax) return false;
    out->smem_fwd = sf;
    out->smem_bwd = sb;
    const int64_t groups_per_cta = bwd_warps >> p.wb;
    const int64_t ctas = (n_batches(p, B) + groups_per_cta - 1) / groups_per_cta;
    // backward: one CTA per SM (255 registers); the narrow variants fit two
    out->grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, (int64_t)c->num_sm * (p.kr < kR ? 2 : 1)));
  }
  out->ws_coef = align256(std::max<size_t>(p.coef_ops.size(), 1) * QNN_COEF_STRIDE * sizeof(float));
  out->ws_diag = align256(((size_t)p.n_big_tables * (((size_t)32 << p.kr) << p.wb) + ((size_t)p.n_small_tables << p.kr) + 1) * sizeof(float2));
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
  if ((rc = up((void**)&p->d_xbase, p->xbase.data(), p->xbase.size() * sizeof(uint32_t)))) return rc;
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
  for (void* d : {(void*)p->d_ops, (void*)p->d_coef_ops, (void*)p->d_terms, (void*)p->d_diags, (void*)p->d_exchanges, (void*)p->d_xbase,
                  (void*)p->d_slots, (void*)p->d_slot_ptr, (void*)p->d_slot_idx})
    if (d) cudaFree(d);
  p->d_ops = nullptr; p->d_coef_ops = nullptr; p->d_terms = nullptr; p->d_diags = nullptr;
  p->d_exchanges = nullptr; p->d_xbase = nullptr; p->d_slots = nullptr; p->d_slot_ptr = nullptr; p->d_slot_idx = nullptr;
}

// The kernel instantiation for (register bits, geometry, encoder).
static int launch_variant(const Program& p, const Plan& pl, bool bwd, bool state_on, const LaunchCfg& lc, const VmArgs& a) {
  if (p.kr == 3) return launch_g0_k3(bwd, state_on, lc, a);
  if (p.kr == 4) return p.wb > 0 ? launch_g1_k4(bwd, state_on, lc, a) : launch_g0_k4(bwd, state_on, lc, a);
  if (pl.cluster > 0) return launch_g2(bwd, state_on, lc, a);
  if (p.wb > 0) return launch_g1(bwd, state_on, lc, a);
  return p.amplitude ? launch_g0a(bwd, state_on, lc, a) : launch_g0(bwd, state_on, lc, a);
}

int forward(const qnn_circuit* c, const Program& p, const Plan& pl, const float* x, const float* w, float* y,
            int64_t B, void* ws, float* state, cudaStream_t st) {
  VmArgs a{};
  fill_args(c, p, pl, ws, false, &a);
  a.x = x; a.y = y; a.B = B; a.state = reinterpret_cast<float4*>(state); a.w = w;
  int rc = launch_prep(p, w, a, st);
  if (rc) return rc;
  int grid = pl.cluster > 0 ? pl.n_clusters_fwd * pl.cluster : pl.grid;
  if (pl.cluster == 0) {
    const int64_t groups_per_cta = (pl.fwd_threads / 32) >> p.wb;
    const int64_t ctas = (n_batches(p, B) + groups_per_cta - 1) / groups_per_cta;
    grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, (int64_t)c->num_sm * (p.kr < kR ? 4 : 1)));
  }
  const LaunchCfg lc{grid, pl.fwd_threads, pl.cluster, pl.smem_fwd, st};
  return launch_variant(p, pl, false, a.state != nullptr, lc, a);
}

int backward(const qnn_circuit* c, const Program& p, const Plan& pl, const float* x, const float* w,
             const float* gy, float* gx, float* gw, int64_t B, void* ws, const float* state, cudaStream_t st) {
  VmArgs a{};
  fill_args(c, p, pl, ws, true, &a);
  a.x = x; a.gy = gy; a.gx = gx; a.B = B; a.state = reinterpret_cast<float4*>(const_cast<float*>(state)); a.w = w;
  int rc = launch_prep(p, w, a, st);
  if (rc) return rc;
  const LaunchCfg lc{pl.grid, kBwdThreads, pl.cluster, pl.smem_bwd, st};
  rc = launch_variant(p, pl, true, a.state != nullptr, lc, a);
  if (rc) return rc;
  if (c->n_weights > 0) {
    // programmatic dependent launch: the finalize grid is set up while the backward kernel still runs (the launch
    // latency of a dependent kernel is a visible share of a small batch's backward pass)
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(c->n_weights); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr; cfg.numAttrs = 1;
    QNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_vm_finalize, (const float*)a.partial, (int)pl.grid, (int)p.n_acc, (const GradSlot*)p.d_slots,
                                      (const int32_t*)p.d_slot_ptr, (const int32_t*)p.d_slot_idx, gw));
    QNN_CUDA_CHECK(cudaGetLastError());
  }
  return QNN_OK;
}

}  // namespace vm
