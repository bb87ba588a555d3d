// Host-side compiler of the register-resident statevector VM (see qnn_vm.h).
//
// Input: the validated gate list of a qnn_circuit.  Output: a linear micro-op program plus the tables
// the kernels need.  Three ideas carry the performance:
//   1. list scheduling over the gate DAG so that as many gates as possible run while their target wire
//      sits on a register bit; when nothing more can run, ONE exchange brings the next wires in;
//   2. CNOTs are free: they only update the GF(2) layout (rowmask[wire] ^= rowmask[control]); the next
//      exchange materialises the accumulated permutation in its addressing;
//   3. diagonal gates (RZ, CZ, and the two RZ factors of U3 = e^{ia} RZ(phi) RY(theta) RZ(lambda)) are
//      collected and applied as one table multiply.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

#include "qnn_vm.h"

namespace vm {
namespace {

enum ItemKind { IT_REAL, IT_RZ, IT_CZ, IT_CP, IT_CNOT, IT_CGEN };

struct Item {
  ItemKind kind;
  int w0, w1;          // w0: target of 1q / control of 2q;  w1: target of 2q
  int gate;            // qnn_gate for coefficient recipes (IT_REAL: RY; IT_CGEN: CU3)
  int32_t src[3];
  float value[3];
  bool done = false;
  bool inline_post = true;   // IT_CGEN: the e^{i phi} phase is applied inside the gate body (else deferred as IT_CP)
};

inline int ctz(uint32_t v) { return __builtin_ctz(v); }
inline bool is_unit(uint32_t v) { return v != 0 && (v & (v - 1)) == 0; }

struct Lcg {
  uint64_t s;
  uint32_t next() { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return (uint32_t)(s >> 33); }
};

// rank of a set of `cnt` vectors of `bits` bits == cnt ?
bool independent(const uint32_t* v, int cnt) {
  uint32_t basis[16];
  int nb = 0;
  for (int i = 0; i < cnt; ++i) {
    uint32_t x = v[i];
    for (int b = 0; b < nb; ++b) x = std::min(x, x ^ basis[b]);
    if (!x) return false;
    basis[nb++] = x;
  }
  return true;
}

// Solve M p = e_w for every wire w, M given by rows rowmask[w] (bit b of rowmask = column b).
bool invert(const uint32_t* rowmask, int n, uint32_t* inv_col) {
  // Augmented elimination on rows: [rowmask | identity(wire)]
  uint32_t a[16], id[16];
  for (int i = 0; i < n; ++i) { a[i] = rowmask[i]; id[i] = 1u << i; }
  int piv_row_of_col[16];
  int r = 0;
  for (int col = 0; col < n; ++col) {
    int p = -1;
    for (int i = r; i < n; ++i) if ((a[i] >> col) & 1u) { p = i; break; }
    if (p < 0) return false;
    std::swap(a[p], a[r]); std::swap(id[p], id[r]);
    for (int i = 0; i < n; ++i) if (i != r && ((a[i] >> col) & 1u)) { a[i] ^= a[r]; id[i] ^= id[r]; }
    piv_row_of_col[col] = r;
    ++r;
  }
  // Now a[piv_row_of_col[col]] == 1<<col and id[...] = combination of wires: p_col = parity(id & k).
  // inv_col[w] = mask of physical bits col with (id[row(col)] >> w) & 1.
  for (int w = 0; w < n; ++w) {
    uint32_t m = 0;
    for (int col = 0; col < n; ++col) if ((id[piv_row_of_col[col]] >> w) & 1u) m |= 1u << col;
    inv_col[w] = m;
  }
  return true;
}

// Inverse of an invertible GF(2) matrix given by its columns (n <= 16).
bool invert_cols(const uint32_t* cols, int n, uint32_t* inv) {
  uint32_t a[16], id[16];          // rows of [M | I]
  for (int r = 0; r < n; ++r) {
    uint32_t row = 0;
    for (int c = 0; c < n; ++c) if ((cols[c] >> r) & 1u) row |= 1u << c;
    a[r] = row; id[r] = 1u << r;
  }
  for (int col = 0; col < n; ++col) {
    int p = -1;
    for (int r = col; r < n; ++r) if ((a[r] >> col) & 1u) { p = r; break; }
    if (p < 0) return false;
    std::swap(a[p], a[col]); std::swap(id[p], id[col]);
    for (int r = 0; r < n; ++r) if (r != col && ((a[r] >> col) & 1u)) { a[r] ^= a[col]; id[r] ^= id[col]; }
  }
  for (int c = 0; c < n; ++c) {     // id rows = rows of M^-1; return its columns
    uint32_t v = 0;
    for (int r = 0; r < n; ++r) if ((id[r] >> c) & 1u) v |= 1u << r;
    inv[c] = v;
  }
  return true;
}
inline uint32_t mat_vec(const uint32_t* cols, int n, uint32_t v) {
  uint32_t r = 0;
  for (int c = 0; c < n; ++c) if ((v >> c) & 1u) r ^= cols[c];
  return r;
}

struct Builder {
  const qnn_circuit* c;
  Program* P;
  int n, tb;
  int kr = kr;             // register bits of this program (5, or 3 / 4 for the narrow variants)
  bool no_local = false;   // cluster geometry: never use the shared-memory (local) exchange
  bool use_dsmem = true;   // cluster geometry: exchanges that cross CTAs go through distributed shared memory (else L2)
  int policy = 0;   // how "apply the pending diagonal" and "exchange" are ordered when both are possible
  std::vector<Item> items;
  uint32_t rowmask[16];
  struct Pending { int item; };
  std::vector<int> pending;          // item indices of RZ / CZ terms waiting for a flush
  std::vector<std::vector<int>> by_wire;
  std::vector<size_t> head;          // per wire: index into by_wire[w] of the first not-done item

  bool lower_items() {
    for (size_t i = 0; i < c->ops.size(); ++i) {
      const qnn_op& o = c->ops[i];
      auto x_src = [](int32_t s) { return s < 0 && s != QNN_SRC_CONST; };
      Item it{};
      for (int k = 0; k < 3; ++k) { it.src[k] = QNN_SRC_CONST; it.value[k] = 0.f; }
      switch (o.kind) {
        case QNN_RY:
          it.kind = IT_REAL; it.w0 = o.wire0; it.w1 = -1; it.gate = QNN_RY; it.src[0] = o.src[0]; it.value[0] = o.value[0];
          items.push_back(it);
          break;
        case QNN_RX: {
          if (x_src(o.src[0])) { P->why_not = "RX with an input-sourced angle"; return false; }
          // RX(t) = RZ(-pi/2) RY(t) RZ(pi/2): the constant RZ factors join the merged diagonals
          Item a = it; a.kind = IT_RZ; a.w0 = o.wire0; a.w1 = -1; a.gate = QNN_RZ; a.value[0] = 1.57079632679489662f;
          Item b = it; b.kind = IT_REAL; b.w0 = o.wire0; b.w1 = -1; b.gate = QNN_RY; b.src[0] = o.src[0]; b.value[0] = o.value[0];
          Item d = it; d.kind = IT_RZ; d.w0 = o.wire0; d.w1 = -1; d.gate = QNN_RZ; d.value[0] = -1.57079632679489662f;
          items.push_back(a); items.push_back(b); items.push_back(d);
          break;
        }
        case QNN_RZ:
          if (x_src(o.src[0])) { P->why_not = "RZ with an input-sourced angle"; return false; }
          it.kind = IT_RZ; it.w0 = o.wire0; it.w1 = -1; it.gate = QNN_RZ; it.src[0] = o.src[0]; it.value[0] = o.value[0];
          items.push_back(it);
          break;
        case QNN_U3: {
          for (int k = 0; k < 3; ++k) if (x_src(o.src[k])) { P->why_not = "U3 with an input-sourced angle"; return false; }
          // U3(t,p,l) = e^{i(p+l)/2} RZ(p) RY(t) RZ(l); the global phase drops out of <Z> and of all gradients
          Item a = it; a.kind = IT_RZ; a.w0 = o.wire0; a.w1 = -1; a.gate = QNN_RZ; a.src[0] = o.src[2]; a.value[0] = o.value[2];
          Item b = it; b.kind = IT_REAL; b.w0 = o.wire0; b.w1 = -1; b.gate = QNN_RY; b.src[0] = o.src[0]; b.value[0] = o.value[0];
          Item d = it; d.kind = IT_RZ; d.w0 = o.wire0; d.w1 = -1; d.gate = QNN_RZ; d.src[0] = o.src[1]; d.value[0] = o.value[1];
          items.push_back(a); items.push_back(b); items.push_back(d);
          break;
        }
        case QNN_CNOT:
          it.kind = IT_CNOT; it.w0 = o.wire0; it.w1 = o.wire1; it.gate = QNN_CNOT;
          items.push_back(it);
          break;
        case QNN_CZ:
          it.kind = IT_CZ; it.w0 = o.wire0; it.w1 = o.wire1; it.gate = QNN_CZ;
          items.push_back(it);
          break;
        case QNN_CU3:
          for (int k = 0; k < 3; ++k) {
            if (x_src(o.src[k])) { P->why_not = "CU3 with an input-sourced angle"; return false; }
            it.src[k] = o.src[k]; it.value[k] = o.value[k];
          }
          it.kind = IT_CGEN; it.w0 = o.wire0; it.w1 = o.wire1; it.gate = QNN_CU3;
          items.push_back(it);
          {   // placeholder for the deferred post-phase e^{i phi} on (control = 1, target = 1); see defer_phases()
            Item cp{};
            for (int k = 0; k < 3; ++k) { cp.src[k] = QNN_SRC_CONST; cp.value[k] = 0.f; }
            cp.kind = IT_CP; cp.w0 = o.wire0; cp.w1 = o.wire1; cp.gate = QNN_CZ;
            cp.src[0] = o.src[1]; cp.value[0] = o.value[1];
            items.push_back(cp);
          }
          break;
        default:
          P->why_not = "unknown gate";
          return false;
      }
    }
    defer_phases();
    by_wire.assign(n, {});
    for (size_t i = 0; i < items.size(); ++i) {
      by_wire[items[i].w0].push_back((int)i);
      if (items[i].w1 >= 0) by_wire[items[i].w1].push_back((int)i);
    }
    head.assign(n, 0);
    return true;
  }

  // CU3 = diag(1, e^{i phi}) R diag(1, e^{i lambda}) on the control = 1 block.  The post-phase is a
  // diagonal on (control, target): it commutes with every later gate that does not TARGET one of the two
  // wires, so it can ride along with a later merged diagonal for free.  Defer it when the next gate that
  // targets either wire is a single-qubit rotation (which flushes a diagonal anyway) or when there is none
  // (a trailing diagonal: no effect on <Z>); keep it inside the gate body when a controlled gate or a CNOT
  // would otherwise force a diagonal of its own.
  void defer_phases() {
    const bool no_defer = false;
    std::vector<Item> out;
    for (size_t i = 0; i < items.size(); ++i) {
      if (items[i].kind == IT_CP) continue;                   // placeholders are re-emitted below
      out.push_back(items[i]);
      if (items[i].kind != IT_CGEN) continue;
      const int c = items[i].w0, t = items[i].w1;
      bool defer = !no_defer;
      for (size_t k = i + 2; defer && k < items.size(); ++k) {
        const Item& nx = items[k];
        const int target = (nx.kind == IT_REAL) ? nx.w0 : (nx.kind == IT_CGEN || nx.kind == IT_CNOT) ? nx.w1 : -1;
        if (target == c || target == t) { defer = nx.kind == IT_REAL; break; }
      }
      out.back().inline_post = !defer;
      if (defer) out.push_back(items[i + 1]);                 // the IT_CP placeholder right behind the gate
    }
    items.swap(out);
  }

  bool ready(int i) const {
    const Item& it = items[i];
    auto first = [&](int w) { return head[w] < by_wire[w].size() ? by_wire[w][head[w]] : -1; };
    if (first(it.w0) != i) return false;
    if (it.w1 >= 0 && first(it.w1) != i) return false;
    return true;
  }
  void mark_done(int i) {
    items[i].done = true;
    for (int w : {items[i].w0, items[i].w1}) {
      if (w < 0) continue;
      while (head[w] < by_wire[w].size() && items[by_wire[w][head[w]]].done) ++head[w];
    }
  }

  bool row_unit(int w) const { return is_unit(rowmask[w]); }
  bool clean(int w) const {   // the wire's physical bit carries this wire and nothing else
    if (!is_unit(rowmask[w])) return false;
    for (int v = 0; v < n; ++v) if (v != w && (rowmask[v] & rowmask[w])) return false;
    return true;
  }
  int bit_of(int w) const { return ctz(rowmask[w]); }
  bool pending_on(int w) const {
    for (int i : pending) if (items[i].w0 == w || items[i].w1 == w) return true;
    return false;
  }

  int new_coef_row(int gate, const int32_t* src, const float* value, int nang) {
    DevOp d{};
    d.kind = gate; d.tpos = 0; d.cpos = -1; d.mode = 1; d.nang = nang; d.slot = -1;
    for (int k = 0; k < 3; ++k) { d.src[k] = src[k]; d.value[k] = value[k]; }
    d.coef = (int)P->coef_ops.size();
    P->coef_ops.push_back(d);
    return d.coef;
  }

  void flush_diag() {
    if (pending.empty()) return;
    DiagDesc D{};
    D.first_term = (int)P->terms.size();
    for (int r = 0; r < kR; ++r) D.acc_reg[r] = -1;
    for (int l = 0; l < kMaxThreadBits; ++l) D.acc_thr[l] = -1;
    bool reg_only = true;
    for (int i : pending) {
      const Item& it = items[i];
      DiagTerm t{};
      t.kind = it.kind == IT_CZ ? 1 : it.kind == IT_CP ? 2 : 0;
      t.src = it.src[0]; t.value = it.value[0];
      t.bit0 = bit_of(it.w0);
      t.bit1 = it.kind == IT_RZ ? -1 : bit_of(it.w1);
      if (t.bit0 >= kr || t.bit1 >= kr) reg_only = false;
      if (t.kind == 0 && t.src >= 0) {
        int32_t& a = t.bit0 < kr ? D.acc_reg[t.bit0] : D.acc_thr[t.bit0 - kr];
        if (a < 0) a = P->n_acc++;
        P->slots.push_back(GradSlot{t.src, a});
      }
      P->terms.push_back(t);
    }
    D.n_terms = (int)pending.size();
    D.is_reg = reg_only ? 1 : 0;
    D.table = reg_only ? P->n_small_tables++ : P->n_big_tables++;
    MicroOp m{};
    m.kind = reg_only ? MOP_DIAG_REG : MOP_DIAG;
    m.idx = D.table; m.aux = (int)P->diags.size(); m.acc = -1;
    P->diags.push_back(D);
    P->ops.push_back(m);
    pending.clear();
  }

  enum Status { ST_OK, ST_NEEDS_FLUSH, ST_BLOCKED };
  // Can item i run in the current layout?  ST_NEEDS_FLUSH: yes, once the pending diagonal is applied.
  Status classify(int i) const {
    const Item& it = items[i];
    switch (it.kind) {
      case IT_RZ: return row_unit(it.w0) ? ST_OK : ST_BLOCKED;
      case IT_CZ:
      case IT_CP: return (row_unit(it.w0) && row_unit(it.w1)) ? ST_OK : ST_BLOCKED;
      case IT_CNOT: return pending_on(it.w1) ? ST_NEEDS_FLUSH : ST_OK;
      case IT_REAL:
        if (!clean(it.w0) || bit_of(it.w0) >= kr) return ST_BLOCKED;
        return pending_on(it.w0) ? ST_NEEDS_FLUSH : ST_OK;
      case IT_CGEN:
        if (!clean(it.w1) || bit_of(it.w1) >= kr || !row_unit(it.w0)) return ST_BLOCKED;
        return pending_on(it.w1) ? ST_NEEDS_FLUSH : ST_OK;
    }
    return ST_BLOCKED;
  }

  // Execute item i in the current layout (flushing the pending diagonal if it must).
  bool try_exec(int i) {
    Item& it = items[i];
    switch (it.kind) {
      case IT_RZ:
        if (!row_unit(it.w0)) return false;
        pending.push_back(i);
        return true;
      case IT_CZ:
      case IT_CP:
        if (!row_unit(it.w0) || !row_unit(it.w1)) return false;
        pending.push_back(i);
        return true;
      case IT_CNOT:
        if (pending_on(it.w1)) flush_diag();
        rowmask[it.w1] ^= rowmask[it.w0];
        return true;
      case IT_REAL: {
        if (!clean(it.w0) || bit_of(it.w0) >= kr) return false;
        if (pending_on(it.w0)) flush_diag();
        MicroOp m{};
        m.a = bit_of(it.w0); m.acc = -1;
        if (it.src[0] < 0 && it.src[0] != QNN_SRC_CONST) {
          m.kind = MOP_ROT_X; m.idx = -1 - it.src[0];
        } else {
          m.kind = MOP_ROT; m.idx = new_coef_row(QNN_RY, it.src, it.value, 1);
          if (it.src[0] >= 0) { m.acc = P->n_acc++; P->slots.push_back(GradSlot{it.src[0], m.acc}); }
        }
        P->ops.push_back(m);
        return true;
      }
      case IT_CGEN: {
        if (!clean(it.w1) || bit_of(it.w1) >= kr || !row_unit(it.w0)) return false;
        if (pending_on(it.w1)) flush_diag();
        MicroOp m{};
        m.a = bit_of(it.w1); m.acc = -1;
        const int cb = bit_of(it.w0);
        if (cb < kr) { m.kind = MOP_CU3_L; m.b = cb; }
        else { m.kind = MOP_CU3_P; m.aux = 1 << (cb - kr); }   // thread-bit mask (lanes, then warps)
        if (it.inline_post) m.b |= kCu3InlinePost;
        m.idx = new_coef_row(QNN_CU3, it.src, it.value, 3);
        bool any_w = false;
        for (int k = 0; k < 3; ++k) any_w |= it.src[k] >= 0;
        if (any_w) {   // three direct accumulators: theta, phi, lambda
          m.acc = P->n_acc; P->n_acc += 3;
          for (int k = 0; k < 3; ++k) if (it.src[k] >= 0) P->slots.push_back(GradSlot{it.src[k], m.acc + k});
        }
        P->ops.push_back(m);
        return true;
      }
    }
    return false;
  }

  // Cluster geometry (n >= 14): bits 13.. of the extended index are the CTA rank.  An exchange that leaves the
  // wires on those bits where they are moves nothing between SMs and can run through the CTA's own shared
  // memory ("local"); only when a wanted wire sits on a CTA bit does the state have to cross the L2 ("global").
  // `global` false: the wanted wires take the register bits, the wires they displace take the vacated bits,
  // every other clean wire stays.  `global` true: the CTA bits receive the wires that are needed last.
  void choose_layout_cluster(int* bit_of_wire, const std::vector<int>& order_in, bool global) {
    std::vector<int> order(order_in.begin(), order_in.begin() + kr);
    if (global && all_clean())
      std::stable_partition(order.begin(), order.end(), [&](int w) { const int b = ctz(rowmask[w]); return b >= kr && b < kr + 5; });
    std::vector<int> nb(n, -1);
    uint32_t taken = 0;
    for (int r = 0; r < kr; ++r) { nb[order[r]] = r; taken |= 1u << r; }
    auto next_use = [&](int w) {
      for (size_t i = 0; i < items.size(); ++i)
        if (!items[i].done && (items[i].kind == IT_REAL || items[i].kind == IT_CGEN) && (items[i].w0 == w || items[i].w1 == w))
          return (int)i;
      return (int)items.size() + 1;
    };
    if (global) {
      std::vector<int> rest;
      for (int w = 0; w < n; ++w) if (nb[w] < 0) rest.push_back(w);
      std::stable_sort(rest.begin(), rest.end(), [&](int a, int b) { return next_use(a) > next_use(b); });
      for (int k = 0; k < n - 13 && k < (int)rest.size(); ++k) { nb[rest[k]] = 13 + k; taken |= 1u << (13 + k); }
    }
    for (int w = 0; w < n; ++w)          // clean wires keep their bit when it is still free
      if (nb[w] < 0 && clean(w)) { const int b = bit_of(w); if (!((taken >> b) & 1u)) { nb[w] = b; taken |= 1u << b; } }
    for (int w = 0; w < n; ++w)
      if (nb[w] < 0) { int b = 0; while ((taken >> b) & 1u) ++b; nb[w] = b; taken |= 1u << b; }
    for (int w = 0; w < n; ++w) bit_of_wire[w] = nb[w];
  }

  // Wires wanted on register bits next, in order of need.
  void choose_layout(int* bit_of_wire, bool global = true) {
    std::vector<int> order;
    auto add = [&](int w) { if (w >= 0 && std::find(order.begin(), order.end(), w) == order.end()) order.push_back(w); };
    for (size_t i = 0; i < items.size() && (int)order.size() < kr; ++i) {
      if (items[i].done) continue;
      if (items[i].kind == IT_REAL) add(items[i].w0);
      else if (items[i].kind == IT_CGEN) add(items[i].w1);
    }
    for (size_t i = 0; i < items.size() && (int)order.size() < kr; ++i)
      if (!items[i].done && items[i].kind == IT_CGEN) add(items[i].w0);
    for (int w = 0; w < n && (int)order.size() < kr; ++w)
      if (is_unit(rowmask[w]) && ctz(rowmask[w]) < kr) add(w);
    for (int w = 0; w < n && (int)order.size() < kr; ++w) add(w);
    if (P->wb > 3) {
      // (global exchanges: wires that now sit on lane bits take the low register bits, so that after the
      // exchange a thread's registers 0..3 are consecutive slots of the L2 buffer, one 32-byte sector)
      choose_layout_cluster(bit_of_wire, order, global);
      return;
    }
    std::vector<bool> used(n, false);
    for (int r = 0; r < kr; ++r) { bit_of_wire[order[r]] = r; used[order[r]] = true; }
    int b = kr;
    for (int w = 0; w < n; ++w) if (!used[w]) bit_of_wire[w] = b++;
  }

  // Cluster geometry, exchange through distributed shared memory.  Every thread stores its amplitudes straight
  // into the shared-memory buffer of the CTA that will hold them (st.shared::cluster) and, after one cluster
  // barrier, loads its new amplitudes from its own CTA's buffer.  The slot of an amplitude is a GF(2)-linear
  // function of its READER-side extended index: the CTA-rank bits (13..) are the reader's CTA rank, the low 13
  // bits are R times the reader's (register, lane, warp) bits, R invertible and chosen so that
  //   * a reading half-warp hits 16 distinct 8-byte banks, and
  //   * where the permutation allows it, the 32 lanes of a writing warp store one contiguous 256-byte run in one
  //     destination CTA (the images of the writer's lane bits go to slot bits 0..4).
  // wimg[e]: reader-side extended index contributed by writer-side bit e (E bits).  Outputs as in emit_exchange:
  // scol[writer bit], rcol[reader bit].
  bool dsmem_maps(const uint32_t* wimg, int E, Lcg& rng, uint32_t* scol, uint32_t* rcol) {
    const int LB = 13;
    const uint32_t lowmask = (1u << LB) - 1u;
    uint32_t u[5];
    for (int i = 0; i < 5; ++i) u[i] = wimg[5 + i] & lowmask;     // low parts of the images of the writer's lane bits
    for (int keep = 5; keep >= 0; --keep) {
      // a maximal independent subset of the first `keep` lane images gets the contiguous-run treatment
      // Lane images that are single reader REGISTER bits go first and so get slot bits 0, 1, ...: a reader's
      // registers that differ in those bits then sit in one 32-byte sector (the backward pass reads lam through
      // an L2 buffer with this map: four of its loads per sector instead of one).
      int sel[5], ns = 0;
      uint32_t chosen[13];
      for (int pass = 0; pass < 2; ++pass)
        for (int i = 0; i < keep; ++i) {
          const bool reg_unit = is_unit(u[i]) && ctz(u[i]) < kR && wimg[5 + i] == u[i];
          if (!u[i] || reg_unit != (pass == 0)) continue;
          chosen[ns] = u[i];
          if (independent(chosen, ns + 1)) sel[ns++] = i;
        }
      for (int attempt = 0; attempt < 4000; ++attempt) {
        uint32_t bm[13], tg[13];
        int nb = 0;
        for (int k = 0; k < ns; ++k) { bm[nb] = u[sel[k]]; tg[nb] = 1u << k; ++nb; }
        int guard = 0;
        while (nb < LB && guard++ < 1000) {                       // complete both to bases with random vectors
          bm[nb] = rng.next() & lowmask;
          if (!independent(bm, nb + 1)) continue;
          int g2 = 0;
          do { tg[nb] = rng.next() & lowmask; } while (!independent(tg, nb + 1) && g2++ < 1000);
          if (g2 >= 1000) break;
          ++nb;
        }
        if (nb < LB) continue;
        uint32_t bminv[13], rlow[13];
        if (!invert_cols(bm, LB, bminv)) continue;
        for (int b = 0; b < LB; ++b) rlow[b] = mat_vec(tg, LB, bminv[b]);   // R = T * B^-1, column b
        // reader side: lanes 0..3 of the reader (extended bits 5..8) must spread over the 16 banks of a half-warp
        uint32_t rl[4];
        for (int i = 0; i < 4; ++i) rl[i] = rlow[5 + i] & 15u;
        if (!independent(rl, 4)) continue;
        for (int b = 0; b < E; ++b) rcol[b] = b < LB ? rlow[b] : (1u << b);
        for (int e = 0; e < E; ++e) scol[e] = mat_vec(rcol, E, wimg[e]);
        return true;
      }
    }
    return false;
  }

  void push_exchange_tables(Exchange& X, const uint32_t* scol, const uint32_t* rcol) {
    const int G = 32 << P->wb;
    X.base = (uint32_t)P->xbase.size();
    for (int v = 0; v < (1 << kr); ++v) {
      uint32_t wo = 0, ro = 0;
      for (int i = 0; i < kr; ++i) if ((v >> i) & 1) { wo ^= scol[i]; ro ^= rcol[i]; }
      X.woff[v] = wo * 8; X.roff[v] = ro * 8;
    }
    for (int side = 0; side < 2; ++side)
      for (int t = 0; t < G; ++t) {
        uint32_t b = 0;
        for (int i = 0; i < 5 + P->wb; ++i) if ((t >> i) & 1) b ^= side ? rcol[kr + i] : scol[kr + i];
        P->xbase.push_back(b * 8);
      }
  }

  bool emit_exchange(const int* new_bit_of_wire, bool allow_global = true, bool* local_out = nullptr) {
    uint32_t inv_col[16];
    if (!invert(rowmask, n, inv_col)) { P->why_not = "internal: singular layout"; return false; }
    // Extended index of an amplitude inside a group's buffer: bits 0..4 register index, 5..9 lane,
    // 10.. warp-in-group.  For n < 10 the lanes above the sample's own are sample-in-warp bits and stay put.
    const int E = kr + 5 + P->wb;
    // acol[b']: old extended index contributed by new extended bit b'
    uint32_t acol[16];
    for (int w = 0; w < n; ++w) acol[new_bit_of_wire[w]] = inv_col[w];
    for (int b = n; b < E; ++b) acol[b] = 1u << b;
    // slot map S (columns scol[e], E bits), random invertible GF(2) matrix, accepted when both the writing
    // and the reading side are well-behaved: the low slot bits must be a bijection of the lane bits
    Lcg rng{0x9E3779B97F4A7C15ULL + P->exchanges.size() * 7919ULL};
    uint32_t scol[16], rcol[16];
    // cluster geometry: the exchange is local when every new CTA bit is the old one and no other new bit
    // depends on an old CTA bit - then an amplitude's slot is an address in its own CTA's 64 KB buffer
    bool local = P->wb > 3 && !no_local;
    if (local)
      for (int b = 0; b < E; ++b) {
        const uint32_t hi = acol[b] >> 13;
        if (hi != (b >= 13 ? 1u << (b - 13) : 0u)) local = false;
      }
    if (local_out) *local_out = local;
    if (P->wb > 3 && !local && !allow_global) return true;     // caller retries with a global layout
    if (P->wb > 3 && !local && use_dsmem) {
      // forward direction: writers hold the old layout, readers the new one; the backward pass runs the inverse
      // permutation and gets its own maps (entry pad[1]) so that its stores, too, are the remote side
      uint32_t ainv[16];
      if (!invert_cols(acol, E, ainv)) { P->why_not = "internal: singular exchange"; return false; }
      Exchange F{}, Bk{};
      bool ok = dsmem_maps(ainv, E, rng, scol, rcol);
      if (ok) {
        push_exchange_tables(F, scol, rcol);
        ok = dsmem_maps(acol, E, rng, scol, rcol);
        if (ok) push_exchange_tables(Bk, scol, rcol);
      }
      if (!ok) { P->why_not = "internal: no conflict-free distributed-shared-memory exchange found"; return false; }
      F.pad[0] = 2u; F.pad[1] = (uint32_t)P->exchanges.size() + 1u;
      Bk.pad[0] = 3u; Bk.pad[1] = (uint32_t)P->exchanges.size();     // 3: backward-direction tables of the entry before
      MicroOp m{};
      m.kind = MOP_XCHG; m.idx = (int)P->exchanges.size(); m.acc = -1; m.aux = 2;
      P->exchanges.push_back(F);
      P->exchanges.push_back(Bk);
      P->ops.push_back(m);
      for (int w = 0; w < n; ++w) rowmask[w] = 1u << new_bit_of_wire[w];
      return true;
    }
    const int ES = local ? 13 : E;                             // slot bits
    const uint32_t emask = (1u << ES) - 1u;
    for (int attempt = 0;; ++attempt) {
      if (attempt > 200000) { P->why_not = "internal: no conflict-free exchange found"; return false; }
      for (int e = 0; e < E; ++e) scol[e] = e < ES ? (rng.next() & emask) : 0u;
      if (P->wb > 3 && !local) {
        // L2 buffer: a warp's 32 lanes write one aligned 256-byte block: the low five slot bits are the lane
        // bits.  (A transposition cannot be coalesced on the reading side as well; but when the wires coming
        // onto register bits 0 and 1 were on lane bits, those lane bits become slot bits 0 and 1 and every
        // reader's registers 0..3 share one 32-byte sector.)
        int slot_of_lane[5] = {-1, -1, -1, -1, -1};
        uint32_t taken = 0;
        for (int b = 0; b < 2; ++b)
          if (is_unit(acol[b])) {
            const int e = ctz(acol[b]);
            if (e >= 5 && e < 10) { slot_of_lane[e - 5] = b; taken |= 1u << b; }
          }
        for (int i = 0; i < 5; ++i)
          if (slot_of_lane[i] < 0) { int b = 0; while ((taken >> b) & 1u) ++b; slot_of_lane[i] = b; taken |= 1u << b; }
        for (int i = 0; i < 5; ++i) scol[5 + i] = 1u << slot_of_lane[i];
      }
      if (!independent(scol, ES)) continue;
      for (int b = 0; b < E; ++b) {
        uint32_t s = 0;
        for (int e = 0; e < E; ++e) if ((acol[b] >> e) & 1u) s ^= scol[e];
        rcol[b] = s;
      }
      if (P->wb > 3 && !local) break;
      // shared-memory buffer: no bank conflicts within a half-warp, writing and reading
      uint32_t wlow[4], rlow[4];
      for (int i = 0; i < 4; ++i) { wlow[i] = scol[kr + i] & 15u; rlow[i] = rcol[kr + i] & 15u; }
      if (independent(wlow, 4) && independent(rlow, 4)) break;
    }
    Exchange X{};
    push_exchange_tables(X, scol, rcol);
    X.pad[0] = local ? 1u : 0u;
    MicroOp m{};
    m.kind = MOP_XCHG; m.idx = (int)P->exchanges.size(); m.acc = -1; m.aux = local ? 1 : 0;
    P->exchanges.push_back(X);
    P->ops.push_back(m);
    for (int w = 0; w < n; ++w) rowmask[w] = 1u << new_bit_of_wire[w];
    return true;
  }

  // Peephole: groups that layered circuits produce become one macro-op (one dispatch instead of 4-5).
  void fuse_groups() {
    std::vector<MicroOp> out;
    const std::vector<MicroOp>& in = P->ops;
    for (size_t i = 0; i < in.size();) {
      auto distinct_bits = [&](int kind, bool need_acc) {
        if (i + kr > in.size()) return false;
        uint32_t seen = 0;
        for (int k = 0; k < kr; ++k) {
          const MicroOp& m = in[i + k];
          if (m.kind != kind || ((seen >> m.a) & 1u)) return false;
          if (need_acc && (m.acc != in[i].acc + k || m.idx != in[i].idx + k)) return false;
          seen |= 1u << m.a;
        }
        return seen == (1u << kr) - 1u;
      };
      if (distinct_bits(MOP_ROT, true)) {
        MicroOp m{};
        m.kind = MOP_ROT5; m.idx = in[i].idx; m.acc = in[i].acc;
        for (int k = 0; k < kr; ++k) m.aux |= k << (3 * in[i + k].a);
        out.push_back(m);
        i += kr;
        continue;
      }
      if (distinct_bits(MOP_ROT_X, false)) {
        MicroOp m{};
        m.kind = MOP_ROTX5; m.acc = -1;
        for (int k = 0; k < kr; ++k) m.idx |= in[i + k].idx << (5 * in[i + k].a);
        out.push_back(m);
        i += kr;
        continue;
      }
      const int steps = kr - 1;                       // (target, control) = (1,0) ... (kr-1, kr-2)
      bool chain = steps >= 2 && i + steps <= in.size();
      for (int k = 0; chain && k < steps; ++k) {
        const MicroOp& m = in[i + k];
        chain = m.kind == MOP_CU3_L && m.a == k + 1 && (m.b & 7) == k && m.idx == in[i].idx + k && in[i].acc >= 0 &&
                m.acc == in[i].acc + 3 * k;
      }
      if (chain) {
        MicroOp m{};
        m.kind = MOP_CU3_CHAIN; m.idx = in[i].idx; m.acc = in[i].acc;
        for (int k = 0; k < steps; ++k) if (in[i + k].b & kCu3InlinePost) m.aux |= 1 << k;   // per-gate inline post-phase
        out.push_back(m);
        i += steps;
        continue;
      }
      out.push_back(in[i++]);
    }
    P->ops.swap(out);
  }

  // One dense handler id per (kind, register bits): the kernel dispatches with a single jump.
  // pad1 = coefficient row to prefetch (0 when the op has none).
  void assign_handlers() {
    for (MicroOp& m : P->ops) {
      m.pad1 = 0;
      switch (m.kind) {
        case MOP_ROT: m.pad0 = H_ROT + m.a; m.pad1 = m.idx; break;
        case MOP_ROT_X: m.pad0 = H_ROTX + m.a; break;
        case MOP_CU3_L: m.pad0 = H_CU3L + m.a * 4 + ((m.b & 7) < m.a ? (m.b & 7) : (m.b & 7) - 1); m.pad1 = m.idx; break;
        case MOP_CU3_P: m.pad0 = H_CU3P + m.a; m.pad1 = m.idx; break;
        case MOP_DIAG: m.pad0 = H_DIAG; break;
        case MOP_DIAG_REG: m.pad0 = H_DIAGR; break;
        case MOP_XCHG: m.pad0 = H_XCHG; break;
        case MOP_ROT5: m.pad0 = H_ROT5; break;
        case MOP_ROTX5: m.pad0 = H_ROTX5; break;
        case MOP_CU3_CHAIN: m.pad0 = H_CHAIN; m.pad1 = m.idx; break;
        default: m.pad0 = H_END; break;
      }
    }
  }

  bool all_clean() const {
    uint32_t seen = 0;
    for (int w = 0; w < n; ++w) {
      if (!is_unit(rowmask[w]) || (seen & rowmask[w])) return false;
      seen |= rowmask[w];
    }
    return true;
  }

  bool run() {
    n = c->n_qubits;
    tb = n - kr;
    P->kr = kr;
    P->n = n; P->tb = tb; P->wb = n > kr + 5 ? n - kr - 5 : 0;
    if (kr == kR && (n < kMinQubits || n > kMaxQubits)) { P->why_not = "n_qubits outside [5, 16]"; return false; }
    // narrow variants (fewer amplitudes per thread): whole samples per warp for kr = 3, up to one CTA per sample for kr = 4
    if (kr != kR && (n < kr || n > (kr == 3 ? 8 : 12))) { P->why_not = "n_qubits outside the narrow variant's range"; return false; }
    P->amplitude = c->encoder == QNN_ENC_AMPLITUDE;
    if (P->amplitude && kr != kR) { P->why_not = "amplitude encoder in a narrow variant"; return false; }
    if (P->amplitude && n > 10) { P->why_not = "amplitude encoder above 10 qubits"; return false; }
    if (!P->amplitude && c->n_inputs > 32) { P->why_not = "more than 32 angle input columns"; return false; }
    if (!lower_items()) return false;
    int bits[16];
    for (int w = 0; w < n; ++w) rowmask[w] = 1u << w;   // provisional, only "unit-ness" matters here
    choose_layout(bits);
    for (int w = 0; w < n; ++w) { rowmask[w] = 1u << bits[w]; P->initial_bit_of_wire[w] = bits[w]; }
    size_t remaining = items.size();
    // One sweep over the ready items; runs those whose status is ST_OK and that `want` accepts.
    auto sweep = [&](auto want) {
      bool progress = false;
      for (size_t i = 0; i < items.size(); ++i) {
        if (items[i].done || !ready((int)i) || !want(items[i]) || classify((int)i) != ST_OK) continue;
        if (try_exec((int)i)) { mark_done((int)i); --remaining; progress = true; }
      }
      return progress;
    };
    auto first_ready = [&](auto pred) {   // program-order index of the first ready item satisfying pred, or -1
      for (size_t i = 0; i < items.size(); ++i)
        if (!items[i].done && ready((int)i) && pred((int)i)) return (int)i;
      return -1;
    };
    while (remaining > 0) {
      // 1. everything that needs neither a flush nor layout bookkeeping
      if (sweep([](const Item& it) { return it.kind != IT_CNOT; })) continue;
      // 2. CNOTs that are free right now (they dirty the layout, so they go after the real work)
      if (sweep([](const Item& it) { return it.kind == IT_CNOT; })) continue;
      // 3./4. a gate that is local and clean but waits for the pending diagonal -> apply the diagonal;
      //       a gate that waits for its wire to come onto a register bit (or to be cleaned) -> exchange.
      //       Whichever waiting gate comes first in program order decides, so that work of the next
      //       layer does not fragment the diagonals of this one.
      const int f = first_ready([&](int i) { return items[i].kind != IT_CNOT && classify(i) == ST_NEEDS_FLUSH; });
      const int b = first_ready([&](int i) { return classify(i) == ST_BLOCKED; });
      bool flush_first = f >= 0 && (b < 0 || f < b);
      if (f >= 0 && policy == 0) flush_first = true;               // always finish local work first
      if (f >= 0 && policy == 2 && items[f].kind == IT_CGEN) {     // ... only for two-qubit gates whose
        bool all_local = true;                                      //     pending terms are register-only
        for (int i : pending) {
          if (bit_of(items[i].w0) >= kr) all_local = false;
          if (items[i].w1 >= 0 && bit_of(items[i].w1) >= kr) all_local = false;
        }
        if (all_local) flush_first = true;
      }
      if (flush_first) { flush_diag(); continue; }
      if (b >= 0) {
        if (P->wb > 3 && !no_local) {          // cluster geometry: a local exchange when it can do the job
          bool local = false;
          choose_layout(bits, false);
          if (!emit_exchange(bits, false, &local)) return false;
          if (local) continue;
        }
        choose_layout(bits);
        if (!emit_exchange(bits)) return false;
        continue;
      }
      // 5. only CNOTs whose target carries a pending diagonal term are left: apply the diagonal
      if (!pending.empty()) { flush_diag(); continue; }
      P->why_not = "internal: scheduler made no progress";
      return false;
    }
    // trailing diagonal terms do not change <Z> and have identically zero gradient: drop them
    pending.clear();
    if (!all_clean()) {
      for (int w = 0; w < n; ++w) bits[w] = -1;
      // keep wires that are clean where they are; give the rest the free bits
      uint32_t taken = 0;
      for (int w = 0; w < n; ++w) if (clean(w)) { bits[w] = bit_of(w); taken |= 1u << bits[w]; }
      for (int w = 0; w < n; ++w) if (bits[w] < 0) { int b = 0; while ((taken >> b) & 1u) ++b; bits[w] = b; taken |= 1u << b; }
      if (!emit_exchange(bits)) return false;
    }
    for (int w = 0; w < n; ++w) P->final_bit_of_wire[w] = bit_of(w);
    fuse_groups();
    assign_handlers();
    MicroOp e{};
    e.kind = MOP_END; e.acc = -1;
    P->ops.push_back(e);
    P->ok = true;
    return true;
  }
};

}  // namespace

// Estimated issue slots per amplitude of one forward pass (the backward pass scales alike).
static double program_cost(const Program& p) {
  double cost = 0;
  for (const MicroOp& m : p.ops) {
    switch (m.kind) {
      case MOP_ROT: case MOP_ROT_X: cost += 1.5; break;
      case MOP_ROT5: case MOP_ROTX5: cost += 7.0; break;
      case MOP_CU3_CHAIN: cost += 11.0; break;
      case MOP_CU3_L: cost += 3.0; break;
      case MOP_CU3_P: cost += 6.0; break;
      case MOP_DIAG_REG: cost += 5.0; break;
      case MOP_DIAG: cost += 6.0; break;
      case MOP_XCHG: cost += p.wb > 3 ? (m.aux == 1 ? 8.0 : m.aux == 2 ? 20.0 : 40.0) : 4.6; break;   // cluster: local / DSMEM / L2
      default: break;
    }
  }
  return cost;
}

void compile(const qnn_circuit* c, Program* out, int kr) {
  // The scheduler is greedy; its one real choice (diagonal first or exchange first) is made by trying
  // each policy and keeping the cheapest program.
  Program best;
  double best_cost = 0;
  // Cluster geometry: shared-memory exchanges (local, and across CTAs through distributed shared memory) need two
  // 64 KB buffers in every CTA next to the program's tables; programs whose tables leave no room for them are
  // compiled with L2 exchanges only.
  for (int no_local = 0; no_local < 2; ++no_local) {
  if (no_local) {
    if (!best.ok || best.wb <= 3) break;
    // backward kernel's tables (qnn_vm.cu: table_layout): program, maps, coefficient rows, small diagonal
    // tables, and eight per-warp copies of the gradient accumulators; 4 KB for the small per-warp scratch
    const size_t tables = best.ops.size() * sizeof(MicroOp) + best.exchanges.size() * sizeof(Exchange) +
                          best.diags.size() * sizeof(DiagDesc) + best.coef_ops.size() * 32 + (size_t)best.n_small_tables * 256 +
                          (size_t)best.n_acc * 32 + 4096;
    if (tables + 2 * 65536 + 64 <= 227 * 1024) break;
    best = Program();
  }
  for (int policy = 0; policy < 3; ++policy) {
    Program cand;
    Builder b{};
    b.c = c; b.P = &cand; b.policy = policy; b.no_local = no_local != 0; b.use_dsmem = no_local == 0; b.kr = kr;
    if (!b.run()) {
      cand.ok = false;
      cand.ops.clear();
      if (!best.ok && best.why_not.empty()) best = cand;
      continue;
    }
    const double cost = program_cost(cand);
    if (!best.ok || cost < best_cost) { best = cand; best_cost = cost; }
  }
  }
  *out = best;
}

std::string describe(const Program& p) {
  std::ostringstream o;
  o.precision(9);
  o << "{\"ok\": " << (p.ok ? "true" : "false") << ", \"why_not\": \"" << p.why_not << "\"";
  if (!p.ok) { o << "}"; return o.str(); }
  o << ", \"amplitude\": " << (p.amplitude ? "true" : "false") << ", \"kr\": " << p.kr << ", \"n\": " << p.n << ", \"tb\": " << p.tb << ", \"wb\": " << p.wb << ", \"n_acc\": " << p.n_acc;
  auto arr = [&](const char* name, const int* v, int cnt) {
    o << ", \"" << name << "\": [";
    for (int i = 0; i < cnt; ++i) o << (i ? ", " : "") << v[i];
    o << "]";
  };
  arr("initial_bit_of_wire", p.initial_bit_of_wire, p.n);
  arr("final_bit_of_wire", p.final_bit_of_wire, p.n);
  o << ", \"ops\": [";
  for (size_t i = 0; i < p.ops.size(); ++i) {
    const MicroOp& m = p.ops[i];
    o << (i ? ", " : "") << "[" << m.kind << ", " << m.a << ", " << m.b << ", " << m.idx << ", " << m.acc << ", " << m.aux << "]";
  }
  o << "], \"coef_ops\": [";
  for (size_t i = 0; i < p.coef_ops.size(); ++i) {
    const DevOp& d = p.coef_ops[i];
    o << (i ? ", " : "") << "{\"kind\": " << d.kind << ", \"src\": [" << d.src[0] << ", " << d.src[1] << ", " << d.src[2]
      << "], \"value\": [" << d.value[0] << ", " << d.value[1] << ", " << d.value[2] << "]}";
  }
  o << "], \"diags\": [";
  for (size_t i = 0; i < p.diags.size(); ++i) {
    const DiagDesc& D = p.diags[i];
    o << (i ? ", " : "") << "{\"is_reg\": " << D.is_reg << ", \"table\": " << D.table << ", \"terms\": [";
    for (int t = 0; t < D.n_terms; ++t) {
      const DiagTerm& T = p.terms[D.first_term + t];
      o << (t ? ", " : "") << "[" << T.kind << ", " << T.src << ", " << T.value << ", " << T.bit0 << ", " << T.bit1 << "]";
    }
    o << "], \"acc_reg\": [";
    for (int r = 0; r < kR; ++r) o << (r ? ", " : "") << D.acc_reg[r];
    o << "], \"acc_thr\": [";
    for (int r = 0; r < kMaxThreadBits; ++r) o << (r ? ", " : "") << D.acc_thr[r];
    o << "]}";
  }
  o << "], \"exchanges\": [";
  const int G = 32 << p.wb;
  for (size_t i = 0; i < p.exchanges.size(); ++i) {
    const Exchange& X = p.exchanges[i];
    o << (i ? ", " : "") << "{";
    auto tab = [&](const char* name, const uint32_t* v, int cnt, bool first) {
      o << (first ? "" : ", ") << "\"" << name << "\": [";
      for (int k = 0; k < cnt; ++k) o << (k ? ", " : "") << v[k];
      o << "]";
    };
    tab("wbase", p.xbase.data() + X.base, G, true);
    tab("woff", X.woff, 1 << p.kr, false);
    tab("rbase", p.xbase.data() + X.base + G, G, false);
    tab("roff", X.roff, 1 << p.kr, false);
    o << ", \"local\": " << X.pad[0] << ", \"link\": " << X.pad[1] << "}";
  }
  o << "], \"slots\": [";
  for (size_t i = 0; i < p.slots.size(); ++i) {
    const GradSlot& s = p.slots[i];
    o << (i ? ", " : "") << "[" << s.weight << ", " << s.acc << "]";
  }
  o << "]}";
  return o.str();
}

}  // namespace vm
