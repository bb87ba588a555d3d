"""numpy emulator of the register-resident VM's micro-op program (csrc/qnn_vm.h), for CPU tests of the
host compiler: layout tracking, merged diagonals, GF(2) exchanges.  Test infrastructure only.

One warp is emulated: 32 lanes x 32 register amplitudes, 32 / 2^tb samples.  Exchanges go through an
emulated 8 KB buffer with the program's XOR addressing, and every access is checked for shared-memory
bank conflicts (64-bit accesses: the 16 lanes of a half-warp must hit 16 distinct 8-byte slots mod 16).
"""
import numpy as np

SRC_CONST = -(2 ** 31)
RX, RY, RZ, U3, CNOT, CZ, CU3 = range(7)
(MOP_END, MOP_G1_REAL, MOP_G1_REAL_X, MOP_G1_GEN, MOP_CG1_L, MOP_CG1_P, MOP_DIAG, MOP_DIAG_REG, MOP_XCHG, MOP_ROT5, MOP_ROTX5, MOP_CU3_CHAIN) = range(12)


def _angle(src, value, w):
    return value if src == SRC_CONST else w[src]


def _matrix(kind, a):
    if kind == RY:
        c, s = np.cos(a[0] / 2), np.sin(a[0] / 2)
        return np.array([[c, -s], [s, c]], complex)
    if kind == RX:
        c, s = np.cos(a[0] / 2), np.sin(a[0] / 2)
        return np.array([[c, -1j * s], [-1j * s, c]], complex)
    t, p, l = a
    c, s = np.cos(t / 2), np.sin(t / 2)
    return np.array([[c, -np.exp(1j * l) * s], [np.exp(1j * p) * s, np.exp(1j * (p + l)) * c]], complex)


def _apply_reg(state, U, rbit, lane_mask=None, ctrl_rbit=None):
    """state[lane, j]; apply U on register bit rbit for the selected lanes / control register bit."""
    j = np.arange(state.shape[1])
    j0 = j[(j >> rbit) & 1 == 0]
    if ctrl_rbit is not None:
        j0 = j0[(j0 >> ctrl_rbit) & 1 == 1]
    j1 = j0 | (1 << rbit)
    lanes = np.arange(state.shape[0]) if lane_mask is None else np.nonzero(lane_mask)[0]
    a0 = state[np.ix_(lanes, j0)].copy()
    a1 = state[np.ix_(lanes, j1)].copy()
    state[np.ix_(lanes, j0)] = U[0, 0] * a0 + U[0, 1] * a1
    state[np.ix_(lanes, j1)] = U[1, 0] * a0 + U[1, 1] * a1


def check_conflict_free(base, off, writing=True, local=False):
    NA = len(off)
    if len(base) > 32 * 8 and not local and NA == 32:              # cluster geometry: L2 buffer; writes must be coalesced
        if writing:
            for j in range(0, 32, 5):
                for w0 in range(0, len(base), 32):
                    blocks = {(base[l] ^ off[j]) // 256 for l in range(w0, w0 + 32)}
                    assert len(blocks) == 1, "a warp's exchange write is not one aligned 256-byte block"
        return
    for j in range(NA):
        for half in range(0, len(base), 16):          # every half-warp of every warp of the group
            slots = {((base[l] ^ off[j]) // 8) % 16 for l in range(half, half + 16)}
            assert len(slots) == 16, "shared-memory bank conflict in an exchange"


def dsmem_exchange(state, X, G, stats):
    """Exchange across the CTAs of a cluster through distributed shared memory: a slot's byte offset holds the
    destination CTA's rank in bits 16.. and the offset in that CTA's 64 KB buffer below.  Every thread must READ
    only its own CTA's buffer, without bank conflicts; a writing warp's 32 stores should fill eight whole 32-byte sectors
    (one 256-byte run, or a few runs in different CTAs when a lane bit's wire moves to a CTA-rank bit): counted."""
    n_cta = G // 256
    buf = {}
    runs = 0
    for jj in range(32):
        for w0 in range(0, G, 32):
            addrs = [X["wbase"][l] ^ X["woff"][jj] for l in range(w0, w0 + 32)]
            runs += len({a // 32 for a in addrs}) == 8      # eight full 32-byte sectors: nothing wasted
            for l, addr in zip(range(w0, w0 + 32), addrs):
                assert addr % 8 == 0 and 0 <= (addr >> 16) < n_cta and addr not in buf
                buf[addr] = state[l, jj]
    new = np.empty_like(state)
    for l in range(G):
        for jj in range(32):
            addr = X["rbase"][l] ^ X["roff"][jj]
            assert addr >> 16 == l // 256, "a thread reads another CTA's buffer"
            new[l, jj] = buf[addr]
    for jj in range(32):
        for half in range(0, G, 16):
            slots = {(((X["rbase"][l] ^ X["roff"][jj]) & 0xFFFF) // 8) % 16 for l in range(half, half + 16)}
            assert len(slots) == 16, "shared-memory bank conflict on the reading side of a DSMEM exchange"
    if stats is not None:
        stats["dsmem_sector_exact_warp_stores"] = stats.get("dsmem_sector_exact_warp_stores", 0) + runs
        stats["dsmem_warp_stores"] = stats.get("dsmem_warp_stores", 0) + 32 * (G // 32)
    return new


def emulate_forward(desc, x, w):
    """x: (samples_per_warp, n_inputs); returns y (samples_per_warp, n)."""
    assert desc["ok"]
    n, tb, wb = desc["n"], desc["tb"], desc.get("wb", 0)
    kr = desc.get("kr", 5)               # register bits: 2^kr amplitudes per thread
    NA = 1 << kr
    G = 32 << wb                          # threads per group (one warp for n <= 10)
    T = 1 << tb                           # threads per sample
    spw = G // T if wb == 0 else 1
    assert x.shape[0] == spw
    state = np.zeros((G, NA), complex)
    if desc.get("amplitude"):
        # x row (first 2^n entries, zero padded), normalised, gathered into the initial layout
        ibw = desc["initial_bit_of_wire"]
        for s in range(spw):
            row = np.zeros(1 << n)
            kk = min(1 << n, x.shape[1])
            row[:kk] = x[s, :kk]
            row /= np.linalg.norm(row)
            for t in range(T):
                for jj in range(NA):
                    p = jj | (t << kr)
                    k = sum(((p >> ibw[w]) & 1) << (n - 1 - w) for w in range(n))
                    state[s * T + t, jj] = row[k]
    else:
        for s in range(spw):
            state[s * T, 0] = 1.0
    lane = np.arange(G)
    lis = lane % T                       # thread in sample
    samp = lane // T
    stats = {"xchg": 0, "diag": 0, "diag_reg": 0, "g1": 0, "cg_l": 0, "cg_p": 0}
    for kind, a, b, idx, acc, aux in desc["ops"]:
        if kind == MOP_END:
            break
        if kind == MOP_G1_REAL or kind == MOP_G1_GEN:
            co = desc["coef_ops"][idx]
            U = _matrix(co["kind"], [_angle(s_, v_, w) for s_, v_ in zip(co["src"], co["value"])])
            _apply_reg(state, U, a)
            stats["g1"] += 1
        elif kind == MOP_G1_REAL_X:
            for s in range(spw):
                _apply_reg(state, _matrix(RY, [x[s, idx]]), a, lane_mask=(samp == s))
            stats["g1"] += 1
        elif kind in (MOP_CG1_L, MOP_CG1_P):
            co = desc["coef_ops"][idx]
            ang = [_angle(s_, v_, w) for s_, v_ in zip(co["src"], co["value"])]
            if not (b & 8):                      # post-phase e^{i phi} deferred into a later diagonal
                ang[1] = 0.0
            U = _matrix(U3, ang)
            if kind == MOP_CG1_L:
                _apply_reg(state, U, a, ctrl_rbit=b & 7)
                stats["cg_l"] += 1
            else:
                par = np.array([bin(int(v) & aux).count("1") & 1 for v in lis], bool)
                _apply_reg(state, U, a, lane_mask=par)
                stats["cg_p"] += 1
        elif kind == MOP_ROT5:                    # rotations on all kr register bits
            for k in range(kr):
                co = desc["coef_ops"][idx + ((aux >> (3 * k)) & 7)]
                _apply_reg(state, _matrix(co["kind"], [_angle(s_, v_, w) for s_, v_ in zip(co["src"], co["value"])]), k)
            stats["g1"] += kr
        elif kind == MOP_ROTX5:
            for k in range(kr):
                col = (idx >> (5 * k)) & 31
                for s in range(spw):
                    _apply_reg(state, _matrix(RY, [x[s, col]]), k, lane_mask=(samp == s))
            stats["g1"] += kr
        elif kind == MOP_CU3_CHAIN:
            for k in range(kr - 1):
                co = desc["coef_ops"][idx + k]
                ang = [_angle(s_, v_, w) for s_, v_ in zip(co["src"], co["value"])]
                if not ((aux >> k) & 1):
                    ang[1] = 0.0
                _apply_reg(state, _matrix(U3, ang), k + 1, ctrl_rbit=k)
            stats["cg_l"] += kr - 1
        elif kind in (MOP_DIAG, MOP_DIAG_REG):
            D = desc["diags"][aux]
            assert bool(D["is_reg"]) == (kind == MOP_DIAG_REG)
            j = np.arange(NA)
            p = j[None, :] | (lis[:, None] << kr)        # physical index of (lane, j) within its sample
            ang = np.zeros((G, NA))
            for tkind, src, value, bit0, bit1 in D["terms"]:
                if tkind == 0:
                    z = 1 - 2 * ((p >> bit0) & 1)
                    ang += -0.5 * _angle(src, value, w) * z
                elif tkind == 1:
                    ang += np.pi * (((p >> bit0) & 1) & ((p >> bit1) & 1))
                else:                              # deferred CU3 post-phase: e^{i phi} where both bits are 1
                    ang += _angle(src, value, w) * (((p >> bit0) & 1) & ((p >> bit1) & 1))
                if kind == MOP_DIAG_REG:
                    assert bit0 < kr and bit1 < kr
            state *= np.exp(1j * ang)
            stats["diag" if kind == MOP_DIAG else "diag_reg"] += 1
        elif kind == MOP_XCHG:
            X = desc["exchanges"][idx]
            mode = int(X.get("local", 0))   # cluster geometry: 1 own CTA's buffer, 2 distributed shared memory, 0 L2 / plain
            assert mode == aux and (mode == 0 or G > 256)
            if mode == 2:
                new = dsmem_exchange(state, X, G, stats)
                # the backward pass runs the inverse permutation with the maps of the linked entry
                Xb = desc["exchanges"][X["link"]]
                assert Xb["local"] == 3 and Xb["link"] == idx
                back = dsmem_exchange(new, Xb, G, None)
                assert np.array_equal(back, state), "backward-direction maps do not invert the exchange"
                state = new
                stats["xchg"] += 1
                continue
            local = mode == 1
            check_conflict_free(X["wbase"], X["woff"], local=local)
            check_conflict_free(X["rbase"], X["roff"], writing=False, local=local)
            size = 65536 if local else 256 * NA * (G // 32)
            cta = (lambda l: l // 256) if local else (lambda l: 0)
            buf = {}
            for l in range(G):
                for jj in range(NA):
                    addr = X["wbase"][l] ^ X["woff"][jj]
                    assert addr % 8 == 0 and 0 <= addr < size and (cta(l), addr) not in buf
                    buf[(cta(l), addr)] = state[l, jj]
            new = np.empty_like(state)
            for l in range(G):
                for jj in range(NA):
                    new[l, jj] = buf[(cta(l), X["rbase"][l] ^ X["roff"][jj])]
            state = new
            stats["xchg"] += 1
        else:
            raise ValueError(f"unknown micro-op {kind}")
    prob = np.abs(state) ** 2
    j = np.arange(NA)
    p = j[None, :] | (lis[:, None] << kr)
    y = np.zeros((spw, n))
    for wire, bit in enumerate(desc["final_bit_of_wire"]):
        z = 1 - 2 * ((p >> bit) & 1)
        for s in range(spw):
            y[s, wire] = (prob * z)[samp == s].sum()
    return y, stats
