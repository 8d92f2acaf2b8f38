"""Pure-Python model of the index arithmetic of csrc/ntt_flat.cu and ntt_small.cu.

It mirrors, line for line, which elements a thread holds in each pass, which table
entry each butterfly reads and in which order the stages run, with Python integers in
place of registers.  It lets the CPU-only test-suite check the *network* (pass
structure, twiddle indexing, swizzle bijectivity, lazy-bound schedule) against the
oracle without a GPU.  It is a model of the kernels, not an implementation anyone ships.
"""
from __future__ import annotations

K_FWD_IN_BOUND = 2


def fwd_xtarget(B, s, logn):
    if s == logn - 1:
        return 2
    return 8 if B + 4 > 16 else B


def fwd_bound_before(s, logn):
    B = K_FWD_IN_BOUND
    for i in range(s):
        B = fwd_xtarget(B, i, logn) + 4
    return B


def deposit3(u, P):
    return ((u >> P) << (P + 3)) | (u & ((1 << P) - 1))


def swz(e):
    chunk = e >> 1
    return ((chunk ^ ((chunk >> 3) & 7)) << 1) | (e & 1)


class Net:
    """Holds tables of one limb: tw/itw lists of (w) in index order tw[m+i]; exact ints."""

    def __init__(self, logn, q, tw, itw, ninv, w1n):
        self.logn, self.q, self.tw, self.itw, self.ninv, self.w1n = logn, q, tw, itw, ninv, w1n
        self.max_bound = 0

    # -- butterflies with explicit lazy ranges, asserting the bound schedule ------------
    def ct(self, x, y, w, s):
        q = self.q
        B = fwd_bound_before(s, self.logn)
        T = min(fwd_xtarget(B, s, self.logn), B)
        assert x < B * q and y < 2 ** 64, (x, B)
        for k in (8, 4, 2, 1):
            if B > k and T <= k:
                if x >= k * q:
                    x -= k * q
        assert x < T * q
        # worst case of the lazy product: [0,4q) with the approximate quotient, [0,2q) in the last stage
        k = 2 if s == self.logn - 1 else 4
        t = (y * w) % q + (k - 1) * q
        yo = x + k * q - ((y * w) % q)
        xo = x + t
        assert xo < 2 ** 64 and yo < 2 ** 64
        self.max_bound = max(self.max_bound, -(-max(xo, yo) // q))
        return xo, yo

    def gs(self, x, y, w):
        q = self.q
        assert x < 4 * q and y < 4 * q
        t = x + 4 * q - y
        s = x + y
        if s >= 4 * q:
            s -= 4 * q
        return s, (t * w) % q + 3 * q          # worst case of the approximate-quotient product

    def fwd_pass(self, v, g0, GB, ACT):
        L = self.logn
        for tb in (2, 1, 0):
            if not (ACT >> tb) & 1:
                continue
            b = GB + tb
            s = L - 1 - b
            base = (1 << s) + (g0 >> (b + 1))
            for jh in range(4 >> tb):
                w = self.tw[base + jh]
                for jl in range(1 << tb):
                    jx = (jh << (tb + 1)) | jl
                    jy = jx | (1 << tb)
                    v[jx], v[jy] = self.ct(v[jx], v[jy], w, s)

    def inv_pass(self, v, g0, GB, ACT):
        L, q = self.logn, self.q
        for tb in (0, 1, 2):
            if not (ACT >> tb) & 1:
                continue
            b = GB + tb
            s = L - 1 - b
            base = (1 << s) + (g0 >> (b + 1))
            for jh in range(4 >> tb):
                w = self.w1n if b == L - 1 else self.itw[base + jh]
                for jl in range(1 << tb):
                    jx = (jh << (tb + 1)) | jl
                    jy = jx | (1 << tb)
                    v[jx], v[jy] = self.gs(v[jx], v[jy], w)
                    if b == L - 1:
                        v[jx] = (v[jx] * self.ninv) % q
                        v[jy] %= q


# ---- flat two-kernel transform ------------------------------------------------------
def col_passes(logn, fwd):
    LOGR = logn - 8
    if fwd:
        P0 = LOGR - 3
        out = [(P0, 7)]
        if P0 > 0:
            P1 = P0 - 3 if P0 > 3 else 0
            out.append((P1, (1 << (P0 - P1)) - 1))
            if P1 > 0:
                P2 = P1 - 3 if P1 > 3 else 0
                assert P2 == 0
                out.append((P2, (1 << (P1 - P2)) - 1))
        return out
    PL = LOGR - 3
    out = [(0, 7)]
    if PL > 0:
        P1 = PL if PL < 3 else 3
        out.append((P1, 7 & ~((1 << (3 - P1)) - 1)))
        if PL > 3:
            P2 = PL if PL < 6 else 6
            out.append((P2, 7 & ~((1 << (P1 + 3 - P2)) - 1)))
    return out


def cols_kernel(net, a, fwd):
    logn = net.logn
    LOGR = logn - 8
    R = 1 << LOGR
    for col in range(0, 256, 37):        # the model samples columns; all columns share code
        pass
    for col in range(256):
        for (P, ACT) in col_passes(logn, fwd):
            for u in range(R // 8):
                row0 = deposit3(u, P)
                idx = [((row0 + (j << P)) << 8) + col for j in range(8)]
                v = [a[i] for i in idx]
                if fwd:
                    net.fwd_pass(v, row0 << 8, 8 + P, ACT)
                else:
                    net.inv_pass(v, row0 << 8, 8 + P, ACT)
                for i, x in zip(idx, v):
                    a[i] = x


def rows_kernel(net, a, fwd):
    logn, q = net.logn, net.q
    R = 1 << (logn - 8)
    for row in range(R):
        gbase = row << 8
        if fwd:
            plan = [(5, 7, lambda lane: lane, 5), (2, 7, lambda lane: ((lane >> 2) << 5) | (lane & 3), 2),
                    (0, 3, lambda lane: lane << 3, 0)]
        else:
            plan = [(0, 7, lambda lane: lane << 3, 0), (3, 7, lambda lane: ((lane >> 3) << 6) | (lane & 7), 3),
                    (5, 6, lambda lane: lane, 5)]
        for (P, ACT, cfun, sh) in plan:
            for lane in range(32):
                c0 = cfun(lane)
                idx = [gbase + c0 + (j << sh) for j in range(8)]
                v = [a[i] for i in idx]
                if fwd:
                    net.fwd_pass(v, gbase | c0, P, ACT)
                else:
                    net.inv_pass(v, gbase | c0, P, ACT)
                for i, x in zip(idx, v):
                    a[i] = x
        if fwd:
            for i in range(gbase, gbase + 256):
                assert a[i] < 4 * q
                a[i] %= q


def flat_forward(net, a):
    a = list(a)
    cols_kernel(net, a, True)
    rows_kernel(net, a, True)
    return a


def flat_inverse(net, a):
    a = list(a)
    rows_kernel(net, a, False)
    cols_kernel(net, a, False)
    return a


# ---- small transform -------------------------------------------------------------------
def bitrev(x, bits):
    r = 0
    for _ in range(bits):
        r = (r << 1) | (x & 1)
        x >>= 1
    return r


def small_forward(net, a, natural):
    logn, q = net.logn, net.q
    n = 1 << logn
    a = list(a)
    P0 = logn - 3
    passes = [(P0, 7)]
    if P0 > 0:
        P1 = P0 - 3 if P0 > 3 else 0
        passes.append((P1, (1 << (P0 - P1)) - 1))
        if P1 > 0:
            passes.append((0, (1 << P1) - 1))
    for (P, ACT) in passes:
        for u in range(n // 8):
            e0 = deposit3(u, P)
            idx = [e0 + (j << P) for j in range(8)]
            v = [a[i] for i in idx]
            net.fwd_pass(v, e0, P, ACT)
            for i, x in zip(idx, v):
                a[i] = x
    a = [x % q for x in a]
    if natural:
        out = [0] * n
        for pos in range(n):
            out[bitrev(pos, logn)] = a[pos]
        a = out
    return a


def small_inverse(net, a, natural):
    logn = net.logn
    n = 1 << logn
    a = list(a)
    if natural:
        t = [0] * n
        for k in range(n):
            t[bitrev(k, logn)] = a[k]
        a = t
    PL = logn - 3
    passes = [(0, 7)]
    if PL > 0:
        P1 = PL if PL < 3 else 3
        passes.append((P1, 7 & ~((1 << (3 - P1)) - 1)))
        if PL > 3:
            P2 = PL if PL < 6 else 6
            passes.append((P2, 7 & ~((1 << (P1 + 3 - P2)) - 1)))
    for (P, ACT) in passes:
        for u in range(n // 8):
            e0 = deposit3(u, P)
            idx = [e0 + (j << P) for j in range(8)]
            v = [a[i] for i in idx]
            net.inv_pass(v, e0, P, ACT)
            for i, x in zip(idx, v):
                a[i] = x
    return a


# ---- table builder (mirror of numth_host.hpp build_tree_tables) -------------------------
def tree_tables(logn, q, psi, M, c0):
    n = 1 << logn
    tw = [1] * n
    itw = [1] * n
    cexp = [0] * (2 * n)
    cexp[1] = c0 % M
    ipsi = pow(psi, q - 2, q)
    for h in range(1, n):
        assert cexp[h] % 2 == 0
        e = cexp[h] // 2
        tw[h] = pow(psi, e, q)
        itw[h] = pow(ipsi, e, q)
        cexp[2 * h] = e
        cexp[2 * h + 1] = (e + M // 2) % M
    ninv = pow(n, q - 2, q)
    return tw, itw, ninv, (itw[1] * ninv) % q
