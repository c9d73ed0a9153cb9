"""Pure-Python model of the warp-per-board bitboard algorithm in muzero_go_b200/csrc/rules.cu
(lane r = row r; same hfill carry trick, same flood/dilate/compute_invalid/step/score structure).
It exists so the ALGORITHM can be checked against the oracle on the CPU-only build box; the CUDA
code itself is checked on the GPU (tests/test_rules_gpu.py).  Test infrastructure only."""
M32 = 0xFFFFFFFF


def brev(x):
    return int('{:032b}'.format(x & M32)[::-1], 2)


def hfill(g, mask):
    up = ((((mask + g) & M32) ^ mask) & mask) | g
    rm, rg = brev(mask), brev(g)
    dn = brev((((rm + rg) & M32) ^ rm) & rm)
    return up | dn


def shup(x):   # value of lane-1, 0 for lane 0
    return [0] + x[:-1]


def shdn(x):
    return x[1:] + [0]


def dilate4(x, bm):
    u, d = shup(x), shdn(x)
    return [(((x[i] << 1) | (x[i] >> 1) | u[i] | d[i]) & bm[i]) & M32 for i in range(32)]


def flood(seed, mask):
    g = [hfill(seed[i] & mask[i], mask[i]) for i in range(32)]
    while True:
        u, d = shup(g), shdn(g)
        v = [(u[i] | d[i]) & mask[i] & ~g[i] & M32 for i in range(32)]
        if not any(v):
            return g
        g = [hfill(g[i] | v[i], mask[i]) for i in range(32)]


def first_point(x):
    for l in range(32):
        if x[l]:
            return l, x[l] & (-x[l]) & M32
    raise ValueError


def count_upto2(x):
    c = sum(bin(v).count('1') for v in x)
    return min(c, 2)


def seed_at(l, bit):
    s = [0] * 32
    s[l] = bit
    return s


def compute_invalid(own, opp, ko, n):
    bm = [((1 << n) - 1) if i < n else 0 for i in range(32)]
    occ = [own[i] | opp[i] for i in range(32)]
    empt = [~occ[i] & bm[i] & M32 for i in range(32)]
    u = [bm[0]] + occ[:-1]
    d = [occ[i + 1] if i < n - 1 else bm[i] for i in range(31)] + [0]
    e1 = [empt[i] & ((occ[i] << 1) | 1) & ((occ[i] >> 1) | (1 << (n - 1))) & u[i] & d[i] & M32 for i in range(32)]
    if not any(e1):
        return [occ[i] | ko[i] for i in range(32)]
    seeds = [a & b for a, b in zip(dilate4(e1, bm), occ)]
    # one flood for the whole board: stones next to a non-surrounded empty point, grown through their own colour
    opn = dilate4([empt[i] & ~e1[i] & M32 for i in range(32)], bm)
    go = [hfill(own[i] & opn[i], own[i]) for i in range(32)]
    gp = [hfill(opp[i] & opn[i], opp[i]) for i in range(32)]
    rem = [seeds[i] & ~(go[i] | gp[i]) & M32 for i in range(32)]
    while any(rem):
        uo, do, up_, dp = shup(go), shdn(go), shup(gp), shdn(gp)
        vo = [(uo[i] | do[i]) & own[i] & ~go[i] & M32 for i in range(32)]
        vp = [(up_[i] | dp[i]) & opp[i] & ~gp[i] & M32 for i in range(32)]
        if not any(vo) and not any(vp):
            break
        go = [hfill(go[i] | vo[i], own[i]) for i in range(32)]
        gp = [hfill(gp[i] | vp[i], opp[i]) for i in range(32)]
        rem = [rem[i] & ~(go[i] | gp[i]) & M32 for i in range(32)]
    atari, multi = [0] * 32, [go[i] | gp[i] for i in range(32)]
    seeds = rem
    while any(seeds):
        l, bit = first_point(seeds)
        is_own = bool(own[l] & bit)
        grp = flood(seed_at(l, bit), own if is_own else opp)
        libs = count_upto2([a & b for a, b in zip(dilate4(grp, bm), empt)])
        if libs == 1:
            atari = [a | b for a, b in zip(atari, grp)]
        elif libs > 1:
            multi = [a | b for a, b in zip(multi, grp)]
        seeds = [s & ~g & M32 for s, g in zip(seeds, grp)]
    possible = dilate4([(own[i] & multi[i]) | (opp[i] & atari[i]) for i in range(32)], bm)
    definite = dilate4([(own[i] & atari[i]) | (opp[i] & multi[i]) for i in range(32)], bm)
    return [occ[i] | (e1[i] & possible[i] & ~definite[i] & M32) | ko[i] for i in range(32)]


class Board:
    def __init__(self, n):
        self.n = n
        self.bl, self.wh, self.inv = [0] * 32, [0] * 32, [0] * 32
        self.turn = self.passed = self.done = 0

    def step(self, action):
        n = self.n
        np_ = n * n
        if self.done:
            return -2
        if action < 0 or action > np_:
            return -1
        bm = [((1 << n) - 1) if i < n else 0 for i in range(32)]
        own = list(self.wh if self.turn else self.bl)
        opp = list(self.bl if self.turn else self.wh)
        ko = [0] * 32
        if action == np_:
            if self.passed:
                self.done = 1
            self.passed = 1
        else:
            r, c = divmod(action, n)
            if self.inv[r] >> c & 1:
                return -1
            self.passed = 0
            stone = seed_at(r, 1 << c)
            own[r] |= 1 << c
            occ = [a | b for a, b in zip(own, opp)]
            adj = dilate4(stone, bm)
            surrounded = not any(a & ~o & M32 for a, o in zip(adj, opp))
            live = dilate4([~o & b & M32 for o, b in zip(occ, bm)], bm)   # next to an empty point: lives
            cand = [a & o & ~v & M32 for a, o, v in zip(adj, opp, live)]
            kill, nk, single, single_row = [0] * 32, 0, False, [0] * 32
            while any(cand):
                l, bit = first_point(cand)
                grp = flood(seed_at(l, bit), opp)
                libs = [a & ~o & M32 for a, o in zip(dilate4(grp, bm), occ)]
                if not any(libs):
                    nk += 1
                    single = count_upto2(grp) == 1
                    single_row = grp
                    kill = [a | b for a, b in zip(kill, grp)]
                cand = [a & ~g & M32 for a, g in zip(cand, grp)]
            opp = [o & ~k & M32 for o, k in zip(opp, kill)]
            if nk == 1 and single and surrounded:
                ko = single_row
        self.inv = compute_invalid(own, opp, ko, n)
        if self.turn:
            self.wh, self.bl = own, opp
        else:
            self.bl, self.wh = own, opp
        self.turn ^= 1
        return 0

    def score(self):
        n = self.n
        bm = [((1 << n) - 1) if i < n else 0 for i in range(32)]
        empt = [~(self.bl[i] | self.wh[i]) & bm[i] & M32 for i in range(32)]
        ba = sum(bin(v).count('1') for v in self.bl)
        wa = sum(bin(v).count('1') for v in self.wh)
        rem = list(empt)
        while any(rem):
            l, bit = first_point(rem)
            region = flood(seed_at(l, bit), empt)
            nb = dilate4(region, bm)
            tb = any(a & b for a, b in zip(nb, self.bl))
            tw = any(a & b for a, b in zip(nb, self.wh))
            size = sum(bin(v).count('1') for v in region)
            if tb and not tw:
                ba += size
            elif tw and not tb:
                wa += size
            rem = [a & ~g & M32 for a, g in zip(rem, region)]
        return ba, wa

    def planes(self):
        import numpy as np
        n = self.n
        st = np.zeros((6, n, n), dtype=np.uint8)
        for r in range(n):
            for c in range(n):
                st[0, r, c] = self.bl[r] >> c & 1
                st[1, r, c] = self.wh[r] >> c & 1
                st[3, r, c] = self.inv[r] >> c & 1
        st[2], st[4], st[5] = self.turn, self.passed, self.done
        return st
