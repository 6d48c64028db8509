#!/usr/bin/env python
"""Generates ska_sdp_func_python_b200/csrc/fft_codelets.inc: straight-line in-register DFTs of
small sizes (the butterflies of the hand-written line FFT in csrc/fft_lines.cuh).

    python tools/gen_fft_codelets.py

Each codelet is  template <int S> HD void dftR(float2 (&v)[R])  computing, in place and in
natural order,  v[u] <- sum_t v[t] * exp(S * 2 pi i * t * u / R)  (S = +1 / -1).
Composite sizes use Cooley-Tukey on registers with literal twiddle constants (trivial factors 1,
-1, +-i are folded away); the prime sizes 2, 3, 5 use the usual symmetric formulas.  Complex
additions and subtractions go through cadd / csub (fft_inplace.cuh): one packed f32x2 instruction
each on sm_100.
"""
import math
import os

SIZES = [2, 3, 4, 5, 8, 9, 10, 15, 16, 20, 25, 32, 40]


class Gen:
    def __init__(self):
        self.lines = []
        self.n = 0

    def tmp(self):
        self.n += 1
        return f"t{self.n}"

    def emit(self, s):
        self.lines.append("    " + s)

    # complex helpers on names (each name is a float2 variable)
    def add(self, a, b):
        t = self.tmp()
        self.emit(f"const float2 {t} = cadd({a}, {b});")
        return t

    def sub(self, a, b):
        t = self.tmp()
        self.emit(f"const float2 {t} = csub({a}, {b});")
        return t

    def scale(self, a, c):
        t = self.tmp()
        self.emit(f"const float2 {t} = make_float2({c!r}f * {a}.x, {c!r}f * {a}.y);")
        return t

    def fma_real(self, a, c, b):
        """b + c * a (c real constant)"""
        t = self.tmp()
        self.emit(f"const float2 {t} = make_float2(fmaf({c!r}f, {a}.x, {b}.x), fmaf({c!r}f, {a}.y, {b}.y));")
        return t

    def add_si(self, m, n):
        """m + S*i*n"""
        t = self.tmp()
        self.emit(f"const float2 {t} = make_float2({m}.x - S * {n}.y, {m}.y + S * {n}.x);")
        return t

    def sub_si(self, m, n):
        """m - S*i*n"""
        t = self.tmp()
        self.emit(f"const float2 {t} = make_float2({m}.x + S * {n}.y, {m}.y - S * {n}.x);")
        return t

    def mul_tw(self, a, num, den):
        """a * exp(S * 2 pi i * num / den) with trivial cases folded."""
        num %= den
        if num == 0:
            return a
        g = math.gcd(num, den)
        num, den = num // g, den // g
        t = self.tmp()
        if den == 2:      # -1
            self.emit(f"const float2 {t} = make_float2(-{a}.x, -{a}.y);")
        elif den == 4 and num == 1:   # S*i
            self.emit(f"const float2 {t} = make_float2(-S * {a}.y, S * {a}.x);")
        elif den == 4 and num == 3:   # -S*i
            self.emit(f"const float2 {t} = make_float2(S * {a}.y, -S * {a}.x);")
        else:
            c = math.cos(2 * math.pi * num / den)
            s = math.sin(2 * math.pi * num / den)
            self.emit(f"const float2 {t} = make_float2(fmaf({c!r}f, {a}.x, -(S * {s!r}f) * {a}.y), "
                      f"fmaf({c!r}f, {a}.y, (S * {s!r}f) * {a}.x));")
        return t

    def dft(self, x):
        """x: list of names; returns list of names of the DFT (natural order)."""
        R = len(x)
        if R == 1:
            return x
        if R == 2:
            return [self.add(x[0], x[1]), self.sub(x[0], x[1])]
        if R == 3:
            t1, t2 = self.add(x[1], x[2]), self.sub(x[1], x[2])
            o0 = self.add(x[0], t1)
            m = self.fma_real(t1, -0.5, x[0])
            n = self.scale(t2, math.sin(2 * math.pi / 3))
            return [o0, self.add_si(m, n), self.sub_si(m, n)]
        if R == 5:
            t1, t2 = self.add(x[1], x[4]), self.add(x[2], x[3])
            t3, t4 = self.sub(x[1], x[4]), self.sub(x[2], x[3])
            c1, c2 = math.cos(2 * math.pi / 5), math.cos(4 * math.pi / 5)
            s1, s2 = math.sin(2 * math.pi / 5), math.sin(4 * math.pi / 5)
            o0 = self.add(self.add(x[0], t1), t2)
            m1 = self.fma_real(t2, c2, self.fma_real(t1, c1, x[0]))
            m2 = self.fma_real(t2, c1, self.fma_real(t1, c2, x[0]))
            n1 = self.fma_real(t4, s2, self.scale(t3, s1))
            n2 = self.fma_real(t4, -s1, self.scale(t3, s2))
            return [o0, self.add_si(m1, n1), self.add_si(m2, n2), self.sub_si(m2, n2), self.sub_si(m1, n1)]
        # composite: R = P * Q, P the smallest prime-ish factor (4 preferred over 2)
        for P in (4, 2, 3, 5):
            if R % P == 0 and (P != 4 or R % 4 == 0):
                break
        if R == 4:
            P = 2
        Q = R // P
        # X[k1 + P k2] = sum_n2 W_R^(n2 k1) [sum_n1 x[Q n1 + n2] W_P^(n1 k1)] W_Q^(n2 k2)
        A = [self.dft([x[Q * n1 + n2] for n1 in range(P)]) for n2 in range(Q)]   # A[n2][k1]
        out = [None] * R
        for k1 in range(P):
            col = [self.mul_tw(A[n2][k1], n2 * k1, R) for n2 in range(Q)]
            res = self.dft(col)
            for k2 in range(Q):
                out[k1 + P * k2] = res[k2]
        return out


def codelet(R):
    g = Gen()
    x = [f"v[{i}]" for i in range(R)]
    # copy inputs to named temporaries first (the outputs overwrite v)
    ins = []
    for i in range(R):
        t = g.tmp()
        g.emit(f"const float2 {t} = v[{i}];")
        ins.append(t)
    out = g.dft(ins)
    for i in range(R):
        g.emit(f"v[{i}] = {out[i]};")
    head = f"template <int S> B2G_HD void dft{R}(float2 (&v)[{R}])\n{{\n"
    return head + "\n".join(g.lines) + "\n}\n"


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    out = os.path.join(os.path.dirname(here), "ska_sdp_func_python_b200", "csrc", "fft_codelets.inc")
    parts = ["// GENERATED by tools/gen_fft_codelets.py -- do not edit.\n"
             "// In-register DFTs: v[u] <- sum_t v[t] exp(S 2 pi i t u / R), natural order.\n"]
    for R in SIZES:
        parts.append(codelet(R))
    parts.append("template <int R, int S> struct SmallDft;\n")
    for R in SIZES:
        parts.append(f"template <int S> struct SmallDft<{R}, S> {{ B2G_HD static void run(float2 (&v)[{R}]) "
                     f"{{ dft{R}<S>(v); }} }};\n")
    with open(out, "w") as fh:
        fh.write("\n".join(parts))
    print("wrote", out, "sizes", SIZES)


if __name__ == "__main__":
    main()
