"""CPU oracle, scheme level -- TEST INFRASTRUCTURE ONLY.

An independent restatement (Python orchestration over the C oracle primitives) of the CKKS scheme operations the
Go shim forwards to Lattigo: key generation, encode/decode, encrypt/decrypt, the evaluator operations of
orion/backend/lattigo/evaluator.go:48-317, and the hoisted BSGS linear transform of lineartransform.go:96-113.
With the same seed and the same order of sampling events it reproduces the device library's keys and ciphertexts
bit for bit.  "parity unpinned" versus Lattigo itself (see ckks_oracle.h).
"""
import ctypes
import math

import numpy as np

from . import Ctx, P, f64p, i64p, ints, lib

LD = np.longdouble


class OCt:
    def __init__(self, data, level, scale, log_slots, degree=1):
        self.data = np.ascontiguousarray(data, dtype=np.uint64)   # [degree+1][level+1][N]
        self.level, self.scale, self.log_slots, self.degree = level, LD(scale), log_slots, degree

    def copy(self):
        return OCt(self.data.copy(), self.level, self.scale, self.log_slots, self.degree)


class OPt:
    def __init__(self, data, level, scale, log_slots):
        self.data = np.ascontiguousarray(data, dtype=np.uint64)   # [level+1][N]
        self.level, self.scale, self.log_slots = level, LD(scale), log_slots


class OScheme:
    def __init__(self, logN, logQ, logP, log_scale, h, seed, moduli=None):
        """moduli = (Q, P): explicit prime chains instead of bit sizes (the bootstrapping scheme extends the residual chain)."""
        from . import gen_moduli
        self.L = lib()
        self.logN, self.N = logN, 1 << logN
        self.Q, self.Pm = moduli if moduli is not None else gen_moduli(logQ, logP, logN + 1)
        self.Q, self.Pm = [int(q) for q in self.Q], [int(q) for q in self.Pm]
        self.nQ, self.nP, self.nmod = len(self.Q), len(self.Pm), len(self.Q) + len(self.Pm)
        self.ctx = Ctx(logN, self.Q + self.Pm, self.nQ, self.nP)
        self.mod = self.Q + self.Pm
        self.h, self.seed, self.next_event = h, seed, 0
        self.max_level, self.log_max_slots = self.nQ - 1, logN - 1
        self.default_scale = LD(2) ** log_scale
        self.sk_coeffs = None
        self.sk = self.pk = self.rlk = None
        self.galois_keys = {}

    # ---- helpers ----
    def new_event(self):
        e = self.next_event * 16384
        self.next_event += 1
        return e

    def qp_mods(self, level, with_p=True):
        return list(range(level + 1)) + ([self.nQ + k for k in range(self.nP)] if with_p else [])

    def _ntt(self, a, mods, inverse=False):
        return self.ctx.ntt(a, mods, inverse)

    def _ew(self, op, a, b, mods):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(len(mods), self.N)
        out = np.empty_like(a)
        if b is not None:
            b = np.ascontiguousarray(b, dtype=np.uint64).reshape(len(mods), self.N)
        for i, m in enumerate(mods):
            if op == "neg":
                self.L.orc_neg(self.ctx.h, m, P(a[i]), P(out[i]))
            else:
                getattr(self.L, "orc_" + op)(self.ctx.h, m, P(a[i]), P(b[i]), P(out[i]))
        return out

    def _mul_scalar(self, a, scalars, mods):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(len(mods), self.N)
        out = np.empty_like(a)
        for i, m in enumerate(mods):
            self.L.orc_mul_scalar(self.ctx.h, m, P(a[i]), int(scalars[i]), P(out[i]))
        return out

    def _add_scalar(self, a, scalars, mods):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(len(mods), self.N)
        out = np.empty_like(a)
        for i, m in enumerate(mods):
            self.L.orc_add_scalar(self.ctx.h, m, P(a[i]), int(scalars[i]), P(out[i]))
        return out

    def _uniform(self, mods, stream0):
        out = np.empty((len(mods), self.N), dtype=np.uint64)
        for t, m in enumerate(mods):
            self.L.orc_sample_uniform(self.ctx.h, m, self.seed, stream0 + t, P(out[t]))
        return out

    def _gaussian_ntt(self, mods, stream):
        small = np.empty(self.N, dtype=np.int64)
        self.L.orc_sample_gaussian_i64(self.N, self.seed, stream, P(small, i64p))
        return self._ntt(self._small_to_rns(small, mods), mods)

    def _small_to_rns(self, small, mods):
        out = np.empty((len(mods), self.N), dtype=np.uint64)
        small = np.ascontiguousarray(small, dtype=np.int64)
        for t, m in enumerate(mods):
            self.L.orc_small_to_rns(self.ctx.h, m, P(small, i64p), P(out[t]))
        return out

    def _ternary(self, stream):
        out = np.empty(self.N, dtype=np.int64)
        self.L.orc_sample_ternary_hw(self.N, self.h, self.seed, stream, P(out, i64p))
        return out

    def _const_rns(self, value, cscale, mods):
        out = np.empty(len(mods), dtype=np.uint64)
        cs = LD(cscale)
        hi = float(cs)
        self.L.orc_scaled_const_rns(self.ctx.h, float(value), hi, float(cs - LD(hi)), ints(mods), len(mods), P(out))
        return out

    def _pmodq(self, level):
        out = []
        for i in range(level + 1):
            p = 1
            for k in range(self.nP):
                p = (p * (self.Pm[k] % self.Q[i])) % self.Q[i]
            out.append(p)
        return out

    # ---- keys ----
    def set_secret_key(self, coeffs):
        self.sk_coeffs = np.array(coeffs, dtype=np.int64)
        full = list(range(self.nmod))
        self.sk = self._ntt(self._small_to_rns(self.sk_coeffs, full), full)

    def gen_secret_key(self):
        self.set_secret_key(self._ternary(self.new_event()))

    def gen_public_key(self):
        ev = self.new_event()
        full = list(range(self.nmod))
        a = self._uniform(full, ev)
        e = self._gaussian_ntt(full, ev + 512)
        b = self._ew("sub", e, self._ew("mul", a, self.sk, full), full)
        self.pk = np.stack([b, a])

    def gen_eval_key(self, sk_in, sk_out):
        ev = self.new_event()
        full = list(range(self.nmod))
        beta = self.ctx.num_digits(self.max_level)
        key = np.empty((beta, 2, self.nmod, self.N), dtype=np.uint64)
        pm = self._pmodq(self.max_level)
        for j in range(beta):
            a = self._uniform(full, ev + j * 128)
            e = self._gaussian_ntt(full, ev + j * 128 + 64)
            b = self._ew("sub", e, self._ew("mul", a, sk_out, full), full)
            lo, hi = j * self.nP, min((j + 1) * self.nP, self.nQ)
            own = list(range(lo, hi))
            term = self._mul_scalar(sk_in[lo:hi], [pm[i] for i in own], own)
            b[lo:hi] = self._ew("add", b[lo:hi], term, own)
            key[j, 0], key[j, 1] = b, a
        return key

    def gen_relin_key(self):
        full = list(range(self.nmod))
        self.rlk = self.gen_eval_key(self._ew("mul", self.sk, self.sk, full), self.sk)

    def galois_key(self, gal_el, generate=True):
        if gal_el not in self.galois_keys:
            if not generate:
                raise KeyError("oracle: Galois key %d was not generated" % gal_el)
            inv = pow(gal_el, -1, 2 * self.N)
            s_out = self.ctx.automorphism_ntt(self.sk, inv)
            self.galois_keys[gal_el] = self.gen_eval_key(self.sk, s_out)
        return self.galois_keys[gal_el]

    # ---- encoder ----
    def encode_coeffs(self, values, scale, mods, log_slots=None):
        """complex/real slot vector -> NTT-domain residues over `mods`."""
        log_slots = self.log_max_slots if log_slots is None else log_slots
        slots = 1 << log_slots
        v = np.zeros(slots, dtype=np.complex128)
        vals = np.asarray(values)
        v[: len(vals)] = vals
        re, im = np.ascontiguousarray(v.real), np.ascontiguousarray(v.imag)
        self.L.orc_special_ifft(P(re, f64p), P(im, f64p), log_slots, 2 * self.N)
        out = np.empty((len(mods), self.N), dtype=np.uint64)
        self.L.orc_encode_round(self.ctx.h, P(re, f64p), P(im, f64p), log_slots, float(LD(scale)), ints(mods), len(mods), P(out))
        return self._ntt(out, mods)

    def encode(self, values, level, scale, log_slots=None):
        log_slots = self.log_max_slots if log_slots is None else log_slots
        return OPt(self.encode_coeffs(values, scale, list(range(level + 1)), log_slots), level, scale, log_slots)

    def decode(self, pt):
        mods = list(range(pt.level + 1))
        coef = self._ntt(pt.data, mods, inverse=True)
        slots = 1 << pt.log_slots
        re, im = np.empty(slots), np.empty(slots)
        self.L.orc_decode_crt(self.ctx.h, P(coef), pt.level, float(pt.scale), pt.log_slots, P(re, f64p), P(im, f64p))
        self.L.orc_special_fft(P(re, f64p), P(im, f64p), pt.log_slots, 2 * self.N)
        return re + 1j * im

    # ---- encryptor ----
    def encrypt(self, pt):
        level, L = pt.level, pt.level + 1
        ev = self.new_event()
        mods = self.qp_mods(level)
        u = self._ntt(self._small_to_rns(self._ternary(ev), mods), mods)
        out = []
        for c in range(2):
            e = self._gaussian_ntt(mods, ev + 1 + c)
            pkc = self.pk[c][mods]
            acc = self._ew("add", self._ew("mul", u, pkc, mods), e, mods)
            out.append(self.ctx.moddown(level, acc) if self.nP else acc)
        out[0] = self._ew("add", out[0], pt.data, list(range(L)))
        return OCt(np.stack(out), level, pt.scale, pt.log_slots)

    def decrypt(self, ct):
        mods = list(range(ct.level + 1))
        m = self._ew("add", self._ew("mul", ct.data[1], self.sk[: ct.level + 1], mods), ct.data[0], mods)
        return OPt(m, ct.level, ct.scale, ct.log_slots)

    # ---- evaluator (orion/backend/lattigo/evaluator.go:48-317 semantics) ----
    @staticmethod
    def _close(a, b):
        r = LD(a) / LD(b)
        return LD(1) - LD(1e-9) < r < LD(1) + LD(1e-9)

    def _addsub(self, a, b_data, b_level, b_scale, sub):
        level = min(a.level, b_level)
        L = level + 1
        mods = list(range(L))
        A = [a.data[p][:L] for p in range(a.degree + 1)]
        B = [b_data[p][:L] for p in range(b_data.shape[0])]
        sa, sb = a.scale, LD(b_scale)
        if not self._close(sa, sb):
            a_small = sa < sb
            ratio = np.floor(sb / sa if a_small else sa / sb)
            if ratio > 1:
                sc = self._const_rns(1.0, ratio, mods)
                if a_small:
                    A = [self._mul_scalar(x, sc, mods) for x in A]
                else:
                    B = [self._mul_scalar(x, sc, mods) for x in B]
        out = []
        for p in range(max(len(A), len(B))):
            if p < len(A) and p < len(B):
                out.append(self._ew("sub" if sub else "add", A[p], B[p], mods))
            elif p < len(A):
                out.append(np.array(A[p]))
            else:
                out.append(self._ew("neg", B[p], None, mods) if sub else np.array(B[p]))
        return OCt(np.stack(out), level, max(sa, sb), a.log_slots, len(out) - 1)

    def add(self, a, b, sub=False):
        if isinstance(b, OPt):
            return self._addsub(a, b.data[None], b.level, b.scale, sub)
        return self._addsub(a, b.data, b.level, b.scale, sub)

    def mul_pt(self, a, pt):
        level = min(a.level, pt.level)
        mods = list(range(level + 1))
        out = [self._ew("mul", a.data[p][: level + 1], pt.data[: level + 1], mods) for p in range(a.degree + 1)]
        return OCt(np.stack(out), level, a.scale * pt.scale, a.log_slots, a.degree)

    def add_const(self, a, c):
        mods = list(range(a.level + 1))
        sc = self._const_rns(c, a.scale, mods)
        out = a.copy()
        out.data[0] = self._add_scalar(a.data[0], sc, mods)
        return out

    def mul_const_scaled(self, a, c, cscale):
        mods = list(range(a.level + 1))
        sc = self._const_rns(c, LD(cscale), mods)
        out = [self._mul_scalar(a.data[p], sc, mods) for p in range(a.degree + 1)]
        return OCt(np.stack(out), a.level, a.scale * LD(cscale), a.log_slots, a.degree)

    def mul_int(self, a, k):
        return self.mul_const_scaled(a, float(k), 1)

    def mul_float(self, a, c):
        if c == math.floor(c) and abs(c) < 9e15:
            return self.mul_int(a, int(c))
        return self.mul_const_scaled(a, c, LD(self.Q[a.level]))

    def keyswitch(self, level, c1, evk):
        return self.ctx.keyswitch(level, c1, evk)

    def mul_relin(self, a, b, relin=True):
        level = min(a.level, b.level)
        L = level + 1
        mods = list(range(L))
        a0, a1, b0, b1 = a.data[0][:L], a.data[1][:L], b.data[0][:L], b.data[1][:L]
        d0 = self._ew("mul", a0, b0, mods)
        d1 = self._ew("add", self._ew("mul", a0, b1, mods), self._ew("mul", a1, b0, mods), mods)
        d2 = self._ew("mul", a1, b1, mods)
        if not relin:
            return OCt(np.stack([d0, d1, d2]), level, a.scale * b.scale, a.log_slots, 2)
        ks = self.keyswitch(level, d2, self.rlk)
        return OCt(np.stack([self._ew("add", ks[0], d0, mods), self._ew("add", ks[1], d1, mods)]), level,
                   a.scale * b.scale, a.log_slots)

    def rescale(self, a):
        out = [self.ctx.rescale(a.level, a.data[p]) for p in range(a.degree + 1)]
        return OCt(np.stack(out), a.level - 1, a.scale / LD(self.Q[a.level]), a.log_slots, a.degree)

    def automorphism(self, a, gal_el, generate=True):
        mods = list(range(a.level + 1))
        ks = self.keyswitch(a.level, a.data[1], self.galois_key(gal_el, generate))
        c0 = self._ew("add", ks[0], a.data[0], mods)
        return OCt(np.stack([self.ctx.automorphism_ntt(c0, gal_el), self.ctx.automorphism_ntt(ks[1], gal_el)]),
                   a.level, a.scale, a.log_slots)

    def rotate(self, a, rot, generate=True):
        g = self.ctx.galois_element(rot)
        return a.copy() if g == 1 else self.automorphism(a, g, generate)

    def conjugate(self, a, generate=True):
        return self.automorphism(a, 2 * self.N - 1, generate)

    def drop_level(self, a, level):
        return OCt(a.data[:, : level + 1].copy(), level, a.scale, a.log_slots, a.degree)

    # ---- linear transforms (orion/backend/lattigo/lineartransform.go:37-113) ----
    @staticmethod
    def bsgs_index(diags, slots, n1):
        index, giants, babies = {}, set(), set()
        for rot in diags:
            rot &= slots - 1
            j, i = ((rot // n1) * n1) & (slots - 1), rot & (n1 - 1)
            index.setdefault(j, []).append(i)
            giants.add(j)
            babies.add(i)
        return {j: sorted(v) for j, v in sorted(index.items())}, sorted(giants), sorted(babies)

    @classmethod
    def find_best_bsgs_ratio(cls, diags, slots, log_max_ratio):
        max_ratio = float(1 << log_max_ratio)
        n1 = 1
        while n1 < slots:
            _, g, b = cls.bsgs_index(diags, slots, n1)
            num, den = float(len(g) - 1), float(len(b) - 1)
            r = (num / den) if den != 0 else (float("nan") if num == 0 else float("inf"))
            if r == max_ratio:
                return n1
            if r < max_ratio:
                return max(n1 // 2, 1)
            n1 <<= 1
        return 1

    def lt_build(self, diagonals, level, scale, log_slots=None, log_bsgs_ratio=0):
        """diagonals: {idx: slots values}.  Returns a dict holding the QP-encoded, pre-rotated diagonals."""
        log_slots = self.log_max_slots if log_slots is None else log_slots
        slots = 1 << log_slots
        idxs = sorted({k & (slots - 1) for k in diagonals})
        n1 = slots if log_bsgs_ratio < 0 else self.find_best_bsgs_ratio(idxs, slots, log_bsgs_ratio)
        index, giants, babies = self.bsgs_index(idxs, slots, n1)
        mods = self.qp_mods(level)
        enc = {}
        by_idx = {k & (slots - 1): np.asarray(v) for k, v in diagonals.items()}
        for d in idxs:
            j = ((d // n1) * n1) & (slots - 1)
            enc[d] = self.encode_coeffs(np.roll(by_idx[d], j), scale, mods, log_slots)
        return dict(level=level, scale=LD(scale), log_slots=log_slots, n1=n1, index=index, giants=giants,
                    babies=babies, diags=enc)

    def lt_galois_elements(self, lt):
        return sorted({self.ctx.galois_element(r) for r in lt["giants"] + lt["babies"] if r})

    def lt_apply(self, lt, ct):
        level = min(ct.level, lt["level"])
        L, K = level + 1, self.nP
        qmods, mods = list(range(L)), self.qp_mods(level)
        lt_mods = self.qp_mods(lt["level"])
        sel = [lt_mods.index(m) for m in mods]
        c0, c1 = ct.data[0][:L], ct.data[1][:L]
        pm = self._pmodq(level)
        zeros_p = np.zeros((K, self.N), dtype=np.uint64)
        lift = lambda x: np.concatenate([self._mul_scalar(x, pm, qmods), zeros_p])
        rots = {}
        digits = None
        for i in lt["babies"]:
            if i == 0:
                rots[0] = (lift(c0), lift(c1))
                continue
            if digits is None:
                digits = self.ctx.decompose(level, c1)
            g = self.ctx.galois_element(i)
            a0, a1 = self.ctx.inner_product(level, digits, self.galois_key(g, False))
            a0[:L] = self._ew("add", a0[:L], self._mul_scalar(c0, pm, qmods), qmods)
            rots[i] = (self.ctx.automorphism_ntt(a0, g), self.ctx.automorphism_ntt(a1, g))
        out = None
        slots = 1 << lt["log_slots"]
        for j, babies in lt["index"].items():
            t0 = t1 = None
            for i in babies:
                d = lt["diags"][(j + i) & (slots - 1)][sel]
                p0, p1 = self._ew("mul", d, rots[i][0], mods), self._ew("mul", d, rots[i][1], mods)
                t0 = p0 if t0 is None else self._ew("add", t0, p0, mods)
                t1 = p1 if t1 is None else self._ew("add", t1, p1, mods)
            if j != 0:
                g = self.ctx.galois_element(j)
                u1 = self.ctx.moddown(level, t1)
                k0, k1 = self.ctx.inner_product(level, self.ctx.decompose(level, u1), self.galois_key(g, False))
                k0 = self._ew("add", k0, t0, mods)
                t0, t1 = self.ctx.automorphism_ntt(k0, g), self.ctx.automorphism_ntt(k1, g)
            out = (t0, t1) if out is None else (self._ew("add", out[0], t0, mods), self._ew("add", out[1], t1, mods))
        res = np.stack([self.ctx.moddown(level, out[0]), self.ctx.moddown(level, out[1])])
        return OCt(res, level, ct.scale * lt["scale"], ct.log_slots)

    # ---- polynomial evaluation (orion/backend/lattigo/polyeval.go:63-84 contract; restates the product's own algorithm,
    #      orion_b200/csrc/polyeval.cu: Paterson-Stockmeyer with backward scale propagation).  Checked on the CPU against
    #      numpy (tests/test_oracle_cpu.py) and word for word against the GPU (tests/test_gpu_scheme.py). ----
    def lin_comb_int(self, ops, ks, c0, level, S):
        """sum_i k_i * a_i + c0 at `level`, declared scale S; k_i, c0 integers carried as long doubles."""
        L = level + 1
        mods = list(range(L))
        polys = ops[0].degree + 1
        acc = None
        for a, k in zip(ops, ks):
            k = LD(k)
            sc = self._const_rns(-1.0 if k < 0 else 1.0, abs(k), mods)
            term = [self._mul_scalar(a.data[p][:L], sc, mods) for p in range(polys)]
            acc = term if acc is None else [self._ew("add", acc[p], term[p], mods) for p in range(polys)]
        c0 = LD(c0)
        if c0 != 0:
            acc[0] = self._add_scalar(acc[0], self._const_rns(-1.0 if c0 < 0 else 1.0, abs(c0), mods), mods)
        return OCt(np.stack(acc), level, S, ops[0].log_slots, polys - 1)

    def lin_comb_scaled(self, ops, cs, c0, level, S):
        S = LD(S)
        ks = []
        for a, c in zip(ops, cs):
            v = np.round(abs(LD(c) * (S / a.scale)))
            ks.append(-v if c < 0 else v)
        v0 = np.round(abs(LD(c0) * S))
        return self.lin_comb_int(ops, ks, -v0 if c0 < 0 else v0, level, S)

    def poly_evaluate(self, ct, coeffs, chebyshev, out_scale, float32_coeffs=True):
        """float32_coeffs: the boundary hands coefficients over as C floats (GenerateChebyshev / GenerateMonomial); the
        bootstrapping circuit's own interpolant (oracle/bootstrap.py) keeps its doubles."""
        def bitlen(x):
            return int(x).bit_length()
        p = [float(np.float32(c)) if float32_coeffs else float(c) for c in coeffs]
        d = len(p) - 1
        while d > 0 and p[d] == 0.0:
            d -= 1
        depth, in_level = bitlen(d), ct.level
        out_level = in_level - depth
        assert d >= 1 and out_level >= 0
        B = 1 << ((depth + 1) // 2)
        P = {1: ct.copy()}
        q_at = lambda l: LD(self.Q[l])

        def power(n):
            if n in P:
                return P[n]
            a = 1 << (bitlen(n - 1) - 1)
            b = n - a
            A, Bc = power(a), power(b)
            prod = self.mul_relin(A, Bc)
            if chebyshev:
                c = a - b
                if c == 0:
                    prod = self.lin_comb_int([prod], [LD(2)], -np.round(prod.scale), prod.level, prod.scale)
                else:
                    Cc = power(c)
                    k = np.round(prod.scale / Cc.scale)
                    prod = self.lin_comb_int([prod, Cc], [LD(2), -k], LD(0), prod.level, prod.scale)
            P[n] = self.rescale(prod)
            return P[n]

        level_of = lambda n: in_level - (0 if n <= 1 else bitlen(n - 1))
        in_basis = lambda n: n < B or (n & (n - 1)) == 0

        def E(p, L, S):
            S = LD(S)
            deg = len(p) - 1
            while deg > 0 and p[deg] == 0.0:
                deg -= 1
            direct = all(p[i] == 0.0 or (in_basis(i) and level_of(i) >= L) for i in range(1, deg + 1))
            if direct:
                ops = [power(i) for i in range(1, deg + 1) if p[i] != 0.0]
                cs = [p[i] for i in range(1, deg + 1) if p[i] != 0.0]
                assert ops, "constant polynomials are not staged in the oracle"
                return self.lin_comb_scaled(ops, cs, p[0], L, S)
            assert deg >= 2
            g = 1 << (bitlen(deg) - 1)
            q = [0.0] * (deg - g + 1)
            r = list(p[:g])
            if chebyshev:
                q[0] = p[g]
                for k in range(1, deg - g + 1):
                    q[k] = 2.0 * p[g + k]
                    r[g - k] -= p[g + k]
            else:
                for k in range(deg - g + 1):
                    q[k] = p[g + k]
            Pg = power(g)
            usePg = self.drop_level(Pg, L) if Pg.level > L else Pg
            if deg == g:
                res = self.mul_const_scaled(usePg, q[0], S / Pg.scale)
            else:
                qS = (S / Pg.scale) * q_at(L + 1)
                A = self.rescale(E(q, L + 1, qS))
                A.scale = S / Pg.scale
                res = self.mul_relin(A, usePg)
            res.scale = S
            if any(v != 0.0 for v in r):
                res = self.add(res, E(r, L, S))
            return res

        res = self.rescale(E(p[: d + 1], out_level + 1, LD(out_scale) * q_at(out_level + 1)))
        res.scale = LD(out_scale)
        return res
