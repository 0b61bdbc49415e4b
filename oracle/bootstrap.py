"""CPU oracle of the bootstrapping circuit -- TEST INFRASTRUCTURE ONLY.

Restates orion_b200/csrc/bootstrap.cu (this backend's own circuit behind the contract of
orion/backend/lattigo/bootstrapper.go:19-80; structure of the reference GPU path:
orion/backend/heongpu/adrianHEonGPU/src/heongpu/lib/host/ckks/operator.cu:3894-4026) on top of oracle.scheme.OScheme:
extended prime chain, merged DFT factor matrices, Chebyshev interpolant of the scaled cosine, ScaleUp, ModRaise, SubSum,
CoeffsToSlots, EvalMod with double-angle steps, SlotsToCoeffs with the shim's sparse-slot post-factor.  With the same seed
the output ciphertext of `OBootstrapper.bootstrap` equals the device library's Bootstrap() word for word
(tests/test_gpu_scheme.py::test_bootstrap_words_equal_the_oracles).

Host floating point is reproduced operation by operation: complex products as four real products (no FMA), libm cos /
cosl / pow / powl through math and numpy.longdouble -- the same glibc routines the library's host code calls.
"parity unpinned" versus Lattigo (different circuit, same contract): see DESIGN.md section 4.
"""
import math

import numpy as np

from . import P, c_u64, lib, u64p
from .scheme import LD, OCt, OScheme

PI = 3.141592653589793238462643383279502884          # the double the C++ literal rounds to
LOG_MSG_RATIO = 10
CTS_DEPTH, STC_DEPTH, CHEB_DEPTH, CHEB_RANGE = 4, 3, 6, 27.0


def gen_ntt_primes(bits, nthroot, count, skip):
    L = lib()
    out = np.zeros(count, dtype=np.uint64)
    sk = np.array(list(skip) if skip else [0], dtype=np.uint64)
    rc = L.orc_gen_ntt_primes(bits, c_u64(nthroot), count, P(sk), len(skip), P(out))
    assert rc == 0, "ran out of NTT-friendly primes of this size"
    return [int(x) for x in out]


def host_roots(M):
    """exp(2 pi i k / M), k = 0..M: first quadrant from cos(), the rest by symmetry (csrc/scheme_kernels.cu host_roots)."""
    quar = M >> 2
    angle = 2.0 * PI / float(M)
    cs = [math.cos(angle * float(k)) for k in range(quar + 1)]
    re, im = np.zeros(M + 1), np.zeros(M + 1)
    for k in range(quar + 1):
        co, si = cs[k], cs[quar - k]
        re[k], im[k] = co, si
        re[2 * quar - k], im[2 * quar - k] = -co, si
        re[2 * quar + k], im[2 * quar + k] = -co, -si
        re[4 * quar - k], im[4 * quar - k] = co, -si
    return re, im


class Diag:
    """A matrix in diagonal form: rotation -> (real part, imaginary part) of the diagonal, float64 arrays of length n."""

    def __init__(self, n):
        self.n, self.d = n, {}

    def get(self, k):
        if k not in self.d:
            self.d[k] = (np.zeros(self.n), np.zeros(self.n))
        return self.d[k]

    def keys(self):
        return sorted(self.d)


def fft_stage(n, ln, M, roots, inverse):
    rr, ri = roots
    lenh, lenq = ln >> 1, ln << 2
    gap = M // lenq
    m = Diag(n)
    d0 = m.get(0); dp = m.get(lenh & (n - 1)); dm = m.get((n - lenh) & (n - 1))
    rot, p = [], 1
    for _ in range(lenh):
        rot.append(p); p = (p * 5) % M
    for s in range(n):
        r = s & (ln - 1)
        if r < lenh:
            idx = (rot[r] % lenq) * gap
            d0[0][s] += 1.0
            if inverse:
                dp[0][s] += 1.0
            else:
                dp[0][s] += rr[idx]; dp[1][s] += ri[idx]
        else:
            j = r - lenh
            idx = (lenq - (rot[j] % lenq)) * gap if inverse else (rot[j] % lenq) * gap
            d0[0][s] += -rr[idx]; d0[1][s] += -ri[idx]
            if inverse:
                dm[0][s] += rr[idx]; dm[1][s] += ri[idx]
            else:
                dm[0][s] += 1.0
    return m


def diag_mul(m2, m1, n):
    """(M2 * M1): diag_{k1+k2}[s] += d2_{k2}[s] * d1_{k1}[s + k2], accumulated in the key order of the C++ maps."""
    out = Diag(n)
    for a in m2.keys():
        ar, ai = m2.d[a]
        for b in m1.keys():
            br, bi = m1.d[b]
            br, bi = np.roll(br, -a), np.roll(bi, -a)
            dr, di = out.get((a + b) & (n - 1))
            dr += ar * br - ai * bi
            di += ar * bi + ai * br
    return out


def dft_factors(log_n, M, depth, inverse, roots):
    n = 1 << log_n
    lens = []
    if inverse:
        ln = n
        while ln >= 2:
            lens.append(ln); ln >>= 1
    else:
        ln = 2
        while ln <= n:
            lens.append(ln); ln <<= 1
    groups = min(depth, log_n)
    out, pos = [], 0
    for g in range(groups):
        cnt = log_n // groups + (1 if g < log_n % groups else 0)
        acc = fft_stage(n, lens[pos], M, roots, inverse)
        for k in range(1, cnt):
            acc = diag_mul(fft_stage(n, lens[pos + k], M, roots, inverse), acc, n)
        pos += cnt
        out.append(acc)
    return out


def to_lt_diags(m, factor):
    out = {}
    for k in m.keys():
        re, im = m.d[k]
        if not np.any(np.hypot(re, im) > 1e-300):
            continue
        out[k] = (re * factor) + 1j * (im * factor)
    return out


def cheb_interp(deg, K1, r):
    n = deg + 1
    f = [math.cos(2.0 * PI * (K1 * math.cos(PI * (k + 0.5) / n) - 0.25) / float(1 << r)) for k in range(n)]
    c = []
    for j in range(n):
        s = LD(0)
        for k in range(n):
            s = s + LD(f[k]) * np.cos(LD(PI) * j * (k + LD(0.5)) / n)
        c.append(float(s * (LD(1.0) if j == 0 else LD(2.0)) / n))
    return c


class OBootstrapper:
    def __init__(self, s, logP, num_slots):
        """s: the OScheme of the residual chain (secret key generated); mirrors boot_new()."""
        self.s = s
        log_slots = 0
        while (1 << log_slots) < num_slots:
            log_slots += 1
        assert (1 << log_slots) == num_slots and 1 <= log_slots <= s.log_max_slots
        self.log_slots = log_slots
        h = min(s.h, s.N)
        K = int(math.ceil(8.0 * math.sqrt(float(h) / 12.0)))
        self.K1, self.r = K + 1, 3
        while 2.0 * 3.14159265358979 * self.K1 / float(1 << self.r) > CHEB_RANGE:
            self.r += 1
        logd = int(round(math.log2(float(s.default_scale))))
        stc_bits = max(39, min(56, logd + 12))
        Q, used, Pb = list(s.Q), list(s.mod), []

        def take(bits, count, dst):
            for q in gen_ntt_primes(bits, 2 * s.N, count, used):
                dst.append(q); used.append(q)
        take(stc_bits, STC_DEPTH, Q)
        take(60, CHEB_DEPTH + self.r, Q)
        take(56, CTS_DEPTH, Q)
        sizes = {}
        for b in logP:
            sizes[b] = sizes.get(b, 0) + 1
        for b in sorted(sizes):
            take(b, sizes[b], Pb)
        log_scale = int(round(math.log2(float(s.default_scale))))
        bs = OScheme(s.logN, None, None, log_scale, s.h, s.seed ^ 0xB007B007 ^ num_slots, moduli=(Q, Pb))
        bs.default_scale = s.default_scale
        bs.set_secret_key(s.sk_coeffs)
        bs.gen_relin_key()
        self.bs, self.top = bs, bs.max_level
        M = 2 * s.N
        roots = host_roots(M)
        self.cts_m = dft_factors(log_slots, M, CTS_DEPTH, True, roots)
        self.stc_m = dft_factors(log_slots, M, STC_DEPTH, False, roots)
        self.cheb = cheb_interp((1 << CHEB_DEPTH) - 1, float(self.K1), self.r)
        self.times_i = bs.encode(np.full(1 << bs.log_max_slots, 1j), self.top, 1, bs.log_max_slots)
        # keys, in the order boot_new() generates them (every generation consumes one sampling event)
        bs.galois_key(2 * bs.N - 1)
        for i in range(log_slots, bs.log_max_slots):
            bs.galois_key(bs.ctx.galois_element(1 << i))
        n = 1 << log_slots
        for ms in (self.cts_m, self.stc_m):
            for m in ms:
                idx = m.keys()
                n1 = OScheme.find_best_bsgs_ratio(idx, n, 0)
                _, gs, bb = OScheme.bsgs_index(idx, n, n1)
                for g in gs:
                    if g:
                        bs.galois_key(bs.ctx.galois_element(g))
                for i in bb:
                    if i:
                        bs.galois_key(bs.ctx.galois_element(i))
        self.built_for = None
        self.cts = self.stc = None

    def bootstrap(self, ct_in, dump=None):
        """dump: optional dict filled with the intermediate ciphertexts (tools/boot_lockstep.py compares them with the
        library's ORION_B200_BOOT_DUMP files)."""
        s, bs = self.s, self.bs
        def D(tag, c):
            if dump is not None:
                dump[tag] = c
        N, n_log, top, r = bs.N, self.log_slots, self.top, self.r
        q_at = lambda l: LD(bs.Q[l])
        q0 = q_at(0)
        # ---- ScaleUp at level 0 ----
        low = s.drop_level(ct_in.copy(), 0)
        kf = np.floor(q0 / (np.ldexp(LD(1.0), LOG_MSG_RATIO) * low.scale) + LD(0.5))
        if kf < 1:
            kf = LD(1)
        if kf > 1:
            low = s.mul_const_scaled(low, 1.0, kf)
        eff = low.scale
        # ---- ModRaise: centred lift of both polynomials from q0 to the whole chain ----
        coef = s.ctx.ntt(low.data[:, 0, :], [0, 0], inverse=True)
        q0i = int(bs.Q[0])
        raised = np.empty((2, top + 1, N), dtype=np.uint64)
        for p in range(2):
            v = coef[p]
            neg = v > np.uint64(q0i >> 1)
            for i in range(top + 1):
                qi = np.uint64(bs.Q[i])
                a = (np.uint64(q0i) - v[neg]) % qi
                out = v % qi
                out[neg] = np.where(a != 0, qi - a, np.uint64(0))
                raised[p, i] = out
        mods = list(range(top + 1))
        ct = OCt(np.stack([bs.ctx.ntt(raised[0], mods), bs.ctx.ntt(raised[1], mods)]), top, q0, n_log)
        D("00_low", low); D("01_raised", ct)
        # ---- SubSum (sparse packing) ----
        g = 1.0
        for i in range(n_log, bs.log_max_slots):
            ct = bs.add(ct, bs.rotate(ct, 1 << i, generate=False))
            g *= 2.0
        # ---- DFT factors for this input scale ----
        cts_groups, stc_groups = len(self.cts_m), len(self.stc_m)
        lvl_after_cts = top - CTS_DEPTH
        lvl_after_mod = lvl_after_cts - (CHEB_DEPTH + r)
        s_mod, s_cheb = q_at(lvl_after_cts), q_at(lvl_after_cts - CHEB_DEPTH)
        if self.built_for is None or self.built_for != eff:
            lam = float(q0 / (LD(2.0) * self.K1 * g * q0))
            n = float(1 << n_log)
            f_cts = math.pow(lam / n, 1.0 / cts_groups)
            cur, self.cts = q0, []
            for k in range(cts_groups):
                lvl = top - k
                ps = q_at(lvl)
                if k == cts_groups - 1:
                    ps = q_at(lvl) * (s_mod / cur)
                self.cts.append(bs.lt_build(to_lt_diags(self.cts_m[k], f_cts), lvl, ps, n_log, 0))
                cur = cur * ps / q_at(lvl)
            s_out = s_cheb
            for k in range(r):
                s_out = s_out * s_out / q_at(lvl_after_cts - CHEB_DEPTH - k)
            mu = float(q0 / (LD(2.0) * PI * eff)) * g
            f_stc = math.pow(mu, 1.0 / stc_groups)
            target = s.default_scale
            per = np.power(target / s_out, LD(1.0) / stc_groups)
            cur, self.stc = s_out, []
            for k in range(stc_groups):
                lvl = lvl_after_mod - k
                ps = q_at(lvl) * per
                if k == stc_groups - 1:
                    ps = q_at(lvl) * (target / cur)
                self.stc.append(bs.lt_build(to_lt_diags(self.stc_m[k], f_stc), lvl, ps, n_log, 0))
                cur = cur * ps / q_at(lvl)
            self.built_for = eff
        D("02_subsum", ct)
        # ---- CoeffsToSlots ----
        for k in range(cts_groups):
            ct = bs.rescale(bs.lt_apply(self.cts[k], ct))
            D("03_cts%d" % k, ct)
        if ct.level > lvl_after_cts:
            ct = bs.drop_level(ct, lvl_after_cts)
        ct.scale = s_mod
        cj = bs.conjugate(ct, generate=False)
        re = bs.add(ct, cj)
        im = bs.add(ct, cj, sub=True)
        im = bs.mul_int(bs.mul_pt(im, self.times_i), -1)
        D("04_re", re); D("05_im", im)

        # ---- EvalMod ----
        def eval_mod(x):
            y = bs.poly_evaluate(x, self.cheb, True, s_cheb, float32_coeffs=False)
            D("06_cheb_re" if x is re else "06_cheb_im", y)
            for _ in range(r):
                sq = bs.mul_relin(y, y)
                sq = bs.lin_comb_int([sq], [LD(2)], -np.round(sq.scale), sq.level, sq.scale)
                y = bs.rescale(sq)
            return y
        vre, vim = eval_mod(re), eval_mod(im)
        ct = bs.add(vre, bs.mul_pt(vim, self.times_i))
        D("07_vre", vre); D("08_evalmod", ct)
        # ---- SlotsToCoeffs ----
        for k in range(stc_groups):
            ct = bs.rescale(bs.lt_apply(self.stc[k], ct))
            D("09_stc%d" % k, ct)
        out_level = s.max_level
        if ct.level > out_level:
            ct = bs.drop_level(ct, out_level)
        assert ct.level == out_level
        return OCt(ct.data.copy(), out_level, s.default_scale, ct_in.log_slots)
