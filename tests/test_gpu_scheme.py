"""GPU parity of the scheme layer (keys, encode, encrypt, evaluator ops, linear transforms) against the CPU oracle,
through the Orion-facing C ABI.  Same seed + same order of sampling events => bit-identical keys and ciphertexts."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PARAMS = dict(logN=12, logQ=[45, 30, 30, 30, 30], logP=[45, 45], log_scale=30, h=64)
SEED = 20261104


@pytest.fixture(scope="module")
def pair(orc, b200lib):
    from oracle.scheme import OScheme
    from orion_b200.bindings import B200Library
    p = PARAMS
    B = B200Library()
    B.B200_SetSeed(SEED)
    B.NewScheme(p["logN"], p["logQ"], p["logP"], p["log_scale"], p["h"], "standard", "", "none")
    O = OScheme(p["logN"], p["logQ"], p["logP"], p["log_scale"], p["h"], SEED)
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey(); B.GenerateEvaluationKeys()
    O.gen_secret_key(); O.gen_public_key(); O.gen_relin_key()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor()
    yield B, O
    B.DeleteScheme()


def export_ct(B, O, ct):
    level, deg = B.GetCiphertextLevel(ct), B.GetCiphertextDegree(ct)
    out = np.empty((deg + 1, level + 1, O.N), dtype=np.uint64)
    B.B200_ExportCiphertext(ct, out)
    return out


def export_pt(B, O, pt):
    level = B.GetPlaintextLevel(pt)
    out = np.empty((level + 1, O.N), dtype=np.uint64)
    B.B200_ExportPlaintext(pt, out)
    return out


def test_keys_bit_exact(pair):
    B, O = pair
    sk = np.empty(O.N, dtype=np.int64)
    B.B200_ExportSecretKey(sk)
    assert np.array_equal(sk, O.sk_coeffs)
    assert int(np.count_nonzero(sk)) == PARAMS["h"]
    pk = np.empty((2, O.nmod, O.N), dtype=np.uint64)
    B.B200_ExportPublicKey(pk)
    assert np.array_equal(pk, O.pk)
    beta = O.ctx.num_digits(O.max_level)
    rlk = np.empty((beta, 2, O.nmod, O.N), dtype=np.uint64)
    B.B200_ExportRelinKey(rlk)
    assert np.array_equal(rlk, O.rlk)
    assert B.GetModuliChain() == O.Q


def test_encode_decode(pair):
    B, O = pair
    rng = np.random.default_rng(1)
    slots = O.N // 2
    v = rng.uniform(-1, 1, slots).astype(np.float32)
    for level in (O.max_level, 1, 0):
        scale = 1 << 30
        pt = B.Encode(v.tolist(), level, scale)
        assert B.GetPlaintextLevel(pt) == level and B.GetPlaintextScale(pt) == scale and B.GetPlaintextSlots(pt) == slots
        opt = O.encode(v.astype(np.float64), level, scale)
        assert np.array_equal(export_pt(B, O, pt), opt.data), "encode level %d" % level
        dec = np.array(B.Decode(pt))
        assert np.max(np.abs(dec - v)) < 1e-6
        odec = O.decode(opt)
        re, im = np.empty(slots), np.empty(slots)
        B.B200_DecodeF64(pt, re, im)
        assert np.array_equal(re, odec.real) and np.array_equal(im, odec.imag)
        B.DeletePlaintext(pt)
    # ragged input: fewer values than slots are zero-padded
    pt = B.Encode(v[:100].tolist(), 2, 1 << 30)
    assert np.array_equal(export_pt(B, O, pt), O.encode(v[:100].astype(np.float64), 2, 1 << 30).data)
    B.DeletePlaintext(pt)


def test_encrypt_ops(pair):
    B, O = pair
    rng = np.random.default_rng(2)
    slots = O.N // 2
    x = rng.uniform(-1, 1, slots).astype(np.float32)
    y = rng.uniform(-1, 1, slots).astype(np.float32)
    scale = 1 << 30
    L = O.max_level
    ptx, pty = B.Encode(x.tolist(), L, scale), B.Encode(y.tolist(), L, scale)
    optx, opty = O.encode(x.astype(np.float64), L, scale), O.encode(y.astype(np.float64), L, scale)
    ctx_, cty = B.Encrypt(ptx), B.Encrypt(pty)
    octx, octy = O.encrypt(optx), O.encrypt(opty)
    assert np.array_equal(export_ct(B, O, ctx_), octx.data)
    assert np.array_equal(export_ct(B, O, cty), octy.data)

    def dec(ct):
        p = B.Decrypt(ct); out = np.array(B.Decode(p)); B.DeletePlaintext(p); return out

    assert np.max(np.abs(dec(ctx_) - x)) < 1e-4
    # add / sub / plaintext ops
    s = B.AddCiphertextNew(ctx_, cty)
    assert np.array_equal(export_ct(B, O, s), O.add(octx, octy).data)
    d = B.SubCiphertextNew(ctx_, cty)
    assert np.array_equal(export_ct(B, O, d), O.add(octx, octy, sub=True).data)
    ap = B.AddPlaintextNew(ctx_, pty)
    assert np.array_equal(export_ct(B, O, ap), O.add(octx, opty).data)
    sp = B.SubPlaintextNew(ctx_, pty)
    assert np.array_equal(export_ct(B, O, sp), O.add(octx, opty, sub=True).data)
    # scalar ops
    a1 = B.AddScalarNew(ctx_, 0.375)
    assert np.array_equal(export_ct(B, O, a1), O.add_const(octx, 0.375).data)
    a2 = B.SubScalarNew(ctx_, 1.5)
    assert np.array_equal(export_ct(B, O, a2), O.add_const(octx, -1.5).data)
    m1 = B.MulScalarIntNew(ctx_, -3)
    assert np.array_equal(export_ct(B, O, m1), O.mul_int(octx, -3).data)
    ng = B.Negate(ctx_)
    assert np.array_equal(export_ct(B, O, ng), O.mul_int(octx, -1).data)
    m2 = B.MulScalarFloatNew(ctx_, 0.3)
    om2 = O.mul_float(octx, float(np.float32(0.3)))
    assert np.array_equal(export_ct(B, O, m2), om2.data)
    m2 = B.Rescale(m2)
    om2 = O.rescale(om2)
    assert np.array_equal(export_ct(B, O, m2), om2.data)
    assert B.GetCiphertextLevel(m2) == L - 1
    assert np.max(np.abs(dec(m2) - np.float32(0.3) * x)) < 1e-4
    # ct x pt, ct x ct (+ relinearisation) and rescale
    mp = B.MulPlaintextNew(ctx_, pty)
    omp = O.mul_pt(octx, opty)
    assert np.array_equal(export_ct(B, O, mp), omp.data)
    mp = B.Rescale(mp); omp = O.rescale(omp)
    assert np.array_equal(export_ct(B, O, mp), omp.data)
    assert np.max(np.abs(dec(mp) - x * y)) < 1e-4
    mc = B.MulRelinCiphertextNew(ctx_, cty)
    omc = O.mul_relin(octx, octy)
    assert np.array_equal(export_ct(B, O, mc), omc.data)
    mc = B.Rescale(mc); omc = O.rescale(omc)
    assert np.array_equal(export_ct(B, O, mc), omc.data)
    assert np.max(np.abs(dec(mc) - x * y)) < 1e-4
    # level mismatch: operands are aligned by dropping limbs
    mix = B.AddCiphertextNew(mc, ctx_)
    assert B.GetCiphertextLevel(mix) == L - 1
    # rotations (keys generated on demand, in the same order on both sides)
    for rot in (1, 7, -2):
        r = B.RotateNew(ctx_, rot)
        orr = O.rotate(octx, rot)
        assert np.array_equal(export_ct(B, O, r), orr.data), "rotate %d" % rot
        assert np.max(np.abs(dec(r) - np.roll(x, -rot))) < 1e-4
        B.DeleteCiphertext(r)
    cj = B.B200_Conjugate(ctx_)
    assert np.array_equal(export_ct(B, O, cj), O.conjugate(octx).data)
    # handle table: smallest free integer is reused
    live = B.GetLiveCiphertexts()
    B.DeleteCiphertext(s)
    t = B.CloneCiphertext(ctx_)
    assert t == s and sorted(B.GetLiveCiphertexts()) == sorted(live)


def test_linear_transform(pair):
    B, O = pair
    rng = np.random.default_rng(3)
    slots = O.N // 2
    level = 3
    scale = 1 << 30
    x = rng.uniform(-1, 1, slots).astype(np.float32)
    diag_idx = list(range(0, 24)) + [slots - 3, slots - 1, 100, 1000]
    diags = {d: rng.uniform(-1, 1, slots).astype(np.float32) for d in diag_idx}
    flat = np.concatenate([diags[d] for d in diag_idx])
    pt = B.Encode(x.tolist(), level, scale)
    ct = B.Encrypt(pt)
    oct_ = O.encrypt(O.encode(x.astype(np.float64), level, scale))
    assert np.array_equal(export_ct(B, O, ct), oct_.data)
    B.NewLinearTransformEvaluator()
    lt = B.GenerateLinearTransform(diag_idx, flat, level, 2.0, "none")
    olt = O.lt_build({d: v.astype(np.float64) for d, v in diags.items()}, level, O.Q[level], log_bsgs_ratio=0)
    gal = B.GetLinearTransformRotationKeys(lt)
    assert sorted(gal) == O.lt_galois_elements(olt)
    for g in sorted(gal):
        B.GenerateLinearTransformRotationKey(g)
        O.galois_key(g)
    out = B.EvaluateLinearTransform(lt, ct)
    oout = O.lt_apply(olt, oct_)
    assert np.array_equal(export_ct(B, O, out), oout.data)
    out = B.Rescale(out)
    p = B.Decrypt(out)
    got = np.array(B.Decode(p))
    want = np.zeros(slots)
    for d, v in diags.items():
        want += v.astype(np.float64) * np.roll(x.astype(np.float64), -d)
    assert np.max(np.abs(got - want)) < 1e-3
    B.DeleteLinearTransform(lt)


def test_polynomial_evaluation(orc, b200lib):
    """Depth-optimal PS evaluation: output level = in - ceil(log2(d+1)), scale == requested, values match numpy."""
    from orion_b200.bindings import B200Library
    B = B200Library()
    B.B200_SetSeed(7)
    logq = [50] + [40] * 7
    B.NewScheme(13, logq, [51, 51], 40, 192, "standard", "", "none")
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewPolynomialEvaluator()
    Q = B.GetModuliChain()
    rng = np.random.default_rng(5)
    slots = 4096
    x = rng.uniform(-1, 1, slots)
    L = len(logq) - 1
    pt = B.B200_EncodeF64(x, slots, L, float(1 << 40))
    ct = B.Encrypt(pt)

    def dec(c):
        p = B.Decrypt(c)
        re, im = np.empty(slots), np.empty(slots)
        B.B200_DecodeF64(p, re, im)
        B.DeletePlaintext(p)
        return re

    for deg, cheb in ((1, False), (3, False), (7, False), (12, False), (15, True), (31, True), (63, True), (27, True)):
        coeffs = rng.uniform(-1, 1, deg + 1) / (1.0 if cheb else np.arange(1, deg + 2))
        poly = B.B200_GeneratePolynomialF64(np.ascontiguousarray(coeffs), deg + 1, 1 if cheb else 0)
        depth = int(np.ceil(np.log2(deg + 1)))
        assert B.GetPolyDepth(poly) == depth
        out_level = L - depth
        out_scale = Q[out_level]                      # what orion's ReLU asks for (activation.py:253-255)
        out = B.EvaluatePolynomial(ct, poly, out_scale)
        assert B.GetCiphertextLevel(out) == out_level
        assert B.GetCiphertextScale(out) == out_scale
        want = np.polynomial.chebyshev.chebval(x, coeffs) if cheb else np.polynomial.polynomial.polyval(x, coeffs)
        err = np.max(np.abs(dec(out) - want))
        assert err < 1e-5, "degree %d cheb=%s err=%g" % (deg, cheb, err)
        B.DeleteCiphertext(out)
    # float32 entry points used by orion (coefficients ascending)
    c32 = [0.5, 0.25, 0.0, -0.125]
    pm = B.GenerateMonomial(c32)
    out = B.EvaluatePolynomial(ct, pm, 1 << 40)
    assert np.max(np.abs(dec(out) - np.polynomial.polynomial.polyval(x, c32))) < 1e-5
    B.DeleteScheme()


@pytest.mark.parametrize("num_slots", [2048, 256])
def test_bootstrap(orc, b200lib, num_slots):
    """Bootstrap contract (lattigo/bootstrapper.go:61-80): any input level -> top level, default scale, same values."""
    from orion_b200.bindings import B200Library
    B = B200Library()
    B.B200_SetSeed(11)
    logq = [55, 40, 40, 40, 40]
    B.NewScheme(12, logq, [56, 56], 40, 64, "standard", "", "none")
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
    B.NewBootstrapper([61, 61, 61, 61], num_slots)
    rng = np.random.default_rng(8)
    slots = 2048
    base = rng.uniform(-1, 1, num_slots)
    # Orion hands Bootstrap a ZERO-PADDED full-slot vector (orion/nn/operations.py:59-60,85-87) and the Go shim multiplies
    # Lattigo's result by MaxSlots/numSlots (LG/bootstrapper.go:73-74): x, 0, ..., 0 comes back as x, x, ..., x
    xin = np.zeros(slots); xin[:num_slots] = base
    x = np.tile(base, slots // num_slots)
    pt = B.B200_EncodeF64(xin, slots, 1, float(1 << 40))
    ct = B.Encrypt(pt)
    out = B.Bootstrap(ct, num_slots)
    assert B.GetCiphertextLevel(out) == len(logq) - 1
    assert B.GetCiphertextScale(out) == 1 << 40
    p = B.Decrypt(out)
    re, im = np.empty(slots), np.empty(slots)
    B.B200_DecodeF64(p, re, im)
    err = np.max(np.abs(re - x))
    tol = 2e-4 * max(1, (slots // num_slots) // 2)        # the post-factor multiplies the bootstrapping error as well
    assert err < tol, err
    assert np.max(np.abs(im)) < tol
    # the refreshed ciphertext is usable: one multiplication + rescale
    sq = B.Rescale(B.MulRelinCiphertextNew(out, out))
    p2 = B.Decrypt(sq)
    B.B200_DecodeF64(p2, re, im)
    assert np.max(np.abs(re - x * x)) < 2.5 * tol
    B.DeleteBootstrappers()
    B.DeleteScheme()


@pytest.mark.parametrize("num_slots,min_bits", [(4096, 18.0), (1024, 16.0)])
def test_bootstrap_precision_at_the_resnet20_parameter_set(orc, b200lib, num_slots, min_bits):
    """The shipped ResNet-20 parameters (configs/resnet.yml: logN=13, LogQ 60+32x30, LogP 60x2, LogScale 30, H=192, boot
    LogP 8x61): a ciphertext at level 1 comes back at level 32 with the default scale and >= 18 bits of precision on full
    slots (max |error| <= 2^-18 for values in [-1, 1]); with 1024 of 4096 slots the Go shim's post-factor MaxSlots/slots = 4
    (LG/bootstrapper.go:73-74) multiplies the error as well: >= 16 bits.  Input as Orion sends it: zero-padded."""
    from orion_b200.bindings import B200Library
    B = B200Library()
    B.B200_SetSeed(11)
    logq = [60] + [30] * 32
    B.NewScheme(13, logq, [60, 60], 30, 192, "standard", "", "none")
    try:
        B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
        B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
        B.NewBootstrapper([61] * 8, num_slots)
        slots = 4096
        rng = np.random.default_rng(8)
        base = rng.uniform(-1, 1, num_slots)
        xin = np.zeros(slots); xin[:num_slots] = base
        want = np.tile(base, slots // num_slots)
        ct = B.Encrypt(B.B200_EncodeF64(xin, slots, 1, float(1 << 30)))
        worst = 0.0
        for _ in range(2):
            out = B.Bootstrap(ct, num_slots)
            assert B.GetCiphertextLevel(out) == 32 and B.GetCiphertextScale(out) == 1 << 30
            re, im = np.empty(slots), np.empty(slots)
            B.B200_DecodeF64(B.Decrypt(out), re, im)
            worst = max(worst, float(np.max(np.abs(re - want))), float(np.max(np.abs(im))))
        bits = -np.log2(worst)
        print("bootstrap at the ResNet-20 chain, %d slots: %.1f bits" % (num_slots, bits))
        assert bits >= min_bits, (bits, worst)
    finally:
        B.DeleteBootstrappers()
        B.DeleteScheme()


@pytest.mark.parametrize("num_slots,boot_p", [(2048, 4), (256, 4), (2048, 8)])
def test_bootstrap_words_equal_the_oracles(orc, b200lib, num_slots, boot_p):
    """a16, word for word: Bootstrap() on the GPU (ScaleUp, ModRaise, SubSum, 4 CoeffsToSlots transforms, EvalMod on both
    halves -- degree-63 Chebyshev + double-angle steps --, 3 SlotsToCoeffs transforms; ~30 rotations keys, ~100 key switches)
    returns the ciphertext oracle/bootstrap.py computes on the CPU from the same input ciphertext and seed.  Full and sparse
    slot counts (the sparse one adds the SubSum rotations and the shim's post-factor); with 8 bootstrapping special primes the
    ModUp / ModDown conversions have 8-source digits, i.e. they run on the INT8 tensor-core kernel (csrc/ks_kernels.cu
    baseconv_tc_kernel), plain and centred -- the shape of the shipped ResNet-20 configuration (boot LogP 8 x 61)."""
    from oracle.bootstrap import OBootstrapper
    from oracle.scheme import OScheme
    from orion_b200.bindings import B200Library
    logn, logq, logp, logscale, h, seed = 12, [55, 40, 40, 40, 40], [56, 56], 40, 64, 11
    B = B200Library()
    B.B200_SetSeed(seed)
    B.NewScheme(logn, logq, logp, logscale, h, "standard", "", "none")
    try:
        B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
        B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
        O = OScheme(logn, logq, logp, logscale, h, seed)
        O.gen_secret_key(); O.gen_public_key(); O.gen_relin_key()
        # NewEvaluator generates the power-of-two rotation keys (LG/evaluator.go:25-31): same sampling events on both sides
        i = 1
        while i < (1 << O.log_max_slots):
            O.galois_key(O.ctx.galois_element(i)); i *= 2
        rng = np.random.default_rng(8)
        slots = 1 << (logn - 1)
        xin = np.zeros(slots); xin[:num_slots] = rng.uniform(-1, 1, num_slots)
        ct = B.Encrypt(B.B200_EncodeF64(xin, slots, 1, float(1 << logscale)))
        oct_ = O.encrypt(O.encode(xin, 1, 2.0 ** logscale))
        assert np.array_equal(export_ct(B, O, ct), oct_.data)
        B.NewBootstrapper([61] * boot_p, num_slots)
        ob = OBootstrapper(O, [61] * boot_p, num_slots)
        got = export_ct(B, O, B.Bootstrap(ct, num_slots))
        want = ob.bootstrap(oct_)
        assert got.shape == want.data.shape
        assert np.array_equal(got, want.data), "bootstrapped ciphertext differs from the oracle's"
        # and it is a bootstrap: the oracle's ciphertext decrypts to replicate(x)
        dec = np.real(O.decode(O.decrypt(want)))
        assert np.max(np.abs(dec - np.tile(xin[:num_slots], slots // num_slots))) < 2e-4 * (slots // num_slots)
    finally:
        B.DeleteBootstrappers()
        B.DeleteScheme()


def test_polynomial_evaluation_words_equal_the_oracles(orc, b200lib):
    """a13, word for word: EvaluatePolynomial (Chebyshev and monomial, degrees 5-31: power basis, Paterson-Stockmeyer
    recursion with backward scale propagation, relinearisations, rescales) gives the ciphertext oracle/scheme.py
    poly_evaluate computes on the CPU from the same input ciphertext."""
    from oracle.scheme import OScheme
    from orion_b200.bindings import B200Library
    logn, logq, logp, logscale, h, seed = 12, [50] + [40] * 7, [56, 56], 40, 64, 7
    B = B200Library()
    B.B200_SetSeed(seed)
    B.NewScheme(logn, logq, logp, logscale, h, "standard", "", "none")
    try:
        B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
        B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewPolynomialEvaluator()
        O = OScheme(logn, logq, logp, logscale, h, seed)
        O.gen_secret_key(); O.gen_public_key(); O.gen_relin_key()
        rng = np.random.default_rng(1)
        x = rng.uniform(-1, 1, 1 << (logn - 1)).astype(np.float32)
        ct = B.Encrypt(B.Encode(x.tolist(), 7, 1 << logscale))
        oct_ = O.encrypt(O.encode(x.astype(np.float64), 7, 2.0 ** logscale))
        assert np.array_equal(export_ct(B, O, ct), oct_.data)
        for deg, cheb in ((7, True), (15, True), (27, True), (31, True), (5, False)):
            c = (rng.uniform(-1, 1, deg + 1) / (1 if cheb else deg)).astype(np.float32)
            poly = B.GenerateChebyshev(c.tolist()) if cheb else B.GenerateMonomial(c.tolist())
            got = export_ct(B, O, B.EvaluatePolynomial(ct, poly, 1 << logscale))
            want = O.poly_evaluate(oct_, c, cheb, 2.0 ** logscale).data
            assert got.shape == want.shape and np.array_equal(got, want), (deg, cheb)
    finally:
        B.DeleteScheme()


def test_bootstrap_graph_replay_is_bit_identical(orc, b200lib):
    """Bootstrap runs as a captured CUDA graph from the second call with the same (slot count, input level, input scale) on
    (csrc/graphs.cu; default for logN <= 14): the first call is eager, later calls capture one instance per concurrently used
    lane and replay it.  Every replay must return the eager call's words, for two different inputs, and an input at another
    level (another key) must not disturb it."""
    from orion_b200.bindings import B200Library
    logn, logq, logp, logscale, h = 12, [55, 40, 40, 40, 40], [56, 56], 40, 64
    B = B200Library()
    B.B200_SetSeed(5)
    B.NewScheme(logn, logq, logp, logscale, h, "standard", "", "none")
    try:
        B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey()
        B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
        slots = 1 << (logn - 1)
        N = 1 << logn
        B.NewBootstrapper([61] * 8, slots)
        rng = np.random.default_rng(3)

        def words(ct):
            out = np.empty(2 * (B.GetCiphertextLevel(ct) + 1) * N, dtype=np.uint64)
            B.B200_ExportCiphertext(ct, out)
            return out

        cts = [B.Encrypt(B.B200_EncodeF64(rng.uniform(-1, 1, slots), slots, 1, float(1 << logscale))) for _ in range(2)]
        other = B.Encrypt(B.B200_EncodeF64(rng.uniform(-1, 1, slots), slots, 2, float(1 << logscale)))     # level 2: another key
        first = [words(B.Bootstrap(c, slots)) for c in cts]                # eager (first sight of the key), then the first capture
        assert not np.array_equal(first[0], first[1])
        other_first = words(B.Bootstrap(other, slots))
        l0 = b200lib.load().B200_LaunchCount()
        for rep in range(6):                                               # replays, on whichever lanes the scheduler picks
            for k, c in enumerate(cts):
                assert np.array_equal(words(B.Bootstrap(c, slots)), first[k]), "replay %d of input %d" % (rep, k)
            if rep == 2:
                assert np.array_equal(words(B.Bootstrap(other, slots)), other_first)
        assert b200lib.load().B200_LaunchCount() - l0 > 13 * 500          # replayed kernel nodes are counted as launches
        # and it is a bootstrap
        re_, im_ = np.empty(slots), np.empty(slots)
        B.B200_DecodeF64(B.Decrypt(B.Bootstrap(cts[0], slots)), re_, im_)
    finally:
        B.DeleteBootstrappers()
        B.DeleteScheme()
