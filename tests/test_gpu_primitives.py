"""GPU parity of the hot-path primitives against the CPU oracle, through the C ABI (B200_*).  Bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [  # (logN, LogQ, LogP)
    (12, [40, 30, 30, 30], [40, 40]),
    (13, [29, 26, 26, 26, 26, 26], [29, 29]),          # configs/mlp.yml
    (14, [60, 30, 30, 30, 30, 30, 30], [60, 60]),      # configs/lola2.yml
    (13, [55, 40, 40, 40, 40, 40, 40], [61, 61, 61]),  # partial last digit (7 = 3+3+1)
    (16, [60, 46, 45, 47, 46], [61, 46]),              # both sides of the FP64 / integer butterfly split (46 bits)
    (15, [46, 46, 46, 46], [46, 46]),                  # FP64 butterflies only, largest admissible primes
    (17, [45, 45], [45]),                              # deepest transform (17 stages) on the FP64 path
    (13, [60] + [30] * 16, [61] * 8),                  # bootstrapping-style chain: 8 special primes, 30-bit digits (raw / 96-bit MAC modes)
    (13, [30] * 9, [61] * 8),                          # every ModUp source below 2^32 (96-bit MAC mode for the whole batch)
]


def make(orc, b200lib, case):
    logN, lq, lp = case
    Q, P = orc.gen_moduli(lq, lp, logN + 1)
    Qg, Pg = b200lib.gen_moduli(lq, lp, logN + 1)
    assert (Q, P) == (Qg, Pg), "prime selection differs between library and oracle"
    oc = orc.Ctx(logN, Q + P, len(Q), len(P))
    gc = b200lib.RawContext(logN, Q, P)
    for i in range(len(Q) + len(P)):
        assert gc.psi(i) == int(oc.L.orc_ctx_psi(oc.h, i))
    return oc, gc


@pytest.mark.parametrize("case", CASES)
def test_ntt_roundtrip_and_parity(orc, b200lib, case):
    oc, gc = make(orc, b200lib, case)
    rng = np.random.default_rng(7)
    idx = list(range(oc.nQ + oc.nP))
    a = oc.uniform_poly(rng, idx)
    f = gc.ntt(a, idx)
    assert np.array_equal(f, oc.ntt(a, idx))
    b = gc.ntt(f, idx, inverse=True)
    assert np.array_equal(b, a)
    assert np.array_equal(gc.ntt(a, idx, inverse=True), oc.ntt(a, idx, inverse=True))
    # edge values: 0, q-1
    e = np.stack([np.full(oc.N, oc.moduli[m] - 1, dtype=np.uint64) for m in idx])
    assert np.array_equal(gc.ntt(e, idx), oc.ntt(e, idx))
    z = np.zeros_like(e)
    assert np.array_equal(gc.ntt(z, idx), z)


@pytest.mark.parametrize("case", CASES)
def test_baseconv(orc, b200lib, case):
    oc, gc = make(orc, b200lib, case)
    rng = np.random.default_rng(8)
    K = oc.nP
    src = list(range(0, min(K, oc.nQ)))
    dst = [i for i in range(oc.nQ + K) if i not in src]
    a = oc.uniform_poly(rng, src)
    assert np.array_equal(gc.baseconv(a, src, dst), oc.baseconv(a, src, dst))
    psrc = list(range(oc.nQ, oc.nQ + K))
    qdst = list(range(oc.nQ))
    b = oc.uniform_poly(rng, psrc)
    assert np.array_equal(gc.baseconv(b, psrc, qdst, centered=True), oc.baseconv(b, psrc, qdst, centered=True))
    # single source limb: copy + reduce
    c = oc.uniform_poly(rng, [0])
    assert np.array_equal(gc.baseconv(c, [0], dst), oc.baseconv(c, [0], dst))


@pytest.mark.parametrize("case", CASES)
def test_elementwise_and_automorphism(orc, b200lib, case):
    oc, gc = make(orc, b200lib, case)
    rng = np.random.default_rng(9)
    idx = list(range(oc.nQ))
    a, b = oc.uniform_poly(rng, idx), oc.uniform_poly(rng, idx)
    for op in ("add", "sub", "mul"):
        assert np.array_equal(gc.elementwise(op, a, b), oc.elementwise(op, a, b)), op
    assert np.array_equal(gc.elementwise("neg", a), oc.elementwise("neg", a))
    for rot in (1, 5, -3, oc.N // 4):
        g = oc.galois_element(rot)
        assert g == gc.galois_element(rot)
        assert np.array_equal(gc.automorphism_ntt(a, g), oc.automorphism_ntt(a, g))
    g = 2 * oc.N - 1
    assert np.array_equal(gc.automorphism_ntt(a, g), oc.automorphism_ntt(a, g))


@pytest.mark.parametrize("case", CASES)
def test_rescale(orc, b200lib, case):
    oc, gc = make(orc, b200lib, case)
    rng = np.random.default_rng(10)
    for level in (oc.nQ - 1, 1):
        a = oc.uniform_poly(rng, list(range(level + 1)))
        b = oc.uniform_poly(rng, list(range(level + 1)))
        out = gc.rescale(level, np.stack([a, b]), polys=2)
        assert np.array_equal(out[0], oc.rescale(level, a))
        assert np.array_equal(out[1], oc.rescale(level, b))


@pytest.mark.parametrize("case", CASES)
def test_keyswitch_stages_and_full(orc, b200lib, case):
    oc, gc = make(orc, b200lib, case)
    rng = np.random.default_rng(11)
    nmod = oc.nQ + oc.nP
    beta_max = oc.num_digits(oc.nQ - 1)
    evk = np.stack([oc.uniform_poly(rng, list(range(nmod))) for _ in range(beta_max * 2)]).reshape(beta_max, 2, nmod, oc.N)
    evk_dev = gc.evk_to_device(evk)
    for level in sorted({oc.nQ - 1, max(0, oc.nQ - 2), 0}):
        c1 = oc.uniform_poly(rng, list(range(level + 1)))
        c0 = oc.uniform_poly(rng, list(range(level + 1)))
        dg = gc.decompose(level, c1)
        assert np.array_equal(dg, oc.decompose(level, c1)), "decompose level %d" % level
        acc = oc.uniform_poly(rng, list(range(level + 1)) + list(range(oc.nQ, nmod)))
        assert np.array_equal(gc.moddown(level, acc), oc.moddown(level, acc)), "moddown level %d" % level
        ref = oc.keyswitch(level, c1, evk)
        out = gc.keyswitch(level, c1, evk_dev)
        assert np.array_equal(out, ref), "keyswitch level %d" % level
        out2 = gc.keyswitch(level, c1, evk_dev, c0=c0)
        ref2 = ref.copy()
        ref2[0] = oc.elementwise("add", ref[0], c0)
        assert np.array_equal(out2, ref2)
    # batched entry points (pipelined host-buffer call and the device-resident call bench.py times): every output
    # ciphertext equals the ORACLE's key switch of that ciphertext.  Under ORION_B200_FUSED_IP=1 (see
    # test_fused_inner_product_switch_*) this is the route into ntt2_ip_kernel / ntt_forward_first / the raw-double
    # ModUp hand-off for every chain in CASES.
    level = oc.nQ - 1
    cts = [np.stack([oc.uniform_poly(rng, list(range(level + 1))) for _ in range(2)]) for _ in range(5)]
    refs = [oc.keyswitch(level, ct[1], evk) for ct in cts]
    refs_c0 = []
    for ct, r in zip(cts, refs):
        r2 = r.copy(); r2[0] = oc.elementwise("add", r[0], ct[0]); refs_c0.append(r2)
    outs = gc.keyswitch_batch(level, cts, [evk_dev] * 5)
    for r, o in zip(refs_c0, outs):
        assert np.array_equal(o, r)
    outs = gc.keyswitch_batch(level, [ct[1] for ct in cts], [evk_dev] * 5, has_c0=False)
    for r, o in zip(refs, outs):
        assert np.array_equal(o, r)
    outs = gc.keyswitch_batch_dev(level, np.stack(cts), [evk_dev] * 5)
    for r, o in zip(refs_c0, outs):
        assert np.array_equal(o, r)
    outs = gc.keyswitch_batch_dev(level, np.stack([ct[1] for ct in cts]), [evk_dev] * 5, has_c0=False)
    for r, o in zip(refs, outs):
        assert np.array_equal(o, r)
    gc.lib.B200_DevFree(evk_dev)


# The configurations bench.py times (and its neighbours), at full size, through the device-resident batch entry point:
#   bench chain            logN=16, L=24 (60 + 23x45 bit), k=3 (61 bit), beta=8: 23 limbs on the FP64 path, 4 on the integer path
#   integer-path chain     50-55-bit scale primes (ordinary for bootstrappable parameter sets): every limb on the integer path
#   beta does not divide L logN=16, L=10, k=3: beta=4 with a single-limb last digit
# (N/1024) * W * n is far above the fused-kernel threshold in all three, so this is ntt2_ip_kernel<ArithF64> and <ArithInt>,
# ntt_forward_first and the raw-double ModUp hand-off, word for word against oracle.Ctx.keyswitch.
BENCH_CASES = [
    (16, [60] + [45] * 23, [61] * 3, 8),
    (16, [60] + [52, 55, 50, 53] * 3, [61] * 3, 6),
    (16, [60] + [45] * 9, [61] * 3, 6),
    (15, [55] + [40] * 12, [61, 61], 12),
]


@pytest.mark.parametrize("case", BENCH_CASES)
def test_keyswitch_batchdev_at_bench_sizes_equals_the_oracle(orc, b200lib, case):
    logN, lq, lp, n = case
    oc, gc = make(orc, b200lib, (logN, lq, lp))
    rng = np.random.default_rng(12)
    nmod = oc.nQ + oc.nP
    level = oc.nQ - 1
    beta = oc.num_digits(level)
    evks = [np.stack([oc.uniform_poly(rng, list(range(nmod))) for _ in range(beta * 2)]).reshape(beta, 2, nmod, oc.N) for _ in range(2)]
    evk_devs = [gc.evk_to_device(e) for e in evks]
    cts = np.stack([np.stack([oc.uniform_poly(rng, list(range(level + 1))) for _ in range(2)]) for _ in range(n)])
    which = [z % 2 for z in range(n)]
    refs = [oc.keyswitch(level, cts[z, 1], evks[which[z]]) for z in range(n)]
    outs = gc.keyswitch_batch_dev(level, cts[:, 1], [evk_devs[w] for w in which], has_c0=False)
    for z in range(n):
        assert np.array_equal(outs[z], refs[z]), "ciphertext %d (no c0)" % z
    outs = gc.keyswitch_batch_dev(level, cts, [evk_devs[w] for w in which], has_c0=True)
    for z in range(n):
        r = refs[z].copy(); r[0] = oc.elementwise("add", refs[z][0], cts[z, 0])
        assert np.array_equal(outs[z], r), "ciphertext %d (+c0)" % z
    # a lower level of the same chain (plans are per level; beta changes)
    lvl = max(1, level - 4)
    outs = gc.keyswitch_batch_dev(lvl, cts[:, 1, : lvl + 1], [evk_devs[w] for w in which], has_c0=False)
    for z in range(0, n, 3):
        assert np.array_equal(outs[z], oc.keyswitch(lvl, cts[z, 1, : lvl + 1], evks[which[z]])), "level %d ciphertext %d" % (lvl, z)
    for e in evk_devs:
        gc.lib.B200_DevFree(e)


def _rerun(env_extra, k):
    import os, subprocess, sys
    env = dict(os.environ, **env_extra)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "tests/test_gpu_primitives.py", "-k", k],
                       cwd=root, env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("fused", ["0", "1"])
def test_fused_base_conversion_switch_gives_the_oracles_words(fused):
    """ORION_B200_FUSED_BC=1 forces the base conversions of the batched key switch into the transforms either side of them
    (csrc/ntt.cu intt2_prep_kernel + ntt1_bc_kernel: ModUp inside INTT-last-pass / NTT-first-pass, ModDown likewise) on every
    chain of CASES and BENCH_CASES -- FP64 and integer contractions, centred and plain, partial digits; =0 forces the
    un-fused kernels.  By default the library picks by size."""
    _rerun({"ORION_B200_FUSED_BC": fused, "ORION_B200_FUSED_IP": "1"}, "keyswitch_stages or batchdev")
    _rerun({"ORION_B200_FUSED_BC": fused, "ORION_B200_FUSED_IP": "0"}, "keyswitch_stages")


@pytest.mark.parametrize("fused", ["0", "1"])
def test_fused_inner_product_switch_gives_the_oracles_words(fused):
    """ORION_B200_FUSED_IP=1 forces the fused second-pass + inner-product kernel (ntt2_ip_kernel) on every chain of CASES,
    =0 forces the un-fused pair of launches; by default the library picks by size.  The switch is read per call, but the
    parity cases run in a child process so the setting cannot leak."""
    _rerun({"ORION_B200_FUSED_IP": fused}, "keyswitch_stages or rescale or batchdev")


def test_cluster_ntt_switch_gives_the_same_words():
    """ORION_B200_CLUSTER_NTT=1: whole-limb NTT in one thread-block-cluster launch (distributed shared memory between the
    strided and the contiguous pass, csrc/ntt.cu ntt_cluster) for 2^12 <= N <= 2^14.  The switch is read once per process,
    so the NTT parity cases (oracle and reference GPU kernel) are re-run in a child process with it on."""
    import os, subprocess, sys
    env = dict(os.environ, ORION_B200_CLUSTER_NTT="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "tests/test_gpu_primitives.py", "tests/test_gpu_refgpu.py",
                        "-k", "ntt_roundtrip or ntt_words or rescale or keyswitch_stages"], cwd=root, env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("setting", [{"ORION_B200_BC_TC": "0"}, {"ORION_B200_BC_TC": "1", "ORION_B200_BC_TC_RT": "1"},
                                     {"ORION_B200_BC_TC": "1", "ORION_B200_BC_TC_RT": "4", "ORION_B200_BC_TC_Z": "1"},
                                     {"ORION_B200_IP_ZFAST": "0"}])
def test_tensor_core_base_conversion_and_cta_order_switches_give_the_oracles_words(setting):
    """ORION_B200_BC_TC=0 forces the scalar base-conversion kernel, =1 (the default) the INT8 tensor-core contraction
    (csrc/ks_kernels.cu baseconv_tc_kernel, plain and centred) with 1 / 2 / 4 coefficient tiles per warp and with / without the
    destination split; ORION_B200_IP_ZFAST=0 restores the tile-fastest CTA order of the fused inner product.  Read once per
    process, so the base-conversion and key-switch cases (chains with 5-8 special primes, partial digits) run in a child process."""
    _rerun(setting, "baseconv or keyswitch_stages or batchdev")
