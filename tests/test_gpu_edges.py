"""Edge cases and the `io_mode: save / load` entry points of the drop-in boundary, through the C ABI.

Covers what the reference's Python layer does around the hot path: key and diagonal blobs written to HDF5 and loaded back
(PY/key_generator.py:17-31, PY/lt_evaluator.py:93-153,283-315 -> LG/keygenerator.go:38-59, LG/lineartransform.go:131-193),
rotations by 0 and by a whole turn, operations at level 0, degenerate plaintexts.  The blobs are this backend's own
little-endian containers; the contract tested here is the reference's: what was
serialised loads back and computes the same ciphertext words.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LOGN, LOGQ, LOGP, LOGSCALE, H, SEED = 12, [45, 30, 30, 30], [45, 45], 30, 64, 21


@pytest.fixture()
def B(b200lib):
    from orion_b200.bindings import B200Library
    B = B200Library()
    B.B200_SetSeed(SEED)
    B.NewScheme(LOGN, LOGQ, LOGP, LOGSCALE, H, "standard", "", "none")
    B.NewKeyGenerator(); B.GenerateSecretKey(); B.GeneratePublicKey(); B.GenerateRelinearizationKey(); B.GenerateEvaluationKeys()
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator(); B.NewLinearTransformEvaluator()
    yield B
    B.DeleteScheme()


def words(B, ct):
    level, deg = B.GetCiphertextLevel(ct), B.GetCiphertextDegree(ct)
    out = np.empty((deg + 1, level + 1, 1 << LOGN), dtype=np.uint64)
    B.B200_ExportCiphertext(ct, out)
    return out


def dec(B, ct):
    p = B.Decrypt(ct)
    out = np.array(B.Decode(p))
    B.DeletePlaintext(p)
    return out


def free_blob(B, ptr):
    B.lib.FreeCArray(ctypes.cast(ptr, ctypes.c_void_p))


def test_rotation_by_zero_and_by_a_whole_turn(B):
    slots = 1 << (LOGN - 1)
    x = np.random.default_rng(1).uniform(-1, 1, slots).astype(np.float32)
    ct = B.Encrypt(B.Encode(x.tolist(), len(LOGQ) - 1, 1 << LOGSCALE))
    r0 = B.RotateNew(ct, 0)
    assert r0 != ct and np.array_equal(words(B, r0), words(B, ct))          # a copy, bit for bit (Galois element 1)
    B.AddRotationKey(slots)                                                  # 5^slots = 1 mod 2N: a whole turn
    rs = B.RotateNew(ct, slots)
    assert np.max(np.abs(dec(B, rs) - x)) < 1e-4
    rm = B.RotateNew(ct, -1)
    r1 = B.RotateNew(rm, 1)
    assert np.max(np.abs(dec(B, r1) - x)) < 1e-4
    assert np.max(np.abs(dec(B, rm) - np.roll(x, 1))) < 1e-4


def test_level_zero_and_degenerate_plaintexts(B):
    slots = 1 << (LOGN - 1)
    rng = np.random.default_rng(2)
    x = rng.uniform(-1, 1, slots).astype(np.float32)
    scale = 1 << LOGSCALE
    ct = B.Encrypt(B.Encode(x.tolist(), 0, scale))                           # a ciphertext born at level 0 (one limb)
    assert B.GetCiphertextLevel(ct) == 0
    y = rng.uniform(-1, 1, slots).astype(np.float32)
    s = B.AddPlaintextNew(ct, B.Encode(y.tolist(), 0, scale))
    assert np.max(np.abs(dec(B, s) - (x + y))) < 1e-4
    a = B.AddScalarNew(ct, -0.25)
    assert np.max(np.abs(dec(B, a) - (x - 0.25))) < 1e-4
    m = B.MulScalarIntNew(ct, 0)                                             # times zero: an encryption of zero, words all 0
    assert not words(B, m).any()
    r = B.RotateNew(ct, 3)                                                   # key switch at level 0: a single digit
    assert np.max(np.abs(dec(B, r) - np.roll(x, -3))) < 1e-4
    # all-zero and extreme plaintexts
    z = B.Encode([0.0] * slots, 2, scale)
    zp = np.empty((3, 1 << LOGN), dtype=np.uint64)
    B.B200_ExportPlaintext(z, zp)
    assert not zp.any()
    big = np.full(slots, 1000.0, dtype=np.float32)
    cb = B.Encrypt(B.Encode(big.tolist(), len(LOGQ) - 1, scale))
    assert np.max(np.abs(dec(B, cb) - big)) < 1e-2
    # rescale all the way down: level 3 -> 0, scale tracked
    top = len(LOGQ) - 1
    c = B.Encrypt(B.Encode(x.tolist(), top, scale))
    for lvl in range(top, 0, -1):
        c = B.MulScalarFloatNew(c, 0.5)                                      # multiplies by round(0.5 * q_level); rescale divides by q_level
        c = B.Rescale(c)
        assert B.GetCiphertextLevel(c) == lvl - 1
    assert np.max(np.abs(dec(B, c) - x * 0.5 ** top)) < 1e-3
    k = B.MulScalarFloatNew(c, 3.0)                                          # integer-valued constants keep the scale (no rescale needed)
    assert B.GetCiphertextScale(k) == B.GetCiphertextScale(c)
    assert np.max(np.abs(dec(B, k) - 3.0 * x * 0.5 ** top)) < 1e-3


def test_secret_key_blob_roundtrip(B):
    slots = 1 << (LOGN - 1)
    x = np.random.default_rng(3).uniform(-1, 1, slots).astype(np.float32)
    ct = B.Encrypt(B.Encode(x.tolist(), 2, 1 << LOGSCALE))
    ref = words(B, ct)
    blob, ptr = B.SerializeSecretKey()
    blob = blob.copy()
    free_blob(B, ptr)
    # a fresh scheme with the same parameters and the loaded key decrypts what the first one encrypted
    B.DeleteScheme()
    B.B200_SetSeed(SEED + 1)                                                 # different randomness: the key must come from the blob
    B.NewScheme(LOGN, LOGQ, LOGP, LOGSCALE, H, "standard", "", "none")
    B.NewKeyGenerator(); B.LoadSecretKey(blob)
    B.NewEncoder(); B.NewEncryptor(); B.NewDecryptor(); B.NewEvaluator()
    ct2 = B.B200_ImportCiphertext(np.ascontiguousarray(ref), 2, 1, float(1 << LOGSCALE), LOGN - 1)
    assert np.max(np.abs(dec(B, ct2) - x)) < 1e-4


def test_rotation_key_and_diagonal_blobs_roundtrip(B):
    """io_mode save -> load: serialise every rotation key and diagonal of a transform, drop them, load them back, and
    require the SAME output words as before."""
    slots = 1 << (LOGN - 1)
    rng = np.random.default_rng(4)
    level, scale = 2, 1 << LOGSCALE
    x = rng.uniform(-1, 1, slots).astype(np.float32)
    diag_idx = [0, 1, 2, 3, 5, 8, 64, slots - 1]
    flat = np.concatenate([rng.uniform(-1, 1, slots).astype(np.float32) for _ in diag_idx])
    ct = B.Encrypt(B.Encode(x.tolist(), level, scale))
    lt = B.GenerateLinearTransform(diag_idx, flat, level, 2.0, "none")
    gal = sorted(B.GetLinearTransformRotationKeys(lt))
    key_blobs = {}
    for g in gal:                                                            # PY/lt_evaluator.py:283-295 (save path)
        blob, ptr = B.GenerateAndSerializeRotationKey(g)
        key_blobs[g] = blob.copy()
        free_blob(B, ptr)
    for g in gal:
        B.LoadRotationKey(key_blobs[g], g)
    want = words(B, B.EvaluateLinearTransform(lt, ct))
    diag_blobs = {}
    for d in diag_idx:                                                       # LG/lineartransform.go:162-176
        blob, ptr = B.SerializeDiagonal(lt, d)
        diag_blobs[d] = blob.copy()
        free_blob(B, ptr)
    B.RemovePlaintextDiagonals(lt)                                           # LG/lineartransform.go:196-201
    B.RemoveRotationKeys()                                                   # :204-211
    for d in diag_idx:
        B.LoadPlaintextDiagonal(diag_blobs[d], lt, d)
    for g in gal:
        B.LoadRotationKey(key_blobs[g], g)
    got = words(B, B.EvaluateLinearTransform(lt, ct))
    assert np.array_equal(got, want)
    # io_mode "load" (LG/lineartransform.go:76-84): a transform created WITHOUT diagonal data -- the reference passes the
    # index list only -- gets its diagonals from the blobs and gives the same words
    lt2 = B.GenerateLinearTransform(diag_idx, np.zeros(0, dtype=np.float32), level, 2.0, "load")
    assert sorted(B.GetLinearTransformRotationKeys(lt2)) == gal
    for d in diag_idx:
        B.LoadPlaintextDiagonal(diag_blobs[d], lt2, d)
    assert np.array_equal(words(B, B.EvaluateLinearTransform(lt2, ct)), want)
    B.DeleteLinearTransform(lt2)
    B.DeleteLinearTransform(lt)
