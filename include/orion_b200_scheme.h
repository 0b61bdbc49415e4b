/*
 * orion_b200_scheme.h -- Orion-level C ABI of liborion_b200.so.
 *
 * Same export names, argument order and conventions as the Go shim the reference binds with ctypes
 * (orion/backend/lattigo/*.go, signatures declared in orion/backend/lattigo/bindings.py:141-746):
 *   - one process-global scheme (lattigo/scheme.go:32), single-threaded by contract;
 *   - plaintexts / ciphertexts / polynomials / linear transforms are int ids from smallest-free-int tables
 *     (lattigo/minheap.go:46-65); in-place ops return the same id, *New ops a fresh id;
 *   - arrays are returned BY VALUE as {Data, Length} (lattigo/utils.go:57-73) and released with FreeCArray;
 *   - float payloads are float32, scales uint64 (lattigo/encoder.go:17-20);
 *   - errors abort the process with a message (the Go shim panics).
 * "LG/" below abbreviates orion/backend/lattigo/.  B200_* entries are additions (seeding, raw limb import/export
 * for parity tests and multi-GPU gathers); everything else replaces the cited Go export one for one.
 */
#ifndef ORION_B200_SCHEME_H
#define ORION_B200_SCHEME_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float* Data; unsigned long Length; } ArrayResultFloat;
typedef struct { double* Data; unsigned long Length; } ArrayResultDouble;
typedef struct { int* Data; unsigned long Length; } ArrayResultInt;
typedef struct { uint64_t* Data; unsigned long Length; } ArrayResultUInt64;
typedef struct { unsigned char* Data; unsigned long Length; } ArrayResultByte;

/* ---- scheme: LG/scheme.go:35-100, LG/utils.go:93 ---- */
void NewScheme(int logN, int* logQ, int lenQ, int* logP, int lenP, int logScale, int h, char* ringType, char* keysPath, char* ioMode);
void DeleteScheme(void);
void FreeCArray(void* p);
void B200_SetSeed(uint64_t seed);      /* deterministic keys / encryption randomness; default: OS randomness */

/* ---- tensors: LG/tensors.go:35-122 ---- */
void DeletePlaintext(int id);
void DeleteCiphertext(int id);
unsigned long GetPlaintextScale(int id);
unsigned long GetCiphertextScale(int id);
void SetPlaintextScale(int id, unsigned long scale);
void SetCiphertextScale(int id, unsigned long scale);
int GetPlaintextLevel(int id);
int GetCiphertextLevel(long id);
int GetPlaintextSlots(long id);
int GetCiphertextSlots(long id);
int GetCiphertextDegree(long id);
ArrayResultUInt64 GetModuliChain(void);
ArrayResultInt GetLivePlaintexts(void);
ArrayResultInt GetLiveCiphertexts(void);

/* ---- key generator: LG/keygenerator.go:13-59 ---- */
void NewKeyGenerator(void);
void GenerateSecretKey(void);
void GeneratePublicKey(void);
void GenerateRelinearizationKey(void);
void GenerateEvaluationKeys(void);
ArrayResultByte SerializeSecretKey(void);
void LoadSecretKey(unsigned char* data, unsigned long len);

/* ---- encoder / encryptor: LG/encoder.go:11-42, LG/encryptor.go:10-38 ---- */
void NewEncoder(void);
int Encode(float* values, int lenValues, int level, unsigned long scale);
ArrayResultFloat Decode(int ptID);
void NewEncryptor(void);
void NewDecryptor(void);
int Encrypt(int ptID);
int Decrypt(int ctID);

/* ---- evaluator: LG/evaluator.go:13-317 ---- */
void NewEvaluator(void);
void AddRotationKey(int rot);
int Negate(int ct);
int Rotate(int ct, int rot);
int RotateNew(int ct, int rot);
int Rescale(int ct);
int RescaleNew(int ct);
int AddScalar(int ct, float c);
int AddScalarNew(int ct, float c);
int SubScalar(int ct, float c);
int SubScalarNew(int ct, float c);
int MulScalarInt(int ct, int k);
int MulScalarIntNew(int ct, int k);
int MulScalarFloat(int ct, float c);
int MulScalarFloatNew(int ct, float c);
int AddPlaintext(int ct, int pt);
int AddPlaintextNew(int ct, int pt);
int SubPlaintext(int ct, int pt);
int SubPlaintextNew(int ct, int pt);
int MulPlaintext(int ct, int pt);
int MulPlaintextNew(int ct, int pt);
int AddCiphertext(int a, int b);
int AddCiphertextNew(int a, int b);
int SubCiphertext(int a, int b);
int SubCiphertextNew(int a, int b);
int MulRelinCiphertext(int a, int b);
int MulRelinCiphertextNew(int a, int b);
/* required by the drifted Python layer, defined only in orion/backend/heongpu/bindings.py:519,960 */
int CloneCiphertext(int ct);
int ModDropCiphertext(int ct);

/* ---- linear transforms: LG/lineartransform.go:26-211 ---- */
void NewLinearTransformEvaluator(void);
int GenerateLinearTransform(int* diagIdx, int nIdx, float* diagData, int nData, int level, float bsgsRatio, char* ioMode);
int EvaluateLinearTransform(int transform, int ct);
void DeleteLinearTransform(int id);
ArrayResultInt GetLinearTransformRotationKeys(int transform);       /* Galois elements */
void GenerateLinearTransformRotationKey(int galEl);
ArrayResultByte GenerateAndSerializeRotationKey(int galEl);
void LoadRotationKey(unsigned char* data, unsigned long len, unsigned long galEl);
ArrayResultByte SerializeDiagonal(int transform, int diag);
void LoadPlaintextDiagonal(unsigned char* data, unsigned long len, int transform, unsigned long diag);
void RemovePlaintextDiagonals(int transform);
void RemoveRotationKeys(void);
void GenerateConsolidatedRotationKeys(int* galEls, int n);           /* heongpu/bindings.py:1879 */

/* ---- polynomial evaluation: LG/polyeval.go:33-167 ---- */
void NewPolynomialEvaluator(void);
int GenerateMonomial(float* coeffs, int n);
int GenerateChebyshev(float* coeffs, int n);
int EvaluatePolynomial(int ct, int poly, unsigned long outScale);
int GetPolyDepth(int poly);                                           /* python/poly_evaluator.py:59 */
int B200_GeneratePolynomialF64(const double* coeffs, int n, int chebyshev);   /* float64 coefficients */
/* GenerateMinimaxSignCoeffs (LG/polyeval.go:91-167) is host-only multi-precision Remez: it is provided by the Python
 * binding layer (orion_b200/minimax.py, B200Library.GenerateMinimaxSignCoeffs), not by this shared library. */

/* ---- bootstrapping: LG/bootstrapper.go:19-93 ---- */
void NewBootstrapper(int* logPs, int n, int numSlots);
int Bootstrap(int ct, int numSlots);
void DeleteBootstrappers(void);

/* ---- additions: raw limbs in/out (NTT domain, [degree+1][level+1][N] uint64) ---- */
int    B200_ImportCiphertext(const uint64_t* data, int level, int degree, double scale, int logSlots);
void   B200_ExportCiphertext(int ct, uint64_t* out);
void*  B200_CiphertextDevicePtr(int ct);
/* multi-GPU exchange on device memory: receive buffer for a remote ciphertext, stream-ordered access for NCCL (no host
 * synchronisation), exact scale as hi + lo doubles.  Replaces the gather of SURVEY.md 8(e) (ncclAllGather of layer outputs). */
int    B200_NewCiphertext(int level, int degree, double scaleHi, double scaleLo, int logSlots);
void*  B200_CiphertextAcquire(int ct, int for_write);
void   B200_GetCiphertextScaleParts(int ct, double* hi, double* lo);
int    B200_GetCiphertextLogSlots(int ct);
double B200_GetCiphertextScaleF64(int ct);
void   B200_SetCiphertextScaleF64(int ct, double scale);
int    B200_ImportPlaintext(const uint64_t* data, int level, double scale);
void   B200_ExportPlaintext(int pt, uint64_t* out);
void   B200_ExportSecretKey(int64_t* coeffs);
void   B200_ImportSecretKey(const int64_t* coeffs);
void   B200_ExportRelinKey(uint64_t* out);                  /* [beta][2][nQ+nP][N], standard residues */
void   B200_ExportGaloisKey(uint64_t galEl, uint64_t* out);
void   B200_ExportPublicKey(uint64_t* out);                 /* [2][nQ+nP][N] */
int    B200_SchemeInfo(int* logN, int* nQ, int* nP, uint64_t* moduli);
int    B200_EncodeF64(const double* values, int n, int level, double scale);
void   B200_DecodeF64(int pt, double* out_re, double* out_im);
int    B200_Conjugate(int ct);
void   B200_NoteEvaluatorKey(uint64_t galEl);

#ifdef __cplusplus
}
#endif
#endif
