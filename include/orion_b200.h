/*
 * orion_b200.h -- C ABI of liborion_b200.so, the B200 (sm_100a) CKKS evaluation backend for Orion.
 *
 * Two groups of entry points:
 *
 *  (1) Primitive level, "B200_*": explicit modulus chains, batched NTT, base conversion, key switching and
 *      rescale on host or device buffers.  This is what the primitive sweep of BASELINE.json (configs[4]:
 *      NTT / rotation / key switch at logN 13-16) and the kernel parity tests drive.  The reference has no
 *      such C surface for Lattigo; the nearest reference interfaces are HEonGPU's C API
 *      (orion/backend/heongpu/adrianHEonGPU/src/heongpu/heongpu_c_api/operator_c_api.h:26-82) and GPU-NTT's
 *      GPU_NTT_Inplace (thirdparty/GPU-NTT/src/include/ntt_merge/ntt.cuh).
 *
 *  (2) Orion level (orion_b200_scheme.h, included below): the exact export names and argument conventions of
 *      the Go shim orion/backend/lattigo/*.go that orion/backend/lattigo/bindings.py binds with ctypes.
 *
 * Plain pointers and sizes only; no C++ or torch types.  Words are uint64_t.  All polynomial buffers are
 * [limb][N] row-major.  Every function aborts the process with a message on error, as the Go shim panics
 * (orion/backend/lattigo/evaluator.go:52-54).  There is no CPU fallback: without a CUDA device any call that
 * computes aborts.
 */
#ifndef ORION_B200_H
#define ORION_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- host-only helpers (usable without a GPU) ---- */
/* Prime chain for (LogQ, LogP): replaces ckks.NewParametersFromLiteral at lattigo/scheme.go:57-65. */
int      B200_GenModuli(const int* logQ, int nQ, const int* logP, int nP, int logNthRoot, uint64_t* Q, uint64_t* P);
uint64_t B200_FindPsi(uint64_t q, uint64_t nthroot);
int      B200_DeviceCount(void);

/* ---- context over an explicit chain Q_0..Q_{nQ-1}, P_0..P_{nP-1} (chain index = position) ---- */
void*    B200_NewContext(int logN, const uint64_t* Q, int nQ, const uint64_t* P, int nP);
void     B200_DeleteContext(void* ctx);
uint64_t B200_ContextPsi(void* ctx, int chain_idx);
uint64_t B200_GaloisElement(void* ctx, int rot);            /* 5^rot mod 2N; lattigo/evaluator.go:34-36 */

/* ---- stream, timing, memory ---- */
void     B200_Sync(void);
unsigned long long B200_LaunchCount(void);                   /* kernels launched by this library so far */
double   B200_AlgorithmicBytes(void);          /* running sum of SURVEY 8(d) algorithmic bytes of the evaluator ops issued so far */
double   B200_DeviceBytes(int peak);           /* device bytes held by the library's buffers now (0) or at their high-water mark (1) */
void     B200_TrimPool(void);                  /* synchronise and hand the pool's unused blocks back to the driver (before another process needs the GPU's memory) */
double   B200_ReclaimedBytes(void);            /* bytes of regenerable state (Galois keys, bootstrap factor sets) dropped under memory pressure so far */
void*    B200_Stream(void);                                  /* cudaStream_t all work is issued on */
void     B200_TimerStart(void);                              /* CUDA events on that stream */
float    B200_TimerStopMs(void);
/* Kernel probe: while enabled, every launch of the key switch's dominant kernel (fused second NTT pass + evaluation-key
 * inner product, FP64-path limbs) is bracketed by CUDA events on its own stream; ReadMs returns the summed kernel time
 * and the launch count since the last read.  Measurement aid for bench.py's roofline object. */
void     B200_HostProfDump(void);                            /* ORION_B200_HOSTPROF=1: print and reset the host-side enqueue timers */
void     B200_ProbeEnable(int mode);              /* 0 off, 1 the dominant kernel only (slot 0), 2 every key-switch stage */
double   B200_ProbeReadMs(long* launches);        /* slot 0 */
int      B200_ProbeSlots(void);
const char* B200_ProbeSlotName(int slot);
double   B200_ProbeReadSlotMs(int slot, long* launches);
void*    B200_DevAlloc(size_t words);
void     B200_DevFree(void* dev);
void     B200_Upload(void* dev, const uint64_t* host, size_t words);
void     B200_UploadAsync(void* dev, const uint64_t* host, size_t words);
void     B200_Download(uint64_t* host, const void* dev, size_t words);
void*    B200_HostAllocPinned(size_t words);
void     B200_HostFreePinned(void* p);

/* ---- negacyclic NTT (natural in, bit-reversed out) / INTT over n limbs; limb i uses chain index mod_idx[i].
 *      Same transform as GPU-NTT's CPU reference, src/lib/ntt_merge/ntt_cpu.cu:81-185. ---- */
void B200_NTT(void* ctx, const uint64_t* in, uint64_t* out, const int* mod_idx, int n, int inverse);
void B200_NTTDev(void* ctx, const void* in, void* out, const int* mod_idx, int n, int inverse);

/* ---- RNS fast base conversion, coefficient domain: in [ns][N] -> out [nd][N]
 *      (replaces HEonGPU base_conversion_DtoQtilde_*: lib/kernel/switchkey.cu:698-872, with Lattigo's rounding). ---- */
void B200_BaseConv(void* ctx, const int* src, int ns, const int* dst, int nd, const uint64_t* in, uint64_t* out, int centered);

/* ---- NTT-domain automorphism X -> X^galEl on n limbs (lattigo/evaluator.go:61-81) ---- */
void B200_AutomorphismNTT(void* ctx, uint64_t galEl, const uint64_t* in, uint64_t* out, int n);
void     B200_AutomorphismNTTDev(void* ctx, uint64_t galEl, const void* in_dev, void* out_dev, int n);   /* device buffers, no copies (primitive sweep) */

/* ---- hybrid key switching (lattigo/evaluator.go:61-81, :297-317; HEonGPU operator.cu:1308-1429).
 *      evk host layout: [beta(max level)][2][nQ+nP][N], NTT domain, standard residues. ---- */
void* B200_EvkToDevice(void* ctx, const uint64_t* evk_host);  /* free with B200_DevFree */
/* out [2][level+1][N] = KeySwitch(c1 [level+1][N]); addend_c0 (may be NULL) is added to polynomial 0 */
void  B200_KeySwitch(void* ctx, int level, const uint64_t* c1, const void* evk_dev, const uint64_t* addend_c0, uint64_t* out);
void  B200_KeySwitchDev(void* ctx, int level, const void* c1, const void* evk_dev, const void* addend_c0, void* out);
/* n independent key switches on HOST buffers in one call, H2D / compute / D2H overlapped on three streams.
 * ct_in[i]: [2][level+1][N] = (c0, c1) when has_c0 (out = KeySwitch(c1) + (c0, 0)), else [level+1][N] = c1.
 * This is the host-buffer entry point a cgo/ctypes caller of lattigo/evaluator.go:61-81 (RotateNew on a batch of
 * ciphertexts) would bind; out[i]: [2][level+1][N]. */
void  B200_KeySwitchBatch(void* ctx, int level, int n, const uint64_t* const* ct_in, int has_c0, const void* const* evk_dev,
                          uint64_t* const* out);
/* The same batch on device buffers: cts [n][2][level+1][N] (or [n][level+1][N] without c0), out [n][2][level+1][N];
 * every kernel of the key switch is launched once for the whole batch. */
void  B200_KeySwitchBatchDev(void* ctx, int level, int n, const void* cts, int has_c0, const void* const* evk_dev, void* out);
void  B200_Decompose(void* ctx, int level, const uint64_t* c1, uint64_t* digits);   /* [beta][level+1+nP][N] */
void  B200_ModDown(void* ctx, int level, const uint64_t* accQP, uint64_t* outQ);

/* ---- rescale by the last prime (lattigo/evaluator.go:84-99; HEonGPU operator.cu:906-980) ---- */
void B200_Rescale(void* ctx, int level, const uint64_t* in, uint64_t* out, int polys);
void B200_RescaleDev(void* ctx, int level, const void* in, void* out, int polys);
/* No counterpart in the reference.  The rescale of B200_RescaleDev for callers
 * that may not allocate (CUDA graph capture, serving loops): its scratch and launch descriptors live in one static device
 * array (32 MiB + 32 KiB per GPU, held until exit), it calls no cudaMalloc / cudaFree / pool routine, and gives the same
 * words.  Limit: polys * (level + 1) * N <= B200_NoAllocMaxWords() (2^22); beyond it returns -1 with a message and launches
 * nothing (0 = ok).  All calls share the array: calls that overlap on the GPU race; calls on the library's stream do not. */
int      B200_RescaleDevNoAlloc(void* ctx, int level, const void* in, void* out, int polys);
size_t   B200_NoAllocMaxWords(void);

/* ---- element-wise limb arithmetic on n limbs (chain index = limb index): 0 add, 1 sub, 2 neg, 3 mul ---- */
void B200_Elementwise(void* ctx, int op, const uint64_t* a, const uint64_t* b, uint64_t* out, int n);

#ifdef __cplusplus
}
#endif
#endif
