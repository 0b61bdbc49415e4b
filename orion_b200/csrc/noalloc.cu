// noalloc.cu -- B200_RescaleDevNoAlloc: the rescale of capi_raw.cu's B200_RescaleDev for callers that may not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// PolyEvaluator::rescale (evaluator.cu) takes two stream-ordered scratch buffers from the pool (the inverse transform of the
// last limb, [polys][N], and its residues modulo the remaining primes, [polys][level][N]) and keeps its launch descriptors in
// pooled device tables built on first use.  Here all three live in ONE file-scope device array of a fixed size:
//
//     g_noalloc_scratch : NOALLOC_MAX_WORDS data words (32 MiB) + NOALLOC_DESC_WORDS descriptor words (32 KiB)
//
// which every GPU that loads the library holds until the process exits.  The entry point calls no cudaMalloc / cudaFree /
// cudaMallocAsync / cudaFreeAsync and no allocating library routine: it launches one descriptor-fill kernel and the same
// four kernels the pooled path launches (ntt_pass x2, rescale_prep_kernel, ntt_pass, ntt2_submul_kernel), so its words are
// those of B200_RescaleDev bit for bit.
//
// Limit: polys * (level + 1) * N <= NOALLOC_MAX_WORDS (= 2^22 words: two polynomials of 32 limbs at logN = 16, or of
// 256 limbs at logN = 13); larger requests are rejected with an error code and a message, nothing is launched.
// Sharing: every call (of every context on that GPU) uses the same array, which keeps the previous call's contents.
// Nothing in it has to start at zero -- each region is fully written by this call before this call reads it -- but two
// calls that overlap on the GPU (different streams, or one re-launched from a graph while another runs) race on it.
// Calls on the library's stream are ordered and safe.
#include "evaluator.h"
#include "../../include/orion_b200.h"

namespace ob {
Context& raw_context(void* handle);        // capi_raw.cu
}
using namespace ob;

namespace {

constexpr size_t NOALLOC_MAX_WORDS = (size_t)1 << 22;     // 32 MiB of 64-bit words
constexpr size_t NOALLOC_DESC_WORDS = (size_t)1 << 12;    // 2048 LimbDesc (16 bytes each)
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_MAX_WORDS + NOALLOC_DESC_WORDS];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address is looked up once per device,
// with that device current, and only that pointer is handed to kernels.
constexpr int MAX_DEVICES = 64;
unsigned long long* g_scratch_ptr[MAX_DEVICES] = {nullptr};

unsigned long long* scratch_for_current_device() {
    int dev = -1;
    OB_CUDA(cudaGetDevice(&dev));
    OB_CHECK(dev >= 0 && dev < MAX_DEVICES, "B200_RescaleDevNoAlloc: device index out of range");
    if (g_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            OB_FATAL(std::string("B200_RescaleDevNoAlloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        g_scratch_ptr[dev] = (unsigned long long*)p;
    }
    return g_scratch_ptr[dev];
}

// Launch descriptors of the three transforms, written by the device (no host table, no copy):
//   d[0 .. polys)                     inverse transform, first pass : limb p*in_stride+level of `in` -> row p of `last`
//   d[polys .. 2 polys)               inverse transform, second pass: row p of `last`, in place
//   d[2 polys .. 2 polys + polys*level) forward first pass           : row p*level+i of `tmp`, in place, modulus i
__global__ void noalloc_fill_desc_kernel(LimbDesc* d, int level, int polys, int in_stride) {
    OB_PDL_SYNC();
    const int n = 2 * polys + polys * level;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        LimbDesc v;
        if (k < polys) v = LimbDesc{(uint32_t)(k * in_stride + level), (uint32_t)k, (uint32_t)level, 0u};
        else if (k < 2 * polys) v = LimbDesc{(uint32_t)(k - polys), (uint32_t)(k - polys), (uint32_t)level, 0u};
        else { const int r = k - 2 * polys; v = LimbDesc{(uint32_t)r, (uint32_t)r, (uint32_t)(r % level), 0u}; }
        d[k] = v;
    }
}

}  // namespace

extern "C" {

size_t B200_NoAllocMaxWords(void) { return NOALLOC_MAX_WORDS; }

// in [polys][level+1][N] -> out [polys][level][N], device buffers owned by the caller.  Returns 0, or -1 (nothing launched)
// when polys * (level+1) * N exceeds B200_NoAllocMaxWords() or an argument is out of range.
int B200_RescaleDevNoAlloc(void* h, int level, const void* in, void* out, int polys) {
    Context& c = raw_context(h);
    const size_t N = c.N;
    if (level < 1 || level >= c.nQ || polys < 1 || polys > 2 || in == nullptr || out == nullptr) {
        std::fprintf(stderr, "B200_RescaleDevNoAlloc: bad arguments (level %d of %d limbs, polys %d)\n", level, c.nQ, polys);
        return -1;
    }
    const size_t need = (size_t)polys * (size_t)(level + 1) * N;
    const size_t ndesc = (size_t)2 * polys + (size_t)polys * level;
    if (need > NOALLOC_MAX_WORDS || ndesc * 2 > NOALLOC_DESC_WORDS) {
        std::fprintf(stderr, "B200_RescaleDevNoAlloc: polys * (level + 1) * N = %zu words exceeds the static scratch of %zu words "
                             "(use B200_RescaleDev, which allocates)\n", need, (size_t)NOALLOC_MAX_WORDS);
        return -1;
    }
    unsigned long long* s = scratch_for_current_device();
    u64* last = (u64*)s;                                   // [polys][N]
    u64* tmp = last + (size_t)polys * N;                   // [polys][level][N]
    LimbDesc* d = (LimbDesc*)(s + NOALLOC_MAX_WORDS);
    // the descriptors are written by a kernel, and the transforms read their descriptors BEFORE griddepcontrol.wait (they are
    // host-built tables everywhere else): this sequence runs as ordinary, fully ordered launches
    PdlSuppress no_pdl;
    OB_LAUNCH(noalloc_fill_desc_kernel, 1, 128, 0, stream(), d, level, polys, level + 1);
    NttPlan inv; inv.first = d; inv.second = d + polys; inv.n = polys;
    NttPlan fwd; fwd.first = d + 2 * polys; fwd.second = d + 2 * polys; fwd.n = polys * level;
    ntt_inverse(c, (const u64*)in, last, inv);
    rescale_prep(c, level, last, tmp, polys);
    ntt_forward_first(c, tmp, tmp, fwd);
    ntt2_submul(c, level, polys, (const u64*)in, level + 1, tmp, c.d_qlinv.as<ulonglong2>() + (size_t)level * c.nQ, nullptr, 0, 0, 0, 0,
                (u64*)out, level);
    return 0;
}

}  // extern "C"
