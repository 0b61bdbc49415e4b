// ref_gpu_shim.cu -- thin extern "C" wrapper around the REFERENCE's own GPU merge NTT
// (GPU-NTT: src/lib/ntt_merge/ntt.cu, gpuntt::GPU_NTT_Inplace, the RNS overload declared at
// src/include/ntt_merge/ntt.cuh:198-201), compiled IN PLACE from /root/reference for sm_100a by
// oracle/Makefile into oracle/_ref/libgpuntt_refgpu.so.  No reference source is copied into this
// repository; this file only calls the reference's public functions and classes.
//
// TEST / MEASUREMENT INFRASTRUCTURE ONLY: (1) a second pin for the NTT row of SURVEY.md section 8 -- the
// product's GPU NTT must give the reference GPU kernel's words on the same (prime, psi) chain; (2) the
// "reference GPU kernel to beat" of SURVEY.md section 8(d), timed with CUDA events in place of nvbench
// (benchmark/bench_merge_ntt.cu needs nvbench, which is not in the image).
#include "ntt.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace gpuntt;

namespace {
struct RefGpu {
    int logn, nmod;
    Root<Data64>* fwd;        // [nmod][N]  psi powers, bit-reversed (the layout mod_index << N_power expects)
    Root<Data64>* inv;        // [nmod][N]
    Modulus<Data64>* mod;     // [nmod]
    Ninverse<Data64>* ninv;   // [nmod]
    std::vector<Data64> primes;
};
#define RG_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
    std::fprintf(stderr, "ref_gpu_shim: %s: %s\n", #x, cudaGetErrorString(e_)); std::abort(); } } while (0)
}  // namespace

extern "C" {

void* refgpu_create(int logn, int nmod, const unsigned long long* primes, const unsigned long long* psis)
{
    RefGpu* h = new RefGpu{logn, nmod, nullptr, nullptr, nullptr, nullptr, {}};
    h->primes.assign(primes, primes + nmod);
    const size_t n = (size_t)1 << logn;
    std::vector<Root<Data64>> f, b;
    std::vector<Modulus<Data64>> m;
    std::vector<Ninverse<Data64>> ni;
    for (int i = 0; i < nmod; ++i) {
        Modulus<Data64> q(primes[i]);
        Data64 omega = OPERATOR<Data64>::mult(psis[i], psis[i], q);
        NTTFactors<Data64> fac(q, omega, psis[i]);
        NTTParameters<Data64> p(logn, fac, ReductionPolynomial::X_N_plus);
        std::vector<Root<Data64>> ft = p.gpu_root_of_unity_table_generator(p.forward_root_of_unity_table);
        std::vector<Root<Data64>> bt = p.gpu_root_of_unity_table_generator(p.inverse_root_of_unity_table);
        if (ft.size() != n || bt.size() != n) { std::fprintf(stderr, "ref_gpu_shim: unexpected table size\n"); std::abort(); }
        f.insert(f.end(), ft.begin(), ft.end());
        b.insert(b.end(), bt.begin(), bt.end());
        m.push_back(q);
        ni.push_back(p.n_inv);
    }
    RG_CUDA(cudaMalloc(&h->fwd, f.size() * sizeof(Root<Data64>)));
    RG_CUDA(cudaMalloc(&h->inv, b.size() * sizeof(Root<Data64>)));
    RG_CUDA(cudaMalloc(&h->mod, m.size() * sizeof(Modulus<Data64>)));
    RG_CUDA(cudaMalloc(&h->ninv, ni.size() * sizeof(Ninverse<Data64>)));
    RG_CUDA(cudaMemcpy(h->fwd, f.data(), f.size() * sizeof(Root<Data64>), cudaMemcpyHostToDevice));
    RG_CUDA(cudaMemcpy(h->inv, b.data(), b.size() * sizeof(Root<Data64>), cudaMemcpyHostToDevice));
    RG_CUDA(cudaMemcpy(h->mod, m.data(), m.size() * sizeof(Modulus<Data64>), cudaMemcpyHostToDevice));
    RG_CUDA(cudaMemcpy(h->ninv, ni.data(), ni.size() * sizeof(Ninverse<Data64>), cudaMemcpyHostToDevice));
    return h;
}

void refgpu_destroy(void* hv)
{
    RefGpu* h = (RefGpu*)hv;
    cudaFree(h->fwd); cudaFree(h->inv); cudaFree(h->mod); cudaFree(h->ninv);
    delete h;
}

// batch polynomials laid out [batch][N]; polynomial i uses modulus i % nmod (ntt.cuh:174-190)
static void run(RefGpu* h, Data64* dev, int batch, int inverse, cudaStream_t s)
{
    ntt_rns_configuration<Data64> cfg = {};
    cfg.n_power = h->logn;
    cfg.ntt_type = inverse ? INVERSE : FORWARD;
    cfg.reduction_poly = ReductionPolynomial::X_N_plus;
    cfg.zero_padding = false;
    cfg.mod_inverse = h->ninv;
    cfg.stream = s;
    GPU_NTT_Inplace(dev, inverse ? h->inv : h->fwd, h->mod, cfg, batch, h->nmod);
}

// host buffers in and out: the parity entry point
void refgpu_ntt_host(void* hv, const unsigned long long* in, unsigned long long* out, int batch, int inverse)
{
    RefGpu* h = (RefGpu*)hv;
    const size_t words = ((size_t)batch) << h->logn;
    Data64* d;
    RG_CUDA(cudaMalloc(&d, words * 8));
    RG_CUDA(cudaMemcpy(d, in, words * 8, cudaMemcpyHostToDevice));
    run(h, d, batch, inverse, 0);
    RG_CUDA(cudaDeviceSynchronize());
    RG_CUDA(cudaMemcpy(out, d, words * 8, cudaMemcpyDeviceToHost));
    RG_CUDA(cudaFree(d));
}

// average milliseconds of one batched in-place transform over `reps` timed runs (3 warm-up runs), CUDA events on the
// launching stream.  `nbuf` distinct buffers are cycled so that consecutive runs do not hit L2 (pass a count whose
// total size exceeds the 126 MB L2).
float refgpu_time_ms(void* hv, int batch, int inverse, int reps, int nbuf)
{
    RefGpu* h = (RefGpu*)hv;
    const size_t words = ((size_t)batch) << h->logn;
    std::vector<Data64*> bufs(nbuf);
    std::vector<Data64> host(words);
    unsigned long long st = 1234;                              // seeded residues in [0, q_i), as bench_merge_ntt.cu:20-24 does
    for (size_t i = 0; i < words; ++i) {
        st = st * 6364136223846793005ull + 1442695040888963407ull;
        host[i] = (st >> 3) % h->primes[(i >> h->logn) % h->nmod];
    }
    for (int i = 0; i < nbuf; ++i) {
        RG_CUDA(cudaMalloc(&bufs[i], words * 8));
        RG_CUDA(cudaMemcpy(bufs[i], host.data(), words * 8, cudaMemcpyHostToDevice));
    }
    cudaStream_t s;
    RG_CUDA(cudaStreamCreate(&s));
    cudaEvent_t e0, e1;
    RG_CUDA(cudaEventCreate(&e0)); RG_CUDA(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) run(h, bufs[i % nbuf], batch, inverse, s);
    RG_CUDA(cudaStreamSynchronize(s));
    RG_CUDA(cudaEventRecord(e0, s));
    for (int i = 0; i < reps; ++i) run(h, bufs[i % nbuf], batch, inverse, s);
    RG_CUDA(cudaEventRecord(e1, s));
    RG_CUDA(cudaStreamSynchronize(s));
    float ms = 0.f;
    RG_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    RG_CUDA(cudaEventDestroy(e0)); RG_CUDA(cudaEventDestroy(e1)); RG_CUDA(cudaStreamDestroy(s));
    for (Data64* p : bufs) cudaFree(p);
    return ms / reps;
}

}  // extern "C"
