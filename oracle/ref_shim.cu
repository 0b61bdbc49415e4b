// ref_shim.cu -- thin extern "C" wrapper around the REFERENCE's own CPU NTT
// (GPU-NTT: src/lib/ntt_merge/ntt_cpu.cu, src/lib/common/nttparameters.cu), compiled IN PLACE from
// /root/reference by oracle/Makefile into oracle/_ref/libgpuntt_ref.so.  No reference source is
// copied into this repository; this file only calls the reference's public classes.
// TEST INFRASTRUCTURE ONLY: used to pin oracle/ckks_oracle.c's NTT and to generate tests/golden/ntt_kat.npz.
#include "ntt_cpu.cuh"
#include <cstring>
using namespace gpuntt;

extern "C" {

// the in-tree known-answer pool for this logN (nttparameters.cu:83-160): prime, omega, psi
void ref_ntt_pool(int logn, unsigned long long* prime, unsigned long long* omega, unsigned long long* psi)
{
    NTTParameters<Data64> p(logn, ReductionPolynomial::X_N_plus);
    *prime = p.modulus.value; *omega = p.omega; *psi = p.psi;
}

static NTTParameters<Data64> make(int logn, unsigned long long prime, unsigned long long psi)
{
    Modulus<Data64> m(prime);
    Data64 omega = OPERATOR<Data64>::mult(psi, psi, m);
    NTTFactors<Data64> f(m, omega, psi);
    return NTTParameters<Data64>(logn, f, ReductionPolynomial::X_N_plus);
}

void ref_ntt_forward(int logn, unsigned long long prime, unsigned long long psi,
                     const unsigned long long* in, unsigned long long* out)
{
    NTTCPU<Data64> cpu(make(logn, prime, psi));
    std::vector<Data64> v(in, in + (1 << logn));
    std::vector<Data64> r = cpu.ntt(v);
    std::memcpy(out, r.data(), sizeof(Data64) * r.size());
}

void ref_ntt_inverse(int logn, unsigned long long prime, unsigned long long psi,
                     const unsigned long long* in, unsigned long long* out)
{
    NTTCPU<Data64> cpu(make(logn, prime, psi));
    std::vector<Data64> v(in, in + (1 << logn));
    std::vector<Data64> r = cpu.intt(v);
    std::memcpy(out, r.data(), sizeof(Data64) * r.size());
}

void ref_schoolbook(int n, unsigned long long prime, const unsigned long long* a,
                    const unsigned long long* b, unsigned long long* out)
{
    Modulus<Data64> m(prime);
    std::vector<Data64> va(a, a + n), vb(b, b + n);
    std::vector<Data64> r = schoolbook_poly_multiplication<Data64>(va, vb, m, ReductionPolynomial::X_N_plus);
    std::memcpy(out, r.data(), sizeof(Data64) * n);
}

}  // extern "C"
