// scheme_kernels.h -- device kernels behind encode / decode / key generation / encryption (scheme_kernels.cu).
#pragma once
#include "context.h"

namespace ob {

// ---- deterministic counter-based sampling (same generator as the oracle: oracle/ckks_oracle.c orc_rng_u64) ----
__host__ __device__ inline u64 rng_mix64(u64 z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL; z ^= z >> 27; z *= 0x94D049BB133111EBULL; z ^= z >> 31; return z;
}
__host__ __device__ inline u64 rng_u64(u64 seed, u64 stream, u64 index) {
    u64 x = rng_mix64(seed ^ (stream * 0xD6E8FEB86659FD93ULL));
    x = rng_mix64(x ^ (index * 0x9E3779B97F4A7C15ULL));
    return rng_mix64(x + 0x9E3779B97F4A7C15ULL);
}

// out[t][x] uniform in [0, q_chain(t)) for the n limbs listed in mods (chain indices); stream = stream0 + t
void sample_uniform(Context& c, u64* out, const std::vector<int>& mods, u64 seed, u64 stream0);
// out[x] = discrete Gaussian (sigma 3.2, |x| <= 19) as int64
void sample_gaussian(Context& c, int64_t* out, u64 seed, u64 stream);
// host: ternary with exactly h non-zeros
void sample_ternary_hw_host(int N, int h, u64 seed, u64 stream, std::vector<int64_t>& out);
// out[t][x] = small[x] mod q_chain(mods[t])   (signed small coefficients -> RNS, coefficient domain)
void small_to_rns(Context& c, const int64_t* small, u64* out, const std::vector<int>& mods);

// ---- canonical-embedding FFT (complex128, "special" FFT over the 5^j rotation group) ----
struct FftTables {
    int log_slots = 0;
    DevBuf roots;       // double2[M+1], M = 4*slots : exp(2*pi*i*k/M)
    DevBuf rot_group;   // int[slots]      : 5^j mod M
};
const FftTables& fft_tables(Context& c, int log_slots);
// in place on `batch` vectors of `slots` complex128 (double2), stride = slots
void special_ifft(Context& c, double2* vals, int log_slots, int batch);
void special_fft(Context& c, double2* vals, int log_slots, int batch);

// ---- encode: complex slots -> RNS coefficients (coefficient domain) ----
// coeffs[b][t][i] for t in mods: real parts to i*gap, imaginary parts to N/2 + i*gap (gap = N/(2*slots)),
// value = round_half_away(x * scale) reduced mod q_t.  vals: [batch][slots] double2 (after special_ifft).
void encode_round_to_rns(Context& c, const double2* vals, int log_slots, double scale, u64* coeffs,
                         const std::vector<int>& mods, int batch);
// ---- decode: RNS coefficients (coefficient domain, limbs 0..level) -> centred value / scale as complex slots ----
void decode_crt_to_complex(Context& c, const u64* coeffs, int level, double scale, double2* vals, int log_slots);

void host_roots(int M, std::vector<double2>& r);
void drop_scheme_tables(Context& c);

}  // namespace ob
