// scheme_kernels.h -- device kernels behind encode / decode / key generation / encryption (scheme_kernels.cu).
#pragma once
#include "context.h"

namespace ob {

// ---- counter-based sampling ----
// Two generators behind one call, rng_u64(rng, stream, index):
//   * ChaCha20 (RFC 8439 block function; 256-bit key, block counter = index, nonce = stream) -- the DEFAULT.  NewScheme
//     draws two independent 256-bit keys from the OS (std::random_device): one for public randomness (the uniform `a`
//     halves of the public / evaluation keys), one for secret material (secret key, errors, encryption masks), so that
//     published words say nothing about the stream the secrets come from.  Lattigo does the same with a blake2b-keyed
//     XOF seeded from crypto/rand.
//   * three chained splitmix64 finalisers keyed by one 64-bit seed -- TEST MODE ONLY, entered through B200_SetSeed: the
//     CPU oracle implements the same function (oracle/ckks_oracle.c orc_rng_u64), which is what makes keys and
//     ciphertexts comparable bit for bit.  It is invertible from a single output word and must never protect real data.
struct Rng {
    u64 seed = 0;                 // test mode
    uint32_t key[8] = {0};        // ChaCha20 key
    uint32_t chacha = 0;          // 0: splitmix test mode, 1: ChaCha20
};
__host__ __device__ inline u64 rng_mix64(u64 z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL; z ^= z >> 27; z *= 0x94D049BB133111EBULL; z ^= z >> 31; return z;
}
__host__ __device__ inline u64 rng_splitmix(u64 seed, u64 stream, u64 index) {
    u64 x = rng_mix64(seed ^ (stream * 0xD6E8FEB86659FD93ULL));
    x = rng_mix64(x ^ (index * 0x9E3779B97F4A7C15ULL));
    return rng_mix64(x + 0x9E3779B97F4A7C15ULL);
}
__host__ __device__ inline uint32_t rng_rotl(uint32_t v, int n) { return (v << n) | (v >> (32 - n)); }
#define OB_CHACHA_QR(a, b, c, d) \
    a += b; d ^= a; d = rng_rotl(d, 16); c += d; b ^= c; b = rng_rotl(b, 12); \
    a += b; d ^= a; d = rng_rotl(d, 8);  c += d; b ^= c; b = rng_rotl(b, 7);
// words 0 and 1 of the ChaCha20 block (key, counter = index, nonce = stream)
__host__ __device__ inline u64 rng_chacha(const uint32_t (&key)[8], u64 stream, u64 index) {
    uint32_t x0 = 0x61707865u, x1 = 0x3320646eu, x2 = 0x79622d32u, x3 = 0x6b206574u;
    uint32_t x4 = key[0], x5 = key[1], x6 = key[2], x7 = key[3], x8 = key[4], x9 = key[5], x10 = key[6], x11 = key[7];
    uint32_t x12 = (uint32_t)index, x13 = (uint32_t)(index >> 32), x14 = (uint32_t)stream, x15 = (uint32_t)(stream >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        OB_CHACHA_QR(x0, x4, x8, x12) OB_CHACHA_QR(x1, x5, x9, x13) OB_CHACHA_QR(x2, x6, x10, x14) OB_CHACHA_QR(x3, x7, x11, x15)
        OB_CHACHA_QR(x0, x5, x10, x15) OB_CHACHA_QR(x1, x6, x11, x12) OB_CHACHA_QR(x2, x7, x8, x13) OB_CHACHA_QR(x3, x4, x9, x14)
    }
    return ((u64)(x1 + 0x3320646eu) << 32) | (u64)(x0 + 0x61707865u);
}
__host__ __device__ inline u64 rng_u64(const Rng& r, u64 stream, u64 index) {
    return r.chacha ? rng_chacha(r.key, stream, index) : rng_splitmix(r.seed, stream, index);
}
// the two generators of a scheme: `pub` for published uniform polynomials, `sec` for sk / errors / encryption masks
struct SchemeRng {
    Rng pub, sec;
    static SchemeRng test_mode(u64 seed);         // both = splitmix(seed): the oracle's generator (B200_SetSeed)
    static SchemeRng from_os();                   // two fresh 256-bit ChaCha20 keys from std::random_device
    SchemeRng derive(u64 tweak) const;            // generator of a dependent scheme (bootstrapping chain): seed ^ tweak in test mode, fresh OS keys otherwise
};

// out[t][x] uniform in [0, q_chain(t)) for the n limbs listed in mods (chain indices); stream = stream0 + t
void sample_uniform(Context& c, u64* out, const std::vector<int>& mods, const Rng& rng, u64 stream0);
// out[x] = discrete Gaussian (sigma 3.2, |x| <= 19) as int64
void sample_gaussian(Context& c, int64_t* out, const Rng& rng, u64 stream);
// host: ternary with exactly h non-zeros
void sample_ternary_hw_host(int N, int h, const Rng& rng, u64 stream, std::vector<int64_t>& out);
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
