// scheme_kernels.cu -- sampling, canonical-embedding FFT, encode rounding and decode CRT kernels.
// Behind LG/encoder.go:16-42 (Encode/Decode), LG/encryptor.go:20-38, LG/keygenerator.go:13-35; HEonGPU analogues:
// lib/kernel/encoding.cu:43-260, lib/kernel/keygeneration.cu:13-776, GPU-FFT src/lib/fft.cu (Special_*Core).
// float64 only, compiled with --fmad=false: every operation is a single IEEE op in a fixed order so that the
// oracle (gcc -ffp-contract=off) reproduces the encodings bit for bit.
#include "scheme_kernels.h"
#include <random>
#include <cmath>

namespace ob {

namespace {

constexpr int TPB = 256;

__constant__ u64 c_gauss_cdt[20] = {
    0x1FEA56814B22FE3AULL, 0x5CB44BA97E1C801CULL, 0x9135C4D2AB6097ACULL, 0xBA57841EC437E30AULL,
    0xD790D29879901FE9ULL, 0xEA65B1358D74FB9DULL, 0xF5672D96E886CF31ULL, 0xFB3C9D22CC87D60FULL,
    0xFE0A92B033D640A2ULL, 0xFF439D38007352AEULL, 0xFFBF67E425D189D1ULL, 0xFFEBCDF3A4629F98ULL,
    0xFFFA3F2F78880D45ULL, 0xFFFE81F40111A23BULL, 0xFFFFA5D06BC38AFBULL, 0xFFFFECA4870DFE3CULL,
    0xFFFFFC3B68ED3B4AULL, 0xFFFFFF58143B9611ULL, 0xFFFFFFE850CE70A6ULL, 0xFFFFFFFFFFFFFFFFULL};

__global__ void __launch_bounds__(TPB)
uniform_kernel(u64* __restrict__ out, int N, const LimbDesc* __restrict__ limbs, const ModConst* __restrict__ mc,
               const Rng rng, u64 stream0) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, x = blockIdx.x * TPB + threadIdx.x;
    const ModConst& m = mc[limbs[t].mod];
    const u64 q = m.q, mask = m.bits >= 64 ? ~0ull : ((1ull << m.bits) - 1);
    u64 r = 0; int a = 0;
    for (; a < 64; ++a) { r = rng_u64(rng, stream0 + t, ((u64)a << 32) | (u64)x) & mask; if (r < q) break; }
    out[(size_t)t * N + x] = a < 64 ? r : r % q;
}

__global__ void __launch_bounds__(TPB)
gaussian_kernel(int64_t* __restrict__ out, const Rng rng, u64 stream) {
    OB_PDL_SYNC();
    const int x = blockIdx.x * TPB + threadIdx.x;
    const u64 r = rng_u64(rng, stream, (u64)x), s = rng_u64(rng, stream, (1ull << 32) | (u64)x);
    int k = 0;
    while (r > c_gauss_cdt[k]) ++k;
    out[x] = (s & 1) ? -(int64_t)k : (int64_t)k;
}

__global__ void __launch_bounds__(TPB)
small_to_rns_kernel(const int64_t* __restrict__ small, u64* __restrict__ out, int N,
                    const LimbDesc* __restrict__ limbs, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int t = blockIdx.y, x = blockIdx.x * TPB + threadIdx.x;
    const ModConst m = mc[limbs[t].mod];
    const int64_t v = small[x];
    const u64 a = barrett_red(0, (u64)(v < 0 ? -v : v), m);
    out[(size_t)t * N + x] = (v < 0 && a != 0) ? m.q - a : a;
}

// ---- special FFT stages ----
__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
    return make_double2(__dsub_rn(__dmul_rn(a.x, b.x), __dmul_rn(a.y, b.y)),
                        __dadd_rn(__dmul_rn(a.x, b.y), __dmul_rn(a.y, b.x)));
}
// inverse stage (decimation in frequency): u = a+b, v = (a-b)*root[(lenq - rot[j] % lenq) * gap]
__global__ void __launch_bounds__(TPB)
ifft_stage_kernel(double2* __restrict__ vals, int slots, int len, int M, const double2* __restrict__ roots,
                  const int* __restrict__ rot) {
    OB_PDL_SYNC();
    const int b = blockIdx.x * TPB + threadIdx.x;
    if (b >= slots / 2) return;
    double2* v = vals + (size_t)blockIdx.y * slots;
    const int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
    const int i = b / lenh, j = b % lenh, k = i * len + j;
    const int idx = (lenq - (rot[j] % lenq)) * gap;
    const double2 x = v[k], y = v[k + lenh];
    v[k] = make_double2(__dadd_rn(x.x, y.x), __dadd_rn(x.y, y.y));
    v[k + lenh] = cmul(make_double2(__dsub_rn(x.x, y.x), __dsub_rn(x.y, y.y)), roots[idx]);
}
// forward stage (decimation in time): v = b*root[(rot[j] % lenq) * gap]; (a+v, a-v)
__global__ void __launch_bounds__(TPB)
fft_stage_kernel(double2* __restrict__ vals, int slots, int len, int M, const double2* __restrict__ roots,
                 const int* __restrict__ rot) {
    OB_PDL_SYNC();
    const int b = blockIdx.x * TPB + threadIdx.x;
    if (b >= slots / 2) return;
    double2* v = vals + (size_t)blockIdx.y * slots;
    const int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
    const int i = b / lenh, j = b % lenh, k = i * len + j;
    const int idx = (rot[j] % lenq) * gap;
    const double2 x = v[k], y = cmul(v[k + lenh], roots[idx]);
    v[k] = make_double2(__dadd_rn(x.x, y.x), __dadd_rn(x.y, y.y));
    v[k + lenh] = make_double2(__dsub_rn(x.x, y.x), __dsub_rn(x.y, y.y));
}
// bit-reversal permutation (in place, pairwise swap), optionally dividing by `div`
__global__ void __launch_bounds__(TPB)
bitrev_scale_kernel(double2* __restrict__ vals, int slots, int log_slots, double div) {
    OB_PDL_SYNC();
    const int i = blockIdx.x * TPB + threadIdx.x;
    if (i >= slots) return;
    double2* v = vals + (size_t)blockIdx.y * slots;
    const int j = (int)(__brev((unsigned)i) >> (32 - log_slots));
    if (j < i) return;
    double2 a = v[i], b = v[j];
    if (div != 0.0) {
        a = make_double2(__ddiv_rn(a.x, div), __ddiv_rn(a.y, div));
        b = make_double2(__ddiv_rn(b.x, div), __ddiv_rn(b.y, div));
    }
    v[i] = b; v[j] = a;
}

// non-negative integer-valued double -> residue mod q
__device__ __forceinline__ u64 double_to_mod(double x, const ModConst& m) {
    if (x < 9223372036854775808.0) return barrett_red(0, (u64)x, m);
    const long long bits = __double_as_longlong(x);
    int e = (int)((bits >> 52) & 0x7ff) - 1075;
    const u64 mant = ((u64)bits & ((1ull << 52) - 1)) | (1ull << 52);
    u64 r = barrett_red(0, mant, m);
    for (; e > 0; --e) { r += r; r = r >= m.q ? r - m.q : r; }
    return r;
}

__global__ void __launch_bounds__(TPB)
encode_round_kernel(const double2* __restrict__ vals, int log_slots, double scale, u64* __restrict__ coeffs,
                    int N, int nlimbs, const LimbDesc* __restrict__ limbs, const ModConst* __restrict__ mc) {
    OB_PDL_SYNC();
    const int c = blockIdx.x * TPB + threadIdx.x;      // coefficient index
    const int bidx = blockIdx.y;                       // vector in the batch
    const int slots = 1 << log_slots, half = N >> 1, gap = half >> log_slots;
    const int cc = c < half ? c : c - half;
    double x = 0.0; bool used = (cc % gap) == 0;
    if (used) {
        const double2 v = vals[(size_t)bidx * slots + cc / gap];
        x = __dmul_rn(c < half ? v.x : v.y, scale);
    }
    const bool neg = x < 0.0;
    const double ax = floor(__dadd_rn(neg ? -x : x, 0.5));
    for (int t = 0; t < nlimbs; ++t) {
        const ModConst& m = mc[limbs[t].mod];
        u64 r = used ? double_to_mod(ax, m) : 0;
        if (neg && r != 0) r = m.q - r;
        coeffs[((size_t)bidx * nlimbs + t) * N + c] = r;
    }
}

// Garner mixed-radix CRT, centred, evaluated in float64 from the top digit down
constexpr int MAX_DEC_LIMBS = 64;
__global__ void __launch_bounds__(128)
decode_crt_kernel(const u64* __restrict__ coeffs, int L, int nQ, int N, double inv_scale_num, double2* __restrict__ vals,
                  int log_slots, const ulonglong2* __restrict__ qinv /* [j][i] = q_j^-1 mod q_i */,
                  const ModConst* __restrict__ mc, const double* __restrict__ pd) {
    OB_PDL_SYNC();
    const int s = blockIdx.x * 128 + threadIdx.x;
    const int slots = 1 << log_slots, half = N >> 1, gap = half >> log_slots;
    if (s >= 2 * slots) return;
    const int c = s < slots ? s * gap : half + (s - slots) * gap;
    u64 a[MAX_DEC_LIMBS];
    for (int i = 0; i < L; ++i) {
        const ModConst m = mc[i];
        u64 t = coeffs[(size_t)i * N + c];
        for (int j = 0; j < i; ++j) {
            const u64 aj = barrett_red(0, a[j], m);
            const ulonglong2 w = qinv[(size_t)j * nQ + i];
            t = shoup_mul(sub_mod(t, aj, m.q), w.x, w.y, m.q);
        }
        a[i] = t;
    }
    bool neg = false;
    for (int i = L - 1; i >= 0; --i) {
        const u64 h = (mc[i].q - 1) >> 1;
        if (a[i] != h) { neg = a[i] > h; break; }
    }
    if (neg) {                                   // digits of Q - x
        bool carry = true;
        for (int i = 0; i < L; ++i) {
            u64 b = mc[i].q - 1 - a[i];
            if (carry) { b += 1; if (b == mc[i].q) b = 0; else carry = false; }
            a[i] = b;
        }
    }
    double acc = 0.0;
    for (int i = L - 1; i >= 0; --i)
        if (a[i] != 0) acc = __dadd_rn(acc, __dmul_rn(__ull2double_rn(a[i]), pd[i]));
    acc = __ddiv_rn(neg ? -acc : acc, inv_scale_num);
    if (s < slots) vals[s].x = acc; else vals[s - slots].y = acc;
}

const LimbDesc* mods_list(Context& c, const std::vector<int>& mods) {
    std::vector<LimbDesc> l(mods.size());
    for (size_t i = 0; i < mods.size(); ++i) l[i] = LimbDesc{(uint32_t)i, (uint32_t)i, (uint32_t)mods[i], 0x51u};
    return c.limb_list(l);
}

}  // namespace

void sample_uniform(Context& c, u64* out, const std::vector<int>& mods, const Rng& rng, u64 stream0) {
    dim3 grid(c.N / TPB, (unsigned)mods.size());
    OB_LAUNCH(uniform_kernel, grid, TPB, 0, stream(), out, c.N, mods_list(c, mods), c.dmc(), rng, stream0);
}
void sample_gaussian(Context& c, int64_t* out, const Rng& rng, u64 strm) {
    OB_LAUNCH(gaussian_kernel, c.N / TPB, TPB, 0, stream(), out, rng, strm);
}
SchemeRng SchemeRng::test_mode(u64 seed) {
    SchemeRng r; r.pub.seed = r.sec.seed = seed; return r;
}
SchemeRng SchemeRng::from_os() {
    SchemeRng r;
    std::random_device rd;                        // /dev/urandom or RDRAND in libstdc++
    for (Rng* g : {&r.pub, &r.sec}) {
        g->chacha = 1;
        for (uint32_t& w : g->key) w = (uint32_t)rd();
    }
    return r;
}
SchemeRng SchemeRng::derive(u64 tweak) const {
    if (pub.chacha) return from_os();
    return test_mode(pub.seed ^ tweak);
}
void sample_ternary_hw_host(int N, int h, const Rng& seed, u64 strm, std::vector<int64_t>& out) {
    std::vector<int> pos(N);
    out.assign(N, 0);
    for (int i = 0; i < N; ++i) pos[i] = i;
    if (h > N) h = N;
    for (int i = 0; i < h; ++i) {
        int j = i + (int)(rng_u64(seed, strm, (u64)i) % (u64)(N - i));
        std::swap(pos[i], pos[j]);
        out[pos[i]] = (rng_u64(seed, strm, (1ull << 32) | (u64)i) & 1) ? -1 : 1;
    }
}
void small_to_rns(Context& c, const int64_t* small, u64* out, const std::vector<int>& mods) {
    dim3 grid(c.N / TPB, (unsigned)mods.size());
    OB_LAUNCH(small_to_rns_kernel, grid, TPB, 0, stream(), small, out, c.N, mods_list(c, mods), c.dmc());
}

// ---- FFT tables: roots of unity exp(2*pi*i*k/M), M = 2N, first quadrant from cos(), the rest by symmetry ----
struct SchemeTables { FftTables fft; DevBuf pd; };
static std::map<const Context*, std::unique_ptr<SchemeTables>>& g_tables = *new std::map<const Context*, std::unique_ptr<SchemeTables>>();   // heap-allocated and never destroyed: no static-destruction-order hazards (and no CUDA calls) at process exit

void host_roots(int M, std::vector<double2>& r) {
    r.assign(M + 1, make_double2(0, 0));
    const int quar = M >> 2;
    const double angle = 2.0 * 3.141592653589793238462643383279502884 / (double)M;
    std::vector<double> cs(quar + 1);
    for (int k = 0; k <= quar; ++k) cs[k] = std::cos(angle * (double)k);
    for (int k = 0; k <= quar; ++k) {
        const double co = cs[k], si = cs[quar - k];
        r[k] = make_double2(co, si);
        r[2 * quar - k] = make_double2(-co, si);
        r[2 * quar + k] = make_double2(-co, -si);
        r[4 * quar - k] = make_double2(co, -si);
    }
}

static SchemeTables& tables(Context& c) {
    auto it = g_tables.find(&c);
    if (it != g_tables.end()) return *it->second;
    auto t = std::make_unique<SchemeTables>();
    const int M = 2 * c.N, slots = c.N / 2;
    std::vector<double2> r; host_roots(M, r);
    std::vector<int> rot(slots);
    long long p = 1;
    for (int j = 0; j < slots; ++j) { rot[j] = (int)p; p = (p * 5) % M; }
    std::vector<double> pd(c.nQ);
    double acc = 1.0;
    for (int i = 0; i < c.nQ; ++i) { pd[i] = acc; acc = acc * (double)c.mod[i]; }
    t->fft.log_slots = c.logN - 1;
    t->fft.roots.alloc(r.size() * sizeof(double2));
    t->fft.rot_group.alloc(rot.size() * sizeof(int));
    t->pd.alloc(pd.size() * sizeof(double));
    OB_CUDA(cudaMemcpyAsync(t->fft.roots.p, r.data(), r.size() * sizeof(double2), cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaMemcpyAsync(t->fft.rot_group.p, rot.data(), rot.size() * sizeof(int), cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaMemcpyAsync(t->pd.p, pd.data(), pd.size() * sizeof(double), cudaMemcpyHostToDevice, stream()));
    OB_CUDA(cudaStreamSynchronize(stream()));
    return *g_tables.emplace(&c, std::move(t)).first->second;
}
void drop_scheme_tables(Context& c) { g_tables.erase(&c); }
const FftTables& fft_tables(Context& c, int) { return tables(c).fft; }

void special_ifft(Context& c, double2* vals, int log_slots, int batch) {
    const FftTables& t = tables(c).fft;
    const int slots = 1 << log_slots, M = 2 * c.N;
    dim3 grid((slots / 2 + TPB - 1) / TPB, batch);
    for (int len = slots; len >= 2; len >>= 1)
        OB_LAUNCH(ifft_stage_kernel, grid, TPB, 0, stream(), vals, slots, len, M, t.roots.as<double2>(), t.rot_group.as<int>());
    dim3 g2((slots + TPB - 1) / TPB, batch);
    OB_LAUNCH(bitrev_scale_kernel, g2, TPB, 0, stream(), vals, slots, log_slots, (double)slots);
}
void special_fft(Context& c, double2* vals, int log_slots, int batch) {
    const FftTables& t = tables(c).fft;
    const int slots = 1 << log_slots, M = 2 * c.N;
    dim3 g2((slots + TPB - 1) / TPB, batch);
    OB_LAUNCH(bitrev_scale_kernel, g2, TPB, 0, stream(), vals, slots, log_slots, 0.0);
    dim3 grid((slots / 2 + TPB - 1) / TPB, batch);
    for (int len = 2; len <= slots; len <<= 1)
        OB_LAUNCH(fft_stage_kernel, grid, TPB, 0, stream(), vals, slots, len, M, t.roots.as<double2>(), t.rot_group.as<int>());
}

void encode_round_to_rns(Context& c, const double2* vals, int log_slots, double scale, u64* coeffs,
                         const std::vector<int>& mods, int batch) {
    dim3 grid(c.N / TPB, batch);
    OB_LAUNCH(encode_round_kernel, grid, TPB, 0, stream(), vals, log_slots, scale, coeffs, c.N, (int)mods.size(),
              mods_list(c, mods), c.dmc());
}

void decode_crt_to_complex(Context& c, const u64* coeffs, int level, double scale, double2* vals, int log_slots) {
    OB_CHECK(level + 1 <= MAX_DEC_LIMBS, "too many limbs to decode");
    const int slots = 1 << log_slots;
    OB_LAUNCH(decode_crt_kernel, (2 * slots + 127) / 128, 128, 0, stream(), coeffs, level + 1, c.nQ, c.N, scale, vals,
              log_slots, c.d_qlinv.as<ulonglong2>(), c.dmc(), tables(c).pd.as<double>());
}

}  // namespace ob
