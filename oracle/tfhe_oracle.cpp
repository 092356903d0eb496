// oracle/tfhe_oracle.cpp -- CPU restatement of the TFHE PBS / keyswitch arithmetic
// behind HELM's gate evaluation.  TEST INFRASTRUCTURE ONLY (see tfhe_oracle.h).
//
// Follows SURVEY.md Appendix A (tfhe-rs 0.4.x semantics; the crate itself is an
// un-vendored dependency, Cargo.toml:19) and, for the HELM-side encodings,
// /root/reference/src/circuit.rs:27-33 (+-1/8 boolean encoding), :948 (decision
// phase < 2^31), src/gates.rs:255-274 (gate -> boolean op), :746-785 (LUT -> PBS).
// "parity unpinned" at ciphertext level; pinned at decrypted level by tests/.
#include "tfhe_oracle.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <type_traits>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ---------------------------------------------------------------- RNG
struct Rng {
    uint64_t s[4];
    static uint64_t splitmix(uint64_t& x) {
        uint64_t z = (x += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    explicit Rng(uint64_t seed) { for (auto& v : s) v = splitmix(seed); }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next() {   // xoshiro256**
        uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    double uniform() { return ((next() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
    double gauss() {    // Box-Muller, one value per call (deterministic, simple)
        double u = uniform(), v = uniform();
        return std::sqrt(-2.0 * std::log(u)) * std::cos(6.283185307179586476925 * v);
    }
};

template <class T> constexpr int bits_of() { return sizeof(T) * 8; }
template <class T> using S = typename std::make_signed<T>::type;

// noise: round(std * 2^BITS * g) as a torus element
template <class T> T noise(Rng& r, double std) {
    double x = r.gauss() * std * std::ldexp(1.0, bits_of<T>());
    return (T)(S<T>)std::llround(x);
}

// ---------------------------------------------------------------- decomposition (App. A.3)
// digits[l-1] = digit of level l (level 1 = most significant), in [-B/2, B/2]
template <class T> inline void decompose(T x, int base_log, int level, S<T>* digits) {
    const int BITS = bits_of<T>();
    const int shift = BITS - base_log * level;
    T state = x;
    if (shift > 0) {
        state = (T)(x >> (shift - 1));
        state = (T)((state + 1) >> 1);           // round to closest representable
    }
    if (base_log * level < BITS) state &= (((T)1) << (base_log * level)) - 1;
    const T mask = (((T)1) << base_log) - 1;
    for (int l = level; l >= 1; --l) {            // least significant level first
        T res = state & mask;
        state >>= base_log;
        T carry = (T)(((T)(res - 1) | state) & res);
        carry >>= (base_log - 1);
        state += carry;
        digits[l - 1] = (S<T>)(T)(res - (T)(carry << base_log));
    }
}

// ---------------------------------------------------------------- f64 FFT (App. A.4)
struct Fft {
    int M = 0;
    std::vector<double> wr, wi;      // e^{+2 pi i j / M}, j < M/2
    std::vector<double> tr, ti;      // twist e^{+i pi j / (2M)} = omega^j, omega = e^{i pi / N}
    std::vector<std::vector<double>> swr, swi;   // per stage (half = M/2, M/4, ... 1): w[j * step], j < half, contiguous
    void init(int M_) {
        M = M_; wr.resize(M / 2); wi.resize(M / 2); tr.resize(M); ti.resize(M);
        for (int j = 0; j < M / 2; ++j) { double a = 2.0 * M_PI * j / M; wr[j] = std::cos(a); wi[j] = std::sin(a); }
        swr.clear(); swi.clear();
        for (int half = M / 2, step = 1; half >= 1; half >>= 1, step <<= 1) {
            std::vector<double> a(half), b(half);
            for (int j = 0; j < half; ++j) { a[j] = wr[j * step]; b[j] = wi[j * step]; }
            swr.push_back(a); swi.push_back(b);
        }
        for (int j = 0; j < M; ++j) { double a = M_PI * j / (2.0 * M); tr[j] = std::cos(a); ti[j] = std::sin(a); }
    }
    // forward: DIF, natural in -> bit-reversed out, kernel e^{+2 pi i jk/M}.  Twiddles are read from one contiguous
    // table per stage (sw*[stage]) so the j loop vectorises; the arithmetic and its order are those of the plain
    // strided form (w = wr[j * step]), results are bit-identical
    void fwd(double* __restrict re, double* __restrict im) const {
        int stage = 0;
        for (int half = M / 2; half >= 1; half >>= 1, ++stage) {
            const double* __restrict twr = swr[stage].data(); const double* __restrict twi = swi[stage].data();
            for (int base = 0; base < M; base += 2 * half) {
                double* __restrict ar = re + base; double* __restrict ai = im + base;
                double* __restrict br = ar + half; double* __restrict bi = ai + half;
#pragma omp simd
                for (int j = 0; j < half; ++j) {
                    double w_r = twr[j], w_i = twi[j];
                    double xr = ar[j] - br[j], xi = ai[j] - bi[j];
                    ar[j] += br[j]; ai[j] += bi[j];
                    br[j] = xr * w_r - xi * w_i; bi[j] = xr * w_i + xi * w_r;
                }
            }
        }
    }
    // inverse (unnormalised): DIT, bit-reversed in -> natural out, kernel e^{-2 pi i jk/M}
    void inv(double* __restrict re, double* __restrict im) const {
        int stage = (int)swr.size() - 1;
        for (int half = 1; half < M; half <<= 1, --stage) {
            const double* __restrict twr = swr[stage].data(); const double* __restrict twi = swi[stage].data();
            for (int base = 0; base < M; base += 2 * half) {
                double* __restrict ar = re + base; double* __restrict ai = im + base;
                double* __restrict br = ar + half; double* __restrict bi = ai + half;
#pragma omp simd
                for (int j = 0; j < half; ++j) {
                    double w_r = twr[j], w_i = -twi[j];
                    double xr = br[j] * w_r - bi[j] * w_i, xi = br[j] * w_i + bi[j] * w_r;
                    br[j] = ar[j] - xr; bi[j] = ai[j] - xi;
                    ar[j] += xr; ai[j] += xi;
                }
            }
        }
    }
};

}  // namespace

// ---------------------------------------------------------------- keys
struct orc_keys {
    orc_params p;
    std::vector<uint8_t> lwe_sk, glwe_sk;
    std::vector<uint8_t> bsk, ksk;       // raw torus words
    // lazily built Fourier BSK: [n][(k+1)l rows][(k+1) polys][M] re / im
    std::vector<double> fbsk_re, fbsk_im;
    Fft fft;
    bool fourier_ready = false;
};

namespace {

template <class T> void lwe_encrypt(const uint8_t* sk, int dim, T pt, double std, Rng& r, T* ct) {
    T b = pt + noise<T>(r, std);
    for (int i = 0; i < dim; ++i) {
        T a = (T)r.next();
        ct[i] = a;
        if (sk[i]) b += a;
    }
    ct[dim] = b;
}

// out += a(X) * s(X) mod X^N+1 for a binary s
template <class T> void poly_mul_binary_acc(const T* a, const uint8_t* s, int N, T* out) {
    for (int j = 0; j < N; ++j) {
        if (!s[j]) continue;
        // a * X^j : coef c -> c+j, negated on wrap
        for (int c = 0; c < N - j; ++c) out[c + j] += a[c];
        for (int c = N - j; c < N; ++c) out[c + j - N] -= a[c];
    }
}

template <class T> void keygen_t(orc_keys* K, uint64_t seed) {
    const orc_params& p = K->p;
    const int n = p.n, k = p.k, N = p.N, l = p.pbs_level, lk = p.ks_level;
    const int BITS = bits_of<T>();
    Rng r(seed);
    K->lwe_sk.resize(n); K->glwe_sk.resize((size_t)k * N);
    for (auto& b : K->lwe_sk) b = r.next() & 1;
    for (auto& b : K->glwe_sk) b = r.next() & 1;
    // BSK: GGSW(s_i), layout [i][level][row][poly][N]; multi-bit (grouping g > 1): GGSW(prod_j (bit_j(q) ? s_j : 1 - s_j)) for
    // every group of g key bits and every pattern q < 2^g, layout [group][q][level][row][poly][N]
    const size_t row_words = (size_t)(k + 1) * N;
    const int grp = p.grouping > 1 ? p.grouping : 1, pats = grp > 1 ? (1 << grp) : 1;
    const int n_ggsw = grp > 1 ? (n / grp) * pats : n;
    K->bsk.assign((size_t)n_ggsw * l * (k + 1) * row_words * sizeof(T), 0);
    T* bsk = (T*)K->bsk.data();
    // independent stream per GGSW so the loop can run in parallel deterministically
    uint64_t base_seed = r.next();
#pragma omp parallel for schedule(dynamic, 4)
    for (int i = 0; i < n_ggsw; ++i) {
        Rng ri(base_seed + 0x1000003ull * (uint64_t)(i + 1));
        int message = K->lwe_sk[grp > 1 ? 0 : i];
        if (grp > 1) {
            const int g0 = (i / pats) * grp, q = i % pats;
            message = 1;
            for (int j = 0; j < grp; ++j) message &= ((q >> j) & 1) ? K->lwe_sk[g0 + j] : 1 - K->lwe_sk[g0 + j];
        }
        for (int lev = 1; lev <= l; ++lev) {
            for (int row = 0; row <= k; ++row) {
                T* g = bsk + (((size_t)i * l + (lev - 1)) * (k + 1) + row) * row_words;
                T* body = g + (size_t)k * N;
                for (int c = 0; c < N; ++c) body[c] = noise<T>(ri, p.glwe_std);
                for (int j = 0; j < k; ++j) {
                    T* a = g + (size_t)j * N;
                    for (int c = 0; c < N; ++c) a[c] = (T)ri.next();
                    poly_mul_binary_acc<T>(a, K->glwe_sk.data() + (size_t)j * N, N, body);
                }
                if (message) {        // + m * q/B^lev on component `row` (after the body was formed from the mask)
                    T gad = ((T)1) << (BITS - p.pbs_base_log * lev);
                    g[(size_t)row * N] += gad;
                    // phase = body - sum A_j S_j must be E - m*g*S_row (row<k) or E + m*g (row=k):
                    // adding to the mask AFTER computing the body gives exactly that.
                }
            }
        }
    }
    // KSK: [i < kN][level][n+1], LWE_small( s_big[i] * q/B_ks^lev )
    const int kN = k * N;
    K->ksk.assign((size_t)kN * lk * (n + 1) * sizeof(T), 0);
    T* ksk = (T*)K->ksk.data();
    uint64_t ks_seed = r.next();
#pragma omp parallel for schedule(dynamic, 16)
    for (int i = 0; i < kN; ++i) {
        Rng ri(ks_seed + 0x2000003ull * (uint64_t)(i + 1));
        for (int lev = 1; lev <= lk; ++lev) {
            T pt = K->glwe_sk[i] ? ((T)1) << (BITS - p.ks_base_log * lev) : 0;
            lwe_encrypt<T>(K->lwe_sk.data(), n, pt, p.lwe_std, ri, ksk + ((size_t)i * lk + (lev - 1)) * (n + 1));
        }
    }
}

}  // namespace

// ================================================================ evaluation
namespace {

template <class T> inline T rot_coef(const T* in, int N, int c, int a /* in [0,2N) */) {
    int t = c - a; if (t < 0) t += 2 * N;
    return t < N ? in[t] : (T)(0 - in[t - N]);
}

template <class T> inline uint32_t modswitch1(T a, int log2N) {
    const int BITS = bits_of<T>();
    T x = a >> (BITS - log2N - 2);
    return (uint32_t)((x + 1) >> 1);          // in [0, 2N]
}

inline int ilog2(int x) { int r = 0; while ((1 << r) < x) ++r; return r; }

template <class T> void build_fourier(orc_keys* K) {
    if (K->fourier_ready) return;
    const orc_params& p = K->p;
    const int N = p.N, M = N / 2;
    K->fft.init(M);
    const size_t polys = K->bsk.size() / sizeof(T) / N;
    K->fbsk_re.resize(polys * M); K->fbsk_im.resize(polys * M);
    const T* bsk = (const T*)K->bsk.data();
    const double scale = std::ldexp(1.0, -bits_of<T>());
#pragma omp parallel for schedule(static)
    for (long q = 0; q < (long)polys; ++q) {
        const T* g = bsk + (size_t)q * N;
        double* re = K->fbsk_re.data() + (size_t)q * M; double* im = K->fbsk_im.data() + (size_t)q * M;
        for (int j = 0; j < M; ++j) {
            double a = (double)(S<T>)g[j] * scale, b = (double)(S<T>)g[j + M] * scale;
            re[j] = a * K->fft.tr[j] - b * K->fft.ti[j];
            im[j] = a * K->fft.ti[j] + b * K->fft.tr[j];
        }
        K->fft.fwd(re, im);
    }
    K->fourier_ready = true;
}

template <class T> inline T from_torus_f64(double x) {   // x in torus-fraction units
    const int BITS = bits_of<T>();
    double f = x - std::nearbyint(x);
    double v = f * std::ldexp(1.0, BITS);
    v = std::nearbyint(v);
    if (v >= std::ldexp(1.0, BITS - 1)) v -= std::ldexp(1.0, BITS);
    return (T)(S<T>)(long long)v;
}

struct Scratch {
    std::vector<double> dre, dim, ore, oim;
    std::vector<uint8_t> diff, dig;
};

// acc += BSK_i [*] (acc * X^a - acc)        (App. A.5 cmux, A.4 external product)
template <class T> void cmux_t(const orc_keys* K, int i, uint32_t a_tilde, T* acc, int mode, Scratch& sc) {
    const orc_params& p = K->p;
    const int k = p.k, N = p.N, l = p.pbs_level, M = N / 2;
    const int a = (int)(a_tilde % (2u * N));
    if (a == 0) return;
    const size_t row_words = (size_t)(k + 1) * N;
    sc.diff.resize(row_words * sizeof(T));
    T* diff = (T*)sc.diff.data();
    for (int j = 0; j <= k; ++j)
        for (int c = 0; c < N; ++c) diff[(size_t)j * N + c] = rot_coef<T>(acc + (size_t)j * N, N, c, a) - acc[(size_t)j * N + c];
    // digits[lev-1][r][N]
    sc.dig.resize((size_t)l * row_words * sizeof(S<T>));
    S<T>* dig = (S<T>*)sc.dig.data();
    S<T> d[64];
    for (size_t c = 0; c < row_words; ++c) {
        decompose<T>(diff[c], p.pbs_base_log, l, d);
        for (int lev = 0; lev < l; ++lev) dig[(size_t)lev * row_words + c] = d[lev];
    }
    if (mode == ORC_EXACT) {
        const T* bsk = (const T*)K->bsk.data();
        for (int lev = 0; lev < l; ++lev)
            for (int r = 0; r <= k; ++r) {
                const S<T>* dp = dig + (size_t)lev * row_words + (size_t)r * N;
                const T* g = bsk + (((size_t)i * l + lev) * (k + 1) + r) * row_words;
                for (int j = 0; j <= k; ++j) {
                    const T* gp = g + (size_t)j * N; T* out = acc + (size_t)j * N;
                    for (int x = 0; x < N; ++x) {
                        const T dx = (T)dp[x];
                        if (dx == 0) continue;
                        for (int y = 0; y < N - x; ++y) out[x + y] += dx * gp[y];
                        for (int y = N - x; y < N; ++y) out[x + y - N] -= dx * gp[y];
                    }
                }
            }
        return;
    }
    const Fft& F = K->fft;
    sc.dre.resize(M); sc.dim.resize(M);
    sc.ore.assign((size_t)(k + 1) * M, 0.0); sc.oim.assign((size_t)(k + 1) * M, 0.0);
    for (int lev = 0; lev < l; ++lev)
        for (int r = 0; r <= k; ++r) {
            const S<T>* dp = dig + (size_t)lev * row_words + (size_t)r * N;
            double* re = sc.dre.data(); double* im = sc.dim.data();
            for (int j = 0; j < M; ++j) {
                double x = (double)dp[j], y = (double)dp[j + M];
                re[j] = x * F.tr[j] - y * F.ti[j]; im[j] = x * F.ti[j] + y * F.tr[j];
            }
            F.fwd(re, im);
            const size_t q0 = (((size_t)i * l + lev) * (k + 1) + r) * (k + 1);
            for (int j = 0; j <= k; ++j) {
                const double* br = K->fbsk_re.data() + (q0 + j) * M; const double* bi = K->fbsk_im.data() + (q0 + j) * M;
                double* orr = sc.ore.data() + (size_t)j * M; double* oi = sc.oim.data() + (size_t)j * M;
                for (int x = 0; x < M; ++x) { orr[x] += re[x] * br[x] - im[x] * bi[x]; oi[x] += re[x] * bi[x] + im[x] * br[x]; }
            }
        }
    const double invM = 1.0 / M;
    for (int j = 0; j <= k; ++j) {
        double* re = sc.ore.data() + (size_t)j * M; double* im = sc.oim.data() + (size_t)j * M;
        F.inv(re, im);
        T* out = acc + (size_t)j * N;
        for (int x = 0; x < M; ++x) {
            double zr = re[x] * invM, zi = im[x] * invM;       // * conj(omega^x)
            double pr = zr * F.tr[x] + zi * F.ti[x], pi = zi * F.tr[x] - zr * F.ti[x];
            out[x] += from_torus_f64<T>(pr); out[x + M] += from_torus_f64<T>(pi);
        }
    }
}

// Multi-bit blind-rotation step (tfhe-rs multi_bit_blind_rotate_assign as recalled in SURVEY.md sections 0 / 8f; the crate is not
// in /root/reference): for group `grp` with switched mask elements a[0..g), pattern q carries the monomial X^{e_q},
// e_q = sum_{j: bit_j(q)} a_j, and   acc <- ExternalProduct( sum_q GGSW_q * X^{e_q}, acc )   -- exactly one pattern encrypts 1,
// so the accumulator is rotated by X^{sum_j a_j s_j} with ONE external product per g key bits.
template <class T> void multibit_step_t(const orc_keys* K, int grp, const uint32_t* a, T* acc, int mode, Scratch& sc) {
    const orc_params& p = K->p;
    const int k = p.k, N = p.N, l = p.pbs_level, M = N / 2, g = p.grouping, pats = 1 << g;
    const size_t row_words = (size_t)(k + 1) * N;
    sc.dig.resize((size_t)l * row_words * sizeof(S<T>));
    S<T>* dig = (S<T>*)sc.dig.data();
    S<T> d[64];
    for (size_t c = 0; c < row_words; ++c) {
        decompose<T>(acc[c], p.pbs_base_log, l, d);
        for (int lev = 0; lev < l; ++lev) dig[(size_t)lev * row_words + c] = d[lev];
    }
    int e[64];
    for (int q = 0; q < pats; ++q) {
        unsigned s = 0;
        for (int j = 0; j < g; ++j) if ((q >> j) & 1) s += a[j];
        e[q] = (int)(s % (2u * N));
    }
    if (mode == ORC_EXACT) {
        std::vector<T> res(row_words, 0);
        const T* bsk = (const T*)K->bsk.data();
        std::vector<T> part(N);
        for (int q = 0; q < pats; ++q)
            for (int j = 0; j <= k; ++j) {
                std::fill(part.begin(), part.end(), (T)0);
                for (int lev = 0; lev < l; ++lev)
                    for (int r = 0; r <= k; ++r) {
                        const S<T>* dp = dig + (size_t)lev * row_words + (size_t)r * N;
                        const T* gp = bsk + (((((size_t)grp * pats + q) * l + lev) * (k + 1) + r) * (k + 1) + j) * N;
                        for (int x = 0; x < N; ++x) {
                            const T dx = (T)dp[x];
                            if (dx == 0) continue;
                            for (int y = 0; y < N - x; ++y) part[x + y] += dx * gp[y];
                            for (int y = N - x; y < N; ++y) part[x + y - N] -= dx * gp[y];
                        }
                    }
                // times X^{e_q}: coefficient c of the product is coefficient c - e of part (negacyclic)
                T* o = res.data() + (size_t)j * N;
                for (int c = 0; c < N; ++c) o[c] += rot_coef<T>(part.data(), N, c, e[q]);
            }
        std::memcpy(acc, res.data(), row_words * sizeof(T));
        return;
    }
    const Fft& F = K->fft;
    const int lgM = ilog2(M);
    std::vector<double> dre((size_t)l * (k + 1) * M), dim((size_t)l * (k + 1) * M);     // spectra of the digit polynomials
    for (int lev = 0; lev < l; ++lev)
        for (int r = 0; r <= k; ++r) {
            const S<T>* dp = dig + (size_t)lev * row_words + (size_t)r * N;
            double* re = dre.data() + ((size_t)lev * (k + 1) + r) * M; double* im = dim.data() + ((size_t)lev * (k + 1) + r) * M;
            for (int j = 0; j < M; ++j) {
                double x = (double)dp[j], y = (double)dp[j + M];
                re[j] = x * F.tr[j] - y * F.ti[j]; im[j] = x * F.ti[j] + y * F.tr[j];
            }
            F.fwd(re, im);
        }
    sc.ore.assign((size_t)(k + 1) * M, 0.0); sc.oim.assign((size_t)(k + 1) * M, 0.0);
    // spectrum of X^e at frequency f (kernel e^{+2 pi i jf/M}, twist e^{+i pi j/N}): exp(i pi e (4f+1) / N); position x of the
    // bit-reversed output order holds frequency brev(x)
    for (int x = 0; x < M; ++x) {
        int f = 0;
        for (int b = 0; b < lgM; ++b) f |= ((x >> b) & 1) << (lgM - 1 - b);
        double mr[64], mi[64];
        for (int q = 0; q < pats; ++q) {
            const long t = ((long)e[q] * (4L * f + 1)) % (2L * N);
            const double ang = M_PI * (double)t / (double)N;
            mr[q] = std::cos(ang); mi[q] = std::sin(ang);
        }
        for (int lev = 0; lev < l; ++lev)
            for (int r = 0; r <= k; ++r)
                for (int j = 0; j <= k; ++j) {
                    double cr = 0, ci = 0;                                  // combined key entry: sum_q G_q * m_q
                    for (int q = 0; q < pats; ++q) {
                        const size_t poly = ((((size_t)grp * pats + q) * l + lev) * (k + 1) + r) * (k + 1) + j;
                        const double br = K->fbsk_re[poly * M + x], bi = K->fbsk_im[poly * M + x];
                        cr += br * mr[q] - bi * mi[q]; ci += br * mi[q] + bi * mr[q];
                    }
                    const double xr = dre[((size_t)lev * (k + 1) + r) * M + x], xi = dim[((size_t)lev * (k + 1) + r) * M + x];
                    sc.ore[(size_t)j * M + x] += xr * cr - xi * ci; sc.oim[(size_t)j * M + x] += xr * ci + xi * cr;
                }
    }
    const double invM = 1.0 / M;
    for (int j = 0; j <= k; ++j) {
        double* re = sc.ore.data() + (size_t)j * M; double* im = sc.oim.data() + (size_t)j * M;
        F.inv(re, im);
        T* out = acc + (size_t)j * N;
        for (int x = 0; x < M; ++x) {
            double zr = re[x] * invM, zi = im[x] * invM;
            double pr = zr * F.tr[x] + zi * F.ti[x], pi = zi * F.tr[x] - zr * F.ti[x];
            out[x] = from_torus_f64<T>(pr); out[x + M] = from_torus_f64<T>(pi);      // acc <- product (not +=)
        }
    }
}

template <class T> void pbs_t(const orc_keys* K, const T* in, const T* tv, T* out, int mode, Scratch& sc) {
    const orc_params& p = K->p;
    const int n = p.n, k = p.k, N = p.N, lg = ilog2(N);
    const size_t row_words = (size_t)(k + 1) * N;
    std::vector<T> acc(row_words);
    const int bt = (int)(modswitch1<T>(in[n], lg) % (2u * N));
    const int a0 = (2 * N - bt) % (2 * N);            // TV * X^{-b~}
    for (int j = 0; j <= k; ++j)
        for (int c = 0; c < N; ++c) acc[(size_t)j * N + c] = rot_coef<T>(tv + (size_t)j * N, N, c, a0);
    if (p.grouping > 1) {
        uint32_t a[16];
        for (int grp = 0; grp < n / p.grouping; ++grp) {
            for (int j = 0; j < p.grouping; ++j) a[j] = modswitch1<T>(in[grp * p.grouping + j], lg) % (2u * N);
            multibit_step_t<T>(K, grp, a, acc.data(), mode, sc);
        }
    } else
    for (int i = 0; i < n; ++i) cmux_t<T>(K, i, modswitch1<T>(in[i], lg), acc.data(), mode, sc);
    // sample extract coefficient 0 (App. A.5)
    for (int j = 0; j < k; ++j) {
        const T* A = acc.data() + (size_t)j * N; T* o = out + (size_t)j * N;
        o[0] = A[0];
        for (int c = 1; c < N; ++c) o[c] = (T)(0 - A[N - c]);
    }
    out[(size_t)k * N] = acc[(size_t)k * N];
}

template <class T> void keyswitch_t(const orc_keys* K, const T* in, T* out) {
    const orc_params& p = K->p;
    const int n = p.n, kN = p.k * p.N, lk = p.ks_level;
    const T* ksk = (const T*)K->ksk.data();
    for (int c = 0; c < n; ++c) out[c] = 0;
    out[n] = in[kN];
    S<T> d[64];
    for (int i = 0; i < kN; ++i) {
        decompose<T>(in[i], p.ks_base_log, lk, d);
        for (int lev = 0; lev < lk; ++lev) {
            const T dv = (T)d[lev];
            if (dv == 0) continue;
            const T* row = ksk + ((size_t)i * lk + lev) * (n + 1);
            for (int c = 0; c <= n; ++c) out[c] -= dv * row[c];
        }
    }
}

// boolean linear part; words = dim+1
template <class T> void bool_linear_t(int op, const T* a, const T* b, T* out, int words) {
    const T eighth = ((T)1) << (bits_of<T>() - 3);
    for (int c = 0; c < words; ++c) {
        T s = a[c] + b[c];
        switch (op) {
            case ORC_AND: case ORC_OR: out[c] = s; break;
            case ORC_NAND: case ORC_NOR: out[c] = (T)(0 - s); break;
            case ORC_XOR: out[c] = (T)(2 * s); break;
            case ORC_XNOR: out[c] = (T)(0 - 2 * s); break;
            default: out[c] = s;
        }
    }
    T& body = out[words - 1];
    switch (op) {
        case ORC_AND: body -= eighth; break;
        case ORC_NAND: body += eighth; break;
        case ORC_OR: body += eighth; break;
        case ORC_NOR: body -= eighth; break;
        case ORC_XOR: body += 2 * eighth; break;
        case ORC_XNOR: body -= 2 * eighth; break;
    }
}

template <class T> void bool_tv(const orc_params& p, std::vector<T>& tv) {
    tv.assign((size_t)(p.k + 1) * p.N, 0);
    const T eighth = ((T)1) << (bits_of<T>() - 3);
    for (int c = 0; c < p.N; ++c) tv[(size_t)p.k * p.N + c] = eighth;
}

template <class T> void bool_gate_t(const orc_keys* K, int op, const T* a, const T* b, const T* c, T* out, int mode, Scratch& sc) {
    const orc_params& p = K->p;
    const int n = p.n, kN = p.k * p.N;
    const T eighth = ((T)1) << (bits_of<T>() - 3);
    if (op == ORC_NOT) { for (int x = 0; x <= n; ++x) out[x] = (T)(0 - a[x]); return; }
    if (op == ORC_COPY) { std::memcpy(out, a, (size_t)(n + 1) * sizeof(T)); return; }
    std::vector<T> tv; bool_tv<T>(p, tv);
    std::vector<T> lin(n + 1), big(kN + 1);
    if (op == ORC_MUX) {   // a = then, b = else, c = cond  (src/gates.rs:265 passes (in[2], in[0], in[1]))
        std::vector<T> big2(kN + 1);
        for (int x = 0; x <= n; ++x) lin[x] = c[x] + a[x];
        lin[n] -= eighth;
        pbs_t<T>(K, lin.data(), tv.data(), big.data(), mode, sc);
        for (int x = 0; x <= n; ++x) lin[x] = b[x] - c[x];
        lin[n] -= eighth;
        pbs_t<T>(K, lin.data(), tv.data(), big2.data(), mode, sc);
        for (int x = 0; x <= kN; ++x) big[x] += big2[x];
        big[kN] += eighth;
    } else {
        bool_linear_t<T>(op, a, b, lin.data(), n + 1);
        pbs_t<T>(K, lin.data(), tv.data(), big.data(), mode, sc);
    }
    keyswitch_t<T>(K, big.data(), out);
}

template <class T> void shortint_acc_t(const orc_keys* K, const uint64_t* table, T* tv) {
    const orc_params& p = K->p;
    const int N = p.N, k = p.k, mc = p.msg_mod * p.carry_mod, box = N / mc, half = box / 2;
    const T delta = ((T)1) << (bits_of<T>() - 1 - ilog2(mc));
    std::vector<T> body(N);
    for (int i = 0; i < mc; ++i) for (int c = 0; c < box; ++c) body[(size_t)i * box + c] = (T)table[i] * delta;
    std::memset(tv, 0, (size_t)(k + 1) * N * sizeof(T));
    T* o = tv + (size_t)k * N;
    for (int c = 0; c < N - half; ++c) o[c] = body[c + half];
    for (int t = 0; t < half; ++t) o[N - half + t] = (T)(0 - body[t]);
}

}  // namespace

// ================================================================ C API
#define DISPATCH(K, CALL32, CALL64) do { if ((K)->p.torus_bits == 32) { CALL32; } else { CALL64; } } while (0)

extern "C" {

orc_keys* orc_keygen(const orc_params* p, uint64_t seed) {
    orc_keys* K = new orc_keys();
    K->p = *p;
    if (p->torus_bits == 32) keygen_t<uint32_t>(K, seed); else keygen_t<uint64_t>(K, seed);
    return K;
}
void orc_keys_free(orc_keys* K) { delete K; }
const orc_params* orc_get_params(const orc_keys* K) { return &K->p; }
const void* orc_bsk(const orc_keys* K, size_t* bytes) { if (bytes) *bytes = K->bsk.size(); return K->bsk.data(); }
const void* orc_ksk(const orc_keys* K, size_t* bytes) { if (bytes) *bytes = K->ksk.size(); return K->ksk.data(); }
const uint8_t* orc_lwe_sk(const orc_keys* K) { return K->lwe_sk.data(); }
const uint8_t* orc_glwe_sk(const orc_keys* K) { return K->glwe_sk.data(); }

void orc_encrypt(const orc_keys* K, int big, uint64_t pt, uint64_t seed, void* ct) {
    Rng r(seed ^ 0xC1F3E5A7ull);
    const int dim = big ? K->p.k * K->p.N : K->p.n;
    const uint8_t* sk = big ? K->glwe_sk.data() : K->lwe_sk.data();
    const double std = big ? K->p.glwe_std : K->p.lwe_std;
    DISPATCH(K, lwe_encrypt<uint32_t>(sk, dim, (uint32_t)pt, std, r, (uint32_t*)ct),
                lwe_encrypt<uint64_t>(sk, dim, pt, std, r, (uint64_t*)ct));
}
void orc_trivial(const orc_keys* K, int big, uint64_t pt, void* ct) {
    const int dim = big ? K->p.k * K->p.N : K->p.n;
    if (K->p.torus_bits == 32) { uint32_t* c = (uint32_t*)ct; std::memset(c, 0, dim * 4); c[dim] = (uint32_t)pt; }
    else { uint64_t* c = (uint64_t*)ct; std::memset(c, 0, dim * 8); c[dim] = pt; }
}
uint64_t orc_phase(const orc_keys* K, int big, const void* ct) {
    const int dim = big ? K->p.k * K->p.N : K->p.n;
    const uint8_t* sk = big ? K->glwe_sk.data() : K->lwe_sk.data();
    if (K->p.torus_bits == 32) {
        const uint32_t* c = (const uint32_t*)ct; uint32_t ph = c[dim];
        for (int i = 0; i < dim; ++i) if (sk[i]) ph -= c[i];
        return ph;
    }
    const uint64_t* c = (const uint64_t*)ct; uint64_t ph = c[dim];
    for (int i = 0; i < dim; ++i) if (sk[i]) ph -= c[i];
    return ph;
}

void orc_encrypt_batch(const orc_keys* K, int big, const uint64_t* pts, uint64_t seed0, uint32_t count, void* out) {
    const size_t words = (size_t)(big ? K->p.k * K->p.N : K->p.n) + 1, ts = K->p.torus_bits / 8;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)count; ++i) orc_encrypt(K, big, pts[i], seed0 + (uint64_t)i, (char*)out + (size_t)i * words * ts);
}
void orc_phase_batch(const orc_keys* K, int big, const void* cts, uint32_t count, uint64_t* out) {
    const size_t words = (size_t)(big ? K->p.k * K->p.N : K->p.n) + 1, ts = K->p.torus_bits / 8;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)count; ++i) out[i] = orc_phase(K, big, (const char*)cts + (size_t)i * words * ts);
}

void orc_decompose(int torus_bits, int base_log, int level, uint64_t x, int64_t* out) {
    if (torus_bits == 32) { int32_t d[64]; decompose<uint32_t>((uint32_t)x, base_log, level, d); for (int i = 0; i < level; ++i) out[i] = d[i]; }
    else { int64_t d[64]; decompose<uint64_t>(x, base_log, level, d); for (int i = 0; i < level; ++i) out[i] = d[i]; }
}
void orc_keyswitch(const orc_keys* K, const void* in, void* out) {
    DISPATCH(K, keyswitch_t<uint32_t>(K, (const uint32_t*)in, (uint32_t*)out),
                keyswitch_t<uint64_t>(K, (const uint64_t*)in, (uint64_t*)out));
}
static void ensure_mode(const orc_keys* K, int mode) {
    if (mode != ORC_FFT) return;
    orc_keys* Km = const_cast<orc_keys*>(K);
#pragma omp critical(orc_fourier)
    { DISPATCH(K, build_fourier<uint32_t>(Km), build_fourier<uint64_t>(Km)); }
}
void orc_pbs(const orc_keys* K, const void* in, const void* tv, void* out, int mode) {
    ensure_mode(K, mode); Scratch sc;
    DISPATCH(K, pbs_t<uint32_t>(K, (const uint32_t*)in, (const uint32_t*)tv, (uint32_t*)out, mode, sc),
                pbs_t<uint64_t>(K, (const uint64_t*)in, (const uint64_t*)tv, (uint64_t*)out, mode, sc));
}
void orc_cmux(const orc_keys* K, int i, uint32_t a_tilde, void* acc, int mode) {
    ensure_mode(K, mode); Scratch sc;
    DISPATCH(K, cmux_t<uint32_t>(K, i, a_tilde, (uint32_t*)acc, mode, sc),
                cmux_t<uint64_t>(K, i, a_tilde, (uint64_t*)acc, mode, sc));
}
void orc_modswitch(const orc_keys* K, const void* ct, uint32_t* out) {
    const int n = K->p.n, lg = ilog2(K->p.N);
    for (int i = 0; i <= n; ++i)
        out[i] = K->p.torus_bits == 32 ? modswitch1<uint32_t>(((const uint32_t*)ct)[i], lg)
                                       : modswitch1<uint64_t>(((const uint64_t*)ct)[i], lg);
}
void orc_bool_gate(const orc_keys* K, int op, const void* a, const void* b, const void* c, void* out, int mode) {
    ensure_mode(K, mode); Scratch sc;
    DISPATCH(K, bool_gate_t<uint32_t>(K, op, (const uint32_t*)a, (const uint32_t*)b, (const uint32_t*)c, (uint32_t*)out, mode, sc),
                bool_gate_t<uint64_t>(K, op, (const uint64_t*)a, (const uint64_t*)b, (const uint64_t*)c, (uint64_t*)out, mode, sc));
}
void orc_bool_linear(const orc_keys* K, int op, const void* a, const void* b, void* out) {
    DISPATCH(K, bool_linear_t<uint32_t>(op, (const uint32_t*)a, (const uint32_t*)b, (uint32_t*)out, K->p.n + 1),
                bool_linear_t<uint64_t>(op, (const uint64_t*)a, (const uint64_t*)b, (uint64_t*)out, K->p.n + 1));
}
void orc_shortint_accumulator(const orc_keys* K, const uint64_t* table, void* tv) {
    DISPATCH(K, shortint_acc_t<uint32_t>(K, table, (uint32_t*)tv), shortint_acc_t<uint64_t>(K, table, (uint64_t*)tv));
}
void orc_shortint_lut(const orc_keys* K, const void* in, const uint64_t* table, void* out, int mode) {
    ensure_mode(K, mode); Scratch sc;
    const size_t tvw = (size_t)(K->p.k + 1) * K->p.N;
    if (K->p.torus_bits == 32) {
        std::vector<uint32_t> tv(tvw), sm(K->p.n + 1);
        shortint_acc_t<uint32_t>(K, table, tv.data());
        keyswitch_t<uint32_t>(K, (const uint32_t*)in, sm.data());
        pbs_t<uint32_t>(K, sm.data(), tv.data(), (uint32_t*)out, mode, sc);
    } else {
        std::vector<uint64_t> tv(tvw), sm(K->p.n + 1);
        shortint_acc_t<uint64_t>(K, table, tv.data());
        keyswitch_t<uint64_t>(K, (const uint64_t*)in, sm.data());
        pbs_t<uint64_t>(K, sm.data(), tv.data(), (uint64_t*)out, mode, sc);
    }
}

// OpenMP team size of the batched entry points.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU arm
// of bench.py (rank 0 alone) sets the team explicitly.  n <= 0: all online cores.  Returns the team size now in effect.
int orc_set_threads(int n) {
#ifdef _OPENMP
    if (n <= 0) n = omp_get_num_procs();
    omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    (void)n; return 1;
#endif
}

int orc_bool_level(const orc_keys* K, void* wires, const uint8_t* ops, const uint32_t* in0, const uint32_t* in1,
                   const uint32_t* in2, const uint32_t* out, uint32_t count, int mode) {
    ensure_mode(K, mode);
    const size_t words = (size_t)K->p.n + 1;
    int threads = 1;
#ifdef _OPENMP
    threads = omp_get_max_threads();
#endif
    // one gate per thread, like rayon par_iter over a level (src/circuit.rs:531)
#pragma omp parallel
    {
        Scratch sc;
#pragma omp for schedule(dynamic, 1)
        for (long g = 0; g < (long)count; ++g) {
            if (K->p.torus_bits == 32) {
                uint32_t* w = (uint32_t*)wires;
                std::vector<uint32_t> tmp(words);
                bool_gate_t<uint32_t>(K, ops[g], w + in0[g] * words, w + in1[g] * words, w + (in2 ? in2[g] : 0) * words, tmp.data(), mode, sc);
                std::memcpy(w + out[g] * words, tmp.data(), words * 4);
            } else {
                uint64_t* w = (uint64_t*)wires;
                std::vector<uint64_t> tmp(words);
                bool_gate_t<uint64_t>(K, ops[g], w + in0[g] * words, w + in1[g] * words, w + (in2 ? in2[g] : 0) * words, tmp.data(), mode, sc);
                std::memcpy(w + out[g] * words, tmp.data(), words * 8);
            }
        }
    }
    return threads;
}

}  // extern "C"
