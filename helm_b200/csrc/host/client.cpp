// See client.hpp.  LWE / GLWE / GGSW objects as in SURVEY.md App. A.2 (tfhe-rs 0.4.x conventions).
#include "client.hpp"
#include <cmath>
#include <cstring>
#include <map>
#include "netlist.hpp"   // helm::Error

namespace helm {

static Params mk(const char* name, int n, int k, int N, int bl, int l, int kbl, int kl, int bits, int order, int msg, int carry,
                 double lwe_std, double glwe_std) {
    Params p; p.name = name;
    p.hb.n = n; p.hb.k = k; p.hb.N = N; p.hb.pbs_base_log = bl; p.hb.pbs_level = l; p.hb.ks_base_log = kbl; p.hb.ks_level = kl;
    p.hb.torus_bits = bits; p.hb.order = order; p.hb.msg_mod = msg; p.hb.carry_mod = carry;
    p.lwe_std = lwe_std; p.glwe_std = glwe_std;
    return p;
}

const Params& params_by_name(const std::string& name) {
    static const std::map<std::string, Params> sets = {
        // tfhe::boolean DEFAULT_PARAMETERS (gen_keys(), reference src/bin/helm.rs:241)
        {"boolean_default", mk("boolean_default", 722, 2, 512, 6, 3, 3, 4, 32, 0, 0, 0, 1.3071021089943935e-5, 4.990272175010415e-8)},
        // PARAM_MESSAGE_1_CARRY_1_KS_PBS (src/bin/helm.rs:301)
        {"shortint_m1c1", mk("shortint_m1c1", 684, 3, 512, 18, 1, 3, 4, 64, 1, 2, 2, 2.0399883307556e-5, 3.45253423096e-13)},
        // PARAM_MESSAGE_2_CARRY_1_KS_PBS (tests/circuit_test.rs:287); same polynomial shape as M1C1, 8 boxes
        {"shortint_m2c1", mk("shortint_m2c1", 708, 3, 512, 18, 1, 3, 4, 64, 1, 4, 2, 1.3292631075564e-5, 3.45253423096e-13)},
        // PARAM_MESSAGE_2_CARRY_2_KS_PBS
        {"shortint_m2c2", mk("shortint_m2c2", 742, 1, 2048, 23, 1, 3, 5, 64, 1, 4, 4, 7.069849454709433e-6, 2.94036403378e-16)},
        // concrete-core GPU path of the reference (src/bin/helm.rs:141-146)
        {"concrete_gpu", mk("concrete_gpu", 512, 1, 1024, 7, 3, 2, 8, 32, 0, 0, 0, 2.9802322387695312e-8, 2.9802322387695312e-8)},
        // small sets for fast tests (NOT secure): same kernel instances as the real ones
        {"toy32", mk("toy32", 48, 2, 512, 6, 3, 3, 4, 32, 0, 0, 0, 9.5367431640625e-7, 3.725290298461914e-9)},
        {"toy64", mk("toy64", 40, 1, 2048, 23, 1, 3, 5, 64, 1, 4, 4, 5.960464477539063e-8, 2.220446049250313e-16)},
        {"toy64b", mk("toy64b", 36, 3, 512, 18, 1, 3, 4, 64, 1, 2, 2, 5.960464477539063e-8, 2.2737367544323206e-13)},
    };
    auto it = sets.find(name);
    if (it == sets.end()) throw Error("unknown parameter set \"" + name + "\"");
    return it->second;
}

// ---------------------------------------------------------------- RNG (xoshiro256**, splitmix64 seeding)
static uint64_t splitmix(uint64_t& x) {
    uint64_t z = (x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
Rng::Rng(uint64_t seed) { for (auto& v : s) v = splitmix(seed); }
uint64_t Rng::next() {
    auto rotl = [](uint64_t x, int k) { return (x << k) | (x >> (64 - k)); };
    const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
}
double Rng::gauss() {
    const double u = ((next() >> 11) + 0.5) * (1.0 / 9007199254740992.0), v = ((next() >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    return std::sqrt(-2.0 * std::log(u)) * std::cos(6.283185307179586476925 * v);
}

template <class T> static T torus_noise(Rng& r, double std) {
    const double x = r.gauss() * std * std::ldexp(1.0, (int)sizeof(T) * 8);
    return (T)(typename std::make_signed<T>::type)std::llround(x);
}

template <class T> static void lwe_encrypt_t(const uint8_t* sk, int dim, T pt, double std, Rng& r, T* ct) {
    T b = pt + torus_noise<T>(r, std);
    for (int i = 0; i < dim; ++i) { const T a = (T)r.next(); ct[i] = a; if (sk[i]) b += a; }
    ct[dim] = b;
}

ClientKey ClientKey::from_raw(const Params& p, const uint8_t* lwe_sk, const uint8_t* glwe_sk, uint64_t enc_seed) {
    ClientKey ck; ck.p = p;
    ck.lwe_sk.assign(lwe_sk, lwe_sk + p.small_dim());
    ck.glwe_sk.assign(glwe_sk, glwe_sk + p.big_dim());
    ck.rng = Rng(enc_seed);
    return ck;
}

void ClientKey::encrypt_raw(bool big, uint64_t pt, void* ct) const {
    const int dim = big ? p.big_dim() : p.small_dim();
    const uint8_t* sk = big ? glwe_sk.data() : lwe_sk.data();
    const double std = big ? p.glwe_std : p.lwe_std;
    if (p.hb.torus_bits == 32) lwe_encrypt_t<uint32_t>(sk, dim, (uint32_t)pt, std, rng, (uint32_t*)ct);
    else lwe_encrypt_t<uint64_t>(sk, dim, pt, std, rng, (uint64_t*)ct);
}

uint64_t ClientKey::phase(bool big, const void* ct) const {
    const int dim = big ? p.big_dim() : p.small_dim();
    const uint8_t* sk = big ? glwe_sk.data() : lwe_sk.data();
    if (p.hb.torus_bits == 32) {
        const uint32_t* c = (const uint32_t*)ct; uint32_t ph = c[dim];
        for (int i = 0; i < dim; ++i) if (sk[i]) ph -= c[i];
        return ph;
    }
    const uint64_t* c = (const uint64_t*)ct; uint64_t ph = c[dim];
    for (int i = 0; i < dim; ++i) if (sk[i]) ph -= c[i];
    return ph;
}

uint64_t ClientKey::trivial_bool_plaintext(bool v) const { return (v ? 1ull : 7ull) << (p.hb.torus_bits - 3); }
void ClientKey::encrypt_bool(bool v, void* ct) const { encrypt_raw(p.hb.order == 1, trivial_bool_plaintext(v), ct); }
bool ClientKey::decrypt_bool(const void* ct) const { return phase(p.hb.order == 1, ct) < (1ull << (p.hb.torus_bits - 1)); }

uint64_t ClientKey::delta() const {
    const int mc = p.hb.msg_mod * p.hb.carry_mod;
    int lg = 0; while ((1 << lg) < mc) ++lg;
    return 1ull << (p.hb.torus_bits - 1 - lg);
}
void ClientKey::encrypt_shortint(uint64_t m, void* ct) const { encrypt_raw(true, (m % (uint64_t)p.hb.msg_mod) * delta(), ct); }
uint64_t ClientKey::decrypt_shortint_full(const void* ct) const {
    const uint64_t d = delta(), mask = p.hb.torus_bits == 64 ? ~0ull : 0xffffffffull;
    const uint64_t ph = (phase(true, ct) + d / 2) & mask;                     // round to the nearest multiple of delta
    return (ph / d) % (uint64_t)(p.hb.msg_mod * p.hb.carry_mod);              // drops the padding bit
}
uint64_t ClientKey::decrypt_shortint(const void* ct) const { return decrypt_shortint_full(ct) % (uint64_t)p.hb.msg_mod; }

// ---------------------------------------------------------------- key generation
// out += a(X) * s(X) mod X^N + 1 for a binary s
template <class T> static void poly_mul_binary_acc(const T* a, const uint8_t* s, int N, T* out) {
    for (int j = 0; j < N; ++j) {
        if (!s[j]) continue;
        for (int c = 0; c < N - j; ++c) out[c + j] += a[c];
        for (int c = N - j; c < N; ++c) out[c + j - N] -= a[c];
    }
}

template <class T> static void gen_server_keys_t(const ClientKey& ck, uint64_t seed, ServerKeyMaterial* sk) {
    const Params& p = ck.p;
    const int n = p.hb.n, k = p.hb.k, N = p.hb.N, l = p.hb.pbs_level, lk = p.hb.ks_level, BITS = (int)sizeof(T) * 8;
    const size_t row_words = (size_t)(k + 1) * N;
    sk->bsk.assign((size_t)n * l * (k + 1) * row_words * sizeof(T), 0);
    T* bsk = (T*)sk->bsk.data();
    // BSK_i = GGSW(s_i): for level lev and row r, a GLWE encryption of 0 plus s_i * q/B^lev on component r (App. A.2)
#pragma omp parallel for schedule(dynamic, 4)
    for (int i = 0; i < n; ++i) {
        Rng r(seed ^ (0x9E3779B97F4A7C15ull * (uint64_t)(i + 1)));
        for (int lev = 1; lev <= l; ++lev)
            for (int row = 0; row <= k; ++row) {
                T* g = bsk + (((size_t)i * l + (lev - 1)) * (k + 1) + row) * row_words;
                T* body = g + (size_t)k * N;
                for (int c = 0; c < N; ++c) body[c] = torus_noise<T>(r, p.glwe_std);
                for (int j = 0; j < k; ++j) {
                    T* a = g + (size_t)j * N;
                    for (int c = 0; c < N; ++c) a[c] = (T)r.next();
                    poly_mul_binary_acc<T>(a, ck.glwe_sk.data() + (size_t)j * N, N, body);
                }
                if (ck.lwe_sk[i]) g[(size_t)row * N] += ((T)1) << (BITS - p.hb.pbs_base_log * lev);
            }
    }
    // KSK[i][lev] = LWE_small( s_big[i] * q / B_ks^lev )
    const int kN = k * N;
    sk->ksk.assign((size_t)kN * lk * (n + 1) * sizeof(T), 0);
    T* ksk = (T*)sk->ksk.data();
#pragma omp parallel for schedule(dynamic, 16)
    for (int i = 0; i < kN; ++i) {
        Rng r(seed * 0x2545F4914F6CDD1Dull + 0xD1B54A32D192ED03ull * (uint64_t)(i + 1));
        for (int lev = 1; lev <= lk; ++lev) {
            const T pt = ck.glwe_sk[i] ? ((T)1) << (BITS - p.hb.ks_base_log * lev) : (T)0;
            lwe_encrypt_t<T>(ck.lwe_sk.data(), n, pt, p.lwe_std, r, ksk + ((size_t)i * lk + (lev - 1)) * (n + 1));
        }
    }
}

void gen_keys(const Params& p, uint64_t seed, ClientKey* ck, ServerKeyMaterial* sk) {
    if (p.hb.torus_bits != 32 && p.hb.torus_bits != 64) throw Error("torus_bits must be 32 or 64");
    Rng r(seed);
    ck->p = p;
    ck->lwe_sk.resize(p.small_dim()); ck->glwe_sk.resize(p.big_dim());
    for (auto& b : ck->lwe_sk) b = r.next() & 1;
    for (auto& b : ck->glwe_sk) b = r.next() & 1;
    ck->rng = Rng(r.next());
    const uint64_t ks = r.next();
    if (sk) { if (p.hb.torus_bits == 32) gen_server_keys_t<uint32_t>(*ck, ks, sk); else gen_server_keys_t<uint64_t>(*ck, ks, sk); }
}

}  // namespace helm
