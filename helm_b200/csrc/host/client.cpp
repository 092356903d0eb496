// See client.hpp.  LWE / GLWE / GGSW objects as in SURVEY.md App. A.2 (tfhe-rs 0.4.x conventions).
#include "client.hpp"
#include <cmath>
#include <sys/random.h>
#include <cstring>
#include <map>
#include "netlist.hpp"   // helm::Error

namespace helm {

static Params mk(const char* name, int n, int k, int N, int bl, int l, int kbl, int kl, int bits, int order, int msg, int carry,
                 double lwe_std, double glwe_std, int grouping = 0) {
    Params p; p.name = name;
    p.hb.n = n; p.hb.k = k; p.hb.N = N; p.hb.pbs_base_log = bl; p.hb.pbs_level = l; p.hb.ks_base_log = kbl; p.hb.ks_level = kl;
    p.hb.torus_bits = bits; p.hb.order = order; p.hb.msg_mod = msg; p.hb.carry_mod = carry; p.hb.grouping = grouping;
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
        // PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_3_KS_PBS: the reference's arithmetic-mode set (src/bin/helm.rs:81-86)
        {"shortint_multibit_m2c2_g3", mk("shortint_multibit_m2c2_g3", 888, 1, 2048, 21, 1, 4, 4, 64, 1, 4, 4, 6.125031601933181e-7, 2.9403601535432533e-16, 3)},
        // concrete-core GPU path of the reference (src/bin/helm.rs:141-146)
        {"concrete_gpu", mk("concrete_gpu", 512, 1, 1024, 7, 3, 2, 8, 32, 0, 0, 0, 2.9802322387695312e-8, 2.9802322387695312e-8)},
        // small sets for fast tests (NOT secure): same kernel instances as the real ones
        {"toy32", mk("toy32", 48, 2, 512, 6, 3, 3, 4, 32, 0, 0, 0, 9.5367431640625e-7, 3.725290298461914e-9)},
        {"toy64", mk("toy64", 40, 1, 2048, 23, 1, 3, 5, 64, 1, 4, 4, 5.960464477539063e-8, 2.220446049250313e-16)},
        {"toy64m", mk("toy64m", 36, 1, 2048, 21, 1, 4, 4, 64, 1, 4, 4, 5.960464477539063e-8, 2.220446049250313e-16, 3)},
        {"toy64b", mk("toy64b", 36, 3, 512, 18, 1, 3, 4, 64, 1, 2, 2, 5.960464477539063e-8, 2.2737367544323206e-13)},
    };
    auto it = sets.find(name);
    if (it == sets.end()) throw Error("unknown parameter set \"" + name + "\"");
    return it->second;
}

// ---------------------------------------------------------------- RNG: ChaCha20 keystream (RFC 8439 block function)
// Every random word the client emits -- secret keys, public LWE / GLWE masks, Gaussian noise -- is ChaCha20 output, so
// seeing masks (they are written into ciphertexts in the clear) reveals nothing about the noise or the keys.  The key is
// 32 bytes from getrandom(2) (Rng::from_os, what hh_session_create uses for seed 0), or is expanded from a 64-bit seed for
// reproducible tests and benchmarks (a 64-bit seed is not a production key: see client.hpp).  `stream` selects an
// independent keystream (nonce word) so that masks and noise never share one.
static uint64_t splitmix(uint64_t& x) {
    uint64_t z = (x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
Rng::Rng(uint64_t seed, uint32_t stream) {
    for (int i = 0; i < 4; ++i) { const uint64_t v = splitmix(seed); key[2 * i] = (uint32_t)v; key[2 * i + 1] = (uint32_t)(v >> 32); }
    nonce = stream;
}
Rng Rng::from_os(uint32_t stream) {
    Rng r(0, stream);
    size_t got = 0;
    while (got < sizeof r.key) {
        const ssize_t n = getrandom((char*)r.key + got, sizeof r.key - got, 0);
        if (n <= 0) throw Error("getrandom failed: no OS entropy for the client keys");
        got += (size_t)n;
    }
    return r;
}
Rng Rng::fork(uint32_t stream) const { Rng r = *this; r.nonce = stream; r.counter = 0; r.have = 0; return r; }
// out = ChaCha20 block function of key / words 12..15 (RFC 8439 section 2.3; known-answer test in tests/test_host_cpu.py)
void chacha20_block(const uint32_t key[8], const uint32_t w12_15[4], uint32_t out[16]) {
    static const uint32_t sigma[4] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u};
    uint32_t st[16], x[16];
    for (int i = 0; i < 4; ++i) st[i] = sigma[i];
    for (int i = 0; i < 8; ++i) st[4 + i] = key[i];
    for (int i = 0; i < 4; ++i) st[12 + i] = w12_15[i];
    for (int i = 0; i < 16; ++i) x[i] = st[i];
    auto rotl = [](uint32_t v, int k) { return (v << k) | (v >> (32 - k)); };
    auto qr = [&](int a, int b, int c, int d) {
        x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 16); x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 12);
        x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 8);  x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 7);
    };
    for (int round = 0; round < 10; ++round) {
        qr(0, 4, 8, 12); qr(1, 5, 9, 13); qr(2, 6, 10, 14); qr(3, 7, 11, 15);
        qr(0, 5, 10, 15); qr(1, 6, 11, 12); qr(2, 7, 8, 13); qr(3, 4, 9, 14);
    }
    for (int i = 0; i < 16; ++i) out[i] = x[i] + st[i];
}
void Rng::refill() {
    const uint32_t w[4] = {(uint32_t)counter, (uint32_t)(counter >> 32), nonce, 0x48454C4Du};
    uint32_t o[16];
    chacha20_block(key, w, o);
    for (int i = 0; i < 8; ++i) buf[i] = (uint64_t)o[2 * i] | ((uint64_t)o[2 * i + 1] << 32);
    ++counter; have = 8;
}
uint64_t Rng::next() { if (!have) refill(); return buf[--have]; }
double Rng::gauss() {
    const double u = ((next() >> 11) + 0.5) * (1.0 / 9007199254740992.0), v = ((next() >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    return std::sqrt(-2.0 * std::log(u)) * std::cos(6.283185307179586476925 * v);
}

template <class T> static T torus_noise(Rng& r, double std) {
    const double x = r.gauss() * std * std::ldexp(1.0, (int)sizeof(T) * 8);
    return (T)(typename std::make_signed<T>::type)std::llround(x);
}

template <class T> static void lwe_encrypt_t(const uint8_t* sk, int dim, T pt, double std, Rng& mask, Rng& noise, T* ct) {
    T b = pt + torus_noise<T>(noise, std);
    for (int i = 0; i < dim; ++i) { const T a = (T)mask.next(); ct[i] = a; if (sk[i]) b += a; }
    ct[dim] = b;
}

ClientKey ClientKey::from_raw(const Params& p, const uint8_t* lwe_sk, const uint8_t* glwe_sk, uint64_t enc_seed) {
    ClientKey ck; ck.p = p;
    ck.lwe_sk.assign(lwe_sk, lwe_sk + p.small_dim());
    ck.glwe_sk.assign(glwe_sk, glwe_sk + p.big_dim());
    const Rng base = enc_seed ? Rng(enc_seed, 0) : Rng::from_os(0);
    ck.rng_mask = base.fork(1); ck.rng_noise = base.fork(2);
    return ck;
}

void ClientKey::encrypt_raw(bool big, uint64_t pt, void* ct) const {
    const int dim = big ? p.big_dim() : p.small_dim();
    const uint8_t* sk = big ? glwe_sk.data() : lwe_sk.data();
    const double std = big ? p.glwe_std : p.lwe_std;
    if (p.hb.torus_bits == 32) lwe_encrypt_t<uint32_t>(sk, dim, (uint32_t)pt, std, rng_mask, rng_noise, (uint32_t*)ct);
    else lwe_encrypt_t<uint64_t>(sk, dim, pt, std, rng_mask, rng_noise, (uint64_t*)ct);
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

template <class T> static void gen_server_keys_t(const ClientKey& ck, const Rng& base, ServerKeyMaterial* sk) {
    const Params& p = ck.p;
    const int n = p.hb.n, k = p.hb.k, N = p.hb.N, l = p.hb.pbs_level, lk = p.hb.ks_level, BITS = (int)sizeof(T) * 8;
    const size_t row_words = (size_t)(k + 1) * N;
    // multi-bit (grouping g > 1): per group of g key bits 2^g GGSWs, pattern q encrypts prod_j (bit_j(q) ? s_j : 1 - s_j)
    const int grp = p.hb.grouping > 1 ? p.hb.grouping : 1, pats = grp > 1 ? (1 << grp) : 1;
    if (grp > 1 && n % grp) throw Error("multi-bit parameter set: n must be a multiple of the grouping factor");
    const int n_ggsw = grp > 1 ? (n / grp) * pats : n;
    sk->bsk.assign((size_t)n_ggsw * l * (k + 1) * row_words * sizeof(T), 0);
    T* bsk = (T*)sk->bsk.data();
    // BSK_i = GGSW(s_i): for level lev and row r, a GLWE encryption of 0 plus s_i * q/B^lev on component r (App. A.2)
#pragma omp parallel for schedule(dynamic, 4)
    for (int i = 0; i < n_ggsw; ++i) {
        int message = ck.lwe_sk[grp > 1 ? 0 : i];
        if (grp > 1) {
            const int g0 = (i / pats) * grp, q = i % pats;
            message = 1;
            for (int j = 0; j < grp; ++j) message &= ((q >> j) & 1) ? ck.lwe_sk[g0 + j] : 1 - ck.lwe_sk[g0 + j];
        }
        Rng r = base.fork(0x10000u + 2u * (uint32_t)i), rn = base.fork(0x10001u + 2u * (uint32_t)i);   // masks, noise
        for (int lev = 1; lev <= l; ++lev)
            for (int row = 0; row <= k; ++row) {
                T* g = bsk + (((size_t)i * l + (lev - 1)) * (k + 1) + row) * row_words;
                T* body = g + (size_t)k * N;
                for (int c = 0; c < N; ++c) body[c] = torus_noise<T>(rn, p.glwe_std);
                for (int j = 0; j < k; ++j) {
                    T* a = g + (size_t)j * N;
                    for (int c = 0; c < N; ++c) a[c] = (T)r.next();
                    poly_mul_binary_acc<T>(a, ck.glwe_sk.data() + (size_t)j * N, N, body);
                }
                if (message) g[(size_t)row * N] += ((T)1) << (BITS - p.hb.pbs_base_log * lev);
            }
    }
    // KSK[i][lev] = LWE_small( s_big[i] * q / B_ks^lev )
    const int kN = k * N;
    sk->ksk.assign((size_t)kN * lk * (n + 1) * sizeof(T), 0);
    T* ksk = (T*)sk->ksk.data();
#pragma omp parallel for schedule(dynamic, 16)
    for (int i = 0; i < kN; ++i) {
        Rng r = base.fork(0x1000000u + 2u * (uint32_t)i), rn = base.fork(0x1000001u + 2u * (uint32_t)i);
        for (int lev = 1; lev <= lk; ++lev) {
            const T pt = ck.glwe_sk[i] ? ((T)1) << (BITS - p.hb.ks_base_log * lev) : (T)0;
            lwe_encrypt_t<T>(ck.lwe_sk.data(), n, pt, p.lwe_std, r, rn, ksk + ((size_t)i * lk + (lev - 1)) * (n + 1));
        }
    }
}

void gen_keys(const Params& p, uint64_t seed, ClientKey* ck, ServerKeyMaterial* sk) {
    if (p.hb.torus_bits != 32 && p.hb.torus_bits != 64) throw Error("torus_bits must be 32 or 64");
    // seed 0: key the generator from the OS (production); any other seed: reproducible keys for tests and benchmarks
    const Rng base = seed ? Rng(seed, 0) : Rng::from_os(0);
    Rng r = base.fork(3);                                    // secret-key stream
    ck->p = p;
    ck->lwe_sk.resize(p.small_dim()); ck->glwe_sk.resize(p.big_dim());
    for (auto& b : ck->lwe_sk) b = r.next() & 1;
    for (auto& b : ck->glwe_sk) b = r.next() & 1;
    ck->rng_mask = base.fork(1); ck->rng_noise = base.fork(2);
    if (sk) { if (p.hb.torus_bits == 32) gen_server_keys_t<uint32_t>(*ck, base, sk); else gen_server_keys_t<uint64_t>(*ck, base, sk); }
}

}  // namespace helm
