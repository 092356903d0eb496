// Client side of the host library: parameter sets, key generation, LWE encryption / decryption.
//
// In the reference these are tfhe-rs calls: boolean::gen_keys() (src/bin/helm.rs:241), shortint gen_keys(PARAM_...)
// (:301), client_key.encrypt / decrypt (src/circuit.rs:460-477, :558; :977-996, :1092).  A HELM build that links
// the engine through the Rust shim keeps using tfhe-rs for all of this and only hands the server key material to
// hb_load_bsk / hb_load_ksk; this file exists so that the C++ harness and CLI are self-contained.
// The GPU engine never sees anything declared here except the two server-key arrays.
//
// Randomness: every key bit, public mask word and noise sample is ChaCha20 keystream (client.cpp), masks and noise on
// separate streams.  With seed 0 (gen_keys / from_raw / hh_session_create) the 256-bit ChaCha key comes from getrandom(2):
// that is the production setting.  A non-zero seed expands 64 bits into the key so that tests and benchmarks are
// reproducible -- 64 bits of entropy, TEST / BENCH USE ONLY.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "../../../include/helm_b200.h"

namespace helm {

struct Params {
    std::string name;
    hb_params hb{};
    double lwe_std = 0, glwe_std = 0;      // noise standard deviations as fractions of the torus
    int small_dim() const { return hb.n; }
    int big_dim() const { return hb.k * hb.N; }
    int ct_dim() const { return hb.order == 0 ? small_dim() : big_dim(); }     // key the wires live under
    size_t word_bytes() const { return hb.torus_bits / 8; }
};
// names: boolean_default, shortint_m1c1, shortint_m2c1 (tests/circuit_test.rs:287), shortint_m2c2, concrete_gpu,
// and the small "toy32"/"toy64"/"toy64b" sets used by fast tests.  Values: SURVEY.md section 8 / App. A.1.
const Params& params_by_name(const std::string& name);

void chacha20_block(const uint32_t key[8], const uint32_t w12_15[4], uint32_t out[16]);

struct Rng {
    uint32_t key[8]; uint32_t nonce = 0; uint64_t counter = 0;
    uint64_t buf[8]; int have = 0;
    explicit Rng(uint64_t seed = 0, uint32_t stream = 0);   // test / bench: key expanded from a 64-bit seed
    static Rng from_os(uint32_t stream);                     // production: key from getrandom(2)
    Rng fork(uint32_t stream) const;                         // same key, independent keystream
    uint64_t next();
    double gauss();
private:
    void refill();
};

struct ServerKeyMaterial {                 // standard-domain keys in the layouts hb_load_bsk / hb_load_ksk take
    std::vector<uint8_t> bsk;              // [n][level][k+1 rows][k+1 polys][N]
    std::vector<uint8_t> ksk;              // [kN][level][n+1]
};

class ClientKey {
public:
    Params p;
    std::vector<uint8_t> lwe_sk, glwe_sk;  // binary; glwe_sk flattened (kN)
    mutable Rng rng_mask{0, 1}, rng_noise{0, 2};   // public masks / secret noise: separate ChaCha20 streams

    static ClientKey from_raw(const Params& p, const uint8_t* lwe_sk, const uint8_t* glwe_sk, uint64_t enc_seed);
    // torus plaintext -> ciphertext words (dim+1) under the small (big = false) or big key
    void encrypt_raw(bool big, uint64_t plaintext, void* ct_out) const;
    uint64_t phase(bool big, const void* ct) const;

    // encodings.  boolean (src/circuit.rs:27-33, :948): true = +1/8, false = -1/8, decision phase < 1/2.
    void encrypt_bool(bool v, void* ct_out) const;                 // under ct_dim()'s key
    bool decrypt_bool(const void* ct) const;
    uint64_t trivial_bool_plaintext(bool v) const;
    // shortint (App. A.8): m * delta, delta = 2^(bits-1) / (msg*carry); decrypt rounds and reduces mod msg
    uint64_t delta() const;
    void encrypt_shortint(uint64_t m, void* ct_out) const;
    uint64_t decrypt_shortint(const void* ct) const;               // message part (mod msg_mod)
    uint64_t decrypt_shortint_full(const void* ct) const;          // message + carry (mod msg*carry)
};

void gen_keys(const Params& p, uint64_t seed, ClientKey* ck, ServerKeyMaterial* sk);

}  // namespace helm
