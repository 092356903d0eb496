/*
 * oracle/tfhe_oracle.h -- CPU restatement of the TFHE arithmetic on HELM's hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under helm_b200/ may include, link or call
 * this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker or as the timed CPU baseline.
 *
 * What it restates: the arithmetic HELM (reference @ fc9f75f) reaches through
 *   src/gates.rs:255-274   boolean::ServerKey::{and,nand,or,nor,xor,xnor,mux,not}
 *   src/gates.rs:754-785   shortint bivariate / lookup-table PBS
 *   src/circuit.rs:450-480, 551-576  encrypt / decrypt of wires
 * That arithmetic lives in the crates.io dependency `tfhe = "0.4.1"`
 * (Cargo.toml:19), which is NOT under /root/reference and cannot be built here
 * (no Rust toolchain, no network).  So this file follows the published TFHE
 * algorithm as tfhe-rs 0.4.x implements it:
 * signed gadget decomposition, folded negacyclic f64 FFT external product,
 * CMUX blind rotation, sample extraction, LWE keyswitch.
 *
 * PARITY STATUS: "parity unpinned" at the ciphertext level -- the reference
 * holds no ciphertext/key golden vectors (encryption is randomised) and the
 * reference cannot run here.  Parity is pinned at the DECRYPTED level only:
 * the reference's own known-answer tests (tests/gates_test.rs:9-108 truth
 * tables, tests/circuit_test.rs:17-94 two-bit adder, :266-312 LUT adder,
 * hdl-benchmarks/test-cases CSV goldens) are restated in tests/ and the
 * oracle must reproduce them.
 *
 * Two interchangeable polynomial back-ends:
 *   ORC_EXACT  negacyclic products in Z/2^BITS (schoolbook), no rounding
 *   ORC_FFT    folded f64 FFT exactly as Appendix A.4 (the timed CPU baseline)
 */
#ifndef TFHE_ORACLE_H
#define TFHE_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t n;            /* small LWE dimension */
    int32_t k;            /* GLWE dimension */
    int32_t N;            /* polynomial size */
    int32_t pbs_base_log, pbs_level;
    int32_t ks_base_log, ks_level;
    int32_t torus_bits;   /* 32 or 64 */
    int32_t order;        /* 0 = PBS then KS (ciphertexts under small key),
                             1 = KS then PBS (ciphertexts under big key) */
    int32_t msg_mod, carry_mod;   /* shortint only; 0 for boolean */
    int32_t grouping;             /* 0 / 1: classic blind rotation (one CMUX per key bit); g > 1: multi-bit blind rotation with
                                     grouping factor g (tfhe-rs multi_bit PBS, reference src/bin/helm.rs:81-86): the BSK holds, per group of
                                     g key bits, 2^g GGSWs -- pattern q encrypts prod_j (bit_j(q) ? s_j : 1 - s_j) -- layout
                                     [n/g][2^g][level][row][poly][N] */
    double lwe_std, glwe_std;     /* noise std-dev as a fraction of the torus */
} orc_params;

enum { ORC_EXACT = 0, ORC_FFT = 1 };
enum { ORC_AND = 0, ORC_NAND, ORC_OR, ORC_NOR, ORC_XOR, ORC_XNOR, ORC_MUX, ORC_NOT, ORC_COPY, ORC_LUT };

typedef struct orc_keys orc_keys;

/* keygen: binary secret keys, standard-domain BSK (GGSW list), KSK; seeded */
orc_keys* orc_keygen(const orc_params* p, uint64_t seed);
void orc_keys_free(orc_keys*);
const orc_params* orc_get_params(const orc_keys*);
/* raw key material (torus words are uint32_t or uint64_t per torus_bits) */
const void* orc_bsk(const orc_keys*, size_t* bytes);      /* [n][level 1..l][row k+1][poly k+1][N] */
const void* orc_ksk(const orc_keys*, size_t* bytes);      /* [kN][level 1..l_ks][n+1] */
const uint8_t* orc_lwe_sk(const orc_keys*);              /* n bits   */
const uint8_t* orc_glwe_sk(const orc_keys*);             /* kN bits (flattened) */

/* client side. big=0: under the small key (dim n); big=1: under the flattened GLWE key (dim kN) */
void orc_encrypt(const orc_keys*, int big, uint64_t plaintext, uint64_t seed, void* ct_out);
void orc_trivial(const orc_keys*, int big, uint64_t plaintext, void* ct_out);
uint64_t orc_phase(const orc_keys*, int big, const void* ct);   /* b - <a,s> */

/* batched client side: ciphertext i uses seed0 + i (identical to orc_encrypt with that seed) */
void orc_encrypt_batch(const orc_keys*, int big, const uint64_t* plaintexts, uint64_t seed0, uint32_t count, void* cts_out);
void orc_phase_batch(const orc_keys*, int big, const void* cts, uint32_t count, uint64_t* phases_out);

/* primitives */
void orc_decompose(int torus_bits, int base_log, int level, uint64_t x, int64_t* digits_out /* [level], index l-1 = level l */);
void orc_keyswitch(const orc_keys*, const void* ct_big, void* ct_small);
/* test vector: (k+1)*N torus words (GLWE).  in: dim n ; out: dim kN */
void orc_pbs(const orc_keys*, const void* ct_small, const void* test_vector, void* ct_big, int mode);
/* one external product step, exposed for golden vectors:
 * acc (k+1)*N words in/out; computes acc += BSK_i [*] (acc*X^a - acc) */
void orc_cmux(const orc_keys*, int i, uint32_t a_tilde, void* acc, int mode);
void orc_modswitch(const orc_keys*, const void* ct_small, uint32_t* out /* n+1, values in [0,2N] */);

/* boolean gates: inputs/outputs under the small key (order 0) */
void orc_bool_gate(const orc_keys*, int op, const void* a, const void* b, const void* c, void* out, int mode);
/* the linear pre-processing alone (for unit parity of the fused prologue) */
void orc_bool_linear(const orc_keys*, int op, const void* a, const void* b, void* out);

/* shortint (order 1): build the boxed accumulator body for f given as table[msg*carry] */
void orc_shortint_accumulator(const orc_keys*, const uint64_t* table, void* test_vector /* (k+1)*N */);
/* ct_out = PBS(KS(ct_in)) with accumulator from table */
void orc_shortint_lut(const orc_keys*, const void* ct_in, const uint64_t* table, void* ct_out, int mode);

/* OpenMP team size of orc_bool_level (n <= 0: all online cores); returns the size now in effect */
int orc_set_threads(int n);

/* batched CPU baseline: evaluates `count` independent boolean gates (OpenMP, one gate per thread).
 * wires: [n_wires][n+1]; gate g reads in0[g], in1[g] (and in2[g] for MUX), writes out[g]. returns #threads used */
int orc_bool_level(const orc_keys*, void* wires, const uint8_t* ops, const uint32_t* in0, const uint32_t* in1,
                   const uint32_t* in2, const uint32_t* out, uint32_t count, int mode);

#ifdef __cplusplus
}
#endif
#endif
