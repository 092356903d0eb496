/*
 * helm_host.h -- C ABI of the host-side mirror of HELM's circuit surface (libhelm_host.so), used by tests, by the
 * `helm` CLI and by bindings.  It wraps helm_b200/csrc/host/: read_verilog_file (reference src/verilog_parser.rs:138),
 * Circuit::{sort_circuit, compute_levels, evaluate} (src/circuit.rs:122-381) and the EvalCircuit implementations
 * GateCircuit / LutCircuit (src/circuit.rs:449-577, 969-1111), which run every level as one hb_eval_level batch.
 * Text results are written into caller buffers as "name=value\n" lines.  All functions return 0 on success;
 * hh_last_error() describes the last failure on this thread (reference panics become errors).
 */
#ifndef HELM_HOST_H
#define HELM_HOST_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct hh_circuit hh_circuit;
typedef struct hh_session hh_session;

const char* hh_last_error(void);

/* netlist + levelization (no GPU needed) */
int hh_circuit_load(const char* verilog_path, int is_arith, int strict_reference_levels, hh_circuit** out);
int hh_circuit_load_text(const char* verilog_text, int is_arith, int strict_reference_levels, hh_circuit** out);
void hh_circuit_free(hh_circuit*);
/* what: 0 gates, 1 wire_set, 2 inputs (with duplicates, DFF outputs appended), 3 outputs, 4 dff_outputs, 5 levels
 * (DFF level included), 6 has_luts, 7 has_arith, 8 bootstrapped gates (PBS gates; MUX counts once), 9 batches of the encrypted gates-mode
 * evaluation: levels with NOT / BUF folded into their consumers (a NOT is a level in the reference, src/circuit.rs:196-212, but no bootstrap) */
int64_t hh_circuit_count(const hh_circuit*, int what);
int64_t hh_circuit_level_width(const hh_circuit*, int level_index);
/* one name per line: what = 2 declared inputs (duplicates kept, DFF outputs appended), 3 outputs, 4 dff_outputs */
int hh_circuit_names(const hh_circuit*, int what, char* buf, size_t cap, size_t* needed);
/* "level:type:name:out:in0,in1,...\n" per gate in level order */
int hh_circuit_describe(const hh_circuit*, char* buf, size_t cap, size_t* needed);
/* plaintext evaluation: inputs as "wire=value\n" lines (empty => all false/0); ptxt_type "bool" | "u8".."u128";
 * cycles > 1 feeds the wire map back like the reference main loop.  Writes every wire with a value. */
int hh_circuit_eval_plain(const hh_circuit*, const char* inputs, const char* ptxt_type, int cycles,
                          char* buf, size_t cap, size_t* needed);
/* Yosys structural Verilog (gate cells) -> HELM netlist text: the hdl-benchmarks preprocessor for gate netlists
 * (reference hdl-benchmarks/src/lib.rs:319-391, 608-810; how c880 / c7552 of BASELINE.json configs[1] are produced) */
int hh_preprocess_yosys(const char* yosys_text, char* buf, size_t cap, size_t* needed);
/* the same with mode bits.  HH_PP_ARITH: the reference's `arith` switch (convert_verilog(..., arith), lib.rs:392-398) --
 * multi-bit nets stay whole and `assign y = a + b;` becomes `add a<id>(a, b, y);`.  HH_PP_EXACT_CONSTS: write the LUT
 * constants Yosys wrote; without it the reference's own reading is reproduced, which cuts `assign`-spelled hex constants
 * at every 'd' digit (8'hd0 -> 0x0; lib.rs:511-514) -- its checked-in *_lut*.v netlists carry those constants. */
enum { HH_PP_ARITH = 1, HH_PP_EXACT_CONSTS = 2 };
int hh_preprocess_yosys_mode(const char* yosys_text, int mode, char* buf, size_t cap, size_t* needed);
/* the per-bit assign dictionary as "dest=source" lines (build_assign_dict, hdl-benchmarks/src/lib.rs:319-391; the
 * reference tests it directly, hdl-benchmarks/tests/preprocessor_test.rs:8-250) */
int hh_preprocess_assign_dict(const char* yosys_text, int arith, char* buf, size_t cap, size_t* needed);
/* CSV helper (reference read_input_wires): file -> "wire=value\n" lines */
int hh_read_input_csv(const char* path, const char* ptxt_type, char* buf, size_t cap, size_t* needed);

/* the ChaCha20 block function (RFC 8439 2.3) behind every random word of the client side (keys, masks, noise): exported
 * for its known-answer test.  key8: 8 words, w12_15: counter / nonce words 12..15, out16: 16 keystream words. */
void hh_chacha20_block(const uint32_t* key8, const uint32_t* w12_15, uint32_t* out16);

/* encrypted evaluation on the GPU engine.  Keys: generated here (raw pointers NULL), or supplied by the caller (tfhe-rs in
 * the drop-in; the oracle in tests): lwe_sk n bytes, glwe_sk kN bytes, bsk/ksk standard domain.
 * seed == 0: the client generator (ChaCha20; secret keys, masks and noise of keys and input ciphertexts) is keyed with 256
 * bits from getrandom(2) -- the production setting.  seed != 0: reproducible keys from 64 bits, tests and benchmarks only. */
int hh_session_create(const char* param_set, int device, uint64_t seed,
                      const uint8_t* lwe_sk, const uint8_t* glwe_sk, const void* bsk, size_t bsk_bytes,
                      const void* ksk, size_t ksk_bytes, hh_session** out);
/* the same over several GPUs of this node (engine: hb_ctx_create_multi): keys and wires replicated per device, levels wider than
 * one gate per SM split across the devices inside hb_eval_level, outputs exchanged over NVLink -- the circuit surface is unchanged */
int hh_session_create_multi(const char* param_set, const int* devices, int n_devices, uint64_t seed,
                            const uint8_t* lwe_sk, const uint8_t* glwe_sk, const void* bsk, size_t bsk_bytes,
                            const void* ksk, size_t ksk_bytes, hh_session** out);
void hh_session_free(hh_session*);
/* encrypt_inputs -> cycles x (evaluate_encrypted [+ evaluate_ready]) -> decrypt_outputs.  Output wires as
 * "wire=value\n"; if all_wires != 0 every wire that carries a value is decrypted (the reference's tests compare every
 * wire with the plaintext evaluation).  seconds[3] = encrypt, evaluate, decrypt wall times (may be NULL). */
int hh_session_run(hh_session*, const hh_circuit*, const char* inputs, int cycles, int all_wires, int verbose,
                   char* buf, size_t cap, size_t* needed, double* seconds);
/* arithmetic mode (reference ArithCircuit, src/circuit.rs:1113-1483): ptxt_type "u8".."u128"; needs a session on a
 * parameter set with msg_mod = carry_mod = 4 (shortint_m2c2 / toy64).  counts[2] = bootstraps, batches (may be NULL). */
int hh_session_run_arith(hh_session*, const hh_circuit*, const char* inputs, const char* ptxt_type, int cycles,
                         int all_wires, int verbose, char* buf, size_t cap, size_t* needed, double* seconds,
                         uint64_t* counts);
/* engine statistics of the session: pbs_count, ks_count, levels, kernel_launches */
int hh_session_stats(hh_session*, uint64_t* out4);
/* multi-device statistics: n_devices, levels / batches split over more than one device, bytes pushed into peer replicas */
int hh_session_stats_multi(hh_session*, uint64_t* out3);

#ifdef __cplusplus
}
#endif
#endif
