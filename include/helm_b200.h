/*
 * helm_b200.h -- C ABI of the B200-native batched TFHE bootstrapping engine behind HELM's
 * gate / LUT evaluation.  Plain pointers and sizes only; no exceptions cross the boundary.
 *
 * What each entry point replaces in the reference (TrustworthyComputing/helm @ fc9f75f):
 *   hb_ctx_create / hb_set_params / hb_load_bsk / hb_load_ksk
 *        key conversion done once per run: src/bin/helm.rs:187-192 (convert_lwe_bootstrap_key,
 *        convert_lwe_keyswitch_key on the reference's own GPU path) and the ServerKey built by
 *        gen_keys(), src/bin/helm.rs:241 / :301.
 *   hb_wires_*   the wire map HashMap<String, Arc<RwLock<Ciphertext>>> of
 *        src/circuit.rs:517-520 (and the per-level H2D/D2H staging of src/circuit.rs:745-761,
 *        877-892, which disappears: wires stay resident in HBM between levels).
 *   hb_eval_level   one iteration of the level loop, src/circuit.rs:524-543 (gates.par_iter_mut()
 *        calling Gate::evaluate_encrypted, src/gates.rs:241-280; LUT mode src/circuit.rs:1050-1077
 *        calling lut(), src/gates.rs:754-785) -- the reference's precedent for this batched shape
 *        is discard_<gate>_lwe_ciphertext_vector(out, in1, in2, bsk, ksk), src/circuit.rs:807-873.
 *   hb_pbs_batch / hb_keyswitch_batch   the two tfhe-rs primitives underneath
 *        (programmable_bootstrap_lwe_ciphertext, keyswitch_lwe_ciphertext).
 *
 * Threading: one hb_ctx is bound to one CUDA device (hb_ctx_create) or to a group of devices (hb_ctx_create_multi) and driven
 * by one host thread at a time; work is stream-ordered on the context's stream(s).  Distinct contexts share no mutable state and
 * may be driven from different threads, on the same or on different devices.  Host buffers are only read/written during the
 * call.  The engine never sees a secret key.
 * There is no CPU fallback: every compute entry point returns HB_ERR_CUDA when no device exists.
 */
#ifndef HELM_B200_H
#define HELM_B200_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t hb_status;
enum { HB_OK = 0, HB_ERR_ARG = 1, HB_ERR_CUDA = 2, HB_ERR_STATE = 3, HB_ERR_UNSUPPORTED = 4 };

typedef struct hb_ctx hb_ctx;
typedef struct hb_wires hb_wires;

typedef struct {
    int32_t n, k, N;                    /* small LWE dim, GLWE dim, polynomial size */
    int32_t pbs_base_log, pbs_level;
    int32_t ks_base_log, ks_level;
    int32_t torus_bits;                 /* 32 | 64 */
    int32_t order;                      /* 0 = PBS then KS (boolean; wires under the small key)
                                           1 = KS then PBS (shortint; wires under the big key) */
    int32_t msg_mod, carry_mod;         /* shortint only */
    int32_t grouping;                   /* 0 / 1: classic blind rotation.  g > 1 (u64, k = 1, N = 2048, pbs_level = 1 instance): multi-bit
                                           blind rotation with grouping factor g -- the reference's arithmetic-mode set
                                           PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_3_KS_PBS (src/bin/helm.rs:81-86): n/g sequential steps, the
                                           key holds 2^g GGSWs per group of g key bits (pattern q encrypts prod_j (bit_j(q) ? s_j : 1 - s_j)) */
} hb_params;

/* gate opcodes (GateType of src/gates.rs:23-45 that reach the FHE library) */
enum { HB_AND = 0, HB_NAND, HB_OR, HB_NOR, HB_XOR, HB_XNOR, HB_MUX, HB_NOT, HB_COPY, HB_LUT };

typedef struct {
    uint8_t op;          /* HB_* */
    uint8_t arity;       /* LUT: number of inputs (1..4) */
    uint16_t flags;      /* bit s: input s is taken NEGATED (boolean parameter sets only): a NOT gate folded into its consumer, so that NOT
                            levels (src/circuit.rs:196-212 counts NOT as a level) cost no extra batch; 0 otherwise */
    uint32_t in[4];      /* wire indices; MUX: in[0]=then, in[1]=else, in[2]=select (src/gates.rs:265) */
    uint32_t out;        /* wire index */
    uint64_t lut;        /* LUT: truth table, bit i = output for index i; index = sum in[j] << (arity-1-j)
                            (first input is the MSB, src/gates.rs:158-181) */
} hb_gate;

typedef struct {
    uint64_t pbs_count, ks_count, levels;
    double pbs_ms, ks_ms;               /* device time of the last hb_eval_level / batch call */
    uint64_t kernel_launches;           /* engine kernels launched since context creation */
    double pbs_ms_total, ks_ms_total;   /* device time summed over every hb_eval_level since creation */
    double gap_ms_total;                /* stream time between consecutive batches' kernels (job upload, launch gaps) */
    uint64_t push_bytes;                /* multi-device contexts: bytes stored into peer replicas over NVLink (level exchange) */
    uint32_t n_devices;                 /* 1, or the size of the hb_ctx_create_multi group */
    uint32_t sharded_levels;            /* levels / batches that were split over more than one device */
} hb_stats_t;

hb_status hb_ctx_create(int32_t device, hb_ctx** out);
/* One context over n_devices GPUs of this node (1 <= n <= 8, distinct, all peer-accessible: NVLink / NVSwitch).  The reference's
 * own GPU backend owns its device handles inside the circuit object the same way (src/circuit.rs:81-91, 650-934).  Keys are
 * replicated on every device (hb_load_bsk / hb_load_ksk load all of them), every hb_wires array has one replica per device, and
 * hb_eval_level / hb_eval_shortint_ops split a wide batch into contiguous slices, one per device; each device stores its output
 * rows straight into every peer's replica (peer-mapped stores right behind the producing kernels) and a per-level event per
 * device orders the next level after the exchange -- no host round trip (the reference gathers every level through host memory,
 * src/circuit.rs:745-761, 877-892) and no collective library.  Batches narrower than min_gates_per_device x 2 stay on the first
 * device (its outputs are still pushed, so the replicas stay coherent).  min_gates_per_device = 0 picks the default (one gate per
 * SM: a narrower slice cannot make a level faster).  Every other entry point takes the group context like an ordinary one;
 * hb_wires_read* read the first replica, hb_wires_write* / hb_wires_set_trivial write all of them, hb_stream / hb_pbs_batch /
 * hb_keyswitch_batch address the first device, hb_wires_attach is not supported. */
hb_status hb_ctx_create_multi(const int32_t* devices, int32_t n_devices, uint32_t min_gates_per_device, hb_ctx** out);
hb_status hb_ctx_destroy(hb_ctx*);
const char* hb_last_error(const hb_ctx*);          /* ctx may be NULL: last creation error */
hb_status hb_set_params(hb_ctx*, const hb_params*);
/* standard-domain GGSW list, layout [n][level 1..l][row 0..k][poly 0..k][N] (multi-bit: [n/g][2^g][level][row][poly][N]);
 * transformed once, resident */
hb_status hb_load_bsk(hb_ctx*, const void* bsk, size_t bytes);
/* layout [kN][level 1..l_ks][n+1]; resident */
hb_status hb_load_ksk(hb_ctx*, const void* ksk, size_t bytes);

/* device-resident wire array [n_wires][stride]; ciphertext words = n+1 (order 0) or kN+1 (order 1) */
hb_status hb_wires_create(hb_ctx*, uint32_t n_wires, hb_wires** out);
/* wrap caller-owned device memory (e.g. a torch tensor used for the per-level NCCL all-gather) */
hb_status hb_wires_attach(hb_ctx*, void* device_ptr, uint32_t n_wires, hb_wires** out);
hb_status hb_wires_destroy(hb_wires*);
uint32_t hb_ct_words(const hb_ctx*);                /* words per ciphertext on the wire */
uint32_t hb_wire_stride(const hb_ctx*);             /* words per wire row in device memory */
void* hb_wires_device_ptr(hb_wires*);
/* host <-> device: packed rows of hb_ct_words() torus words */
hb_status hb_wires_write(hb_wires*, const uint32_t* idx, uint32_t cnt, const void* lwe_host);
hb_status hb_wires_read(hb_wires*, const uint32_t* idx, uint32_t cnt, void* lwe_host);
/* contiguous rows [first, first+cnt): one strided copy straight from/to the caller's (ideally pinned) buffer */
hb_status hb_wires_write_rows(hb_wires*, uint32_t first, uint32_t cnt, const void* lwe_host);
hb_status hb_wires_read_rows(hb_wires*, uint32_t first, uint32_t cnt, void* lwe_host);
/* The same row IO, asynchronous on the context's copy stream, for pipelined serving loops (the reference stages every level synchronously
 * through host memory, src/circuit.rs:745-761, 877-892): the upload of batch s+1 and the download of batch s-1 overlap the evaluation of batch
 * s.  write: every hb_eval_* call issued AFTER it sees the rows; read: returns the rows as of the work enqueued BEFORE it.  Host buffers
 * (pinned for real overlap) must stay valid, and read results are complete, after hb_io_sync.  The rows of an upload in flight must not be read or
 * written by a batch issued earlier that is still running.  Single-device contexts only. */
hb_status hb_wires_write_rows_async(hb_wires*, uint32_t first, uint32_t cnt, const void* lwe_host);
hb_status hb_wires_read_rows_async(hb_wires*, uint32_t first, uint32_t cnt, void* lwe_host);
hb_status hb_io_sync(hb_ctx*);
/* the upload stream of the context (cudaStream_t as void*) and a fence: work a caller enqueues on that stream after an asynchronous upload
 * (one process per GPU: the all-gather of the ranks' input shares) is ordered before every hb_eval_* call issued after hb_io_fence */
void* hb_io_stream(hb_ctx*);
hb_status hb_io_fence(hb_ctx*);
hb_status hb_wires_set_trivial(hb_wires*, const uint32_t* idx, uint32_t cnt, const uint64_t* plaintext);

/* evaluate every gate of one level as one batched launch sequence (asynchronous on the ctx stream).
 * No gate of the batch may read a wire written by another gate of the same batch. */
hb_status hb_eval_level(hb_ctx*, hb_wires*, const hb_gate* gates, uint32_t n_gates);
/* shortint block operations (order 1 parameter sets), the batched form of what tfhe-rs's integer layer issues block
 * by block underneath FheUintN `+ - * << >>` (reached from reference src/gates.rs:306-702 through the high-level
 * API): unchecked_add / unchecked_scalar_add / unchecked_sub (kind HB_SOP_LINEAR, no bootstrap) and
 * apply_lookup_table / unchecked_apply_lookup_table_bivariate (kind HB_SOP_LUT: one keyswitch + one PBS).
 * v = sum_s coef[s] * in[s] + cst (cst in message units, i.e. multiples of delta); LINEAR: out = v;
 * LUT: out = encoding of table entry (v mod msg_mod*carry_mod), 4 bits per entry, entry i at bits [4i, 4i+4).
 * Same batch rule as hb_eval_level: no op may read a wire another op of the batch writes (out == own input is fine
 * for LINEAR). */
enum { HB_SOP_LINEAR = 0, HB_SOP_LUT = 1 };
typedef struct {
    uint8_t kind;        /* HB_SOP_* */
    uint8_t n_in;        /* 0..4 (0: LINEAR writes the trivial encryption of cst) */
    int8_t coef[4];
    int16_t cst;
    uint32_t in[4];
    uint32_t out;
    uint64_t table;
} hb_sop;
hb_status hb_eval_shortint_ops(hb_ctx*, hb_wires*, const hb_sop* ops, uint32_t n_ops);
hb_status hb_sync(hb_ctx*);
hb_status hb_stats(hb_ctx*, hb_stats_t* out);
/* CUDA-graph capture of a sequence of batches (a cycle of a sequential circuit issues the same batches every time; the reference
 * re-issues every gate per cycle, src/bin/helm.rs:256-270).  Between hb_graph_begin and hb_graph_end the hb_eval_level /
 * hb_eval_shortint_ops / hb_wires_set_trivial calls on the context are RECORDED, not executed (hb_sync is a no-op; no other entry point
 * may be used); job_bytes bounds the descriptors of the recorded batches (~ 200 bytes per gate is ample).  The same batches must have run
 * eagerly once before, so that the context's scratch buffers have their final size.  hb_graph_launch replays the sequence on the
 * context's stream (asynchronous; any number of times); it fails with HB_ERR_STATE if a later, larger batch reallocated the scratch
 * buffers the graph points into.  Single-device contexts only. */
typedef struct hb_graph hb_graph;
hb_status hb_graph_begin(hb_ctx*, size_t job_bytes);
hb_status hb_graph_end(hb_ctx*, hb_graph** out);
hb_status hb_graph_launch(hb_graph*);
hb_status hb_graph_destroy(hb_graph*);
/* CUDA stream of the context (cudaStream_t as void*), for callers that interleave collectives */
void* hb_stream(hb_ctx*);

/* primitives on host buffers (unit parity with the oracle):
 * lwe_in: count x (n+1); test_vector_bodies: count x N body polynomials (1 x N if shared_tv != 0; the
 * mask polynomials of the accumulator are zero, as in every caller on this path); out: count x (kN+1) */
hb_status hb_pbs_batch(hb_ctx*, const void* lwe_in, const void* test_vector_bodies, int32_t shared_tv,
                       uint32_t count, void* lwe_out);
/* lwe_in: count x (kN+1); out: count x (n+1) */
hb_status hb_keyswitch_batch(hb_ctx*, const void* lwe_in, uint32_t count, void* lwe_out);
/* Keyswitch of `count` rows that allocates NOTHING (no cudaMalloc / cudaFree / allocating library call), for callers under CUDA
 * graph capture, in tight serving loops or in harnesses that forbid allocation.  No counterpart in the reference.
 * lwe_in_dev: count x (kN+1) packed torus words in DEVICE memory; lwe_out_dev: count x (n+1) in
 * DEVICE memory (the caller's buffers; the call is asynchronous on the context's stream -- hb_sync before reading, and make the
 * inputs visible to that stream first).  Same arithmetic as hb_keyswitch_batch (bit-identical).  The scratch of the tensor-core
 * keyswitch (job descriptors + int8 digit matrix) lives in one static device array of HB_NOALLOC_SCRATCH_BYTES that every GPU
 * loading the library holds until exit: count is limited to floor((HB_NOALLOC_SCRATCH_BYTES - 64 KiB) / (kN * ks_level)) rounded
 * down to a multiple of 64 and to at most 1344 rows (boolean default: 1344, shortint M2C2: 768); larger counts return HB_ERR_ARG with
 * the limit in hb_last_error.  All calls on one device share that array: calls that overlap on the GPU (two contexts / streams) race.
 * Parameter sets without the tensor-core keyswitch return HB_ERR_UNSUPPORTED. */
#define HB_NOALLOC_SCRATCH_BYTES (8u << 20)
hb_status hb_keyswitch_batch_noalloc(hb_ctx*, const void* lwe_in_dev, uint32_t count, void* lwe_out_dev);
/* measured FP64 FMA throughput of this device in TFLOP/s (register-resident DFMA loop) */
hb_status hb_measure_fp64_peak(hb_ctx*, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
