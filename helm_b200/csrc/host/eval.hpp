// Encrypted circuit evaluation on the B200 engine: the C++ mirror of the reference's trait EvalCircuit<C>
// (/root/reference/src/circuit.rs:35-58) and of its implementations GateCircuit (449-577) and LutCircuit (969-1111).
//
// What changes underneath: the reference walks the level map and runs gates.par_iter_mut() on CPU threads, one
// tfhe-rs call per gate (circuit.rs:524-543, 1050-1077).  Here every level is ONE hb_eval_level call (include/
// helm_b200.h) on a device-resident wire array; ciphertexts cross PCIe only in encrypt_inputs / decrypt_outputs.
#pragma once
#include <memory>
#include "circuit.hpp"
#include "client.hpp"

namespace helm {

// Server key = engine context with the bootstrapping / keyswitching keys resident in HBM.
class ServerKey {
public:
    ServerKey(const Params& p, const ServerKeyMaterial& keys, int device = 0);
    ServerKey(const Params& p, const void* bsk, size_t bsk_bytes, const void* ksk, size_t ksk_bytes, int device = 0);
    // one server key over several GPUs of the node (hb_ctx_create_multi): keys replicated, wide levels sharded inside the engine
    ServerKey(const Params& p, const void* bsk, size_t bsk_bytes, const void* ksk, size_t ksk_bytes, const std::vector<int>& devices);
    ~ServerKey();
    ServerKey(const ServerKey&) = delete;
    hb_ctx* ctx() const { return ctx_; }
    const Params& params() const { return p_; }
    hb_stats_t stats() const;
private:
    void init(const void* bsk, size_t bsk_bytes, const void* ksk, size_t ksk_bytes, const std::vector<int>& devices);
    Params p_;
    hb_ctx* ctx_ = nullptr;
};

// The reference's HashMap<String, Ciphertext>: names -> rows of one device array.
class EncWireMap {
public:
    EncWireMap(const ServerKey& sk, const std::vector<std::string>& names, uint32_t extra_rows = 0);
    ~EncWireMap();
    EncWireMap(const EncWireMap&) = delete;
    bool contains(const std::string& w) const { return index_.count(w) != 0; }
    uint32_t at(const std::string& w) const;
    uint32_t size() const { return n_rows_; }
    hb_wires* handle() const { return w_; }
    void write(const std::vector<uint32_t>& idx, const void* cts);
    void read(const std::vector<uint32_t>& idx, void* cts) const;
    void set_trivial(const std::vector<uint32_t>& idx, uint64_t plaintext);
    const std::unordered_map<std::string, uint32_t>& index() const { return index_; }
private:
    const ServerKey& sk_;
    hb_wires* w_ = nullptr;
    std::unordered_map<std::string, uint32_t> index_;
    uint32_t n_rows_ = 0;
};

class EvalCircuit {
public:
    virtual ~EvalCircuit() = default;
    // every wire := trivial false/0, inputs := encrypt(value) (all false when `inputs` is empty or holds "dummy"),
    // DFF outputs := encrypt(false)            (circuit.rs:450-480 / 970-1000)
    virtual std::unique_ptr<EncWireMap> encrypt_inputs(const std::set<std::string>& wire_set, const WireValues& inputs) = 0;
    // one pass over the level map; wires are updated in place on the device and the same map is returned
    virtual EncWireMap& evaluate_encrypted(EncWireMap& enc, size_t cycle, const std::string& ptxt_type) = 0;
    // READY-latched outputs of sequential circuits (circuit.rs:482-504 / 1002-1030): rows after the named wires
    virtual void init_ready(EncWireMap& enc) = 0;
    virtual void evaluate_ready(EncWireMap& enc) = 0;
    virtual WireValues decrypt_outputs(const EncWireMap& enc, bool verbose, bool latched = false) = 0;
    bool quiet = false;                       // suppress the reference's progress lines
};

class BatchedCircuit : public EvalCircuit {
public:
    BatchedCircuit(const ClientKey& ck, const ServerKey& sk, const Circuit& c, bool lut_mode);
    std::unique_ptr<EncWireMap> encrypt_inputs(const std::set<std::string>& wire_set, const WireValues& inputs) override;
    EncWireMap& evaluate_encrypted(EncWireMap& enc, size_t cycle, const std::string& ptxt_type) override;
    void init_ready(EncWireMap& enc) override;
    void evaluate_ready(EncWireMap& enc) override;
    WireValues decrypt_outputs(const EncWireMap& enc, bool verbose, bool latched = false) override;
    uint64_t gates_bootstrapped() const { return n_pbs_gates_; }
    // Gates mode: NOT / BUF cost no bootstrap (ciphertext negation / copy), yet the reference's levelizer gives them a level of their own
    // (src/circuit.rs:196-212), i.e. one more sequential batch.  With fold_linear (default; HELM_NO_FOLD=1 turns it off) a consumer
    // reads the NOT's source wire with a negated coefficient (hb_gate.flags), so NOT / BUF add no depth; they are still evaluated (their
    // wires are observable), as linear jobs of the batch after their source.  Results are bit-identical: -a + b is the same wrapping sum.
    bool fold_linear = true;
    size_t batches_per_cycle() const { return lowered_.size(); }
protected:
    std::vector<hb_gate> lower_level(const EncWireMap& enc, const std::vector<Gate>& gates) const;
    void lower_folded(const EncWireMap& enc);
    const ClientKey& ck_;
    const ServerKey& sk_;
    const Circuit& circuit_;
    bool lut_mode_;
    uint32_t latch_base_ = 0;                 // first of the |outputs| latched rows (+ 2 temporaries each in LUT mode)
    std::vector<std::vector<hb_gate>> lowered_;
    const EncWireMap* lowered_for_ = nullptr;
    uint64_t n_pbs_gates_ = 0;
};

// GateCircuit: tfhe::boolean (u32 torus, PBS then KS).  LutCircuit: tfhe::shortint (u64 torus, KS then PBS).
struct GateCircuit : BatchedCircuit {
    GateCircuit(const ClientKey& ck, const ServerKey& sk, const Circuit& c) : BatchedCircuit(ck, sk, c, false) {}
};
struct LutCircuit : BatchedCircuit {
    LutCircuit(const ClientKey& ck, const ServerKey& sk, const Circuit& c) : BatchedCircuit(ck, sk, c, true) {}
};

}  // namespace helm
