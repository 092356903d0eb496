// Circuit: topological sort, levelization and plaintext evaluation of a HELM netlist -- the C++ mirror of
// /root/reference/src/circuit.rs:104-381 (Circuit::{new, sort_circuit, compute_levels, initialize_wire_map,
// evaluate}).  The level map it produces is what the B200 engine consumes: one hb_eval_level batch per level.
#pragma once
#include "netlist.hpp"

namespace helm {

bool is_numeric_string(const std::string& s);                 // src/circuit.rs:100-102

class Circuit {
public:
    // strict_reference_levels = true reproduces the reference's treatment of numeric wire names as constants
    // (src/circuit.rs:152, :202-203), which mis-levelizes netlists whose internal wires are bare
    // numbers.  The default treats a numeric token as a constant only if nothing drives it.
    explicit Circuit(const Netlist& nl, bool strict_reference_levels = false);

    void sort_circuit();
    void compute_levels();
    const std::vector<Gate>& ordered_gates() const { return ordered_; }
    // levels 1..depth hold combinational gates; DFFs sit in one extra final level (src/circuit.rs:225-234)
    const std::vector<std::vector<Gate>>& levels() const { return levels_; }
    size_t dff_level() const { return dff_level_; }            // index into levels(), or npos when there is none
    static constexpr size_t npos = (size_t)-1;

    WireValues initialize_wire_map(const WireValues& user_inputs, const std::string& ptxt_type) const;
    // plaintext reference semantics (src/gates.rs:151-239); int_bits = width of arithmetic wires (u8..u128)
    WireValues evaluate(const WireValues& wire_map, unsigned int_bits = 0) const;

    // Gates-mode schedule with NOT / BUF contracted (they cost no bootstrap, yet the reference's levels count them, src/circuit.rs:196-212):
    // every input of a gate is resolved through chains of NOT / BUF to (source wire, negated?); a bootstrapped gate sits in batch
    // 1 + the deepest batch of its sources; a NOT / BUF is evaluated from its resolved source in the batch after that source's.  The
    // DFF level stays last (sources as written).  gate points into levels().
    struct FoldedGate { const Gate* gate; std::vector<std::string> src; std::vector<bool> neg; };
    std::vector<std::vector<FoldedGate>> folded_batches() const;

    const Netlist& netlist() const { return nl_; }
    bool is_constant_operand(const std::string& wire) const;

private:
    Netlist nl_;
    bool strict_;
    std::set<std::string> driven_;                             // inputs + gate outputs
    std::vector<Gate> ordered_;
    std::vector<std::vector<Gate>> levels_;
    size_t dff_level_ = npos;
    bool sorted_ = false, levelled_ = false;
};

Ptxt evaluate_gate(const Gate& g, const std::vector<Ptxt>& in, unsigned int_bits);

}  // namespace helm
