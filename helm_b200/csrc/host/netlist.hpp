// HELM netlist front-end for the B200 engine's host side (C++ mirror of the reference's Rust surface).
//
// Mirrors, by name and behaviour, /root/reference/src/verilog_parser.rs:
//   read_verilog_file (138-276), parse_gate (31-120), read_input_wires (278-317), write_output_wires (319-348)
// and the Gate / GateType of src/gates.rs:23-86.  Errors that are panics in the reference are exceptions
// (helm::Error) here; the C API and the CLI turn them into status codes / exit code 101 like a Rust panic.
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace helm {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

enum class GateType : uint8_t {
    And, Dff, Lut, Mux, Nand, Nor, Not, Or, Xnor, Xor, Buf, ConstOne, ConstZero,
    Add, Mult, Div, Sub, Shl, Shr, Copy
};
const char* gate_type_name(GateType t);
inline bool is_arith_type(GateType t) { return t >= GateType::Add; }

struct Gate {
    std::string name;
    GateType type = GateType::And;
    std::vector<std::string> inputs;
    std::vector<uint64_t> lut_const;   // LUT: 1 << arity entries, entry i = bit i of the constant (LSB first)
    std::string output;
    size_t level = 0;
};

// value of a plaintext wire: Bool or an unsigned integer of the --arithmetic width (src/lib.rs:13-22)
struct Ptxt {
    enum Kind : uint8_t { None, Bool, Int } kind = None;
    unsigned __int128 v = 0;
    static Ptxt boolean(bool b) { Ptxt p; p.kind = Bool; p.v = b; return p; }
    static Ptxt integer(unsigned __int128 x) { Ptxt p; p.kind = Int; p.v = x; return p; }
    bool operator==(const Ptxt& o) const { return kind == o.kind && v == o.v; }
};
std::string ptxt_to_string(const Ptxt& p);
unsigned ptxt_type_bits(const std::string& ptxt_type);     // "bool" -> 1, "u8" -> 8, ... "u128" -> 128

struct Netlist {
    std::vector<Gate> gates;                 // HashSet<Gate> keyed by name in the reference: duplicates collapse
    std::set<std::string> wire_set;          // every gate output
    std::vector<std::string> inputs;         // declared inputs + DFF outputs (appended, :225-227); may hold duplicates
    std::vector<std::string> outputs;
    std::vector<std::string> dff_outputs;
    bool has_luts = false, has_arith = false;
};

Netlist read_verilog_file(const std::string& file_name, bool is_arith);
Netlist read_verilog_string(const std::string& text, bool is_arith);
Gate parse_gate(const std::vector<std::string>& tokens);

using WireValues = std::map<std::string, Ptxt>;
WireValues read_input_wires(const std::string& file_name, const std::string& ptxt_type);
Ptxt parse_input_wire(const std::string& text, const std::string& ptxt_type);     // src/lib.rs:90-111
std::string hex_to_bitstring(const std::string& hex);                              // src/lib.rs:183-194
void expand_bus_value(const std::string& wire, const std::string& hex, size_t width, WireValues* out);
void write_output_wires(const std::string& file_name, const WireValues& values);
// Yosys netlist -> HELM netlist text (preprocess.cpp; reference hdl-benchmarks/src/lib.rs): gate cells, both LUT
// spellings, multi-bit assigns; arith = true keeps multi-bit nets whole and turns `assign y = a + b` into `add` cells
std::string preprocess_yosys(const std::string& yosys_text, bool arith, bool exact_consts = false);
std::string preprocess_yosys_gates(const std::string& yosys_text);       // = preprocess_yosys(text, false)
// per-bit assign dictionary, "dest=source" lines (reference build_assign_dict, lib.rs:319-391)
std::string preprocess_assign_dict(const std::string& yosys_text, bool arith);

}  // namespace helm
