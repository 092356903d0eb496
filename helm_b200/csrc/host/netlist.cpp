// See netlist.hpp.  Grammar of a processed HELM netlist (reference src/verilog_parser.rs:161-246):
// one statement per line; tokens split on ',' and ' '; `input/output [hi:lo] name;` expand to name[i] in the
// boolean modes; any other first token is a cell `<type> <inst>(<args>);`.
#include "netlist.hpp"
#include <algorithm>
#include <fstream>
#include <sstream>

namespace helm {

static const char* kTypeNames[] = {"and", "dff", "lut", "mux", "nand", "nor", "not", "or", "xnor", "xor", "buf",
                                   "cone", "czero", "add", "mult", "div", "sub", "shl", "shr", "copy"};
const char* gate_type_name(GateType t) { return kTypeNames[(int)t]; }

static GateType gate_type_from(const std::string& s) {
    for (int i = 0; i < 20; ++i)
        if (s == kTypeNames[i]) return (GateType)i;
    throw Error("Invalid gate type \"" + s + "\"");
}

static std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}
static std::string trim_end(std::string s, char c) { while (!s.empty() && s.back() == c) s.pop_back(); return s; }
static std::string trim_both(std::string s, char c) {
    while (!s.empty() && s.front() == c) s.erase(s.begin());
    return trim_end(s, c);
}
// strips the statement tail of an argument token: trailing ';' then trailing ')'
static std::string strip_tail(const std::string& s) { return trim_end(trim_end(s, ';'), ')'); }
static std::string strip_list(const std::string& s) { return trim_end(trim_both(s, ','), ';'); }

static std::vector<std::string> tokenize(const std::string& line) {
    std::vector<std::string> out; std::string cur;
    for (char c : line) {
        if (c == ',' || c == ' ') { if (!cur.empty()) out.push_back(cur); cur.clear(); }
        else cur.push_back(c);
    }
    if (!cur.empty()) out.push_back(cur);
    return out;
}

static bool parse_usize(const std::string& s, unsigned long long* v, int base = 10) {
    if (s.empty()) return false;
    size_t pos = 0;
    try { *v = std::stoull(s, &pos, base); } catch (...) { return false; }
    return pos == s.size() && s[0] != '-' && s[0] != '+';
}

// `[hi:lo]` or `[i]` -> (min, max); anything else: no range (reference parse_range, :122-136)
static bool parse_range(const std::string& tok, size_t* start, size_t* end) {
    std::string t = tok;
    while (!t.empty() && (t.front() == '[' || t.front() == ']')) t.erase(t.begin());
    while (!t.empty() && (t.back() == '[' || t.back() == ']')) t.pop_back();
    size_t colon = t.find(':');
    unsigned long long a, b;
    if (!parse_usize(colon == std::string::npos ? t : t.substr(0, colon), &a)) return false;
    b = a;
    if (colon != std::string::npos) {
        std::string rest = t.substr(colon + 1);
        size_t c2 = rest.find(':');
        if (!parse_usize(c2 == std::string::npos ? rest : rest.substr(0, c2), &b)) b = a;
    }
    *start = std::min(a, b); *end = std::max(a, b);
    return true;
}

Gate parse_gate(const std::vector<std::string>& tokens) {
    if (tokens.size() < 2) throw Error("Malformed cell statement");
    Gate g;
    g.type = gate_type_from(tokens[0]);
    // tokens[1] = "inst(arg0" : split on '(' and ','
    std::vector<std::string> head; { std::string cur;
        for (char c : tokens[1]) { if (c == '(' || c == ',') { if (!trim(cur).empty()) head.push_back(cur); cur.clear(); } else cur.push_back(c); }
        if (!trim(cur).empty()) head.push_back(cur); }
    if (head.empty()) throw Error("Malformed cell statement: " + tokens[1]);
    g.name = head[0];
    auto need = [&](size_t n) { if (tokens.size() < n) throw Error("Too few arguments for cell " + g.name); };
    switch (g.type) {
        case GateType::Not: case GateType::Dff: case GateType::Buf: case GateType::Copy:
            need(3); if (head.size() < 2) throw Error("Missing input of cell " + g.name);
            g.inputs = {trim(head[1])};
            g.output = strip_tail(tokens[2]);
            break;
        case GateType::Mux: case GateType::Lut:
            need(3); if (head.size() < 2) throw Error("Missing input of cell " + g.name);
            g.inputs.push_back(head[1]);
            for (size_t i = 2; i + 1 < tokens.size(); ++i) g.inputs.push_back(trim(trim_end(tokens[i], ',')));
            g.output = strip_tail(tokens.back());
            break;
        case GateType::ConstOne: case GateType::ConstZero: {
            size_t a = tokens[1].find('('), b = tokens[1].find(')', a == std::string::npos ? 0 : a);
            if (a == std::string::npos) throw Error("Opening parenthesis not found");
            if (b == std::string::npos) throw Error("Closing parenthesis not found");
            g.output = tokens[1].substr(a + 1, b - a - 1);
            break;
        }
        default:
            need(4); if (head.size() < 2) throw Error("Missing input of cell " + g.name);
            g.inputs = {head[1], trim(trim_end(tokens[2], ','))};
            g.output = strip_tail(tokens[3]);
    }
    if (g.type == GateType::Lut) {
        if (g.inputs.size() < 2) throw Error("LUT cell without inputs: " + g.name);
        std::string c = g.inputs.front();
        g.inputs.erase(g.inputs.begin());
        unsigned long long v = 0;
        if (c.rfind("0x", 0) == 0) { if (!parse_usize(c.substr(2), &v, 16)) throw Error("Failed to parse hex"); }
        else if (!parse_usize(c, &v)) throw Error("Failed to parse integer");
        if (g.inputs.size() > 6) throw Error("LUT arity above 6 is not representable (constant is 64-bit): " + g.name);
        const size_t entries = (size_t)1 << g.inputs.size();
        for (size_t i = 0; i < entries; ++i) g.lut_const.push_back((v >> i) & 1);
    }
    return g;
}

Netlist read_verilog_string(const std::string& text, bool is_arith) {
    Netlist nl;
    std::unordered_map<std::string, size_t> by_name;      // HashSet<Gate> hashes by gate name only (src/gates.rs:62-86)
    std::istringstream in(text);
    std::string raw;
    while (std::getline(in, raw)) {
        const std::string line = trim(raw);
        if (line.empty() || line.rfind("module", 0) == 0 || line.rfind("endmodule", 0) == 0 || line.rfind("//", 0) == 0) continue;
        const std::vector<std::string> tokens = tokenize(line);
        if (tokens.empty()) continue;
        if (tokens[0] == "input" || tokens[0] == "output") {
            std::vector<std::string>& dst = tokens[0] == "input" ? nl.inputs : nl.outputs;
            size_t lo, hi;
            if (tokens.size() > 2 && parse_range(tokens[1], &lo, &hi)) {
                if (is_arith) for (size_t i = 2; i < tokens.size(); ++i) dst.push_back(strip_list(tokens[i]));
                else { const std::string name = strip_list(tokens[2]);
                       for (size_t i = lo; i <= hi; ++i) dst.push_back(name + "[" + std::to_string(i) + "]"); }
            } else {
                for (size_t i = 1; i < tokens.size(); ++i) dst.push_back(strip_list(tokens[i]));
            }
        } else if (tokens[0] == "wire") {
            continue;                                      // declared wires are not used (:217-221)
        } else {
            Gate g = parse_gate(tokens);
            if (g.type == GateType::Dff) { nl.inputs.push_back(g.output); nl.dff_outputs.push_back(g.output); }
            else if (g.type == GateType::Lut) nl.has_luts = true;
            else if (is_arith_type(g.type)) nl.has_arith = true;
            nl.wire_set.insert(g.output);
            if (!by_name.count(g.name)) { by_name[g.name] = nl.gates.size(); nl.gates.push_back(std::move(g)); }
        }
    }
    if (nl.gates.empty())
        throw Error(nl.has_arith ? "[!] Parser error, no arithmetic gates detected."
                                 : "[!] Parser error, no gates detected. Make sure to use the 'no-expr' flag in Yosys.");
    if (nl.has_arith && nl.has_luts) throw Error("Can't mix LUTs with arithmetic operators!");
    return nl;
}

Netlist read_verilog_file(const std::string& file_name, bool is_arith) {
    std::ifstream f(file_name);
    if (!f) throw Error("Failed to open file: " + file_name);
    std::stringstream ss; ss << f.rdbuf();
    return read_verilog_string(ss.str(), is_arith);
}

// ---------------------------------------------------------------- values, CSV
unsigned ptxt_type_bits(const std::string& t) {
    static const std::pair<const char*, unsigned> kTypes[] = {{"bool", 1}, {"u8", 8}, {"u16", 16}, {"u32", 32}, {"u64", 64}, {"u128", 128}};
    for (const auto& kv : kTypes)
        if (t == kv.first) return kv.second;
    throw Error("unknown plaintext type " + t);
}

static std::string u128_to_string(unsigned __int128 v) {
    if (v == 0) return "0";
    std::string s; while (v) { s.push_back((char)('0' + (int)(v % 10))); v /= 10; }
    std::reverse(s.begin(), s.end()); return s;
}
std::string ptxt_to_string(const Ptxt& p) {
    if (p.kind == Ptxt::Bool) return p.v ? "true" : "false";
    if (p.kind == Ptxt::Int) return u128_to_string(p.v);
    return "None";
}

Ptxt parse_input_wire(const std::string& text, const std::string& ptxt_type) {
    const std::string s = trim(text);
    if (ptxt_type == "bool") return Ptxt::boolean(s == "1" || s == "true");     // anything else is false (src/lib.rs:92-97)
    const unsigned bits = ptxt_type_bits(ptxt_type);
    // integers are parsed before trimming in the reference (wire.parse::<uN>()): surrounding blanks are an error
    if (text.empty() || text.find_first_not_of("0123456789") != std::string::npos) throw Error("invalid integer input \"" + text + "\"");
    unsigned __int128 v = 0;
    for (char c : text) {
        const unsigned __int128 nv = v * 10 + (unsigned)(c - '0');
        if (nv / 10 != v) throw Error("integer input out of range: " + text);
        v = nv;
    }
    if (bits < 128 && (v >> bits) != 0) throw Error("integer input out of range for " + ptxt_type + ": " + text);
    return Ptxt::integer(v);
}

std::string hex_to_bitstring(const std::string& hex) {
    std::string out;
    for (char c : hex) {
        int d = c >= '0' && c <= '9' ? c - '0' : c >= 'a' && c <= 'f' ? c - 'a' + 10 : c >= 'A' && c <= 'F' ? c - 'A' + 10 : -1;
        if (d < 0) throw Error(std::string("invalid hex digit '") + c + "'");
        for (int b = 3; b >= 0; --b) out.push_back(((d >> b) & 1) ? '1' : '0');
    }
    return out;
}

// expands `wire, hexvalue, width` into wire[0..width) LSB first, zero padded (src/verilog_parser.rs:290-305)
void expand_bus_value(const std::string& wire, const std::string& hex, size_t width, WireValues* out) {
    std::string bits = hex_to_bitstring(trim(hex));
    std::reverse(bits.begin(), bits.end());
    for (size_t i = 0; i < width; ++i)
        (*out)[wire + "[" + std::to_string(i) + "]"] = Ptxt::boolean(i < bits.size() && bits[i] == '1');
}

WireValues read_input_wires(const std::string& file_name, const std::string& ptxt_type) {
    std::ifstream f(file_name);
    if (!f) throw Error("Failed to open CSV file: " + file_name);
    WireValues m;
    std::string line;
    bool header = true;                                     // csv::Reader skips the first record as the header
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) continue;
        if (header) { header = false; continue; }
        std::vector<std::string> rec; { std::string cur; for (char c : line) { if (c == ',') { rec.push_back(cur); cur.clear(); } else cur.push_back(c); } rec.push_back(cur); }
        const std::string name = trim(rec[0]);
        if (rec.size() == 2) m[name] = parse_input_wire(trim(rec[1]), ptxt_type);
        else if (rec.size() == 3 && ptxt_type == "bool") {
            unsigned long long width = 0;
            if (!parse_usize(trim(rec[2]), &width)) throw Error("bad width in CSV record: " + line);
            if (width > 1) expand_bus_value(name, rec[1], (size_t)width, &m);
            else m[name] = parse_input_wire(trim(rec[1]), ptxt_type);
        } else throw Error("The CSV should contain either two or three columns");
    }
    return m;
}

void write_output_wires(const std::string& file_name, const WireValues& values) {
    std::ofstream f(file_name);
    if (!f) throw Error("Failed to create CSV file: " + file_name);
    for (const auto& kv : values) f << kv.first << ", " << ptxt_to_string(kv.second) << "\n";
}

}  // namespace helm
