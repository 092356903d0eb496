// See circuit.hpp.  Follows /root/reference/src/circuit.rs:122-239 (sort + levels), :245-333 (wire map),
// :348-381 (plaintext evaluate) and src/gates.rs:151-239 (gate semantics).
#include "circuit.hpp"
#include <algorithm>

namespace helm {

bool is_numeric_string(const std::string& s) {
    return std::all_of(s.begin(), s.end(), [](char c) { return c >= '0' && c <= '9'; });
}
static bool parses_as_u32(const std::string& s) {             // wire.parse::<u32>().is_ok()
    if (s.empty() || s.size() > 11) return false;
    size_t i = s[0] == '+' ? 1 : 0;
    if (i == s.size()) return false;
    unsigned long long v = 0;
    for (; i < s.size(); ++i) { if (s[i] < '0' || s[i] > '9') return false; v = v * 10 + (s[i] - '0'); if (v > 0xffffffffull) return false; }
    return true;
}

Circuit::Circuit(const Netlist& nl, bool strict) : nl_(nl), strict_(strict) {
    for (const auto& w : nl_.inputs) driven_.insert(w);
    for (const auto& g : nl_.gates) driven_.insert(g.output);
}

bool Circuit::is_constant_operand(const std::string& wire) const {
    if (!parses_as_u32(wire)) return false;
    return strict_ || !driven_.count(wire);
}

void Circuit::sort_circuit() {
    if (nl_.gates.empty()) throw Error("sort_circuit: no gates");
    if (sorted_) throw Error("sort_circuit: already sorted");
    std::set<std::string> ready_wires(nl_.inputs.begin(), nl_.inputs.end());
    std::vector<Gate> pending = nl_.gates, dffs;
    while (!pending.empty()) {
        std::vector<Gate> level, rest;
        std::set<std::string> next;
        for (auto& g : pending) {
            if (g.type == GateType::Dff) { next.insert(g.output); dffs.push_back(g); continue; }
            // cone/czero cells are collected but never scheduled by the reference (:131, :142-147; SURVEY C-7)
            if (g.type == GateType::ConstOne || g.type == GateType::ConstZero) { next.insert(g.output); continue; }
            bool ok = true;
            for (const auto& w : g.inputs) if (!ready_wires.count(w) && !is_constant_operand(w)) { ok = false; break; }
            if (ok) { next.insert(g.output); level.push_back(g); } else rest.push_back(g);
        }
        if (level.empty() && next.empty() && !rest.empty())
            throw Error("sort_circuit: gate \"" + rest.front().name + "\" can never become ready (undriven input or combinational loop)");
        ready_wires.insert(next.begin(), next.end());
        std::sort(level.begin(), level.end(), [](const Gate& a, const Gate& b) { return a.name < b.name; });
        ordered_.insert(ordered_.end(), level.begin(), level.end());
        pending.swap(rest);
    }
    ordered_.insert(ordered_.end(), dffs.begin(), dffs.end());
    sorted_ = true;
}

void Circuit::compute_levels() {
    if (!sorted_) throw Error("compute_levels: sort_circuit has not run");
    if (levelled_) throw Error("compute_levels: already done");
    std::unordered_map<std::string, size_t> wire_level;
    for (const auto& w : nl_.inputs) wire_level[w] = 0;
    std::map<size_t, std::vector<Gate>> by_level;
    std::vector<Gate> dffs;
    for (auto& g : ordered_) {
        if (g.type == GateType::Dff) { dffs.push_back(g); continue; }
        size_t depth = 0;
        for (const auto& w : g.inputs) {
            size_t d;
            auto it = wire_level.find(w);
            if (it != wire_level.end()) d = it->second;
            else if (is_constant_operand(w)) d = 0;
            else throw Error("Input " + w + " not found in wire_levels");
            depth = std::max(depth, d + 1);
        }
        g.level = depth;
        by_level[depth].push_back(g);
        wire_level[g.output] = depth;
    }
    levels_.clear();
    for (auto& kv : by_level) levels_.push_back(std::move(kv.second));
    if (!dffs.empty()) {
        dff_level_ = levels_.size();
        for (auto& g : dffs) g.level = by_level.size();      // the reference re-keys the DFF level to level_map.len()
        levels_.push_back(std::move(dffs));
    }
    levelled_ = true;
}

WireValues Circuit::initialize_wire_map(const WireValues& user, const std::string& ptxt_type) const {
    WireValues m;
    for (const auto& w : nl_.wire_set) m[w] = Ptxt();          // intermediate wires: None
    const bool is_bool = ptxt_type == "bool";
    ptxt_type_bits(ptxt_type);
    for (const auto& w : nl_.inputs) {
        if (user.empty()) m[w] = is_bool ? Ptxt::boolean(false) : Ptxt::integer(0);
        else {
            auto it = user.find(w);
            if (it == user.end()) throw Error("\n Input wire \"" + w + "\" not in input wires!");
            m[w] = it->second;
        }
    }
    for (const auto& w : nl_.dff_outputs) m[w] = is_bool ? Ptxt::boolean(false) : Ptxt::integer(0);
    return m;
}

static bool as_bool(const Ptxt& p) {
    if (p.kind != Ptxt::Bool) throw Error("Expected PtxtType::Bool variant");
    return p.v != 0;
}

Ptxt evaluate_gate(const Gate& g, const std::vector<Ptxt>& in, unsigned int_bits) {
    auto count_true = [&] { size_t c = 0; for (const auto& p : in) c += as_bool(p); return c; };
    const unsigned __int128 mask = int_bits >= 128 ? ~(unsigned __int128)0 : (((unsigned __int128)1 << int_bits) - 1);
    switch (g.type) {
        case GateType::Dff: case GateType::Buf: case GateType::Copy: return in.at(0);
        case GateType::And: return Ptxt::boolean(count_true() == in.size());
        case GateType::Nand: return Ptxt::boolean(count_true() != in.size());
        case GateType::Or: return Ptxt::boolean(count_true() > 0);
        case GateType::Nor: return Ptxt::boolean(count_true() == 0);
        case GateType::Xor: return Ptxt::boolean(count_true() % 2 == 1);
        case GateType::Xnor: return Ptxt::boolean(count_true() % 2 != 1);
        case GateType::Not: return Ptxt::boolean(!as_bool(in.at(0)));
        case GateType::Mux: return Ptxt::boolean(as_bool(in.at(2)) ? as_bool(in.at(0)) : as_bool(in.at(1)));   // S ? A : B
        case GateType::ConstOne: return Ptxt::boolean(true);
        case GateType::ConstZero: return Ptxt::boolean(false);
        case GateType::Lut: {
            size_t idx = 0;                                   // first input is the MSB (src/gates.rs:158-181)
            for (size_t i = 0; i < in.size(); ++i) if (as_bool(in[i])) idx += (size_t)1 << (in.size() - 1 - i);
            if (g.lut_const.empty()) throw Error("Lut const not provided");
            return Ptxt::boolean(g.lut_const.at(idx) & 1);
        }
        default: break;
    }
    // arithmetic cells: wrapping unsigned arithmetic at the --arithmetic width (what FheUintN computes).  The
    // reference's plaintext evaluate() is unreachable!() for these; this is the model its goldens follow.
    if (in.size() < 2 || in[0].kind != Ptxt::Int || in[1].kind != Ptxt::Int) throw Error("arithmetic cell needs two integer operands");
    const unsigned __int128 a = in[0].v & mask, b = in[1].v & mask;
    switch (g.type) {
        case GateType::Add: return Ptxt::integer((a + b) & mask);
        case GateType::Sub: return Ptxt::integer((a - b) & mask);
        case GateType::Mult: return Ptxt::integer((a * b) & mask);
        case GateType::Div: return Ptxt::integer(b == 0 ? mask : a / b);     // tfhe-rs: division by zero yields all ones
        case GateType::Shl: return Ptxt::integer((a << (unsigned)(b % int_bits)) & mask);
        case GateType::Shr: return Ptxt::integer(a >> (unsigned)(b % int_bits));
        default: throw Error("unsupported gate type in evaluate");
    }
}

WireValues Circuit::evaluate(const WireValues& wire_map, unsigned int_bits) const {
    if (!levelled_) throw Error("evaluate: compute_levels has not run");
    WireValues v = wire_map;
    for (const auto& level : levels_) {
        std::vector<std::pair<const Gate*, Ptxt>> results;     // gates of a level read the pre-level state
        for (const auto& g : level) {
            std::vector<Ptxt> in;
            for (const auto& w : g.inputs) {
                auto it = v.find(w);
                if (it != v.end()) in.push_back(it->second);
                else if (is_constant_operand(w) && is_arith_type(g.type)) in.push_back(parse_input_wire(w, "u128"));
                else throw Error("evaluate: wire \"" + w + "\" has no value");
            }
            results.emplace_back(&g, evaluate_gate(g, in, int_bits ? int_bits : 128));
        }
        for (auto& r : results) v[r.first->output] = r.second;
    }
    return v;
}

}  // namespace helm

namespace helm {

std::vector<std::vector<Circuit::FoldedGate>> Circuit::folded_batches() const {
    if (!levelled_) throw Error("folded_batches: compute_levels has not run");
    std::unordered_map<std::string, const Gate*> producer;
    for (size_t li = 0; li < levels_.size(); ++li)
        if (li != dff_level_) for (const Gate& g : levels_[li]) producer[g.output] = &g;
    auto resolve = [&](const std::string& w0, bool* neg_out) {
        std::string w = w0; bool neg = false;
        for (;;) {
            auto p = producer.find(w);
            if (p == producer.end() || (p->second->type != GateType::Not && p->second->type != GateType::Buf)) break;
            if (p->second->type == GateType::Not) neg = !neg;
            w = p->second->inputs.at(0);
        }
        *neg_out = neg;
        return w;
    };
    std::unordered_map<std::string, size_t> depth;             // of wires driven by bootstrapped gates; anything else is 0
    auto wire_depth = [&](const std::string& w) { auto it = depth.find(w); return it == depth.end() ? (size_t)0 : it->second; };
    std::vector<std::vector<FoldedGate>> out;
    auto at_batch = [&](size_t d) -> std::vector<FoldedGate>& { if (out.size() < d) out.resize(d); return out[d - 1]; };
    for (size_t li = 0; li < levels_.size(); ++li) {           // the level order is a topological order
        if (li == dff_level_) continue;
        for (const Gate& g : levels_[li]) {
            FoldedGate f; f.gate = &g;
            const bool linear = g.type == GateType::Not || g.type == GateType::Buf;
            size_t d = 0;
            if (linear) {
                bool neg; const std::string src = resolve(g.output, &neg);
                f.src.push_back(src); f.neg.push_back(neg);
                d = wire_depth(src);
            } else {
                for (const auto& w : g.inputs) {
                    bool neg; const std::string src = resolve(w, &neg);
                    f.src.push_back(src); f.neg.push_back(neg);
                    d = std::max(d, wire_depth(src));
                }
                depth[g.output] = d + 1;
            }
            at_batch(d + 1).push_back(std::move(f));
        }
    }
    if (dff_level_ != npos) {
        out.emplace_back();
        for (const Gate& g : levels_[dff_level_]) {
            FoldedGate f; f.gate = &g; f.src = g.inputs; f.neg.assign(g.inputs.size(), false);
            out.back().push_back(std::move(f));
        }
    }
    return out;
}

}  // namespace helm
