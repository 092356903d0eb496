// Yosys structural Verilog -> HELM netlist, gate cells only: the restatement of the hdl-benchmarks preprocessor
// (/root/reference/hdl-benchmarks/src/lib.rs:319-391 build_assign_dict, :392-900 convert_verilog; recipe in
// RUNNING.md:220-238) for what BASELINE.json configs[1] needs -- ISCAS'85 c880 / c7552 exist in the reference only as
// Yosys output (hdl-benchmarks/netlists/*.v).
//
// What it does, as the reference does it:
//   * `assign a = b;` between single-bit nets builds a substitution: an internal net assigned FROM a port is replaced
//     by the port, an internal net assigned TO an output port is replaced by that output; chains are followed until a
//     non-substituted name or an output port is reached (lib.rs:777-783);
//   * every `\$_AND_  _17_ ( .A(x), .B(y), .Y(z) );` cell becomes `and g_17_(x', y', z');` with ports in the order
//     they are listed, so `\$_MUX_` (.A,.B,.S,.Y) becomes `mux g(A, B, S, Y)` -- NOTE: HELM evaluates that as
//     S ? A : B (src/gates.rs:189-194, :265) while Yosys means S ? B : A; parity is with the reference, so the order
//     is kept;
//   * `\$_DFF_P_` cells drop the clock port (lib.rs:791-806);
//   * an input (or another output) assigned straight to an output becomes a `buf` (lib.rs:868-885).
// LUT cells ($lut / assign ... >> {...}) and arithmetic assigns are NOT restated here.
#include <algorithm>
#include <cctype>
#include <map>
#include <set>
#include <sstream>
#include "netlist.hpp"

namespace helm {

static std::string pp_trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) ++a;
    while (b > a && std::isspace((unsigned char)s[b - 1])) --b;
    return s.substr(a, b - a);
}
static std::vector<std::string> pp_names(const std::string& decl) {      // "input a, b;" -> {a, b}
    std::vector<std::string> out;
    std::string cur;
    const size_t sp = decl.find(' ');
    for (char ch : decl.substr(sp == std::string::npos ? decl.size() : sp + 1)) {
        if (ch == ',' || ch == ';') { cur = pp_trim(cur); if (!cur.empty()) out.push_back(cur); cur.clear(); }
        else cur.push_back(ch);
    }
    cur = pp_trim(cur);
    if (!cur.empty()) out.push_back(cur);
    return out;
}

std::string preprocess_yosys_gates(const std::string& text) {
    std::vector<std::string> lines;
    {
        std::istringstream in(text);
        std::string l;
        while (std::getline(in, l)) {
            l = pp_trim(l);
            if (l.empty() || l.rfind("(*", 0) == 0 || l.rfind("/*", 0) == 0 || l.rfind("//", 0) == 0) continue;
            lines.push_back(l);
        }
    }
    std::string module_line;
    std::vector<std::string> wires, inputs, outputs;
    std::set<std::string> in_set, out_set;
    std::map<std::string, std::string> subst;
    std::vector<std::pair<std::string, std::string>> direct;              // input -> output
    // pass 1: declarations and assigns
    for (const auto& l : lines) {
        if (l.rfind("module", 0) == 0) module_line = l;
        else if (l.rfind("wire ", 0) == 0) { for (auto& n : pp_names(l)) wires.push_back(n); }
        else if (l.rfind("input ", 0) == 0) { for (auto& n : pp_names(l)) { inputs.push_back(n); in_set.insert(n); } }
        else if (l.rfind("output ", 0) == 0) { for (auto& n : pp_names(l)) { outputs.push_back(n); out_set.insert(n); } }
    }
    for (const auto& l : lines) {
        if (l.rfind("assign ", 0) != 0) continue;
        if (l.find_first_of("+-*/{>~&|^?") != std::string::npos)
            throw Error("preprocess_yosys_gates: only net-to-net assigns are supported: " + l);
        const size_t eq = l.find('=');
        if (eq == std::string::npos) continue;
        std::string lhs = pp_trim(l.substr(7, eq - 7)), rhs = pp_trim(l.substr(eq + 1));
        if (!rhs.empty() && rhs.back() == ';') rhs = pp_trim(rhs.substr(0, rhs.size() - 1));
        if (out_set.count(lhs) && (in_set.count(rhs) || out_set.count(rhs))) direct.emplace_back(rhs, lhs);   // port -> output: buf
        else if (out_set.count(lhs)) {
            auto it = subst.find(rhs);
            if (it != subst.end() && out_set.count(it->second)) direct.emplace_back(it->second, lhs);   // second output of one net
            else subst[rhs] = lhs;                                       // internal net drives an output port
        } else subst[lhs] = rhs;                                         // internal net is an alias of rhs
    }
    auto resolve = [&](std::string w) {
        for (int guard = 0; guard < 64; ++guard) {
            auto it = subst.find(w);
            if (it == subst.end()) break;
            w = it->second;
            if (out_set.count(w)) break;
        }
        return w;
    };
    // pass 2: cells
    std::vector<std::string> gates;
    for (size_t i = 0; i < lines.size(); ++i) {
        const std::string& l = lines[i];
        if (l.rfind("\\$_", 0) != 0) continue;
        std::istringstream ts(l);
        std::string cell, inst;
        ts >> cell >> inst;
        std::string type = cell.substr(3, cell.size() - 4);               // \$_AND_ -> AND
        std::transform(type.begin(), type.end(), type.begin(), [](unsigned char c) { return (char)std::tolower(c); });
        const bool is_dff = type.find("dff") != std::string::npos;
        if (is_dff) type = "dff";
        std::vector<std::pair<std::string, std::string>> ports;
        for (++i; i < lines.size() && lines[i].rfind(");", 0) != 0; ++i) {
            const std::string& pl = lines[i];
            const size_t dot = pl.find('.'), op = pl.find('('), cp = pl.find(')');
            if (dot == std::string::npos || op == std::string::npos || cp == std::string::npos) continue;
            ports.emplace_back(pl.substr(dot + 1, op - dot - 1), pp_trim(pl.substr(op + 1, cp - op - 1)));
        }
        std::string g = type + " g" + inst + "(";
        bool first = true;
        for (const auto& p : ports) {
            if (is_dff && p.first == "C") continue;                       // the clock is implicit in HELM
            if (!first) g += ", ";
            g += resolve(p.second);
            first = false;
        }
        g += ");";
        gates.push_back(g);
    }
    if (module_line.empty() || gates.empty()) throw Error("preprocess_yosys_gates: no module or no gate cells found");
    auto join = [](const std::vector<std::string>& v) {
        std::string s;
        for (size_t i = 0; i < v.size(); ++i) { if (i) s += ", "; s += v[i]; }
        return s;
    };
    std::string out = module_line + "\n";
    out += "  wire " + join(wires) + ";\n";
    if (!inputs.empty()) out += "  input " + join(inputs) + ";\n";
    if (!outputs.empty()) out += "  output " + join(outputs) + ";\n";
    for (const auto& g : gates) out += "  " + g + "\n";
    for (size_t b = 0; b < direct.size(); ++b)
        out += "  buf b_" + std::to_string(b) + "_(" + direct[b].first + ", " + direct[b].second + ");\n";
    out += "\nendmodule\n";
    return out;
}

}  // namespace helm
