// See eval.hpp.
#include "eval.hpp"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace helm {

static void ck_status(hb_ctx* c, hb_status s, const char* what) {
    if (s != HB_OK) throw Error(std::string(what) + " failed (" + std::to_string(s) + "): " + hb_last_error(c));
}

// ---------------------------------------------------------------- ServerKey
void ServerKey::init(const void* bsk, size_t bsk_bytes, const void* ksk, size_t ksk_bytes, const std::vector<int>& devices) {
    if (devices.empty()) throw Error("ServerKey: no device");
    std::vector<int32_t> dv(devices.begin(), devices.end());
    hb_status s = dv.size() == 1 ? hb_ctx_create(dv[0], &ctx_) : hb_ctx_create_multi(dv.data(), (int32_t)dv.size(), 0, &ctx_);
    if (s != HB_OK) throw Error(std::string("hb_ctx_create failed: ") + hb_last_error(nullptr));
    try {
        ck_status(ctx_, hb_set_params(ctx_, &p_.hb), "hb_set_params");
        ck_status(ctx_, hb_load_bsk(ctx_, bsk, bsk_bytes), "hb_load_bsk");
        ck_status(ctx_, hb_load_ksk(ctx_, ksk, ksk_bytes), "hb_load_ksk");
    } catch (...) { hb_ctx_destroy(ctx_); ctx_ = nullptr; throw; }
}
ServerKey::ServerKey(const Params& p, const ServerKeyMaterial& k, int device) : p_(p) {
    init(k.bsk.data(), k.bsk.size(), k.ksk.data(), k.ksk.size(), {device});
}
ServerKey::ServerKey(const Params& p, const void* bsk, size_t bb, const void* ksk, size_t kb, int device) : p_(p) {
    init(bsk, bb, ksk, kb, {device});
}
ServerKey::ServerKey(const Params& p, const void* bsk, size_t bb, const void* ksk, size_t kb, const std::vector<int>& devices) : p_(p) {
    init(bsk, bb, ksk, kb, devices);
}
ServerKey::~ServerKey() { if (ctx_) hb_ctx_destroy(ctx_); }
hb_stats_t ServerKey::stats() const { hb_stats_t st{}; ck_status(ctx_, hb_stats(ctx_, &st), "hb_stats"); return st; }

// ---------------------------------------------------------------- EncWireMap
EncWireMap::EncWireMap(const ServerKey& sk, const std::vector<std::string>& names, uint32_t extra) : sk_(sk) {
    for (const auto& n : names)
        if (!index_.count(n)) { const uint32_t i = (uint32_t)index_.size(); index_[n] = i; }
    n_rows_ = (uint32_t)index_.size() + extra;
    ck_status(sk.ctx(), hb_wires_create(sk.ctx(), n_rows_ ? n_rows_ : 1, &w_), "hb_wires_create");
}
EncWireMap::~EncWireMap() { if (w_) hb_wires_destroy(w_); }
uint32_t EncWireMap::at(const std::string& w) const {
    auto it = index_.find(w);
    if (it == index_.end()) throw Error("wire \"" + w + "\" is not in the encrypted wire map");
    return it->second;
}
void EncWireMap::write(const std::vector<uint32_t>& idx, const void* cts) {
    ck_status(sk_.ctx(), hb_wires_write(w_, idx.data(), (uint32_t)idx.size(), cts), "hb_wires_write");
}
void EncWireMap::read(const std::vector<uint32_t>& idx, void* cts) const {
    ck_status(sk_.ctx(), hb_wires_read(w_, idx.data(), (uint32_t)idx.size(), cts), "hb_wires_read");
}
void EncWireMap::set_trivial(const std::vector<uint32_t>& idx, uint64_t pt) {
    if (idx.empty()) return;
    std::vector<uint64_t> pts(idx.size(), pt);
    ck_status(sk_.ctx(), hb_wires_set_trivial(w_, idx.data(), (uint32_t)idx.size(), pts.data()), "hb_wires_set_trivial");
}

// ---------------------------------------------------------------- BatchedCircuit
BatchedCircuit::BatchedCircuit(const ClientKey& ck, const ServerKey& sk, const Circuit& c, bool lut)
    : ck_(ck), sk_(sk), circuit_(c), lut_mode_(lut) {
    if ((sk.params().hb.order == 1) != lut) throw Error(lut ? "LutCircuit needs shortint (KS->PBS) parameters" : "GateCircuit needs boolean (PBS->KS) parameters");
}

std::unique_ptr<EncWireMap> BatchedCircuit::encrypt_inputs(const std::set<std::string>& wire_set, const WireValues& inputs) {
    const Netlist& nl = circuit_.netlist();
    std::vector<std::string> names(wire_set.begin(), wire_set.end());
    names.insert(names.end(), nl.inputs.begin(), nl.inputs.end());
    names.insert(names.end(), nl.dff_outputs.begin(), nl.dff_outputs.end());
    const uint32_t per_out = lut_mode_ ? 3 : 1;               // latched value (+ 2 temporaries in LUT mode)
    auto enc = std::make_unique<EncWireMap>(sk_, names, (uint32_t)nl.outputs.size() * per_out);
    latch_base_ = (uint32_t)enc->index().size();
    // every wire starts as trivial false / 0
    std::vector<uint32_t> all(enc->size());
    for (uint32_t i = 0; i < enc->size(); ++i) all[i] = i;
    enc->set_trivial(all, lut_mode_ ? 0 : ck_.trivial_bool_plaintext(false));
    // client-side encryption of the inputs and of the DFF state
    const bool all_false = inputs.empty() || inputs.count("dummy");
    const size_t words = (size_t)sk_.params().ct_dim() + 1, wb = sk_.params().word_bytes();
    std::vector<uint32_t> idx; std::vector<uint8_t> cts;
    auto push = [&](const std::string& w, bool v) {
        idx.push_back(enc->at(w));
        cts.resize(cts.size() + words * wb);
        void* dst = cts.data() + cts.size() - words * wb;
        if (lut_mode_) ck_.encrypt_shortint(v ? 1 : 0, dst); else ck_.encrypt_bool(v, dst);
    };
    std::set<std::string> seen;
    for (const auto& w : nl.inputs) {
        if (!seen.insert(w).second) continue;                 // the input list may hold duplicates (bus + bits)
        bool v = false;
        if (!all_false) {
            auto it = inputs.find(w);
            if (it == inputs.end()) throw Error("\n Input wire \"" + w + "\" not in input wires!");
            if (it->second.kind != Ptxt::Bool) throw Error("boolean modes need Bool inputs");
            v = it->second.v != 0;
        }
        push(w, v);
    }
    for (const auto& w : nl.dff_outputs) {                    // overrides an input of the same name: encrypt(false)
        auto pos = std::find(idx.begin(), idx.end(), enc->at(w));
        if (pos != idx.end()) {
            void* dst = cts.data() + (size_t)(pos - idx.begin()) * words * wb;
            if (lut_mode_) ck_.encrypt_shortint(0, dst); else ck_.encrypt_bool(false, dst);
        } else push(w, false);
    }
    if (!idx.empty()) enc->write(idx, cts.data());
    return enc;
}

// Gate -> hb_gate (what Gate::evaluate_encrypted, src/gates.rs:241-280, and lut(), :754-785, dispatch per gate)
std::vector<hb_gate> BatchedCircuit::lower_level(const EncWireMap& enc, const std::vector<Gate>& gates) const {
    std::vector<hb_gate> out;
    out.reserve(gates.size());
    const int mc = sk_.params().hb.msg_mod * sk_.params().hb.carry_mod;
    for (const Gate& g : gates) {
        hb_gate h{};
        h.out = enc.at(g.output);
        auto in = [&](size_t i) { if (i >= g.inputs.size()) throw Error("gate " + g.name + ": missing input"); return enc.at(g.inputs[i]); };
        switch (g.type) {
            case GateType::And: h.op = HB_AND; break;   case GateType::Nand: h.op = HB_NAND; break;
            case GateType::Or: h.op = HB_OR; break;     case GateType::Nor: h.op = HB_NOR; break;
            case GateType::Xor: h.op = HB_XOR; break;   case GateType::Xnor: h.op = HB_XNOR; break;
            case GateType::Mux: h.op = HB_MUX; break;   case GateType::Not: h.op = HB_NOT; break;
            case GateType::Dff: case GateType::Buf: h.op = HB_COPY; break;
            case GateType::Lut: h.op = HB_LUT; break;
            default: throw Error(std::string("gate type \"") + gate_type_name(g.type) + "\" cannot be evaluated in " + (lut_mode_ ? "LUT" : "gates") + " mode");
        }
        if (h.op == HB_LUT) {
            if (!lut_mode_) throw Error("Can't mix LUTs with Boolean gates!");     // src/gates.rs:273-276
            const size_t a = g.inputs.size();
            if (a < 1 || a > 4 || (1 << a) > mc)
                throw Error("LUT " + g.name + " has " + std::to_string(a) + " inputs: needs msg_mod*carry_mod >= " + std::to_string(1 << a) +
                            " (parameter set " + sk_.params().name + " has " + std::to_string(mc) + "; SURVEY App. C-4)");
            h.arity = (uint8_t)a;
            for (size_t i = 0; i < a; ++i) h.in[i] = in(i);
            for (size_t i = 0; i < g.lut_const.size(); ++i) if (g.lut_const[i] & 1) h.lut |= 1ull << i;
        } else if (h.op == HB_MUX) { h.in[0] = in(0); h.in[1] = in(1); h.in[2] = in(2); }
        else if (h.op == HB_NOT || h.op == HB_COPY) h.in[0] = in(0);
        else { h.in[0] = in(0); h.in[1] = in(1); }            // the encrypted path uses exactly inputs 0 and 1 (:255-271)
        out.push_back(h);
    }
    return out;
}

// Batches with NOT / BUF folded into their consumers (see eval.hpp).  Every combinational wire is resolved to (source wire, negated?)
// through chains of NOT / BUF; a bootstrapped gate's depth is 1 + the largest depth of its resolved sources; a NOT / BUF is evaluated
// from its resolved source in the batch after that source's.  The DFF level stays last.
void BatchedCircuit::lower_folded(const EncWireMap& enc) {
    const auto batches = circuit_.folded_batches();
    const size_t dffl = circuit_.dff_level();
    std::vector<std::vector<hb_gate>> out;
    for (size_t bi = 0; bi < batches.size(); ++bi) {
        if (dffl != Circuit::npos && bi + 1 == batches.size()) { out.push_back(lower_level(enc, circuit_.levels()[dffl])); break; }
        std::vector<hb_gate> hg;
        for (const auto& f : batches[bi]) {
            const Gate& g = *f.gate;
            hb_gate h{};
            if (g.type == GateType::Not || g.type == GateType::Buf) {
                h.op = f.neg.at(0) ? HB_NOT : HB_COPY; h.in[0] = enc.at(f.src.at(0)); h.out = enc.at(g.output);
            } else {
                h = lower_level(enc, std::vector<Gate>{g}).at(0);
                const size_t nin = h.op == HB_MUX ? 3 : 2;        // the encrypted path uses exactly these inputs (src/gates.rs:255-271)
                for (size_t s = 0; s < nin; ++s) {
                    h.in[s] = enc.at(f.src.at(s));
                    if (f.neg.at(s)) h.flags |= (uint16_t)(1u << s);
                }
            }
            hg.push_back(h);
        }
        out.push_back(std::move(hg));
    }
    lowered_ = std::move(out);
}

EncWireMap& BatchedCircuit::evaluate_encrypted(EncWireMap& enc, size_t /*cycle*/, const std::string& /*ptxt_type*/) {
    const auto& levels = circuit_.levels();
    if (levels.empty()) throw Error("evaluate_encrypted: compute_levels has not run");
    const bool fold = fold_linear && !lut_mode_ && !std::getenv("HELM_NO_FOLD");
    if (lowered_for_ != &enc) {                                // wire indices are per map: lower once, reuse every cycle
        lowered_.clear();
        if (fold) lower_folded(enc);
        else for (const auto& lv : levels) lowered_.push_back(lower_level(enc, lv));
        lowered_for_ = &enc;
    }
    const size_t total = lowered_.size();
    for (size_t li = 0; li < total; ++li) {
        const auto& hg = lowered_[li];
        ck_status(sk_.ctx(), hb_eval_level(sk_.ctx(), enc.handle(), hg.data(), (uint32_t)hg.size()), "hb_eval_level");
        for (const auto& h : hg) n_pbs_gates_ += (h.op != HB_NOT && h.op != HB_COPY);
        if (!quiet) {
            const size_t key = fold ? li + 1 : li == circuit_.dff_level() ? total : levels[li].front().level;
            std::printf("  Evaluated gates in level [%zu/%zu]\n", key, total);
        }
    }
    ck_status(sk_.ctx(), hb_sync(sk_.ctx()), "hb_sync");
    return enc;
}

void BatchedCircuit::init_ready(EncWireMap& enc) {
    const size_t n_out = circuit_.netlist().outputs.size();
    std::vector<uint32_t> idx;
    for (uint32_t i = 0; i < n_out * (lut_mode_ ? 3u : 1u); ++i) idx.push_back(latch_base_ + i);
    enc.set_trivial(idx, lut_mode_ ? 0 : ck_.trivial_bool_plaintext(false));
}

// valid[o] = READY ? wire[o] : valid[o]   for every output wire (circuit.rs:492-504; LUT mode :1012-1030)
void BatchedCircuit::evaluate_ready(EncWireMap& enc) {
    const auto& outs = circuit_.netlist().outputs;
    const uint32_t ready = enc.at("READY"), n_out = (uint32_t)outs.size();
    std::vector<hb_gate> l1, l2;
    for (uint32_t o = 0; o < n_out; ++o) {
        if (!enc.contains(outs[o])) continue;
        const uint32_t cur = enc.at(outs[o]), valid = latch_base_ + o;
        if (!lut_mode_) {
            hb_gate h{}; h.op = HB_MUX; h.in[0] = cur; h.in[1] = valid; h.in[2] = ready; h.out = valid;
            l1.push_back(h);
        } else {   // new*READY + old*(1-READY) as three bivariate LUTs (fits every message space >= 2 bits)
            const uint32_t t1 = latch_base_ + n_out + 2 * o, t2 = t1 + 1;
            hb_gate a{}; a.op = HB_LUT; a.arity = 2; a.in[0] = cur; a.in[1] = ready; a.lut = 0x8; a.out = t1;     // cur & READY
            hb_gate b{}; b.op = HB_LUT; b.arity = 2; b.in[0] = valid; b.in[1] = ready; b.lut = 0x4; b.out = t2;   // valid & !READY
            hb_gate c{}; c.op = HB_LUT; c.arity = 2; c.in[0] = t1; c.in[1] = t2; c.lut = 0xE; c.out = valid;      // or
            l1.push_back(a); l1.push_back(b); l2.push_back(c);
        }
    }
    ck_status(sk_.ctx(), hb_eval_level(sk_.ctx(), enc.handle(), l1.data(), (uint32_t)l1.size()), "hb_eval_level");
    if (!l2.empty()) ck_status(sk_.ctx(), hb_eval_level(sk_.ctx(), enc.handle(), l2.data(), (uint32_t)l2.size()), "hb_eval_level");
    ck_status(sk_.ctx(), hb_sync(sk_.ctx()), "hb_sync");
}

WireValues BatchedCircuit::decrypt_outputs(const EncWireMap& enc, bool verbose, bool latched) {
    const auto& outs = circuit_.netlist().outputs;
    const size_t words = (size_t)sk_.params().ct_dim() + 1, wb = sk_.params().word_bytes();
    std::vector<uint32_t> idx;
    for (uint32_t o = 0; o < outs.size(); ++o) idx.push_back(latched ? latch_base_ + o : enc.at(outs[o]));
    std::vector<uint8_t> cts(idx.size() * words * wb);
    if (!idx.empty()) enc.read(idx, cts.data());
    WireValues res;
    for (size_t o = 0; o < outs.size(); ++o) {
        const void* ct = cts.data() + o * words * wb;
        res[outs[o]] = lut_mode_ ? Ptxt::integer(ck_.decrypt_shortint(ct)) : Ptxt::boolean(ck_.decrypt_bool(ct));
    }
    if (!quiet) {
        size_t i = 0;
        for (const auto& kv : res) {                           // sorted by name, as the reference prints them
            if (i++ > 10 && !verbose) { std::printf("[!] More than ten output_wires, pass `--verbose` to see output.\n"); break; }
            std::printf(" %s: %s\n", kv.first.c_str(), ptxt_to_string(kv.second).c_str());
        }
    }
    return res;
}

}  // namespace helm
