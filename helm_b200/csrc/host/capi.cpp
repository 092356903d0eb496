// C ABI of the host library (include/helm_host.h).
#include "../../../include/helm_host.h"
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include "eval.hpp"
#include "arith.hpp"

using namespace helm;

struct hh_circuit { std::unique_ptr<Circuit> c; bool is_arith = false; };
struct hh_session {
    Params p; ClientKey ck; std::unique_ptr<ServerKey> sk;
};

static thread_local std::string g_err;
const char* hh_last_error(void) { return g_err.c_str(); }

template <class F> static int guarded(F f) {
    try { f(); return 0; }
    catch (const std::exception& e) { g_err = e.what(); return 1; }
    catch (...) { g_err = "unknown error"; return 1; }
}

static int emit(const std::string& s, char* buf, size_t cap, size_t* needed) {
    if (needed) *needed = s.size() + 1;
    if (buf && cap) { const size_t m = std::min(cap - 1, s.size()); memcpy(buf, s.data(), m); buf[m] = 0; }
    if (s.size() + 1 > cap) { g_err = "output buffer too small"; return 2; }
    return 0;
}

static WireValues parse_values(const char* text, const std::string& ptxt_type) {
    WireValues m;
    if (!text) return m;
    std::istringstream in(text);
    std::string line;
    while (std::getline(in, line)) {
        if (line.empty()) continue;
        const size_t eq = line.find('=');
        if (eq == std::string::npos) throw Error("bad input line (want wire=value): " + line);
        m[line.substr(0, eq)] = parse_input_wire(line.substr(eq + 1), ptxt_type);
    }
    return m;
}
static std::string format_values(const WireValues& v) {
    std::string s;
    for (const auto& kv : v) if (kv.second.kind != Ptxt::None) s += kv.first + "=" + ptxt_to_string(kv.second) + "\n";
    return s;
}

static int make_circuit(Netlist nl, int is_arith, int strict, hh_circuit** out) {
    auto h = std::make_unique<hh_circuit>();
    h->is_arith = is_arith != 0;
    h->c = std::make_unique<Circuit>(nl, strict != 0);
    h->c->sort_circuit();
    h->c->compute_levels();
    *out = h.release();
    return 0;
}

extern "C" {

int hh_circuit_load(const char* path, int is_arith, int strict, hh_circuit** out) {
    return guarded([&] { make_circuit(read_verilog_file(path, is_arith != 0), is_arith, strict, out); });
}
int hh_circuit_load_text(const char* text, int is_arith, int strict, hh_circuit** out) {
    return guarded([&] { make_circuit(read_verilog_string(text, is_arith != 0), is_arith, strict, out); });
}
void hh_circuit_free(hh_circuit* c) { delete c; }

int64_t hh_circuit_count(const hh_circuit* c, int what) {
    const Netlist& nl = c->c->netlist();
    switch (what) {
        case 0: return (int64_t)nl.gates.size();
        case 1: return (int64_t)nl.wire_set.size();
        case 2: return (int64_t)nl.inputs.size();
        case 3: return (int64_t)nl.outputs.size();
        case 4: return (int64_t)nl.dff_outputs.size();
        case 5: return (int64_t)c->c->levels().size();
        case 6: return nl.has_luts;
        case 7: return nl.has_arith;
        case 8: { int64_t n = 0;
                  for (const auto& lv : c->c->levels()) for (const auto& g : lv)
                      n += !(g.type == GateType::Not || g.type == GateType::Buf || g.type == GateType::Dff || g.type == GateType::Copy);
                  return n; }
        case 9: return c->c->netlist().has_luts || c->c->netlist().has_arith ? (int64_t)c->c->levels().size() : (int64_t)c->c->folded_batches().size();
    }
    return -1;
}
int64_t hh_circuit_level_width(const hh_circuit* c, int li) {
    const auto& lv = c->c->levels();
    return li < 0 || (size_t)li >= lv.size() ? -1 : (int64_t)lv[li].size();
}
int hh_circuit_names(const hh_circuit* c, int what, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] {
        const Netlist& nl = c->c->netlist();
        const std::vector<std::string>* v = what == 2 ? &nl.inputs : what == 3 ? &nl.outputs : what == 4 ? &nl.dff_outputs : nullptr;
        if (!v) throw Error("hh_circuit_names: what must be 2 (inputs), 3 (outputs) or 4 (dff_outputs)");
        for (const auto& w : *v) s += w + "\n";
    }) || emit(s, buf, cap, needed);
}
int hh_circuit_describe(const hh_circuit* c, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] {
        for (const auto& lv : c->c->levels()) for (const auto& g : lv) {
            s += std::to_string(g.level) + ":" + gate_type_name(g.type) + ":" + g.name + ":" + g.output + ":";
            for (size_t i = 0; i < g.inputs.size(); ++i) s += (i ? "," : "") + g.inputs[i];
            s += "\n";
        }
    }) || emit(s, buf, cap, needed);
}

int hh_circuit_eval_plain(const hh_circuit* c, const char* inputs, const char* ptxt_type, int cycles, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] {
        const std::string t = ptxt_type ? ptxt_type : "bool";
        WireValues m = c->c->initialize_wire_map(parse_values(inputs, t), t);
        const unsigned bits = t == "bool" ? 0 : ptxt_type_bits(t);
        for (int cy = 0; cy < (cycles < 1 ? 1 : cycles); ++cy) m = c->c->evaluate(m, bits);
        s = format_values(m);
    }) || emit(s, buf, cap, needed);
}

int hh_preprocess_yosys(const char* yosys_text, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] { s = preprocess_yosys_gates(yosys_text ? yosys_text : ""); }) || emit(s, buf, cap, needed);
}

int hh_preprocess_yosys_mode(const char* yosys_text, int mode, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] { s = preprocess_yosys(yosys_text ? yosys_text : "", (mode & 1) != 0, (mode & 2) != 0); }) || emit(s, buf, cap, needed);
}

int hh_preprocess_assign_dict(const char* yosys_text, int arith, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] { s = preprocess_assign_dict(yosys_text ? yosys_text : "", arith != 0); }) || emit(s, buf, cap, needed);
}

void hh_chacha20_block(const uint32_t* key8, const uint32_t* w12_15, uint32_t* out16) { chacha20_block(key8, w12_15, out16); }

int hh_read_input_csv(const char* path, const char* ptxt_type, char* buf, size_t cap, size_t* needed) {
    std::string s;
    return guarded([&] { s = format_values(read_input_wires(path, ptxt_type ? ptxt_type : "bool")); }) || emit(s, buf, cap, needed);
}

}  // extern "C"

// ---------------------------------------------------------------- encrypted sessions
extern "C" {

int hh_session_create_multi(const char* param_set, const int* devices, int n_devices, uint64_t seed, const uint8_t* lwe_sk,
                            const uint8_t* glwe_sk, const void* bsk, size_t bsk_bytes, const void* ksk, size_t ksk_bytes, hh_session** out) {
    return guarded([&] {
        if (!devices || n_devices < 1) throw Error("hh_session_create_multi: no devices");
        const std::vector<int> dv(devices, devices + n_devices);
        auto s = std::make_unique<hh_session>();
        s->p = params_by_name(param_set);
        if (lwe_sk && glwe_sk && bsk && ksk) {
            s->ck = ClientKey::from_raw(s->p, lwe_sk, glwe_sk, seed);
            s->sk = std::make_unique<ServerKey>(s->p, bsk, bsk_bytes, ksk, ksk_bytes, dv);
        } else {
            ServerKeyMaterial m;
            gen_keys(s->p, seed, &s->ck, &m);
            s->sk = std::make_unique<ServerKey>(s->p, m.bsk.data(), m.bsk.size(), m.ksk.data(), m.ksk.size(), dv);
        }
        *out = s.release();
    });
}
int hh_session_create(const char* param_set, int device, uint64_t seed, const uint8_t* lwe_sk, const uint8_t* glwe_sk,
                      const void* bsk, size_t bsk_bytes, const void* ksk, size_t ksk_bytes, hh_session** out) {
    return hh_session_create_multi(param_set, &device, 1, seed, lwe_sk, glwe_sk, bsk, bsk_bytes, ksk, ksk_bytes, out);
}
void hh_session_free(hh_session* s) { delete s; }

int hh_session_run(hh_session* s, const hh_circuit* c, const char* inputs, int cycles, int all_wires, int verbose,
                   char* buf, size_t cap, size_t* needed, double* seconds) {
    std::string text;
    return guarded([&] {
        using clk = std::chrono::steady_clock;
        auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
        const Netlist& nl = c->c->netlist();
        const bool lut = nl.has_luts;
        if (nl.has_arith) throw Error("arithmetic circuits go through the integer layer (hh_arith_*), not hh_session_run");
        BatchedCircuit ev(s->ck, *s->sk, *c->c, lut);
        ev.quiet = !verbose;
        const WireValues in = parse_values(inputs, "bool");
        const auto t0 = clk::now();
        auto enc = ev.encrypt_inputs(nl.wire_set, in);
        const bool ready = cycles > 1 && enc->contains("READY");
        if (ready) ev.init_ready(*enc);
        const auto t1 = clk::now();
        // Sequential circuits: every cycle issues the same batches.  Cycle 0 runs eagerly (it also sizes the engine's scratch buffers),
        // cycle 1 is recorded into a CUDA graph (hb_graph_begin / hb_graph_end) and that graph is launched for cycle 1 and every later
        // cycle: one launch per cycle instead of ~3 launches + a descriptor upload per level.  HELM_NO_GRAPH=1 keeps every cycle eager.
        hb_ctx* ctx = s->sk->ctx();
        hb_stats_t st0{}; hb_stats(ctx, &st0);
        const bool use_graph = cycles >= 3 && st0.n_devices == 1 && !std::getenv("HELM_NO_GRAPH");
        hb_graph* graph = nullptr;
        bool in_capture = false;
        try {
            for (int cy = 0; cy < (cycles < 1 ? 1 : cycles); ++cy) {
                if (graph) {
                    if (hb_graph_launch(graph) != HB_OK) throw Error(std::string("hb_graph_launch failed: ") + hb_last_error(ctx));
                    continue;
                }
                const bool capture = use_graph && cy == 1;
                if (capture) {
                    const size_t bytes = 256 * (nl.gates.size() + 4 * nl.outputs.size()) + 1024 * (c->c->levels().size() + 4) + 65536;
                    if (hb_graph_begin(ctx, bytes) != HB_OK) throw Error(std::string("hb_graph_begin failed: ") + hb_last_error(ctx));
                    in_capture = true;
                }
                ev.evaluate_encrypted(*enc, 1, "bool");        // the reference passes the literal 1 (helm.rs:258-263)
                if (ready) ev.evaluate_ready(*enc);
                if (capture) {
                    in_capture = false;
                    if (hb_graph_end(ctx, &graph) != HB_OK) throw Error(std::string("hb_graph_end failed: ") + hb_last_error(ctx));
                    if (hb_graph_launch(graph) != HB_OK) throw Error(std::string("hb_graph_launch failed: ") + hb_last_error(ctx));
                }
            }
            if (hb_sync(ctx) != HB_OK) throw Error(std::string("hb_sync failed: ") + hb_last_error(ctx));
        } catch (...) {
            if (in_capture) { hb_graph* dead = nullptr; hb_graph_end(ctx, &dead); if (dead) hb_graph_destroy(dead); }   // abort the capture
            if (graph) hb_graph_destroy(graph);
            throw;
        }
        if (graph) hb_graph_destroy(graph);
        const auto t2 = clk::now();
        WireValues res = ev.decrypt_outputs(*enc, verbose != 0, ready);
        if (all_wires) {                                       // every wire that was driven or is an input
            std::vector<std::string> names; std::vector<uint32_t> idx;
            for (const auto& kv : enc->index()) { names.push_back(kv.first); idx.push_back(kv.second); }
            const size_t words = (size_t)s->p.ct_dim() + 1, wb = s->p.word_bytes();
            std::vector<uint8_t> cts(idx.size() * words * wb);
            if (!idx.empty()) enc->read(idx, cts.data());
            for (size_t i = 0; i < names.size(); ++i) {
                const void* ct = cts.data() + i * words * wb;
                res[names[i]] = lut ? Ptxt::integer(s->ck.decrypt_shortint(ct)) : Ptxt::boolean(s->ck.decrypt_bool(ct));
            }
        }
        const auto t3 = clk::now();
        if (seconds) { seconds[0] = secs(t0, t1); seconds[1] = secs(t1, t2); seconds[2] = secs(t2, t3); }
        text = format_values(res);
    }) || emit(text, buf, cap, needed);
}

int hh_session_run_arith(hh_session* s, const hh_circuit* c, const char* inputs, const char* ptxt_type, int cycles,
                         int all_wires, int verbose, char* buf, size_t cap, size_t* needed, double* seconds, uint64_t* counts) {
    std::string text;
    return guarded([&] {
        using clk = std::chrono::steady_clock;
        auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
        const Netlist& nl = c->c->netlist();
        if (!nl.has_arith) throw Error("hh_session_run_arith needs an arithmetic netlist");
        const std::string pt = ptxt_type ? ptxt_type : "u32";
        ArithCircuit ev(s->ck, *s->sk, *c->c, ptxt_type_bits(pt));
        ev.quiet = !verbose;
        const WireValues in = parse_values(inputs, pt);
        const auto t0 = clk::now();
        auto enc = ev.encrypt_inputs(nl.wire_set, in);
        const bool ready = cycles > 1 && enc->contains("READY");
        if (ready) ev.init_ready(*enc);
        const auto t1 = clk::now();
        for (int cy = 0; cy < (cycles < 1 ? 1 : cycles); ++cy) {
            ev.evaluate_encrypted(*enc, 1, pt);
            if (ready) ev.evaluate_ready(*enc);
        }
        const auto t2 = clk::now();
        WireValues res = ev.decrypt_outputs(*enc, verbose != 0, ready);
        if (all_wires) {
            std::vector<std::string> names(nl.wire_set.begin(), nl.wire_set.end());
            for (const auto& w : nl.inputs) if (!nl.wire_set.count(w)) names.push_back(w);
            std::sort(names.begin(), names.end()); names.erase(std::unique(names.begin(), names.end()), names.end());
            for (const auto& kv : ev.decrypt_wires(*enc, names)) res[kv.first] = kv.second;
        }
        const auto t3 = clk::now();
        if (seconds) { seconds[0] = secs(t0, t1); seconds[1] = secs(t1, t2); seconds[2] = secs(t2, t3); }
        if (counts) { counts[0] = ev.bootstraps(); counts[1] = ev.batches(); }
        text = format_values(res);
    }) || emit(text, buf, cap, needed);
}

int hh_session_stats(hh_session* s, uint64_t* out4) {
    return guarded([&] {
        const hb_stats_t st = s->sk->stats();
        out4[0] = st.pbs_count; out4[1] = st.ks_count; out4[2] = st.levels; out4[3] = st.kernel_launches;
    });
}
int hh_session_stats_multi(hh_session* s, uint64_t* out3) {
    return guarded([&] {
        const hb_stats_t st = s->sk->stats();
        out3[0] = st.n_devices; out3[1] = st.sharded_levels; out3[2] = st.push_bytes;
    });
}

}  // extern "C"
