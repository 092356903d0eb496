// See arith.hpp.
#include "arith.hpp"
#include <algorithm>
#include <cstdio>
#include <cstring>

namespace helm {

static void ck_status(hb_ctx* c, hb_status s, const char* what) {
    if (s != HB_OK) throw Error(std::string(what) + " failed (" + std::to_string(s) + "): " + hb_last_error(c));
}

// ---------------------------------------------------------------- RadixProgram
static hb_sop make_sop(int kind, const BlockExpr& e, uint64_t table, uint32_t out) {
    hb_sop o{};
    o.kind = (uint8_t)kind; o.n_in = (uint8_t)e.n; o.cst = (int16_t)e.cst; o.out = out; o.table = table;
    for (int s = 0; s < e.n; ++s) { o.in[s] = e.row[s]; o.coef[s] = (int8_t)e.coef[s]; }
    return o;
}
void RadixProgram::lut(size_t stage, const BlockExpr& e, uint64_t table, uint32_t out) {
    if (stages.size() <= stage) stages.resize(stage + 1);
    stages[stage].push_back(make_sop(HB_SOP_LUT, e, table, out));
}
void RadixProgram::lin(size_t stage, const BlockExpr& e, uint32_t out) {
    if (stages.size() <= stage) stages.resize(stage + 1);
    stages[stage].push_back(make_sop(HB_SOP_LINEAR, e, 0, out));
}
size_t RadixProgram::bootstraps() const {
    size_t n = 0;
    for (const auto& st : stages) for (const auto& o : st) n += o.kind == HB_SOP_LUT;
    return n;
}

// ---------------------------------------------------------------- RadixBuilder
RadixBuilder::RadixBuilder(int msg, int carry, int blocks, uint32_t first, uint32_t rows)
    : msg_(msg), carry_(carry), mc_(msg * carry), nb_(blocks), first_(first), rows_(rows), next_(first) {
    if (msg_ != 4 || carry_ != 4) throw Error("arithmetic mode needs a parameter set with msg_mod = carry_mod = 4 (shortint_m2c2)");
}
uint32_t RadixBuilder::tmp() {
    if (next_ - first_ >= rows_) throw Error("arithmetic mode: scratch rows exhausted");
    return next_++;
}

// Carry propagation.  state_i = 2 (block i generates a carry), 1 (propagates an incoming one), 0 (absorbs it);
// an inclusive prefix scan with combine(prev, cur) = (cur == 1 ? prev : cur) gives for every block whether a carry
// leaves it; result_i = (sum_i + carry_in_i) mod msg.
size_t RadixBuilder::propagate(RadixProgram& p, size_t st, const std::vector<BlockExpr>& sums, uint32_t out, uint32_t* carry_out) {
    const int m = msg_;
    const uint64_t t_msg = table([m](int v) { return v % m; });
    const uint64_t t_state = table([m](int v) { return v >= m ? 2 : (v == m - 1 ? 1 : 0); });
    const uint64_t t_comb = table([m](int v) { const int prev = v / m, cur = v % m; return cur == 1 ? prev : cur; });
    const uint64_t t_fin = table([m](int v) { const int prev = v / m, ms = v % m; return (ms + (prev == 2)) % m; });
    const uint64_t t_cout = table([](int v) { return v == 2 ? 1 : 0; });
    const int ns = carry_out ? nb_ : nb_ - 1;                          // blocks whose outgoing carry is needed
    if (ns == 0) { p.lut(st, sums[0], t_msg, out); return st + 1; }
    std::vector<uint32_t> msgr(nb_), state(ns);
    p.lut(st, sums[0], t_msg, out);                                   // block 0 has no incoming carry
    for (int i = 1; i < nb_; ++i) { msgr[i] = tmp(); p.lut(st, sums[i], t_msg, msgr[i]); }
    for (int i = 0; i < ns; ++i) { state[i] = tmp(); p.lut(st, sums[i], t_state, state[i]); }
    size_t s = st + 1;
    for (int d = 1; d < ns; d *= 2, ++s) {
        std::vector<uint32_t> nxt(state);
        for (int i = d; i < ns; ++i) {
            nxt[i] = tmp();
            p.lut(s, BlockExpr::of(state[i]).plus(state[i - d], m), t_comb, nxt[i]);
        }
        state.swap(nxt);
    }
    for (int i = 1; i < nb_; ++i) p.lut(s, BlockExpr::of(msgr[i]).plus(state[i - 1], m), t_fin, out + i);
    if (carry_out) { *carry_out = tmp(); p.lut(s, BlockExpr::of(state[nb_ - 1]), t_cout, *carry_out); }
    return s + 1;
}

uint32_t RadixBuilder::zero_row(RadixProgram& p) {
    if (zero_ == 0xffffffffu) { zero_ = tmp(); p.lin(0, BlockExpr::constant(0), zero_); }
    return zero_;
}

size_t RadixBuilder::mux_rows(RadixProgram& p, size_t st, uint32_t sel, const std::vector<uint32_t>& a, const std::vector<uint32_t>& b,
                              const std::vector<uint32_t>& out) {
    const int m = msg_;
    const uint64_t t_if = table([m](int v) { return (v / m) ? v % m : 0; });       // index = sel*msg + block
    const uint64_t t_else = table([m](int v) { return (v / m) ? 0 : v % m; });
    for (size_t j = 0; j < out.size(); ++j) {
        const uint32_t t1 = tmp(), t2 = tmp();
        p.lut(st, BlockExpr::of(a[j]).plus(sel, m), t_if, t1);
        p.lut(st, BlockExpr::of(b[j]).plus(sel, m), t_else, t2);
        p.lin(st + 1, BlockExpr::of(t1).plus(t2), out[j]);
    }
    return st + 2;
}

void RadixBuilder::add(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out) {
    std::vector<BlockExpr> sums;
    for (int i = 0; i < nb_; ++i) sums.push_back(BlockExpr::of(a + i).plus(b + i));
    propagate(p, 0, sums, out);
}
// a - b = a + ~b + 1 (two's complement over msg^blocks): block sums a_i + (msg-1 - b_i), +1 on block 0
void RadixBuilder::sub(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out) {
    std::vector<BlockExpr> sums;
    for (int i = 0; i < nb_; ++i) {
        BlockExpr e = BlockExpr::of(a + i).plus(b + i, -1);
        e.cst = msg_ - 1 + (i == 0);
        sums.push_back(e);
    }
    propagate(p, 0, sums, out);
}
void RadixBuilder::scalar_add(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out) {
    std::vector<BlockExpr> sums;
    for (int i = 0; i < nb_; ++i) {
        BlockExpr e = BlockExpr::of(a + i);
        e.cst = (int)(k % msg_); k /= msg_;
        sums.push_back(e);
    }
    propagate(p, 0, sums, out);
}
void RadixBuilder::scalar_sub(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out) {
    scalar_add(p, a, (unsigned __int128)0 - k, out);                  // digits beyond `blocks` are never read
}
void RadixBuilder::copy(RadixProgram& p, uint32_t a, uint32_t out) {
    for (int i = 0; i < nb_; ++i) p.lin(0, BlockExpr::of(a + i), out + i);
}

// Column compression.  Every term is a clean block (< msg).  A chunk of up to `carry` terms of one column sums to
// < msg*carry, so two bootstraps split it into a clean block of the same column and a clean carry for the next one.
// When no column holds more than two terms the remaining sums go through the carry propagation.
void RadixBuilder::compress(RadixProgram& p, size_t st, std::vector<std::vector<uint32_t>> cols, uint32_t out) {
    const int m = msg_;
    const uint64_t t_msg = table([m](int v) { return v % m; });
    const uint64_t t_carry = table([m](int v) { return v / m; });
    cols.resize(nb_);
    const size_t chunk = (size_t)((mc_ - 1) / (m - 1));               // terms whose sum stays below msg*carry (5 for 4x4)
    const size_t take = std::min<size_t>(chunk, 4);                   // a block expression holds 4 rows
    for (;; ++st) {
        size_t high = 0;
        for (const auto& c : cols) high = std::max(high, c.size());
        if (high <= 2) break;
        std::vector<std::vector<uint32_t>> nxt(nb_);
        for (int k = 0; k < nb_; ++k) {
            const auto& c = cols[k];
            for (size_t i = 0; i < c.size(); i += take) {
                const size_t cnt = std::min(take, c.size() - i);
                if (cnt == 1) { nxt[k].push_back(c[i]); continue; }
                BlockExpr e;
                for (size_t j = 0; j < cnt; ++j) e.plus(c[i + j]);
                const uint32_t lo = tmp();
                p.lut(st, e, t_msg, lo);
                nxt[k].push_back(lo);
                if (k + 1 < nb_) { const uint32_t hi = tmp(); p.lut(st, e, t_carry, hi); nxt[k + 1].push_back(hi); }
            }
        }
        cols.swap(nxt);
    }
    std::vector<BlockExpr> sums;
    for (int k = 0; k < nb_; ++k) {
        BlockExpr e;
        for (uint32_t r : cols[k]) e.plus(r);
        sums.push_back(e);
    }
    propagate(p, st, sums, out);
}

void RadixBuilder::mul(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out) {
    const int m = msg_;
    const uint64_t t_lo = table([m](int v) { return ((v / m) * (v % m)) % m; });
    const uint64_t t_hi = table([m](int v) { return ((v / m) * (v % m)) / m; });
    std::vector<std::vector<uint32_t>> cols(nb_);
    for (int i = 0; i < nb_; ++i)
        for (int j = 0; i + j < nb_; ++j) {
            const BlockExpr e = BlockExpr::of(b + j).plus(a + i, m);  // index = a_i * msg + b_j
            const uint32_t lo = tmp();
            p.lut(0, e, t_lo, lo);
            cols[i + j].push_back(lo);
            if (i + j + 1 < nb_) { const uint32_t hi = tmp(); p.lut(0, e, t_hi, hi); cols[i + j + 1].push_back(hi); }
        }
    compress(p, 1, cols, out);
}

void RadixBuilder::scalar_mul(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out) {
    const int m = msg_;
    std::vector<std::vector<uint32_t>> cols(nb_);
    bool any_lut = false;
    for (int j = 0; j < nb_ && k != 0; ++j, k /= m) {
        const int d = (int)(k % m);
        if (d == 0) continue;
        for (int i = 0; i + j < nb_; ++i) {
            if (d == 1) { cols[i + j].push_back(a + i); continue; }
            const uint64_t t_lo = table([m, d](int v) { return ((v % m) * d) % m; });
            const uint64_t t_hi = table([m, d](int v) { return ((v % m) * d) / m; });
            const uint32_t lo = tmp();
            p.lut(0, BlockExpr::of(a + i), t_lo, lo);
            cols[i + j].push_back(lo);
            if (i + j + 1 < nb_) { const uint32_t hi = tmp(); p.lut(0, BlockExpr::of(a + i), t_hi, hi); cols[i + j + 1].push_back(hi); }
            any_lut = true;
        }
    }
    compress(p, any_lut ? 1 : 0, cols, out);
}

// shift by a clear amount: block shift = relabelling, the remaining bit shift is one (bi)variate LUT per block
void RadixBuilder::scalar_shift(RadixProgram& p, uint32_t a, unsigned k, bool left, uint32_t out) {
    const int m = msg_;
    int bits_per_block = 0; while ((1 << bits_per_block) < m) ++bits_per_block;
    const int kb = (int)(k / bits_per_block), s = (int)(k % bits_per_block);
    for (int i = 0; i < nb_; ++i) {
        const int main = left ? i - kb : i + kb;                       // block that supplies the shifted-in-place bits
        const int nbr = left ? main - 1 : main + 1;                    // neighbour that supplies the bits crossing over
        const bool has_main = main >= 0 && main < nb_, has_nbr = s != 0 && nbr >= 0 && nbr < nb_;
        if (!has_main && !has_nbr) { p.lin(0, BlockExpr::constant(0), out + i); continue; }
        if (s == 0) { p.lin(0, BlockExpr::of(a + main), out + i); continue; }
        // index = main * msg + nbr (absent operand = 0)
        const uint64_t t = left ? table([m, s, bits_per_block](int v) { return (((v / m) << s) % m) | ((v % m) >> (bits_per_block - s)); })
                                : table([m, s, bits_per_block](int v) { return ((v / m) >> s) | (((v % m) << (bits_per_block - s)) % m); });
        BlockExpr e;
        if (has_main) e.plus(a + main, m);
        if (has_nbr) e.plus(a + nbr, 1);
        p.lut(0, e, t, out + i);
    }
}

// Restoring division.  Per quotient bit (MSB first): R <- (R << 1) | next dividend bit; D <- R - b with carry-out c
// (c = 1 iff R >= b); R <- c ? D : R; quotient bit = c.  With b = 0 every c is 1: the quotient is all ones (tfhe-rs).
void RadixBuilder::div(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out) {
    const int m = msg_;
    int bpb = 0; while ((1 << bpb) < m) ++bpb;                          // bits per block
    const int bits = nb_ * bpb;
    std::vector<uint32_t> R(nb_);
    for (int j = 0; j < nb_; ++j) { R[j] = tmp(); p.lin(0, BlockExpr::constant(0), R[j]); }
    for (int j = 0; j < nb_; ++j) p.lin(0, BlockExpr::constant(0), out + j);
    size_t st = 1;
    for (int i = bits - 1; i >= 0; --i) {
        const int ab = i / bpb, sh = i % bpb;
        // shift the remainder left by one bit, bringing in bit i of the dividend
        std::vector<uint32_t> S(nb_);
        const uint64_t t_in = table([m, sh](int v) { return (((v / m) << 1) % m) | (((v % m) >> sh) & 1); });   // R_0*m + a_blk
        const uint64_t t_sh = table([m, bpb](int v) { return (((v / m) << 1) % m) | ((v % m) >> (bpb - 1)); }); // R_j*m + R_{j-1}
        for (int j = 0; j < nb_; ++j) {
            S[j] = tmp();
            if (j == 0) p.lut(st, BlockExpr::of(a + ab).plus(R[0], m), t_in, S[0]);
            else p.lut(st, BlockExpr::of(R[j - 1]).plus(R[j], m), t_sh, S[j]);
        }
        // D = S - b, c = carry out
        std::vector<BlockExpr> sums;
        for (int j = 0; j < nb_; ++j) {
            BlockExpr e = BlockExpr::of(S[j]).plus(b + j, -1);
            e.cst = m - 1 + (j == 0);
            sums.push_back(e);
        }
        const uint32_t D = next_;
        for (int j = 0; j < nb_; ++j) tmp();
        uint32_t c = 0;
        st = propagate(p, st + 1, sums, D, &c);
        // R = c ? D : S ; quotient bit
        std::vector<uint32_t> Dv(nb_), Rn(nb_);
        for (int j = 0; j < nb_; ++j) { Dv[j] = D + (uint32_t)j; Rn[j] = tmp(); }
        p.lin(st, BlockExpr::of(out + ab).plus(c, 1 << sh), out + ab);
        st = mux_rows(p, st, c, Dv, S, Rn);
        R.swap(Rn);
    }
}

void RadixBuilder::scalar_div(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out) {
    const uint32_t kb = next_;
    for (int j = 0; j < nb_; ++j) { const uint32_t r = tmp(); p.lin(0, BlockExpr::constant((int)(k % msg_)), r); k /= msg_; }
    RadixProgram q;                                                     // the divisor rows are written at stage 0; shift the division by one stage
    div(q, a, kb, out);
    for (size_t s = 0; s < q.stages.size(); ++s) {
        if (p.stages.size() <= s + 1) p.stages.resize(s + 2);
        p.stages[s + 1].insert(p.stages[s + 1].end(), q.stages[s].begin(), q.stages[s].end());
    }
}

// Shift by an encrypted amount s (mod the bit width): for every bit t of s, cur <- s_t ? (cur shifted by 2^t) : cur.
void RadixBuilder::shift(RadixProgram& p, uint32_t a, uint32_t amount, bool left, uint32_t out) {
    const int m = msg_;
    int bpb = 0; while ((1 << bpb) < m) ++bpb;
    const int bits = nb_ * bpb;
    int nsel = 0; while ((1 << nsel) < bits) ++nsel;
    const uint32_t zero = zero_row(p);
    std::vector<uint32_t> sel(nsel);
    for (int t = 0; t < nsel; ++t) {
        const int sh = t % bpb;
        sel[t] = tmp();
        p.lut(0, BlockExpr::of(amount + t / bpb), table([sh](int v) { return (v >> sh) & 1; }), sel[t]);
    }
    std::vector<uint32_t> cur(nb_);
    for (int j = 0; j < nb_; ++j) cur[j] = a + (uint32_t)j;
    size_t st = 1;
    for (int t = 0; t < nsel; ++t) {
        const int k = 1 << t, kb = k / bpb, s = k % bpb;
        std::vector<uint32_t> sh_rows(nb_), nxt(nb_);
        for (int i = 0; i < nb_; ++i) {
            const int main = left ? i - kb : i + kb, nbr = left ? main - 1 : main + 1;
            const bool has_main = main >= 0 && main < nb_, has_nbr = s != 0 && nbr >= 0 && nbr < nb_;
            if (s == 0) { sh_rows[i] = has_main ? cur[main] : zero; continue; }
            if (!has_main && !has_nbr) { sh_rows[i] = zero; continue; }
            const uint64_t tb = left ? table([m, s, bpb](int v) { return (((v / m) << s) % m) | ((v % m) >> (bpb - s)); })
                                     : table([m, s, bpb](int v) { return ((v / m) >> s) | (((v % m) << (bpb - s)) % m); });
            BlockExpr e;
            if (has_main) e.plus(cur[main], m);
            if (has_nbr) e.plus(cur[nbr], 1);
            sh_rows[i] = tmp();
            p.lut(st, e, tb, sh_rows[i]);
        }
        if (s != 0) ++st;
        const bool last = t + 1 == nsel;
        for (int i = 0; i < nb_; ++i) nxt[i] = last ? out + (uint32_t)i : tmp();
        st = mux_rows(p, st, sel[t], sh_rows, cur, nxt);
        cur.swap(nxt);
    }
}

// ---------------------------------------------------------------- ArithCircuit
ArithCircuit::ArithCircuit(const ClientKey& ck, const ServerKey& sk, const Circuit& c, unsigned int_bits)
    : ck_(ck), sk_(sk), circuit_(c), bits_(int_bits) {
    const hb_params& hp = sk.params().hb;
    if (hp.order != 1 || hp.msg_mod != 4 || hp.carry_mod != 4)
        throw Error("ArithCircuit needs shortint parameters with msg_mod = carry_mod = 4 (shortint_m2c2)");
    if (int_bits != 8 && int_bits != 16 && int_bits != 32 && int_bits != 64 && int_bits != 128)
        throw Error("arithmetic width must be one of u8, u16, u32, u64, u128");
    nb_ = (int)int_bits / 2;
}

void ArithCircuit::encrypt_value(unsigned __int128 v, uint8_t* dst) const {
    const size_t row_bytes = ((size_t)sk_.params().ct_dim() + 1) * sk_.params().word_bytes();
    for (int i = 0; i < nb_; ++i) ck_.encrypt_shortint((uint64_t)((v >> (2 * i)) & 3), dst + (size_t)i * row_bytes);
}

static unsigned __int128 scalar_of(const std::string& w) {
    unsigned __int128 v = 0;
    for (char ch : w) { if (ch < '0' || ch > '9') return 0; v = v * 10 + (unsigned)(ch - '0'); }   // parse failure -> 0 (unwrap_or(0))
    return v;
}

// One cell -> block program (dispatch of circuit.rs:1326-1431: a numeric operand selects the *_plain form, which is
// always "ciphertext op scalar" whatever the operand order in the netlist)
void ArithCircuit::build_gate(RadixBuilder& rb, RadixProgram& p, const RowOf& row_of, const Gate& g) const {
    auto base_row = [&](const RowOf&, const std::string& w) { return row_of(w); };
    const RowOf& enc = row_of;
    const uint32_t out = base_row(enc, g.output);
    if (g.type == GateType::Dff || g.type == GateType::Buf || g.type == GateType::Copy) { rb.copy(p, base_row(enc, g.inputs.at(0)), out); return; }
    if (!is_arith_type(g.type)) throw Error(std::string("gate type \"") + gate_type_name(g.type) + "\" cannot be evaluated in arithmetic mode");
    if (g.inputs.size() < 2) throw Error("arithmetic cell " + g.name + " needs two operands");
    const bool c0 = circuit_.is_constant_operand(g.inputs[0]), c1 = circuit_.is_constant_operand(g.inputs[1]);
    const unsigned __int128 mask = bits_ >= 128 ? ~(unsigned __int128)0 : (((unsigned __int128)1 << bits_) - 1);
    if (c0 || c1) {
        if (c0 && c1) throw Error("Empty ctxt operand!");                      // circuit.rs:1356
        const uint32_t a = base_row(enc, g.inputs[c0 ? 1 : 0]);
        const unsigned __int128 k = scalar_of(g.inputs[c0 ? 0 : 1]) & mask;
        switch (g.type) {
            case GateType::Add: rb.scalar_add(p, a, k, out); return;
            case GateType::Sub: rb.scalar_sub(p, a, k, out); return;
            case GateType::Mult: rb.scalar_mul(p, a, k, out); return;
            case GateType::Shl: rb.scalar_shift(p, a, (unsigned)(k % bits_), true, out); return;
            case GateType::Shr: rb.scalar_shift(p, a, (unsigned)(k % bits_), false, out); return;
            case GateType::Div: rb.scalar_div(p, a, k, out); return;
            default: throw Error("arithmetic cell " + g.name + ": unsupported type");
        }
    }
    const uint32_t a = base_row(enc, g.inputs[0]), b = base_row(enc, g.inputs[1]);
    switch (g.type) {
        case GateType::Add: rb.add(p, a, b, out); return;
        case GateType::Sub: rb.sub(p, a, b, out); return;
        case GateType::Mult: rb.mul(p, a, b, out); return;
        case GateType::Div: rb.div(p, a, b, out); return;
        case GateType::Shl: rb.shift(p, a, b, true, out); return;
        case GateType::Shr: rb.shift(p, a, b, false, out); return;
        default: throw Error("arithmetic cell " + g.name + ": unsupported type");
    }
}

std::unique_ptr<EncWireMap> ArithCircuit::encrypt_inputs(const std::set<std::string>& wire_set, const WireValues& inputs) {
    const Netlist& nl = circuit_.netlist();
    std::vector<std::string> names(wire_set.begin(), wire_set.end());
    names.insert(names.end(), nl.inputs.begin(), nl.inputs.end());
    names.insert(names.end(), nl.dff_outputs.begin(), nl.dff_outputs.end());
    names.push_back("__helm_one"); names.push_back("__helm_not_ready");       // READY latching temporaries
    for (const auto& o : nl.outputs) { names.push_back("__helm_valid_" + o); names.push_back("__helm_t1_" + o); names.push_back("__helm_t2_" + o); }
    std::set<std::string> uniq(names.begin(), names.end());
    named_ = (uint32_t)uniq.size();
    scratch_first_ = named_ * (uint32_t)nb_;
    // size the scratch area by building every level's programs (and the READY-latching ones) for real against an unbounded
    // allocator: the row numbers of the named wires are those the wire map will assign (first occurrence in `names`)
    uint32_t need = 0;
    {
        std::unordered_map<std::string, uint32_t> index;
        for (const auto& n : names) if (!index.count(n)) { const uint32_t i = (uint32_t)index.size(); index[n] = i; }
        const RowOf row_of = [&](const std::string& w) {
            auto it = index.find(w);
            if (it == index.end()) throw Error("wire \"" + w + "\" is not in the encrypted wire map");
            return it->second * (uint32_t)nb_;
        };
        RadixBuilder rb(4, 4, nb_, scratch_first_, 0x40000000u);
        for (const auto& level : circuit_.levels()) {
            rb.reset_scratch();
            for (const Gate& g : level) { RadixProgram p; build_gate(rb, p, row_of, g); }
            need = std::max(need, rb.scratch_used());
        }
        if (index.count("READY")) need = std::max(need, build_ready(rb, row_of, nullptr));
    }
    scratch_rows_ = need;
    auto enc = std::make_unique<EncWireMap>(sk_, names, scratch_first_ + scratch_rows_ - named_);
    std::vector<uint32_t> all(scratch_first_);
    for (uint32_t i = 0; i < scratch_first_; ++i) all[i] = i;
    enc->set_trivial(all, 0);                                                  // FheType::None wires read as 0 here
    const bool all_zero = inputs.empty() || inputs.count("dummy");
    const size_t row_bytes = ((size_t)sk_.params().ct_dim() + 1) * sk_.params().word_bytes();
    std::vector<uint32_t> idx; std::vector<uint8_t> cts;
    auto push = [&](const std::string& w, unsigned __int128 v) {
        const uint32_t base = base_row(*enc, w);
        for (int i = 0; i < nb_; ++i) idx.push_back(base + (uint32_t)i);
        cts.resize(cts.size() + (size_t)nb_ * row_bytes);
        encrypt_value(v, cts.data() + cts.size() - (size_t)nb_ * row_bytes);
    };
    const unsigned __int128 mask = bits_ >= 128 ? ~(unsigned __int128)0 : (((unsigned __int128)1 << bits_) - 1);
    std::set<std::string> seen;
    for (const auto& w : nl.inputs) {
        if (!seen.insert(w).second) continue;
        if (std::find(nl.dff_outputs.begin(), nl.dff_outputs.end(), w) != nl.dff_outputs.end()) continue;
        unsigned __int128 v = 0;
        if (!all_zero) {
            auto it = inputs.find(w);
            if (it == inputs.end()) throw Error("\n Input wire \"" + w + "\" not found in input wires!");
            v = it->second.v & mask;
        }
        push(w, v);
    }
    for (const auto& w : nl.dff_outputs) push(w, 0);
    if (!idx.empty()) enc->write(idx, cts.data());
    return enc;
}

// lock-step execution: stage s of every program is one batch
void ArithCircuit::run(EncWireMap& enc, std::vector<RadixProgram>& progs) {
    size_t depth = 0;
    for (const auto& p : progs) depth = std::max(depth, p.stages.size());
    std::vector<hb_sop> batch;
    for (size_t s = 0; s < depth; ++s) {
        batch.clear();
        for (const auto& p : progs)
            if (s < p.stages.size()) batch.insert(batch.end(), p.stages[s].begin(), p.stages[s].end());
        if (batch.empty()) continue;
        ck_status(sk_.ctx(), hb_eval_shortint_ops(sk_.ctx(), enc.handle(), batch.data(), (uint32_t)batch.size()), "hb_eval_shortint_ops");
        ++n_batches_;
    }
    for (const auto& p : progs) n_pbs_ += p.bootstraps();
}

EncWireMap& ArithCircuit::evaluate_encrypted(EncWireMap& enc, size_t /*cycle*/, const std::string& /*ptxt_type*/) {
    const auto& levels = circuit_.levels();
    if (levels.empty()) throw Error("evaluate_encrypted: compute_levels has not run");
    RadixBuilder rb(4, 4, nb_, scratch_first_, scratch_rows_);
    const size_t total = levels.size();
    for (size_t li = 0; li < total; ++li) {
        rb.reset_scratch();
        std::vector<RadixProgram> progs(levels[li].size());
        for (size_t gi = 0; gi < levels[li].size(); ++gi) {
            build_gate(rb, progs[gi], [&](const std::string& w) { return base_row(enc, w); }, levels[li][gi]);
        }
        run(enc, progs);
        if (!quiet) {
            const size_t key = li == circuit_.dff_level() ? total : levels[li].front().level;
            std::printf("  Evaluated gates in level [%zu/%zu]\n", key, total);
        }
    }
    ck_status(sk_.ctx(), hb_sync(sk_.ctx()), "hb_sync");
    return enc;
}

void ArithCircuit::init_ready(EncWireMap& enc) {
    std::vector<uint32_t> idx;
    for (const auto& o : circuit_.netlist().outputs) {
        const uint32_t b = base_row(enc, "__helm_valid_" + o);
        for (int i = 0; i < nb_; ++i) idx.push_back(b + (uint32_t)i);
    }
    enc.set_trivial(idx, 0);
    // trivial one: block 0 = 1 * delta
    enc.set_trivial({base_row(enc, "__helm_one")}, ck_.delta());
}

// valid = wire * READY + valid * (1 - READY) for every output (circuit.rs:1209-1297): three lock-step phases.
// Returns the largest scratch use of a phase; with phases == nullptr it only measures.
uint32_t ArithCircuit::build_ready(RadixBuilder& rb, const RowOf& row, std::vector<std::vector<RadixProgram>>* phases) const {
    const auto& outs = circuit_.netlist().outputs;
    const uint32_t ready = row("READY"), one = row("__helm_one"), nr = row("__helm_not_ready");
    uint32_t need = 0;
    std::vector<RadixProgram> sub(1), mul(2 * outs.size()), add(outs.size());
    rb.reset_scratch();
    rb.sub(sub[0], one, ready, nr);
    need = std::max(need, rb.scratch_used());
    rb.reset_scratch();
    for (size_t o = 0; o < outs.size(); ++o) {
        rb.mul(mul[2 * o], row(outs[o]), ready, row("__helm_t1_" + outs[o]));
        rb.mul(mul[2 * o + 1], row("__helm_valid_" + outs[o]), nr, row("__helm_t2_" + outs[o]));
    }
    need = std::max(need, rb.scratch_used());
    rb.reset_scratch();
    for (size_t o = 0; o < outs.size(); ++o)
        rb.add(add[o], row("__helm_t1_" + outs[o]), row("__helm_t2_" + outs[o]), row("__helm_valid_" + outs[o]));
    need = std::max(need, rb.scratch_used());
    if (phases) { phases->clear(); phases->push_back(std::move(sub)); phases->push_back(std::move(mul)); phases->push_back(std::move(add)); }
    return need;
}

void ArithCircuit::evaluate_ready(EncWireMap& enc) {
    RadixBuilder rb(4, 4, nb_, scratch_first_, scratch_rows_);
    std::vector<std::vector<RadixProgram>> phases;
    build_ready(rb, [&](const std::string& w) { return base_row(enc, w); }, &phases);
    for (auto& ph : phases) run(enc, ph);
    ck_status(sk_.ctx(), hb_sync(sk_.ctx()), "hb_sync");
}

WireValues ArithCircuit::decrypt_wires(const EncWireMap& enc, const std::vector<std::string>& wires) const {
    const size_t row_bytes = ((size_t)sk_.params().ct_dim() + 1) * sk_.params().word_bytes();
    std::vector<uint32_t> idx;
    for (const auto& w : wires) {
        const uint32_t b = base_row(enc, w);
        for (int i = 0; i < nb_; ++i) idx.push_back(b + (uint32_t)i);
    }
    std::vector<uint8_t> cts(idx.size() * row_bytes);
    if (!idx.empty()) enc.read(idx, cts.data());
    WireValues res;
    for (size_t o = 0; o < wires.size(); ++o) {
        unsigned __int128 v = 0;
        for (int i = 0; i < nb_; ++i)
            v |= (unsigned __int128)(ck_.decrypt_shortint(cts.data() + (o * nb_ + i) * row_bytes) & 3) << (2 * i);
        res[wires[o]] = Ptxt::integer(v);
    }
    return res;
}

WireValues ArithCircuit::decrypt_outputs(const EncWireMap& enc, bool verbose, bool latched) {
    const auto& outs = circuit_.netlist().outputs;
    std::vector<std::string> names;
    for (const auto& o : outs) names.push_back(latched ? "__helm_valid_" + o : o);
    const WireValues raw = decrypt_wires(enc, names);
    WireValues res;
    for (size_t o = 0; o < outs.size(); ++o) res[outs[o]] = raw.at(names[o]);
    if (!quiet) {
        size_t i = 0;
        for (const auto& kv : res) {
            if (i++ > 10 && !verbose) { std::printf("[!] More than ten output_wires, pass `--verbose` to see output.\n"); break; }
            std::printf(" %s: %s\n", kv.first.c_str(), ptxt_to_string(kv.second).c_str());
        }
    }
    return res;
}

}  // namespace helm
