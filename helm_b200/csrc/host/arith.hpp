// Arithmetic mode on the B200 engine: the C++ mirror of the reference's ArithCircuit (/root/reference/src/circuit.rs:
// 1113-1483) and of Gate::evaluate_encrypted_{mul,add,sub,shift,copy}_block[_plain] (src/gates.rs:306-702).
//
// In the reference every arithmetic cell is ONE call into tfhe-rs's high-level integer API (FheUint32 `*`, `+`, `-`,
// `<<`, `>>` ...), which underneath is a radix decomposition into shortint blocks (2 message bits + 2 carry bits under
// PARAM_MESSAGE_2_CARRY_2) and a sequence of block-level linear operations and programmable bootstraps.  Here that
// block-level sequence is produced explicitly (RadixProgram: stages of hb_sop operations) and the programs of ALL
// cells of a netlist level advance in lock-step: stage s of every cell is one hb_eval_shortint_ops batch, so a level
// of W cells with P bootstraps each is a handful of launches of ~W*P bootstraps instead of W*P library calls.
//
// Algorithms (decrypted results equal wrapping unsigned arithmetic at the chosen width, which is what FheUintN gives):
//   add / sub / scalar add : blockwise sum, then carry propagation as a Hillis-Steele prefix scan over
//                            generate/propagate states (depth 2 + ceil(log2(blocks-1)) bootstraps)
//   mul (ct x ct)          : block products lo/hi by bivariate LUT, column compression 4 -> (msg, carry) per
//                            bootstrap pair until two terms per column remain, then the carry propagation above
//   mul by scalar          : the same compression over per-digit scaled copies (digit 1 costs nothing)
//   shl / shr by scalar    : one bivariate LUT per block (block shift is a relabelling)
//   copy                   : row copies
//   div                    : restoring long division (shift-in, subtract with carry-out, select), ct / ct and ct / scalar
//   shl / shr by ciphertext: barrel shifter, one select per bit of the amount
#pragma once
#include <functional>
#include "eval.hpp"

namespace helm {

// sum_s coef[s]*row[s] + cst : the argument of a LUT (or a LINEAR op), at most 4 rows
struct BlockExpr {
    int n = 0;
    uint32_t row[4] = {0, 0, 0, 0};
    int coef[4] = {0, 0, 0, 0};
    int cst = 0;
    static BlockExpr of(uint32_t r) { BlockExpr e; e.n = 1; e.row[0] = r; e.coef[0] = 1; return e; }
    static BlockExpr constant(int c) { BlockExpr e; e.cst = c; return e; }
    BlockExpr& plus(uint32_t r, int c = 1) { row[n] = r; coef[n] = c; ++n; return *this; }
};

struct RadixProgram {
    std::vector<std::vector<hb_sop>> stages;
    void lut(size_t stage, const BlockExpr& e, uint64_t table, uint32_t out);
    void lin(size_t stage, const BlockExpr& e, uint32_t out);
    size_t bootstraps() const;
};

// Builds block-level programs; temporaries come from a bump allocator over a scratch row range.
class RadixBuilder {
public:
    RadixBuilder(int msg_mod, int carry_mod, int blocks, uint32_t scratch_first, uint32_t scratch_rows);
    void reset_scratch() { next_ = first_; zero_ = 0xffffffffu; }
    uint32_t scratch_used() const { return next_ - first_; }
    int blocks() const { return nb_; }
    // operands and results are `blocks` consecutive rows starting at the given row (little-endian blocks)
    void add(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out);
    void sub(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out);
    void scalar_add(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out);
    void scalar_sub(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out);      // a - k
    void mul(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out);
    void scalar_mul(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out);
    void scalar_shift(RadixProgram& p, uint32_t a, unsigned k, bool left, uint32_t out);
    void copy(RadixProgram& p, uint32_t a, uint32_t out);
    // restoring long division, one quotient bit per iteration (quotient of x / 0 is all ones, as tfhe-rs defines it)
    void div(RadixProgram& p, uint32_t a, uint32_t b, uint32_t out);
    void scalar_div(RadixProgram& p, uint32_t a, unsigned __int128 k, uint32_t out);
    // shift by an encrypted amount (taken modulo the bit width): barrel shifter over the bits of the amount
    void shift(RadixProgram& p, uint32_t a, uint32_t amount, bool left, uint32_t out);
    // table helpers: entry i = f(i) for i in [0, msg*carry)
    template <class F> uint64_t table(F f) const {
        uint64_t t = 0;
        for (int i = 0; i < mc_; ++i) t |= (uint64_t)(f(i) & 15) << (4 * i);
        return t;
    }
private:
    uint32_t tmp();
    // sums[i] < msg*carry with true value <= 2*msg - 1; writes the propagated result blocks to out..out+nb-1,
    // starting at stage `st`
    // returns the first free stage; if carry_out != nullptr a clean 0/1 block with the carry out of the top block is
    // produced as well (available at the returned stage)
    size_t propagate(RadixProgram& p, size_t st, const std::vector<BlockExpr>& sums, uint32_t out, uint32_t* carry_out = nullptr);
    // out_j = sel ? a_j : b_j for block row lists (sel: clean 0/1 block); two stages
    size_t mux_rows(RadixProgram& p, size_t st, uint32_t sel, const std::vector<uint32_t>& a, const std::vector<uint32_t>& b,
                    const std::vector<uint32_t>& out);
    uint32_t zero_row(RadixProgram& p);
    // columns[k] = rows of clean blocks (value < msg) of weight msg^k; starts at stage `st`
    void compress(RadixProgram& p, size_t st, std::vector<std::vector<uint32_t>> columns, uint32_t out);
    int msg_, carry_, mc_, nb_;
    uint32_t first_, rows_, next_;
    uint32_t zero_ = 0xffffffffu;
};

// ArithCircuit: tfhe high-level integers (FheUint8..128) -> radix blocks of a shortint parameter set with
// msg_mod = carry_mod = 4.  One named wire = `blocks` consecutive device rows.
class ArithCircuit : public EvalCircuit {
public:
    ArithCircuit(const ClientKey& ck, const ServerKey& sk, const Circuit& c, unsigned int_bits);
    std::unique_ptr<EncWireMap> encrypt_inputs(const std::set<std::string>& wire_set, const WireValues& inputs) override;
    EncWireMap& evaluate_encrypted(EncWireMap& enc, size_t cycle, const std::string& ptxt_type) override;
    void init_ready(EncWireMap& enc) override;
    void evaluate_ready(EncWireMap& enc) override;
    WireValues decrypt_outputs(const EncWireMap& enc, bool verbose, bool latched = false) override;
    WireValues decrypt_wires(const EncWireMap& enc, const std::vector<std::string>& wires) const;
    uint64_t bootstraps() const { return n_pbs_; }
    uint64_t batches() const { return n_batches_; }
private:
    uint32_t base_row(const EncWireMap& enc, const std::string& wire) const { return enc.at(wire) * (uint32_t)nb_; }
    using RowOf = std::function<uint32_t(const std::string&)>;               // wire name -> first block row
    void build_gate(RadixBuilder& rb, RadixProgram& p, const RowOf& row_of, const Gate& g) const;
    uint32_t build_ready(RadixBuilder& rb, const RowOf& row_of, std::vector<std::vector<RadixProgram>>* phases) const;
    void run(EncWireMap& enc, std::vector<RadixProgram>& progs);
    void encrypt_value(unsigned __int128 v, uint8_t* dst) const;
    const ClientKey& ck_;
    const ServerKey& sk_;
    const Circuit& circuit_;
    unsigned bits_;
    int nb_;
    uint32_t named_ = 0, latch_first_ = 0, scratch_first_ = 0, scratch_rows_ = 0;
    uint64_t n_pbs_ = 0, n_batches_ = 0;
};

}  // namespace helm
