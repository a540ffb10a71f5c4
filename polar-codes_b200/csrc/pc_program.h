// Host-side "tree compiler": turns (N, frozen set) into the flat node program the decode kernels
// interpret, one instruction per tree action. It replaces the reference's object tree built by
// FastSscFip::createDecoder (reference src/polarcode/decoding/fastssc_fip_char.cpp:496-569): the
// pattern-match ORDER and the 32-lane "short" thresholds are kept because the leaf type decides the
// saturation order and the SPC tie rules, i.e. they are part of the bit-exact semantics.
#ifndef PC_PROGRAM_H
#define PC_PROGRAM_H
#include <stdint.h>
#include <vector>

#if defined(__CUDACC__)
#define PC_PHD __host__ __device__ __forceinline__
#else
#define PC_PHD inline
#endif

namespace pc {

// instruction word: op[5:0] | s[10:6] (log2 of the node length) | off[31:11] (element offset of the node)
enum Op : uint32_t {
    OP_F = 0,     // child llr = f(lo, hi)                         RateRNode::decode :401-405
    OP_G = 1,     // child llr = g(lo, hi, left bits)              :406-408
    OP_G0 = 2,    // left child rate-0: child llr = sat(lo + hi)   ZeroRNode::decode :476-483
    OP_G1 = 3,    // right child rate-1: fused g + hard decision + combine   ROneNode :427-449
    OP_COMB = 4,  // bits[lo] ^= bits[hi]                          CombineInPlace fip_char.h:164-171
    OP_COMB0 = 5, // bits[lo] = bits[hi]                           Combine_0R fip_char.h:209-220
    OP_R0 = 6,    // RateZeroDecoder :202-208
    OP_R1 = 7,    // RateOneDecoder :210-215
    OP_REP = 8,   // RepetitionDecoder / ShortRepetitionDecoder :225-272
    OP_DREP = 9,  // DoubleRepetitionDecoder :249-263
    OP_SPC = 10,  // SpcDecoder (n > 32) / ShortSpcDecoder (n <= 32) :274-319
    OP_S01 = 11,  // ShortZeroOneDecoder :390-399
    OP_0SPC = 12, // ZeroSpcDecoder (n >= 64) / ShortZeroSpcDecoder :321-388
    OP_FLUSH = 13, // store the register-held 32-bit block of hard decisions
    OP_END = 14,
    OP_COUNT
};

PC_PHD uint32_t mk_instr(uint32_t op, uint32_t s, uint32_t off) { return op | (s << 6) | (off << 11); }
PC_PHD uint32_t instr_op(uint32_t i) { return i & 63u; }
PC_PHD uint32_t instr_s(uint32_t i) { return (i >> 6) & 31u; }
PC_PHD uint32_t instr_off(uint32_t i) { return i >> 11; }

// One piece of the info-bit gather (CharContainer::getPackedInformationBits, bitcontainer.cpp:536-564):
// take `len` consecutive hard decisions starting at bit `sshift` of 32-bit block `sword` and append
// them at bit `dshift` of the current output word; `last` closes the output word.
// packed: sword[10:0] | sshift[15:11] | (len-1)[20:16] | dshift[25:21] | last[26]
PC_PHD uint32_t mk_run(uint32_t sword, uint32_t sshift, uint32_t len, uint32_t dshift, uint32_t last)
{
    return sword | (sshift << 11) | ((len - 1) << 16) | (dshift << 21) | (last << 26);
}

struct FastSscProgram {
    int N = 0, K = 0, S = 0;
    bool staged_root = false;  // channel LLRs are first copied into the on-chip tree (N < 32 or root is a leaf)
    bool root_clean = false;
    std::vector<uint32_t> code; // node program, terminated by OP_END
    std::vector<uint32_t> runs; // info-bit gather
    std::vector<uint32_t> info_mask; // N/32 words, bit i%32 of word i/32 set for non-frozen positions
    std::vector<uint32_t> lut;  // K info positions, ascending (BitContainer::calculateLUT)
    int census[OP_COUNT] = { 0 };
};

// SCL node program (reference SclFip::createDecoder, src/polarcode/decoding/scl_fip_char.cpp:729-755): leaves are
// R0 / R1 / REP / SPC only, in that match order; every other node splits. Instruction word:
// op[5:0] | s[10:6] | side[11] (1: this node's decisions go to the "right" buffer of its stage, 0: "left")
enum SclOp : uint32_t { SCL_F = 0, SCL_G = 1, SCL_COMB = 2, SCL_R0 = 3, SCL_R1 = 4, SCL_REP = 5, SCL_SPC = 6, SCL_END = 7 };
PC_PHD uint32_t scl_instr(uint32_t op, uint32_t s, uint32_t side) { return op | (s << 6) | (side << 11); }
struct SclProgram {
    int N = 0, K = 0, S = 0;
    bool root_leaf = false;       // the root itself is a leaf: the channel row is staged into stage S
    std::vector<uint32_t> code;
    int census[8] = { 0 };
};
// float_rules: the leaf set of SclAvx::createDecoder (scl_avx_float.cpp:624-651): REP leaves only for n < 8
bool compile_scl(int N, const unsigned* frozen, int nf, SclProgram& out, const char** msg, bool float_rules = false);

// fp32 Fast-SSC node program (FastSscAvx::createDecoder, src/polarcode/decoding/fastssc_avx_float.cpp:797-896);
// opcodes FOp in pc_fastssc_f32_core.h. Fails like the reference (std::invalid_argument) on the frozen patterns its
// DoubleRepetition / TripleRepetition checks reject.
struct FloatProgram {
    int N = 0, K = 0, S = 0;
    bool staged_root = false;
    std::vector<uint32_t> code;
    int census[32] = { 0 };
};
bool compile_fastssc_f32(int N, const unsigned* frozen, int nf, FloatProgram& out, const char** msg);

// returns false (and leaves msg) on invalid input
bool compile_fastssc_i8(int N, const unsigned* frozen, int nf, FastSscProgram& out, const char** msg);

} // namespace pc
#endif
