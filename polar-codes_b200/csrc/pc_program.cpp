#include "pc_program.h"

#include <algorithm>

namespace pc {

namespace {

constexpr int kV = 32; // lanes of the reference's fipv register (include/polarcode/avxconvenience.h:52-54)

struct Builder {
    FastSscProgram& p;
    int blockS; // nodes of this size own the register-held block of hard decisions

    // frozen: sorted indices relative to the node; mirrors splitFrozenBits (src/polarcode/polarcode.cpp:14-34)
    static void split(const std::vector<unsigned>& fz, unsigned half, std::vector<unsigned>& l, std::vector<unsigned>& r)
    {
        for (unsigned v : fz) {
            if (v < half)
                l.push_back(v);
            else
                r.push_back(v - half);
        }
    }

    void emit(uint32_t op, int s, unsigned off)
    {
        p.code.push_back(mk_instr(op, (uint32_t)s, off));
        p.census[op]++;
    }

    void node(int s, unsigned off, const std::vector<unsigned>& fz)
    {
        const unsigned n = 1u << s, h = n >> 1;
        const size_t nf = fz.size();
        bool done = true;
        // order of FastSscFip::createDecoder (fastssc_fip_char.cpp:496-569)
        if (nf == n)
            emit(OP_R0, s, off);
        else if (nf == 0)
            emit(OP_R1, s, off);
        else if (nf == n - 1)
            emit(OP_REP, s, off);
        else if (nf == 1)
            emit(OP_SPC, s, off);
        else if (nf == n - 2 && n >= (unsigned)kV)
            emit(OP_DREP, s, off);
        else
            done = false;
        if (!done) {
            std::vector<unsigned> lf, rf;
            split(fz, h, lf, rf);
            const bool leftZero = lf.size() == h, rightOne = rf.empty();
            if (n <= (unsigned)kV && leftZero && rightOne) {
                emit(OP_S01, s, off);
            } else if (leftZero && rf.size() == 1) {
                emit(OP_0SPC, s, off); // short variant for n <= 32, long for n >= 64: decided by s in the kernel
            } else if (rightOne) {
                emit(OP_F, s, off);
                node(s - 1, off, lf);
                emit(OP_G1, s, off);
            } else if (leftZero) {
                emit(OP_G0, s, off);
                node(s - 1, off + h, rf);
                emit(OP_COMB0, s, off);
            } else {
                emit(OP_F, s, off);
                node(s - 1, off, lf);
                emit(OP_G, s, off);
                node(s - 1, off + h, rf);
                emit(OP_COMB, s, off);
            }
        }
        if (s == blockS) emit(OP_FLUSH, s, off);
    }
};

} // namespace

bool compile_fastssc_i8(int N, const unsigned* frozen, int nf, FastSscProgram& out, const char** msg)
{
    static const char* e_n = "block length must be a power of two in [2, 65536]";
    static const char* e_f = "frozen indices must be unique and < N";
    int S = 0;
    while ((1 << S) < N) S++;
    if (N < 2 || (1 << S) != N || N > 65536) {
        if (msg) *msg = e_n;
        return false;
    }
    std::vector<unsigned> fz(frozen, frozen + nf);
    std::sort(fz.begin(), fz.end());
    for (size_t i = 0; i < fz.size(); ++i) {
        if (fz[i] >= (unsigned)N || (i && fz[i] == fz[i - 1])) {
            if (msg) *msg = e_f;
            return false;
        }
    }
    out = FastSscProgram();
    out.N = N;
    out.K = N - nf;
    out.S = S;
    Builder b{ out, S < 5 ? S : 5 };
    b.node(S, 0, fz);
    b.emit(OP_END, 0, 0);
    // the root's LLRs stay in HBM unless the root itself is a leaf or narrower than one 16-byte chunk pair
    uint32_t rootop = instr_op(out.code[0]);
    out.staged_root = (S < 5) || !(rootop == OP_F || rootop == OP_G0);

    // info positions, mask and gather runs
    out.info_mask.assign((N + 31) / 32, 0u);
    {
        size_t fc = 0;
        for (int i = 0; i < N; ++i) {
            if (fc < fz.size() && fz[fc] == (unsigned)i) {
                fc++;
            } else {
                out.lut.push_back((uint32_t)i);
                out.info_mask[i >> 5] |= 1u << (i & 31);
            }
        }
    }
    const int K = out.K;
    int k = 0;
    while (k < K) {
        // maximal run of consecutive positions inside one source word and one destination word
        uint32_t pos = out.lut[k];
        int len = 1;
        while (k + len < K && out.lut[k + len] == pos + len && ((pos + len) >> 5) == (pos >> 5) &&
               ((k + len) >> 5) == (k >> 5))
            len++;
        int kend = k + len;
        uint32_t last = (kend == K || (kend & 31) == 0) ? 1u : 0u;
        out.runs.push_back(mk_run(pos >> 5, pos & 31, (uint32_t)len, (uint32_t)(k & 31), last));
        k = kend;
    }
    return true;
}


namespace {
void scl_node(SclProgram& p, int s, uint32_t side, const std::vector<unsigned>& fz, bool float_rules)
{
    const unsigned n = 1u << s, h = n >> 1;
    const size_t nf = fz.size();
    uint32_t leaf = 0xFFu;
    if (nf == n)
        leaf = SCL_R0;
    else if (nf == 0)
        leaf = SCL_R1;
    else if (nf == n - 1 && (!float_rules || n < 8)) // SclAvx builds REP leaves only below one vector (scl_avx_float.cpp:637-639)
        leaf = SCL_REP;
    else if (nf == 1)
        leaf = SCL_SPC;
    if (leaf != 0xFFu) {
        p.code.push_back(scl_instr(leaf, (uint32_t)s, side));
        p.census[leaf]++;
        return;
    }
    std::vector<unsigned> lf, rf;
    for (unsigned v : fz) {
        if (v < h)
            lf.push_back(v);
        else
            rf.push_back(v - h);
    }
    p.code.push_back(scl_instr(SCL_F, (uint32_t)s, side));
    scl_node(p, s - 1, 0, lf, float_rules);
    p.code.push_back(scl_instr(SCL_G, (uint32_t)s, side));
    scl_node(p, s - 1, 1, rf, float_rules);
    p.code.push_back(scl_instr(SCL_COMB, (uint32_t)s, side));
    p.census[SCL_F]++;
}
} // namespace

bool compile_scl(int N, const unsigned* frozen, int nf, SclProgram& out, const char** msg, bool float_rules)
{
    int S = 0;
    while ((1 << S) < N) S++;
    if (N < 2 || (1 << S) != N || N > 65536) {
        if (msg) *msg = "block length must be a power of two in [2, 65536]";
        return false;
    }
    std::vector<unsigned> fz(frozen, frozen + nf);
    std::sort(fz.begin(), fz.end());
    out = SclProgram();
    out.N = N;
    out.K = N - nf;
    out.S = S;
    scl_node(out, S, 1, fz, float_rules);
    out.root_leaf = instr_op(out.code[0]) != SCL_F;
    out.code.push_back(scl_instr(SCL_END, 0, 0));
    return true;
}

namespace {
// opcode values of pc_fastssc_f32_core.h (FOp)
enum { F_F = 0, F_G = 1, F_G0 = 2, F_G1 = 3, F_COMB = 4, F_COMB0 = 5, F_R0 = 6, F_R1 = 7, F_REP = 8, F_SPC = 9, F_DREP = 10,
       F_DSPC = 11, F_DSPC8 = 12, F_ZSPC = 13, F_ZSPC8 = 14, F_TREP = 15, F_T5 = 16, F_REPONE8 = 17, F_END = 18 };

struct FBuilder {
    FloatProgram& p;
    bool bad = false;
    void emit(uint32_t op, int s, unsigned off)
    {
        p.code.push_back(mk_instr(op, (uint32_t)s, off));
        p.census[op]++;
    }
    static bool is_prefix(const std::vector<unsigned>& fz)
    {
        for (size_t i = 0; i < fz.size(); ++i)
            if (fz[i] != i) return false;
        return true;
    }
    void node(int s, unsigned off, const std::vector<unsigned>& fz)
    {
        const unsigned n = 1u << s, h = n >> 1;
        const size_t nf = fz.size();
        if (nf == n) return emit(F_R0, s, off);
        if (nf == 0) return emit(F_R1, s, off);
        if (nf == n - 1) return emit(F_REP, s, off);
        if (nf == 1) return emit(F_SPC, s, off);
        if (nf == n - 2) {
            if (!is_prefix(fz)) bad = true;
            return emit(F_DREP, s, off);
        }
        if (nf == 2 && fz[0] == 0 && fz[1] == 1) return emit(n == 8 ? F_DSPC8 : F_DSPC, s, off);
        if (nf == n - 3 && n > 8 && fz[nf - 1] == n - 4) {
            if (!is_prefix(fz)) bad = true;
            return emit(F_TREP, s, off);
        }
        if (nf == n - 4 && fz[nf - 1] == n - 4 && fz[nf - 2] == n - 6) return emit(F_T5, s, off);
        if (n == 8 && nf == 3 && fz[0] == 0 && fz[1] == 1 && fz[2] == 2) return emit(F_REPONE8, s, off);
        if (n == 8 && nf == 5 && fz[nf - 1] == n - 4 && fz[nf - 2] == n - 5) return emit(F_ZSPC8, s, off);
        std::vector<unsigned> lf, rf;
        for (unsigned v : fz) {
            if (v < h)
                lf.push_back(v);
            else
                rf.push_back(v - h);
        }
        if (n > 8) {
            if (lf.size() == h && rf.size() == 1) return emit(F_ZSPC, s, off);
            if (rf.empty()) {
                emit(F_F, s, off);
                node(s - 1, off, lf);
                return emit(F_G1, s, off);
            }
            if (lf.size() == h) {
                emit(F_G0, s, off);
                node(s - 1, off + h, rf);
                return emit(F_COMB0, s, off);
            }
        }
        emit(F_F, s, off);
        node(s - 1, off, lf);
        emit(F_G, s, off);
        node(s - 1, off + h, rf);
        emit(F_COMB, s, off);
    }
};
} // namespace

bool compile_fastssc_f32(int N, const unsigned* frozen, int nf, FloatProgram& out, const char** msg)
{
    int S = 0;
    while ((1 << S) < N) S++;
    if (N < 2 || (1 << S) != N || N > 65536) {
        if (msg) *msg = "block length must be a power of two in [2, 65536]";
        return false;
    }
    std::vector<unsigned> fz(frozen, frozen + nf);
    std::sort(fz.begin(), fz.end());
    out = FloatProgram();
    out.N = N;
    out.K = N - nf;
    out.S = S;
    FBuilder b{ out };
    b.node(S, 0, fz);
    b.emit(F_END, 0, 0);
    if (b.bad) {
        if (msg) *msg = "frozen pattern rejected by the float decoder (DoubleRepetition / TripleRepetition need a frozen prefix)";
        return false;
    }
    const uint32_t rootop = instr_op(out.code[0]);
    out.staged_root = !(rootop == F_F || rootop == F_G0);
    return true;
}

} // namespace pc
