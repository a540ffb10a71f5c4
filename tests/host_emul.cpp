// DEBUG HARNESS (tests only): compiles the device decoder core (polar-codes_b200/csrc/pc_fastssc_i8_core.h,
// all __host__ __device__) with g++ and runs it for ONE emulated thread per codeword on the CPU, so that the
// node program, the SWAR arithmetic and the info-bit gather can be diffed against the oracle without a GPU.
// It is not a product path: nothing under polar-codes_b200/ references it and the C-ABI has no CPU fallback.
#include <climits>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../polar-codes_b200/csrc/pc_fastssc_f32_core.h"
#include "../polar-codes_b200/csrc/pc_fastssc_i8_core.h"
#include "../polar-codes_b200/csrc/pc_program.h"

using namespace pc;

namespace pc {
thread_local uint32_t* g_tm_emul = nullptr;
}

extern "C" {

// mode: 0 all stages in "shared memory", 1 top stage + decision words in emulated tensor memory,
//       2 top stage + decision words in the "arena", 3 top stage in tensor memory and the next one in the arena
// returns 0 on success; ok[i] = detector verdict
int emul_fastssc_i8_decode_mode(int N, const unsigned* frozen, int nf, int systematic, int crc_kind, const int8_t* llr,
                                long B, uint8_t* out, uint8_t* ok, int mode)
{
    FastSscProgram p;
    const char* msg = nullptr;
    if (!compile_fastssc_i8(N, frozen, nf, p, &msg)) return 1;
    CrcSpec crc;
    if (!crc_spec(crc_kind, crc)) return 2;
    uint32_t table[256];
    crc_build_table(crc, table);
    const int S = p.S;
    const int top = p.staged_root ? S : S - 1;
    const int nchunks = (top >= 4 ? (1 << (top - 3)) : 1) + 1;
    const int nbw = (N + 31) / 32;
    std::vector<u4> rowbuf((N + 15) / 16 + 1);
    std::vector<u4> smem(nchunks), arena(nchunks);
    std::vector<uint32_t> tmem(4 * nchunks + nbw + 8), small(4), bits(nbw), place(32, 0u);
    for (int s = 4; s <= top; ++s) {
        uint32_t kind = PK_SMEM, base = (1u << (s - 4)) - 1u;
        if (s >= 6) {
            if ((mode == 1 || mode == 3) && s == top) { kind = PK_TMEM; base = 0; }
            if (mode == 2 && s == top) { kind = PK_ARENA; base = 0; }
            if (mode == 3 && s == top - 1) { kind = PK_ARENA; base = 0; }
        }
        place[s] = mk_place(kind, base);
    }
    const int kb = (p.K + 7) / 8;
    std::vector<uint8_t> outrow(kb + 8);
    g_tm_emul = tmem.data();
    for (long i = 0; i < B; ++i) {
        memcpy(rowbuf.data(), llr + i * N, N);
        Ctx c;
        c.S = S;
        c.place = place.data();
        c.small = small.data();
        c.ws = 1;
        c.sm = reinterpret_cast<uint32_t*>(smem.data());
        c.scs = 4;
        c.ar = reinterpret_cast<uint32_t*>(arena.data());
        c.acs = 4;
        c.tm = 0;
        c.bits_kind = mode == 1 ? PK_TMEM : PK_SMEM;
        c.bits = bits.data();
        c.bws = 1;
        c.tbits = (uint32_t)(4 * nchunks);
        fastssc_i8_run(c, p.code.data(), reinterpret_cast<const int8_t*>(rowbuf.data()), p.staged_root);
        if (!systematic) transform_bits_inplace(c);
        bool res = extract_info(c, p.runs.data(), (int)p.runs.size(), p.K, outrow.data(), crc, table);
        memcpy(out + i * kb, outrow.data(), kb);
        if (ok) ok[i] = res ? 1 : 0;
    }
    return 0;
}

int emul_fastssc_i8_decode(int N, const unsigned* frozen, int nf, int systematic, int crc_kind, const int8_t* llr,
                           long B, uint8_t* out, uint8_t* ok)
{
    return emul_fastssc_i8_decode_mode(N, frozen, nf, systematic, crc_kind, llr, B, out, ok, 0);
}

// fp32 Fast-SSC core on the host (one emulated thread); returns 0 on success, 3 when the frozen pattern is rejected
int emul_fastssc_f32_decode(int N, const unsigned* frozen, int nf, int systematic, int crc_kind, const float* llr, long B,
                            uint8_t* out, uint8_t* ok)
{
    FloatProgram fp;
    FastSscProgram ip; // gather runs
    const char* msg = nullptr;
    if (!compile_fastssc_f32(N, frozen, nf, fp, &msg)) return 3;
    if (!compile_fastssc_i8(N, frozen, nf, ip, &msg)) return 1;
    CrcSpec crc;
    if (!crc_spec(crc_kind, crc)) return 2;
    uint32_t table[256];
    crc_build_table(crc, table);
    const int S = fp.S;
    std::vector<int> off(S + 2, 0);
    int w = 0;
    for (int s = 0; s <= S; ++s) {
        off[s] = w;
        w += 1 << s;
    }
    std::vector<float> llrs(w + 8);
    std::vector<uint32_t> bits((N + 31) / 32 + 1);
    const int kb = (fp.K + 7) / 8;
    std::vector<uint8_t> outrow(kb + 8);
    for (long i = 0; i < B; ++i) {
        FCtx c;
        c.S = S;
        c.llr_off = off.data();
        c.llr = llrs.data();
        c.ls = 1;
        c.bits = bits.data();
        c.bs = 1;
        std::fill(bits.begin(), bits.end(), 0u);
        fastssc_f32_run(c, fp.code.data(), llr + i * N, fp.staged_root);
        Ctx ic; // reuse the int8 core's bit transform / gather on the same decision words
        ic.S = S;
        ic.bits_kind = PK_SMEM;
        ic.bits = bits.data();
        ic.bws = 1;
        if (!systematic) transform_bits_inplace(ic);
        bool res = extract_info(ic, ip.runs.data(), (int)ip.runs.size(), ip.K, outrow.data(), crc, table);
        memcpy(out + i * kb, outrow.data(), kb);
        if (ok) ok[i] = res ? 1 : 0;
    }
    return 0;
}

// the fp32 core with the soft-output context (the float bit container of Decoder::getSoftCodeword) on the host
int emul_fastssc_f32_soft(int N, const unsigned* frozen, int nf, const float* llr, long B, uint32_t* soft_code)
{
    FloatProgram fp;
    const char* msg = nullptr;
    if (!compile_fastssc_f32(N, frozen, nf, fp, &msg)) return 3;
    const int S = fp.S;
    std::vector<int> off(S + 2, 0);
    int w = 0;
    for (int s = 0; s <= S; ++s) {
        off[s] = w;
        w += 1 << s;
    }
    std::vector<float> llrs(w + 8);
    std::vector<uint32_t> bits((N + 31) / 32 + 1), soft(N);
    for (long i = 0; i < B; ++i) {
        FSoftCtx c;
        c.S = S;
        c.llr_off = off.data();
        c.llr = llrs.data();
        c.ls = 1;
        c.bits = bits.data();
        c.bs = 1;
        c.soft = soft.data();
        c.ss = 1;
        std::fill(bits.begin(), bits.end(), 0u);
        const float* row = llr + i * N;
        if (fp.staged_root)
            for (int k = 0; k < N; ++k) c.st(S, k, row[k]);
        for (size_t pc = 0;; ++pc) {
            const uint32_t ins = fp.code[pc];
            const uint32_t op = instr_op(ins);
            const int s = (int)instr_s(ins), o = (int)instr_off(ins);
            if (op == FOP_END) break;
            const FView in = (s == S && !fp.staged_root) ? FView{ row, 1 } : c.stage(s);
            fastssc_f32_step(c, in, op, s, o);
        }
        memcpy(soft_code + i * N, soft.data(), 4 * (size_t)N);
    }
    return 0;
}

int emul_program(int N, const unsigned* frozen, int nf, uint32_t* code, int cap)
{
    FastSscProgram p;
    const char* msg = nullptr;
    if (!compile_fastssc_i8(N, frozen, nf, p, &msg)) return -1;
    int n = (int)p.code.size();
    for (int i = 0; i < n && i < cap; ++i) code[i] = p.code[i];
    return n;
}
}
