// DEBUG HARNESS (tests only): compiles the device decoder core (polar-codes_b200/csrc/pc_fastssc_i8_core.h,
// all __host__ __device__) with g++ and runs it for ONE emulated thread per codeword on the CPU, so that the
// node program, the SWAR arithmetic and the info-bit gather can be diffed against the oracle without a GPU.
// It is not a product path: nothing under polar-codes_b200/ references it and the C-ABI has no CPU fallback.
#include <climits>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../polar-codes_b200/csrc/pc_fastssc_i8_core.h"
#include "../polar-codes_b200/csrc/pc_program.h"

using namespace pc;

extern "C" {

// returns 0 on success; ok[i] = detector verdict
int emul_fastssc_i8_decode(int N, const unsigned* frozen, int nf, int systematic, int crc_kind, const int8_t* llr,
                           long B, uint8_t* out, uint8_t* ok)
{
    FastSscProgram p;
    const char* msg = nullptr;
    if (!compile_fastssc_i8(N, frozen, nf, p, &msg)) return 1;
    CrcSpec crc;
    if (!crc_spec(crc_kind, crc)) return 2;
    uint32_t table[256];
    crc_build_table(crc, table);
    const int S = p.S;
    const int nchunks = (S >= 4 ? (1 << (S - 3)) : 1) + 1;
    const int nbw = (N + 31) / 32;
    // 16-byte aligned row copy (rows of N < 16 bytes are unaligned in the batch)
    std::vector<u4> rowbuf((N + 15) / 16 + 1);
    std::vector<u4> chunks(nchunks);
    std::vector<uint32_t> small(4), bits(nbw);
    const int kb = (p.K + 7) / 8;
    std::vector<uint8_t> outrow(kb + 8);
    for (long i = 0; i < B; ++i) {
        memcpy(rowbuf.data(), llr + i * N, N);
        Ctx c;
        c.m.small = small.data();
        c.m.chunks = reinterpret_cast<uint32_t*>(chunks.data());
        c.m.bits = bits.data();
        c.m.ws = 1;
        c.m.bws = 1;
        c.m.cs = 4;
        c.m.gchunks = nullptr;
        c.m.gcs = 0;
        c.m.gsplit = INT_MAX;
        c.S = S;
        fastssc_i8_run(c, p.code.data(), reinterpret_cast<const int8_t*>(rowbuf.data()), p.staged_root);
        if (!systematic) transform_bits_inplace(c);
        bool res = extract_info(c, p.runs.data(), (int)p.runs.size(), p.K, outrow.data(), crc, table);
        memcpy(out + i * kb, outrow.data(), kb);
        if (ok) ok[i] = res ? 1 : 0;
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
