// Fast-SSC int8 batch decoder kernel for sm_100a: one thread per codeword, LLR tree in shared memory
// (thread-interleaved 16-byte chunks, conflict-free LDS.128/STS.128), SWAR int8 arithmetic, hard decisions
// on bit-packed words. Replaces the per-codeword FastSscFipChar::decode of the reference
// (src/polarcode/decoding/fastssc_fip_char.cpp:616-630) for B codewords per launch.
//
// Why a thread (not a warp) per codeword: the pruned Fast-SSC tree of a (1024,512) code has 117 nodes of
// which 90 are <= 32 elements wide; a warp per codeword would leave 24-31 lanes idle in those, while all
// threads of a CTA run the SAME node program, so a thread per codeword has no divergence except the
// data-dependent SPC flip. Throughput is then bounded by the integer pipes, not by lane utilisation.
#include <climits>

#include "pc_fastssc_i8_core.h"
#include "pc_kernels.cuh"

namespace pc {

__constant__ uint32_t c_pool[kConstWords];

cudaError_t const_pool_upload(int offset_words, const uint32_t* h_src, int nwords, cudaStream_t st)
{
    return cudaMemcpyToSymbolAsync(c_pool, h_src, (size_t)nwords * 4, (size_t)offset_words * 4,
                                   cudaMemcpyHostToDevice, st);
}

struct KParams {
    const int8_t* llr;
    uint8_t* info;
    uint8_t* ok;
    unsigned long long B;
    int S, K, nruns, staged_root, systematic;
    CrcSpec crc;
    int const_code, const_runs;
    const uint32_t* g_code;
    const uint32_t* g_runs;
    const uint32_t* g_crc_table;
    int smem_chunks, gsplit, g_chunks, bits_in_smem, nbw;
    uint32_t* arena;
};

template <bool HYB, bool CONSTPROG>
__global__ void __launch_bounds__(256) k_fastssc_i8(const KParams p)
{
    extern __shared__ __align__(16) uint32_t smem[];
    const int T = blockDim.x, t = threadIdx.x;
    uint32_t* crc_table = smem;                        // 256 words
    uint32_t* small = smem + 256;                      // 4*T words
    uint32_t* chunks = small + 4 * T;                  // smem_chunks*4*T words
    uint32_t* sbits = chunks + (size_t)p.smem_chunks * 4 * T;
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.g_crc_table[i];
    __syncthreads();

    Ctx c;
    c.S = p.S;
    c.m.small = small + t;
    c.m.chunks = chunks + 4 * t;
    c.m.ws = T;
    c.m.bws = T;
    c.m.cs = 4 * T;
    c.m.bits = sbits + t;
    c.m.gchunks = nullptr;
    c.m.gcs = 0;
    c.m.gsplit = INT_MAX;
    if (HYB) {
        const size_t GT = (size_t)gridDim.x * T, gt = (size_t)blockIdx.x * T + t;
        c.m.gchunks = p.arena + 4 * gt;
        c.m.gcs = (int)(4 * GT);
        c.m.gsplit = p.gsplit;
        if (!p.bits_in_smem) {
            c.m.bits = p.arena + (size_t)p.g_chunks * 4 * GT + gt;
            c.m.bws = (int)GT;
        }
    }
    const uint32_t* code = CONSTPROG ? (c_pool + p.const_code) : p.g_code;
    const uint32_t* runs = CONSTPROG ? (c_pool + p.const_runs) : p.g_runs;
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    const unsigned long long stride = (unsigned long long)gridDim.x * T;
    for (unsigned long long f = (unsigned long long)blockIdx.x * T + t; f < p.B; f += stride) {
        fastssc_i8_run(c, code, p.llr + f * N, p.staged_root != 0);
        if (!p.systematic) transform_bits_inplace(c);
        bool res = extract_info(c, runs, p.nruns, p.K, p.info + f * kb, p.crc, crc_table);
        if (p.ok) p.ok[f] = res ? 1 : 0;
    }
}

template <bool HYB, bool CONSTPROG>
static cudaError_t launch_t(const FastSscLaunch& L, const KParams& kp, cudaStream_t st)
{
    auto kern = k_fastssc_i8<HYB, CONSTPROG>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<L.grid, L.threads, L.smem_bytes, st>>>(kp);
    return cudaGetLastError();
}

cudaError_t launch_fastssc_i8(const FastSscLaunch& L, const int8_t* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                              cudaStream_t st)
{
    KParams kp;
    kp.llr = d_llr;
    kp.info = d_info;
    kp.ok = d_ok;
    kp.B = B;
    kp.S = L.S;
    kp.K = L.K;
    kp.nruns = L.nruns;
    kp.staged_root = L.staged_root;
    kp.systematic = L.systematic;
    kp.crc = L.crc;
    kp.const_code = L.const_code;
    kp.const_runs = L.const_runs;
    kp.g_code = L.g_code;
    kp.g_runs = L.g_runs;
    kp.g_crc_table = L.g_crc_table;
    kp.smem_chunks = L.smem_chunks;
    kp.gsplit = L.gsplit;
    kp.g_chunks = L.g_chunks;
    kp.bits_in_smem = L.bits_in_smem;
    kp.nbw = L.nbw;
    kp.arena = L.arena;
    const bool hyb = L.gsplit != INT_MAX || !L.bits_in_smem;
    const bool cp = L.const_code >= 0 && L.const_runs >= 0;
    if (hyb) return cp ? launch_t<true, true>(L, kp, st) : launch_t<true, false>(L, kp, st);
    return cp ? launch_t<false, true>(L, kp, st) : launch_t<false, false>(L, kp, st);
}

} // namespace pc
