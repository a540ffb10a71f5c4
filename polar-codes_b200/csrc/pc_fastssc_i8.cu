// Fast-SSC int8 batch decoder kernel for sm_100a: one thread per codeword, LLR tree stages in shared memory and
// TENSOR MEMORY (used as 256 KB/SM of private per-thread storage through tcgen05.ld/st 32x32b -- no MMA involved),
// SWAR int8 arithmetic built on VABSDIFF4/PRMT/LOP3, hard decisions on bit-packed words. Replaces the per-codeword
// FastSscFipChar::decode of the reference (src/polarcode/decoding/fastssc_fip_char.cpp:616-630) for B codewords
// per launch.
//
// Why a thread (not a warp) per codeword: the pruned Fast-SSC tree of a (1024,512) code has 117 nodes of
// which 90 are <= 32 elements wide; a warp per codeword would leave 24-31 lanes idle in those, while all
// threads of a CTA run the SAME node program, so a thread per codeword has no divergence except the
// data-dependent SPC flip. Throughput is then bounded by the integer pipes, not by lane utilisation; the number
// of resident warps is bounded by on-chip bytes per codeword (about N + N/8), hence shared + tensor memory.
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
    uint32_t place[24];
    int smem_chunks, tmem_alloc_cols, tmem_cols_thread, arena_chunks, bits_kind, bits_base, nbw;
    uint32_t* arena;
};

template <bool CONSTPROG>
__global__ void __launch_bounds__(1024) k_fastssc_i8(const __grid_constant__ KParams p)
{
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ uint32_t tm_base_s;
    const int T = blockDim.x, t = threadIdx.x, warp = t >> 5;
    uint32_t* crc_table = smem;                        // 256 words
    uint32_t* small = smem + 256;                      // 4*T words
    uint32_t* chunks = small + 4 * T;                  // smem_chunks*4*T words
    uint32_t* sbits = chunks + (size_t)p.smem_chunks * 4 * T;
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.g_crc_table[i];
    if (p.tmem_alloc_cols) {
        if (warp == 0) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             (uint32_t)__cvta_generic_to_shared(&tm_base_s)),
                         "r"((uint32_t)p.tmem_alloc_cols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
    }
    __syncthreads();
    if (p.tmem_alloc_cols) asm volatile("tcgen05.fence::after_thread_sync;");

    Ctx c;
    c.S = p.S;
    c.cur = 0;
    c.place = p.place;
    c.small = small + t;
    c.ws = T;
    c.sm = chunks + 4 * t;
    c.scs = 4 * T;
    const size_t GT = (size_t)gridDim.x * T, gt = (size_t)blockIdx.x * T + t;
    c.ar = p.arena + 4 * gt;
    c.acs = 4 * GT;
    // a warp reaches the 32 tensor-memory lanes of its quarter (warp % 4); warps sharing a quarter are stacked along columns
    c.tm = p.tmem_alloc_cols ? tm_base_s + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * p.tmem_cols_thread)
                             : 0u;
    c.bits_kind = (uint32_t)p.bits_kind;
    c.tbits = c.tm + (uint32_t)p.bits_base;
    if (p.bits_kind == PK_ARENA) {
        c.bits = p.arena + (size_t)p.arena_chunks * 4 * GT + gt;
        c.bws = GT;
    } else {
        c.bits = sbits + t;
        c.bws = (size_t)T;
    }
    const uint32_t* code = CONSTPROG ? (c_pool + p.const_code) : p.g_code;
    const uint32_t* runs = CONSTPROG ? (c_pool + p.const_runs) : p.g_runs;
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    const unsigned long long stride = GT;
    // warp-uniform trip count: lanes past the end of the batch decode the last frame again and store nothing
    for (unsigned long long base = (unsigned long long)blockIdx.x * T + (t & ~31); base < p.B; base += stride) {
        const unsigned long long fl = base + (t & 31);
        const bool live = fl < p.B;
        const unsigned long long f = live ? fl : p.B - 1;
        fastssc_i8_run(c, code, p.llr + f * N, p.staged_root != 0);
        if (!p.systematic) transform_bits_inplace(c);
        bool res = extract_info(c, runs, p.nruns, p.K, p.info + f * kb, p.crc, crc_table, live);
        if (p.ok && live) p.ok[f] = res ? 1 : 0;
    }
    if (p.tmem_alloc_cols) {
        asm volatile("tcgen05.fence::before_thread_sync;");
        __syncthreads();
        if (warp == 0)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm_base_s),
                         "r"((uint32_t)p.tmem_alloc_cols));
    }
}

template <bool CONSTPROG>
static cudaError_t launch_t(const FastSscLaunch& L, const KParams& kp, cudaStream_t st)
{
    auto kern = k_fastssc_i8<CONSTPROG>;
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
    for (int i = 0; i < 24; ++i) kp.place[i] = L.place[i];
    kp.smem_chunks = L.smem_chunks;
    kp.tmem_alloc_cols = L.tmem_alloc_cols;
    kp.tmem_cols_thread = L.tmem_cols_thread;
    kp.arena_chunks = L.arena_chunks;
    kp.bits_kind = L.bits_kind;
    kp.bits_base = L.bits_base;
    kp.nbw = L.nbw;
    kp.arena = L.arena;
    const bool cp = L.const_code >= 0 && L.const_runs >= 0;
    return cp ? launch_t<true>(L, kp, st) : launch_t<false>(L, kp, st);
}

} // namespace pc
