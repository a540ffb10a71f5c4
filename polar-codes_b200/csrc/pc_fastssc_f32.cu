// fp32 Fast-SSC batch decoder kernel for sm_100a: one thread per codeword interpreting the float node program
// (pc_fastssc_f32_core.h). Replaces FastSscAvxFloat::decode (reference src/polarcode/decoding/fastssc_avx_float.cpp:940-955).
// The float LLR tree of a codeword is 4 N bytes, far beyond shared memory at useful occupancy, so the per-thread
// stages and the decision words live in a thread-interleaved arena (every warp access to one element index is one
// coalesced 128-byte transaction). With the f / g loops loading eight element pairs before they store, the kernel is
// DRAM-bandwidth bound on that arena (profiles/r01_fastssc_f32_full_v2.txt: 3.7 TB/s, 47 KB of traffic per codeword).
// Moving the small stages to shared memory through generic pointers measured SLOWER both before and after that change
// (stages <= 3 / 4 / 5 on chip: 37 / 33 / 28 Gbit/s against 41 all-arena, N=1024); the way down is a warp per codeword
// with the whole tree in shared memory, next round.
// Compiled without FMA contraction (-fmad=false is applied through __fadd_rn in the core).
#include "pc_fastssc_f32_core.h"
#include "pc_kernels.cuh"

namespace pc {

__global__ void __launch_bounds__(256) k_fastssc_f32(const __grid_constant__ F32Params p)
{
    __shared__ uint32_t crc_table[256];
    const int T = blockDim.x, t = threadIdx.x;
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.crc_table[i];
    __syncthreads();
    const size_t GT = (size_t)gridDim.x * T, gt = (size_t)blockIdx.x * T + t, AS = p.arena_stride;
    FCtx c;
    c.S = p.S;
    c.llr_off = p.llr_off;
    c.llr = p.arena + gt;
    c.ls = AS;
    c.bits = p.bits_arena + gt;
    c.bs = AS;
    Ctx ic; // the int8 core's decision-word helpers (transform, gather) on the same words
    ic.S = p.S;
    ic.bits_kind = PK_ARENA;
    ic.bits = c.bits;
    ic.bws = AS;
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    for (unsigned long long f = gt; f < p.B; f += GT) {
        fastssc_f32_run(c, p.code, p.llr + f * N, p.staged_root != 0);
        if (!p.systematic) transform_bits_inplace(ic);
        const bool res = extract_info(ic, p.runs, p.nruns, p.K, p.info + f * kb, p.crc, crc_table, true);
        if (p.ok) p.ok[f] = res ? 1 : 0;
    }
}

// FloatContainer::insertLlr(const char*) (reference src/polarcode/bitcontainer.cpp:202-207): int8 -> float
__global__ void __launch_bounds__(256) k_i8_to_f32(const int8_t* __restrict__ in, float* __restrict__ out, unsigned long long n)
{
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x)
        out[i] = (float)in[i];
}

cudaError_t launch_i8_to_f32(const int8_t* d_in, float* d_out, size_t n, cudaStream_t st)
{
    size_t g = (n + 255) / 256;
    if (g > 148 * 32) g = 148 * 32;
    k_i8_to_f32<<<(int)g, 256, 0, st>>>(d_in, d_out, n);
    return cudaGetLastError();
}

cudaError_t launch_fastssc_f32(const F32Launch& L, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, cudaStream_t st)
{
    F32Params p = L.p;
    p.llr = d_llr;
    p.info = d_info;
    p.ok = d_ok;
    p.B = B;
    size_t ctas = (B + L.threads - 1) / L.threads;
    const int grid = (int)(ctas < (size_t)L.grid ? ctas : (size_t)L.grid);
    // the arena is laid out for L.grid * L.threads threads; a smaller grid only uses a prefix of every row
    k_fastssc_f32<<<grid, L.threads, 0, st>>>(p);
    return cudaGetLastError();
}

} // namespace pc
