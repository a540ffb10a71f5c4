// pc_decode_batch_f32_static: the fp32 Fast-SSC batch decode (thread-per-codeword arena form, pc_fastssc_f32.cu
// k_fastssc_f32) with its per-thread LLR stages and decision words in a FILE-SCOPE device array instead of a cudaMalloc'ed
// arena, for callers that must not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc. No counterpart in the reference.
//
// The array is kScratchWords * 4 bytes = 32 MiB of every GPU that loads the library, until exit. Its host-side name is a
// stand-in, not a device address: the address comes from cudaGetSymbolAddress, asked once per device with that device
// current (scratch_for_device). Every call shares the array and it keeps the last call's contents; nothing in it has to
// start at zero (each frame's stages and decision words are written before they are read), so nothing is cleared.
// Calls that overlap on the GPU (two streams, two host threads) race on it.
#include <mutex>

#include "pc_fastssc_f32_core.h"
#include "pc_internal.h"

namespace pc {

constexpr size_t kScratchWords = (size_t)8 << 20; // 32 MiB
constexpr int kStaticMaxN = 4096;                // largest block length the entry point accepts
constexpr int kStaticThreads = 128;

static __device__ __align__(256) uint32_t g_static_scratch[kScratchWords];

// one thread per codeword, grid-stride over the batch; the same core functions as k_fastssc_f32, hence the same bits
__global__ void __launch_bounds__(kStaticThreads) k_fastssc_f32_static(const __grid_constant__ F32Params p)
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

// device address of the array on the CURRENT device, looked up once per device
static int scratch_for_device(int device, uint32_t** out)
{
    static std::mutex mu;
    static uint32_t* ptrs[64] = { nullptr };
    if (device < 0 || device >= 64) return fail(PC_ERR_ARG, "device index out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(static scratch)");
        ptrs[device] = static_cast<uint32_t*>(p);
    }
    *out = ptrs[device];
    return PC_OK;
}

} // namespace pc

using namespace pc;

extern "C" int pc_decode_batch_f32_static(pc_code* c, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, void* stream)
{
    if (!c || !d_llr || !d_info) return fail(PC_ERR_ARG, "null argument");
    if (c->kind != PC_DEC_FASTSSC_F32)
        return fail(PC_ERR_UNSUPPORTED, "pc_decode_batch_f32_static: the code must be a 'float' Fast-SSC decoder (PC_DEC_FASTSSC_F32)");
    if (c->N > kStaticMaxN)
        return fail(PC_ERR_UNSUPPORTED, "pc_decode_batch_f32_static: block length above the static scratch limit (N <= 4096)");
    if (B == 0) return PC_OK;
    if (int s = check_code_device(c)) return s;
    if (int a = check_aligned(d_llr, 4, "d_llr")) return a;
    uint32_t* scratch = nullptr;
    if (int s = scratch_for_device(c->device, &scratch)) return s;
    const FloatProgram& g = c->f32_prog;
    F32Params p = {};
    p.llr = d_llr;
    p.info = d_info;
    p.ok = d_ok;
    p.B = B;
    p.S = g.S;
    p.K = g.K;
    p.nruns = (int)c->prog.runs.size();
    p.staged_root = g.staged_root ? 1 : 0;
    p.systematic = c->systematic;
    crc_spec(c->crc_kind, p.crc);
    p.code = c->d_f32_code;
    p.runs = c->d_runs;
    p.crc_table = c->d_crc_table;
    const int top = g.staged_root ? g.S : g.S - 1;
    int w = 0;
    for (int st = 0; st <= top; ++st) {
        p.llr_off[st] = w;
        w += 1 << st;
    }
    p.elems = w;
    p.nbw = (c->N + 31) / 32;
    // as many CTAs as the array holds (words per thread = LLR stages + decision words), at most four per SM
    const size_t per_cta = (size_t)(p.elems + p.nbw) * kStaticThreads;
    size_t ctas = kScratchWords / per_cta; // >= 3 for N <= 4096
    if (ctas < 1) return fail(PC_ERR_UNSUPPORTED, "pc_decode_batch_f32_static: code too long for the static scratch");
    if (ctas > 148 * 4) ctas = 148 * 4;
    const size_t want = (B + kStaticThreads - 1) / kStaticThreads;
    if (ctas > want) ctas = want;
    p.arena_stride = ctas * kStaticThreads;
    p.arena = reinterpret_cast<float*>(scratch);
    p.bits_arena = scratch + (size_t)p.elems * p.arena_stride;
    k_fastssc_f32_static<<<(int)ctas, kStaticThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "fp32 Fast-SSC (static scratch) launch");
}
