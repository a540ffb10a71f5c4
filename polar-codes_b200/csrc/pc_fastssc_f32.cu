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

// ---- group-of-8 form: EIGHT LANES per codeword, the whole float tree in shared memory ---------------------------------
// The thread-per-codeword kernel above is DRAM-bound on its 4 N-byte-per-thread arena (11x the algorithmic traffic).
// Here a codeword's LLR stages (4 N bytes) and decision words sit in shared memory and the 8 lanes of a group -- the
// width of the reference's AVX vectors -- share the element-wise steps: lane k handles elements k, k + 8, ... of every
// f / g / g0 / g1 / combine step with at least 8 elements per half, which is where 88 % of the element operations of a
// (1024, 512) tree are. Leaves and the steps of nodes with fewer than 16 elements run on lane 0 of the group with
// the SAME core functions as the other kernel (their 8-lane summation / tie orders are already restated on arrays),
// so results are identical by construction. Codeword stride = 8 mod 32 words: the four groups of a warp use
// disjoint banks. About 48 codewords are resident per SM (12 warps).
template <class Op>
__device__ __forceinline__ void g8_map(const FCtx& c, const FView& in, int s, int g, const Op& op)
{
    const int h = 1 << (s - 1);
    float* dst = c.llr + c.llr_off[s - 1];
    for (int i0 = 0; i0 < h; i0 += 32) {
        float a[4], b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + 8 * u + g;
            if (i < h) a[u] = in[i], b[u] = in[i + h];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + 8 * u + g;
            if (i < h) dst[i] = op(a[u], b[u], i);
        }
    }
}

__device__ __forceinline__ uint32_t g8_or(uint32_t v)
{
    v |= __shfl_xor_sync(0xFFFFFFFFu, v, 1);
    v |= __shfl_xor_sync(0xFFFFFFFFu, v, 2);
    v |= __shfl_xor_sync(0xFFFFFFFFu, v, 4);
    return v;
}

// ROneNode::rightDecode with the 8 lanes computing the signs of g (same element-wise operations as fop_g1)
__device__ __forceinline__ void g8_g1(const FCtx& c, const FView& in, int s, int off, int g)
{
    const int h = 1 << (s - 1);
    if (h >= 32) {
        for (int w = 0; w < (h >> 5); ++w) {
            const uint32_t left = c.bld((off >> 5) + w);
            uint32_t r = 0;
            float a[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] = in[32 * w + 8 * u + g], b[u] = in[32 * w + 8 * u + g + h];
#pragma unroll
            for (int u = 0; u < 4; ++u) r |= fsign(g_f32(a[u], b[u], (left >> (8 * u + g)) & 1u)) << (8 * u + g);
            r = g8_or(r);
            if (g == 0) {
                c.bst((off >> 5) + w, left ^ r);
                c.bst(((off + h) >> 5) + w, r);
            }
        }
    } else { // h = 8 or 16: one word holds both halves
        const int sh = off & 31;
        uint32_t word = c.bld(off >> 5);
        uint32_t r = 0;
        for (int k = g; k < h; k += 8) r |= fsign(g_f32(in[k], in[k + h], (word >> (sh + k)) & 1u)) << k;
        r = g8_or(r);
        if (g == 0) {
            word ^= r << sh;
            word = (word & ~(lowmask(h) << (sh + h))) | (r << (sh + h));
            c.bst(off >> 5, word);
        }
    }
}

__device__ __forceinline__ void fastssc_f32_run_g8(const FCtx& c, int g, const uint32_t* code, const float* row, bool staged_root)
{
    const int S = c.S;
    if (staged_root) {
        for (int i = g; i < (1 << S); i += 8) c.st(S, i, row[i]);
        __syncwarp();
    }
    for (int pc = 0;; ++pc) {
        const uint32_t ins = __ldg(code + pc);
        const uint32_t op = instr_op(ins);
        const int s = (int)instr_s(ins), off = (int)instr_off(ins);
        if (op == FOP_END) break;
        const FView in = (s == S && !staged_root) ? FView{ row, 1 } : c.stage(s);
        const int h = 1 << (s - 1);
        if (op <= FOP_COMB0 && h >= 8 && (op < FOP_COMB || h >= 32)) {
            switch (op) {
            case FOP_F: g8_map(c, in, s, g, [](float a, float b, int) { return f_f32(a, b); }); break;
            case FOP_G0: g8_map(c, in, s, g, [](float a, float b, int) { return fadd(a, b); }); break;
            case FOP_G: {
                const uint32_t* bits = c.bits;
                g8_map(c, in, s, g, [bits, off](float a, float b, int i) { return g_f32(a, b, (bits[(off + i) >> 5] >> ((off + i) & 31)) & 1u); });
                break;
            }
            case FOP_G1: g8_g1(c, in, s, off, g); break;
            default: // FOP_COMB / FOP_COMB0 with whole words per half
                for (int w = g; w < (h >> 5); w += 8) {
                    const uint32_t r = c.bld(((off + h) >> 5) + w);
                    c.bst((off >> 5) + w, op == FOP_COMB0 ? r : (c.bld((off >> 5) + w) ^ r));
                }
                break;
            }
        } else if (g == 0) {
            switch (op) {
            case FOP_F: fop_f(c, in, s); break;
            case FOP_G: fop_g(c, in, s, off); break;
            case FOP_G0: fop_g0(c, in, s); break;
            case FOP_G1: fop_g1(c, in, s, off); break;
            case FOP_COMB: fop_comb(c, s, off, false); break;
            case FOP_COMB0: fop_comb(c, s, off, true); break;
            case FOP_R0: fleaf_fill(c, s, off, 0x7F800000u); break;
            case FOP_R1: fleaf_r1(c, in, s, off); break;
            case FOP_REP: fleaf_rep(c, in, s, off); break;
            case FOP_SPC: fleaf_spc(c, in, s, off); break;
            case FOP_DREP: fleaf_drep(c, in, s, off); break;
            case FOP_DSPC: fleaf_dspc(c, in, s, off); break;
            case FOP_DSPC8: fleaf_dspc8(c, in, off); break;
            case FOP_ZSPC: fleaf_zspc(c, in, s, off); break;
            default: {
                float v[8];
                if (op == FOP_ZSPC8 || op == FOP_REPONE8) {
                    for (int k = 0; k < 8; ++k) v[k] = in[k];
                } else {
                    accumulate8(in, 1 << s, v);
                }
                if (op == FOP_ZSPC8 || op == FOP_TREP)
                    fleaf_sum4_spc(c, v, s, off);
                else
                    fleaf_rep_then(c, v, s, off, op == FOP_T5);
                break;
            }
            }
        }
        __syncwarp(); // the next step reads what other lanes of the group wrote
    }
}

__global__ void __launch_bounds__(128) k_fastssc_f32_g8(const __grid_constant__ F32Params p)
{
    extern __shared__ __align__(16) uint32_t smem_g8[];
    const int T = blockDim.x, t = threadIdx.x, g = t & 7, slot = t >> 3, cpc = T >> 3;
    uint32_t* crc_table = smem_g8; // 256 words
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.crc_table[i];
    __syncthreads();
    float* mine = reinterpret_cast<float*>(smem_g8 + 256) + (size_t)slot * p.g8_cw_words;
    FCtx c;
    c.S = p.S;
    c.llr_off = p.llr_off;
    c.llr = mine;
    c.ls = 1;
    c.bits = reinterpret_cast<uint32_t*>(mine + p.elems);
    c.bs = 1;
    Ctx ic; // the int8 core's decision-word helpers (transform, gather) on the same words
    ic.S = p.S;
    ic.bits_kind = PK_SMEM;
    ic.bits = c.bits;
    ic.bws = 1;
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    const unsigned long long nbatch = (p.B + cpc - 1) / cpc;
    for (unsigned long long bi = blockIdx.x; bi < nbatch; bi += gridDim.x) {
        const unsigned long long fl = bi * cpc + slot;
        const bool live = fl < p.B;
        const unsigned long long f = live ? fl : p.B - 1;
        fastssc_f32_run_g8(c, g, p.code, p.llr + f * N, p.staged_root != 0);
        if (g == 0) {
            if (!p.systematic) transform_bits_inplace(ic);
            const bool res = extract_info(ic, p.runs, p.nruns, p.K, p.info + f * kb, p.crc, crc_table, live);
            if (p.ok && live) p.ok[f] = res ? 1 : 0;
        }
        __syncwarp();
    }
}

// ---- on-chip form: one thread per codeword, FEW threads per SM, the whole LLR tree in shared + tensor memory ---------
// The arena kernel above runs at 25 % issue utilisation behind DRAM latency: 3.4 k warp instructions and 47 KB of DRAM
// traffic per codeword. Here a codeword's tree never leaves the SM. Per thread:
//   tensor memory  stage S-1 (N / 2 floats, one column each; tcgen05.ld/st 32x32b.x8 move eight consecutive elements per
//                  instruction, and all lanes of a warp use the same element index, which is what the instruction wants)
//   shared memory  stages 0 .. S-2 (N / 2 - 1 floats) and the N / 32 decision words, thread-interleaved: a warp access is
//                  one conflict-free 128-byte row and, T being a compile-time constant, `base + immediate`
//   channel row    read in place with 16-byte loads, 32 element pairs in flight per thread
// Capacity decides the CTA: N = 1024 needs 512 columns + 543 words per thread -> 96 threads per SM (tensor memory has
// 128 lanes x 512 columns, shared memory 227 KB); N = 512: 2 CTAs x 96; N = 256: 3 CTAs x 128 (plan_f32). Codes whose
// stage S-1 does not fit 512 columns (N > 1024) stay on the arena kernel.
// Same core functions (pc_fastssc_f32_core.h) on different views, so the results are the arena kernel's bit for bit.
__device__ __forceinline__ float lds_f32(uint32_t a)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds_w(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_w(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void tm_ld8_f32(uint32_t taddr, float* a)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(a[0]), "=f"(a[1]), "=f"(a[2]), "=f"(a[3]), "=f"(a[4]), "=f"(a[5]), "=f"(a[6]), "=f"(a[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tm_st8_f32(uint32_t taddr, const float* a)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "f"(a[0]),
                 "f"(a[1]), "f"(a[2]), "f"(a[3]), "f"(a[4]), "f"(a[5]), "f"(a[6]), "f"(a[7])
                 : "memory");
}

__device__ __forceinline__ void lds4_f32(uint32_t a, float* v)
{
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(a));
}
__device__ __forceinline__ void sts4_f32(uint32_t a, const float* v)
{
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
}
// QUAD layout of the shared-memory stages: elements 4q .. 4q+3 of a thread are 16 contiguous bytes, quad q of thread t at
// byte 16 (q T + t) of the stage, so the f / g loops move four elements per LDS.128 / STS.128 (a warp access is 512
// contiguous bytes, conflict-free) and a scalar access costs one shift / mask more. Stages 0 and 1 take a quad each,
// stage s >= 2 takes 2^(s-2): quad offset of stage s = sm_quad_off(s).
__device__ __forceinline__ uint32_t sm_quad_off(int s) { return s < 2 ? (uint32_t)s : (1u << (s - 2)) + 1u; }

template <int T, bool QUAD>
struct SmView { // a stage in shared memory (QUAD: the quad layout above; else element i of a thread at byte 4 T i)
    static constexpr int kBatch = 16;
    uint32_t a; // shared address of this thread's element 0 of the stage
    __device__ __forceinline__ float operator[](int i) const
    {
        if (QUAD) return lds_f32(a + 16u * T * ((uint32_t)i >> 2) + 4u * ((uint32_t)i & 3u));
        return lds_f32(a + 4u * T * (uint32_t)i);
    }
    __device__ __forceinline__ void ld8(int i, float* v) const // i is a multiple of 8
    {
        if (QUAD) {
            const uint32_t q = a + 16u * T * ((uint32_t)i >> 2);
            lds4_f32(q, v);
            lds4_f32(q + 16u * T, v + 4);
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = lds_f32(a + 4u * T * (uint32_t)(i + k));
        }
    }
    __device__ __forceinline__ void wait() const {}
};
struct TmView { // stage S-1 in tensor memory
    static constexpr int kBatch = 32;
    uint32_t t; // tensor-memory address of this warp's lane quarter, column of element 0
    __device__ __forceinline__ float operator[](int i) const
    {
        uint32_t v;
        __syncwarp(); // leaves branch on data; the load is warp-collective
        tm_ld1_nowait(t + (uint32_t)i, v);
        tm_wait_ld();
        return __uint_as_float(v);
    }
    __device__ __forceinline__ void ld8(int i, float* v) const { tm_ld8_f32(t + (uint32_t)i, v); }
    __device__ __forceinline__ void wait() const { tm_wait_ld(); }
};
struct RowView { // the channel row in global memory, contiguous and 16-byte aligned
    static constexpr int kBatch = 32;
    const float* p;
    __device__ __forceinline__ float operator[](int i) const { return __ldg(p + i); }
    __device__ __forceinline__ void ld8(int i, float* v) const
    {
        const float4 x = __ldg(reinterpret_cast<const float4*>(p + i)), y = __ldg(reinterpret_cast<const float4*>(p + i + 4));
        v[0] = x.x, v[1] = x.y, v[2] = x.z, v[3] = x.w, v[4] = y.x, v[5] = y.y, v[6] = y.z, v[7] = y.w;
    }
    __device__ __forceinline__ void wait() const {}
};

// DST_TM: the step's destination stage (s - 1) is stage S-1 in tensor memory, else a shared-memory stage
template <int T, bool DST_TM, bool QUAD>
struct SmCtx {
    static constexpr bool kSoft = false;
    uint32_t sm_llr;  // shared address of this thread's element 0 of stage 0 (stage s starts sm_quad_off(s) quads /
                      // 2^s - 1 elements further)
    uint32_t sm_bits; // shared address of this thread's decision word 0
    uint32_t tm;      // tensor-memory address of column 0 of stage S-1 in this warp's lane quarter
    __device__ __forceinline__ uint32_t sbase(int s) const
    {
        return QUAD ? sm_llr + 16u * T * sm_quad_off(s) : sm_llr + 4u * T * ((1u << s) - 1u);
    }
    __device__ __forceinline__ SmView<T, QUAD> sview(int s) const { return SmView<T, QUAD>{ sbase(s) }; }
    __device__ __forceinline__ void st(int s, int i, float v) const
    {
        if (DST_TM) {
            __syncwarp();
            tm_st1(tm + (uint32_t)i, __float_as_uint(v));
        } else {
            if (QUAD)
                sts_f32(sbase(s) + 16u * T * ((uint32_t)i >> 2) + 4u * ((uint32_t)i & 3u), v);
            else
                sts_f32(sbase(s) + 4u * T * (uint32_t)i, v);
        }
    }
    __device__ __forceinline__ void st8(int s, int i, const float* v) const
    {
        if (DST_TM) {
            tm_st8_f32(tm + (uint32_t)i, v);
        } else {
            if (QUAD) {
                const uint32_t a = sbase(s) + 16u * T * ((uint32_t)i >> 2); // i is a multiple of 8
                sts4_f32(a, v);
                sts4_f32(a + 16u * T, v + 4);
            } else {
                const uint32_t a = sbase(s) + 4u * T * (uint32_t)i;
#pragma unroll
                for (int k = 0; k < 8; ++k) sts_f32(a + 4u * T * k, v[k]);
            }
        }
    }
    __device__ __forceinline__ uint32_t bld(int w) const { return lds_w(sm_bits + 4u * T * (uint32_t)w); }
    __device__ __forceinline__ void bst(int w, uint32_t v) const { sts_w(sm_bits + 4u * T * (uint32_t)w, v); }
    __device__ __forceinline__ uint32_t bit(int i) const { return (bld(i >> 5) >> (i & 31)) & 1u; }
    template <class Gen>
    __device__ __forceinline__ void put_bits(int off, int n, const Gen& gen) const
    {
        if (n >= 32) {
            for (int w = 0; w < (n >> 5); ++w) {
                uint32_t v = 0;
#pragma unroll 8
                for (int k = 0; k < 32; ++k) v |= (gen(32 * w + k) & 1u) << k;
                bst((off >> 5) + w, v);
            }
        } else {
            uint32_t v = 0;
            for (int k = 0; k < n; ++k) v |= (gen(k) & 1u) << k;
            const int sh = off & 31;
            const uint32_t m = lowmask(n) << sh;
            bst(off >> 5, (bld(off >> 5) & ~m) | (v << sh));
        }
    }
    template <class Gen>
    __device__ __forceinline__ void put_vals(int off, int n, const Gen& gen) const
    {
        put_bits(off, n, [&gen](int i) { return gen(i) >> 31; });
    }
};

// HAS_TM: stage S-1 lives in tensor memory (p.sm_top == S - 2); else every stage is in shared memory (p.sm_top == S - 1)
// QUAD: quad layout of the shared-memory stages. Measured (profiles/r02_c24_perf.txt): N = 256 68.2 -> 78.8 Gbit/s,
// N = 128 80.6 -> 91.4, N = 64 77.8 -> 87.5; with the 16-element rounds also N = 1024 next to the tensor-memory stage
// (51.9 -> 53.3, profiles/r02_c66_perf.txt), so every plan uses it (PC_F32_SM_QUAD=0 selects the plain layout).
template <int T, bool HAS_TM, bool QUAD>
__global__ void __launch_bounds__(T, 1) k_fastssc_f32_sm(const __grid_constant__ F32Params p)
{
    extern __shared__ __align__(16) uint32_t smem_sm[];
    __shared__ uint32_t tm_base_s;
    const int t = threadIdx.x;
    uint32_t* crc_table = smem_sm; // 256 words
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.crc_table[i];
    if (HAS_TM) {
        if (t < 32) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             (uint32_t)__cvta_generic_to_shared(&tm_base_s)),
                         "r"((uint32_t)p.sm_tm_cols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
    }
    __syncthreads();
    if (HAS_TM) asm volatile("tcgen05.fence::after_thread_sync;");
    const int S = p.S, top = p.sm_top, g0 = 4 * ((1 << (top - 1)) + 1); // float words per thread (the quad layout's count, either layout)
    uint32_t* bits_gen = smem_sm + 256 + (size_t)g0 * T + t;
    SmCtx<T, false, QUAD> cs;
    cs.sm_llr = (uint32_t)__cvta_generic_to_shared(smem_sm + 256 + (QUAD ? 4 * t : t));
    cs.sm_bits = (uint32_t)__cvta_generic_to_shared(bits_gen);
    cs.tm = HAS_TM ? tm_base_s + ((uint32_t)((t >> 5) & 3) << 21) : 0u; // lane field = 32 * (warp % 4), bits 31:16
    SmCtx<T, true, QUAD> ct;
    ct.sm_llr = cs.sm_llr, ct.sm_bits = cs.sm_bits, ct.tm = cs.tm;
    Ctx ic; // the int8 core's decision-word helpers (transform, gather) on the same words
    ic.S = S;
    ic.bits_kind = PK_SMEM;
    ic.bits = bits_gen;
    ic.bws = T;
    const int N = 1 << S, kb = (p.K + 7) >> 3;
    // warp-uniform frame loop: the tensor-memory accesses are warp-collective, so the lanes past the end of the batch
    // decode the last frame again and keep the result to themselves
    const unsigned long long w0 = (unsigned long long)blockIdx.x * T + (t & ~31);
    for (unsigned long long fw = w0; fw < p.B; fw += (unsigned long long)gridDim.x * T) {
        const unsigned long long fl = fw + (t & 31);
        const bool live = fl < p.B;
        const unsigned long long f = live ? fl : p.B - 1;
        const float* row = p.llr + f * N;
        for (int pc = 0;; ++pc) {
            const uint32_t ins = p.code[pc];
            const uint32_t op = instr_op(ins);
            const int s = (int)instr_s(ins), off = (int)instr_off(ins);
            if (op == FOP_END) break;
            if (s <= top) {
                fastssc_f32_step(cs, cs.sview(s), op, s, off);
            } else if (s == S) {
                // the root is never a leaf here (plan_f32 keeps staged roots on the arena kernel): internal steps only,
                // which spares this instance of the view the whole leaf zoo (a third of the kernel's code)
                const RowView in{ row };
                if (HAS_TM) {
                    switch (op) {
                    case FOP_F: fop_f(ct, in, s); break;
                    case FOP_G: fop_g(ct, in, s, off); break;
                    case FOP_G0: fop_g0(ct, in, s); break;
                    case FOP_G1: fop_g1(cs, in, s, off); break;
                    case FOP_COMB: fop_comb(cs, s, off, false); break;
                    default: fop_comb(cs, s, off, true); break;
                    }
                    tm_wait_st();
                } else {
                    switch (op) {
                    case FOP_F: fop_f(cs, in, s); break;
                    case FOP_G: fop_g(cs, in, s, off); break;
                    case FOP_G0: fop_g0(cs, in, s); break;
                    case FOP_G1: fop_g1(cs, in, s, off); break;
                    case FOP_COMB: fop_comb(cs, s, off, false); break;
                    default: fop_comb(cs, s, off, true); break;
                    }
                }
            } else if (HAS_TM) {
                fastssc_f32_step(cs, TmView{ cs.tm }, op, s, off);
            }
            __syncwarp();
        }
        if (!p.systematic) transform_bits_inplace(ic);
        const bool res = extract_info(ic, p.runs, p.nruns, p.K, p.info + f * kb, p.crc, crc_table, live);
        if (p.ok && live) p.ok[f] = res ? 1 : 0;
    }
    if (HAS_TM) {
        asm volatile("tcgen05.fence::before_thread_sync;");
        __syncthreads();
        if (t < 32)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm_base_s), "r"((uint32_t)p.sm_tm_cols));
    }
}

// ---- soft outputs: Decoder::getSoftCodeword / getSoftInformation of the float decoder (decoder.cpp:147-152) -----------
// FloatContainer::getSoftBits (bitcontainer.cpp:294-297) hands out the decoder's bit container: N floats whose SIGN is the
// decision. Leaves write LLR-valued floats (R0 +inf, R1 the LLRs, REP the sum, SPC the LLRs with the flipped one negated,
// the 8-lane specials their own patterns), and every combine step XORs the 32-bit PATTERNS (avx_float.h:190-236), so the
// words of an inner node are not meaningful numbers (they can be NaN patterns) -- they are reproduced bit for bit anyway.
// The arena kernel's thread-per-codeword form with N more words per thread for that container; the packed decision
// words are kept next to it (the g steps read them), so the hard decisions are those of the other kernels.
__global__ void __launch_bounds__(128) k_fastssc_f32_soft(const __grid_constant__ F32Params p, float* __restrict__ soft_code,
                                                          float* __restrict__ soft_info, const uint32_t* __restrict__ lut)
{
    const int T = blockDim.x, t = threadIdx.x;
    const size_t GT = (size_t)gridDim.x * T, gt = (size_t)blockIdx.x * T + t, AS = p.arena_stride;
    FSoftCtx c;
    c.S = p.S;
    c.llr_off = p.llr_off;
    c.llr = p.arena + gt;
    c.ls = AS;
    c.bits = p.bits_arena + gt;
    c.bs = AS;
    const int N = 1 << p.S;
    c.soft = p.bits_arena + (size_t)p.nbw * AS + gt;
    c.ss = AS;
    for (unsigned long long f = gt; f < p.B; f += GT) {
        const float* row = p.llr + f * N;
        if (p.staged_root)
            for (int i = 0; i < N; ++i) c.st(p.S, i, row[i]);
        for (int pc = 0;; ++pc) {
            const uint32_t ins = p.code[pc];
            const uint32_t op = instr_op(ins);
            const int s = (int)instr_s(ins), off = (int)instr_off(ins);
            if (op == FOP_END) break;
            const FView in = (s == p.S && !p.staged_root) ? FView{ row, 1 } : c.stage(s);
            fastssc_f32_step(c, in, op, s, off);
        }
        if (soft_code) {
            uint32_t* out = reinterpret_cast<uint32_t*>(soft_code) + f * (unsigned long long)N;
            for (int i = 0; i < N; ++i) out[i] = c.soft[(size_t)i * AS];
        }
        if (soft_info) { // FloatContainer::getSoftInformation bitcontainer.cpp:331-339
            uint32_t* out = reinterpret_cast<uint32_t*>(soft_info) + f * (unsigned long long)p.K;
            for (int k = 0; k < p.K; ++k) out[k] = c.soft[(size_t)__ldg(lut + k) * AS];
        }
    }
}

cudaError_t launch_fastssc_f32_soft(const F32Params& p, int grid, int threads, float* d_soft_code, float* d_soft_info, const uint32_t* d_lut,
                                    cudaStream_t st)
{
    k_fastssc_f32_soft<<<grid, threads, 0, st>>>(p, d_soft_code, d_soft_info, d_lut);
    return cudaGetLastError();
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

// the instantiation a plan names: CTA size 96 / 128, stage S-1 in tensor memory or not, quad layout or not
template <class F>
static auto sm_dispatch(const F32Launch& L, F&& f)
{
    const bool tm = L.p.sm_tm_cols != 0, quad = L.sm_quad != 0;
    if (L.sm_threads == 128) {
        if (tm) return quad ? f(k_fastssc_f32_sm<128, true, true>) : f(k_fastssc_f32_sm<128, true, false>);
        return quad ? f(k_fastssc_f32_sm<128, false, true>) : f(k_fastssc_f32_sm<128, false, false>);
    }
    if (tm) return quad ? f(k_fastssc_f32_sm<96, true, true>) : f(k_fastssc_f32_sm<96, true, false>);
    return quad ? f(k_fastssc_f32_sm<96, false, true>) : f(k_fastssc_f32_sm<96, false, false>);
}

template <class K>
static int sm_occ(K kern, const F32Launch& L)
{
    int n = 0;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.sm_smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, L.sm_threads, L.sm_smem);
    return e == cudaSuccess ? n : -1;
}
int fastssc_f32_sm_occupancy(const F32Launch& L)
{
    return sm_dispatch(L, [&](auto kern) { return sm_occ(kern, L); });
}

cudaError_t launch_fastssc_f32(const F32Launch& L, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, cudaStream_t st)
{
    F32Params p = L.p;
    p.llr = d_llr;
    p.info = d_info;
    p.ok = d_ok;
    p.B = B;
    if (L.g8) {
        const int cpc = L.g8_threads / 8;
        size_t nb = (B + cpc - 1) / cpc;
        const int grid = (int)(nb < (size_t)L.g8_grid ? nb : (size_t)L.g8_grid);
        cudaError_t e = cudaFuncSetAttribute(k_fastssc_f32_g8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.g8_smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(k_fastssc_f32_g8, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        k_fastssc_f32_g8<<<grid, L.g8_threads, L.g8_smem, st>>>(p);
        return cudaGetLastError();
    }
    if (L.sm) {
        size_t nb = (B + L.sm_threads - 1) / L.sm_threads;
        const int grid = (int)(nb < (size_t)L.sm_grid ? nb : (size_t)L.sm_grid);
        auto go = [&](auto kern) -> cudaError_t {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.sm_smem);
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
            if (e != cudaSuccess) return e;
            kern<<<grid, L.sm_threads, L.sm_smem, st>>>(p);
            return cudaGetLastError();
        };
        return sm_dispatch(L, go);
    }
    size_t ctas = (B + L.threads - 1) / L.threads;
    const int grid = (int)(ctas < (size_t)L.grid ? ctas : (size_t)L.grid);
    // the arena is laid out for L.grid * L.threads threads; a smaller grid only uses a prefix of every row
    k_fastssc_f32<<<grid, L.threads, 0, st>>>(p);
    return cudaGetLastError();
}

} // namespace pc
