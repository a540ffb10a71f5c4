// CRC-aided SCL fp32 batch decoder for sm_100a: same organisation as pc_scl_i8.cu (a group of G = 2^ceil(log2 L)
// lanes owns one codeword, one lane per list path, owner tables instead of the ref-counted block tree, emulated
// selection sorts). Replaces SclAvxFloat::decode (reference src/polarcode/decoding/scl_avx_float.cpp:690-751).
//
// * f / g are the exact IEEE operations of _mm256_polarf_ps / _mm256_polarg_ps (avx_float.h:55-81, pc_fastssc_f32_core.h);
//   path metrics are floats, larger = better, and every sum that feeds a metric is taken in the reference's order:
//   8 per-lane accumulators over the node's vectors, then reduce_add_ps left to right (avxconvenience.h:256-272).
// * Leaves: R1 -> R0 -> REP (only for n < 8) -> SPC, anything else splits (createDecoder scl_avx_float.cpp:624-651).
// * Hard decisions are one bit per element (the reference's bit floats are only consumed through their sign bit).
// * Storage: a float LLR tree per path is 4 N bytes, so only the small stages (s <= smem_top) and the decision words
//   live in shared memory; the large stages live in a thread-interleaved, L2-resident arena
//   ([element][lane], the lanes of a group are adjacent, so a step of all lanes is one coalesced transaction).
// * Optional indirection (frame_idx / count) lets the adaptive decoder run this kernel over the frames whose first
//   pass failed, without a host round trip (AdaptiveFloat::decode adaptive_float.cpp:33-45).
#include <cstdio>
#include <cstdlib>

#include "pc_fastssc_f32_core.h"
#include "pc_kernels.cuh"
#include "pc_scl_common.cuh"

namespace pc {

__device__ __forceinline__ float lds_f32(uint32_t a)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }

// BM_ARENA: some decision words live in the arena (long codes); BM_TMEM: in tensor memory; 0 = no flag tests.
// TT: threads per CTA as a compile-time constant (0 = run-time): the lane-interleaved shared-memory layout puts element i
// of a stage at byte offset 4 TT i of the lane's column, which is an immediate of the load when i and TT are known
template <int BM, int TT>
struct FLaneT {
    static constexpr int kTT = TT;
    uint32_t* sm;   // smem base + tid (word w at sm[w * T])
    float* ar;      // arena base + global lane id (element e at ar[e * AS])
    size_t AS;
    size_t gl3;     // 3 x global lane id (quad layout of the arena stages, see View)
    int aq;         // arena LLR stages in the quad layout
    uint32_t tm;    // tensor-memory address of this thread's column 0
    int T, G, l;    // threads per CTA, lanes per group, lane index inside the group
    int P;          // live paths (uniform)
    int smem_top;
    int root_right; // virtual top stage: 0 while the left half of the root is decoded, 1 after the root's g step
    float metric;
    Tbl own_llr, own_L, own_R;
    // decision words: shared memory, tensor memory (warp-collective, see pc_scl_common.cuh kTmem), or the arena for the
    // largest stages of long codes (flag in the offset)
    __device__ __forceinline__ uint32_t ld(int word, uint32_t owner) const
    {
        if ((BM & BM_TMEM) && (word & kTmem)) {
            uint32_t v;
            tm_ld1_nowait(tm + (uint32_t)(word & 0xFFFF), v);
            tm_wait_ld();
            return __shfl_sync(0xFFFFFFFFu, v, (int)owner, G);
        }
        if ((BM & BM_ARENA) && (word & kArena)) return reinterpret_cast<const uint32_t*>(ar)[(size_t)(word & (kArena - 1)) * AS + ((int)owner - l)];
        return sm[(size_t)word * T + ((int)owner - l)];
    }
    __device__ __forceinline__ void st(int word, uint32_t v) const
    {
        if ((BM & BM_TMEM) && (word & kTmem)) {
            tm_st1(tm + (uint32_t)(word & 0xFFFF), v); // completion is awaited by st_fence(), once per step
        } else if ((BM & BM_ARENA) && (word & kArena))
            reinterpret_cast<uint32_t*>(ar)[(size_t)(word & (kArena - 1)) * AS] = v;
        else
            sm[(size_t)word * T] = v;
    }
    // tensor-memory stores complete: called before a step re-reads columns it wrote (scl_take_bits after the leaf's own
    // decisions) and at the end of every node-program step
    __device__ __forceinline__ void st_fence() const
    {
        if (BM & BM_TMEM) tm_wait_st();
    }
    // LLR stage s (off = p.llr_off[s]) in lane `owner`'s space, resolved once per node step
    // quad: an arena stage in the QUAD layout (p.arena_quad): elements 4q .. 4q+3 of a lane are 16 contiguous bytes, quad q
    // of lane L at float (q AS + L) * 4 of the stage, so the f / g loops move four elements per 16-byte access (a warp
    // access is 512 contiguous bytes); `stride` is then the distance between a lane's quads (4 AS).
    struct View {
        float* p;
        uint32_t stride; // element index * stride stays below 2^32 (the planner bounds the arena at 1 GiB)
        uint32_t quad;
        __device__ __forceinline__ float operator[](int i) const
        {
            return quad ? p[((uint32_t)i >> 2) * stride + ((uint32_t)i & 3u)] : p[(uint32_t)i * stride];
        }
        __device__ __forceinline__ void set(int i, float v) const
        {
            if (quad)
                p[((uint32_t)i >> 2) * stride + ((uint32_t)i & 3u)] = v;
            else
                p[(uint32_t)i * stride] = v;
        }
        // four elements starting at i (a multiple of 4) of a quad view
        __device__ __forceinline__ float4 ld4(int i) const { return *reinterpret_cast<const float4*>(p + ((uint32_t)i >> 2) * stride); }
        __device__ __forceinline__ void st4(int i, const float4& v) const { *reinterpret_cast<float4*>(p + ((uint32_t)i >> 2) * stride) = v; }
    };
    __device__ __forceinline__ View view(int s, int off, uint32_t owner) const
    {
        const ptrdiff_t d = (int)owner - l;
        if (s <= smem_top) return View{ reinterpret_cast<float*>(sm) + (size_t)off * T + d, (uint32_t)T, 0u };
        if (aq) return View{ ar + (size_t)off * AS + gl3 + 4 * d, 4u * (uint32_t)AS, 1u };
        return View{ ar + (size_t)off * AS + d, (uint32_t)AS, 0u };
    }
    // the lane's OWN stage in shared memory with the compile-time CTA size: 32-bit shared address, element i at + 4 TT i
    struct SView {
        uint32_t a;
        __device__ __forceinline__ float operator[](int i) const { return lds_f32(a + 4u * (uint32_t)(TT ? TT : 1) * (uint32_t)i); }
        __device__ __forceinline__ void set(int i, float v) const { sts_f32(a + 4u * (uint32_t)(TT ? TT : 1) * (uint32_t)i, v); }
    };
    template <class F>
    __device__ __forceinline__ void with_own_view(int s, int off, F&& f) const
    {
        if (TT != 0 && s <= smem_top)
            f(SView{ (uint32_t)__cvta_generic_to_shared(sm) + 4u * (uint32_t)(TT ? TT : 1) * (uint32_t)off });
        else
            f(view(s, off, (uint32_t)l));
    }
    // LLR element i of stage s (off = p.llr_off[s]) in lane `owner`'s space
    __device__ __forceinline__ float lld(int s, int off, int i, uint32_t owner) const
    {
        const ptrdiff_t d = (int)owner - l;
        if (s <= smem_top) return __uint_as_float(sm[(size_t)(off + i) * T + d]);
        return view(s, off, owner)[i];
    }
    __device__ __forceinline__ void lst(int s, int off, int i, float v) const
    {
        if (s <= smem_top)
            sm[(size_t)(off + i) * T] = __float_as_uint(v);
        else
            view(s, off, (uint32_t)l).set(i, v);
    }
};

// _mm256_max_ps lane: returns the second operand unless a > b
__device__ __forceinline__ float fmax_ps(float a, float b) { return a > b ? a : b; }

// Source of an f / g step: a stored stage (view), the channel row (the root's own step), or -- with the VIRTUAL top
// stage -- stage S-1 recomputed from the channel row on the fly: element i of stage S-1 is f(row[i], row[i + N/2]) while
// the left half of the root is decoded and g(row[i], row[i + N/2], root-left decision i) afterwards. Stage S-1 is the
// largest LLR buffer of a path (2 N bytes of the 4 N); not storing it halves the arena, which is what lets the arena
// stay L2-resident (ncu r01/r02: 58x the algorithmic DRAM traffic with all stages >= 4 in a 233 MB arena).
template <class FLane>
struct FgSrc {
    const FLane& c;
    typename FLane::View v;
    const float* row;
    int mode;   // 0 view, 1 row, 2 virtual/left (f), 3 virtual/right (g)
    int half;   // N / 2 (virtual modes)
    int lbv;    // root-left decision words (virtual/right)
    uint32_t olv;
    uint32_t cache_w[2];
    int cache_i[2];
    __device__ __forceinline__ float get(int i, int which)
    {
        if (mode == 0) return v[i];
        if (mode == 1) return __ldg(row + i);
        const float a = __ldg(row + i), b = __ldg(row + i + half);
        if (mode == 2) return f_f32(a, b);
        return g_f32(a, b, (cache_w[which] >> (i & 31)) & 1u);
    }
    // warp-collective: fetch the root-left decision word that covers elements [i, i + 32) for stream `which`
    __device__ __forceinline__ void bits_for(int i, int which)
    {
        if (mode == 3 && (i >> 5) != cache_i[which]) {
            cache_i[which] = i >> 5;
            cache_w[which] = c.ld(lbv + (i >> 5), olv);
        }
    }
};

template <bool IS_G, int UNR, class FLane>
__device__ __forceinline__ void fg_block(const FLane& c, FgSrc<FLane>& src, const typename FLane::View& dst, bool active, int h, int lb, uint32_t ol)
{
    const bool row16 = src.mode == 1 && ((reinterpret_cast<uintptr_t>(src.row) & 15u) == 0);
    uint32_t bw = 0;
    for (int j = 0; j < h; j += UNR) {
        float a[UNR], b[UNR], r[UNR];
        if (row16) {
#pragma unroll
            for (int u = 0; u < UNR; u += 4) {
                const float4 va = __ldg(reinterpret_cast<const float4*>(src.row + j + u)), vb = __ldg(reinterpret_cast<const float4*>(src.row + h + j + u));
                a[u] = va.x, a[u + 1] = va.y, a[u + 2] = va.z, a[u + 3] = va.w;
                b[u] = vb.x, b[u + 1] = vb.y, b[u + 2] = vb.z, b[u + 3] = vb.w;
            }
        } else if (src.mode >= 2 && ((reinterpret_cast<uintptr_t>(src.row) & 15u) == 0)) {
            // virtual stage S-1: four 16-byte row loads per four elements of each stream, then the root's f / g
            src.bits_for(j, 0); // (uniform; no-ops before the root's g step)
            src.bits_for(h + j, 1);
#pragma unroll
            for (int u = 0; u < UNR; u += 4) {
                const float4 x0 = __ldg(reinterpret_cast<const float4*>(src.row + j + u));
                const float4 x1 = __ldg(reinterpret_cast<const float4*>(src.row + j + u + src.half));
                const float4 y0 = __ldg(reinterpret_cast<const float4*>(src.row + h + j + u));
                const float4 y1 = __ldg(reinterpret_cast<const float4*>(src.row + h + j + u + src.half));
                const float xa[4] = { x0.x, x0.y, x0.z, x0.w }, xb[4] = { x1.x, x1.y, x1.z, x1.w };
                const float ya[4] = { y0.x, y0.y, y0.z, y0.w }, yb[4] = { y1.x, y1.y, y1.z, y1.w };
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (src.mode == 2) {
                        a[u + k] = f_f32(xa[k], xb[k]);
                        b[u + k] = f_f32(ya[k], yb[k]);
                    } else {
                        a[u + k] = g_f32(xa[k], xb[k], (src.cache_w[0] >> ((j + u + k) & 31)) & 1u);
                        b[u + k] = g_f32(ya[k], yb[k], (src.cache_w[1] >> ((h + j + u + k) & 31)) & 1u);
                    }
                }
            }
        } else if (src.mode == 0 && src.v.quad) {
            // a stored arena stage in the quad layout: one 16-byte load per four elements
#pragma unroll
            for (int u = 0; u < UNR; u += 4) {
                const float4 va = src.v.ld4(j + u), vb = src.v.ld4(h + j + u);
                a[u] = va.x, a[u + 1] = va.y, a[u + 2] = va.z, a[u + 3] = va.w;
                b[u] = vb.x, b[u + 1] = vb.y, b[u + 2] = vb.z, b[u + 3] = vb.w;
            }
        } else {
            src.bits_for(j, 0);
            src.bits_for(h + j, 1);
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                a[u] = src.get(j + u, 0);
                b[u] = src.get(h + j + u, 1);
            }
        }
        if (IS_G) {
            if ((j & 31) == 0) bw = c.ld(lb + (j >> 5), ol);
#pragma unroll
            for (int u = 0; u < UNR; ++u) r[u] = g_f32(a[u], b[u], (bw >> ((j & 31) + u)) & 1u);
        } else {
#pragma unroll
            for (int u = 0; u < UNR; ++u) r[u] = f_f32(a[u], b[u]);
        }
        if (active) {
            if (dst.quad) {
#pragma unroll
                for (int u = 0; u < UNR; u += 4) dst.st4(j + u, make_float4(r[u], r[u + 1], r[u + 2], r[u + 3]));
            } else {
#pragma unroll
                for (int u = 0; u < UNR; ++u) dst.set(j + u, r[u]);
            }
        }
    }
}

// The same step while ONE path is alive (the frozen prefix of the tree: G - 1 lanes of a group would only watch): all G
// lanes share path 0's step, eight elements per lane and round, reading through lane 0's tables and writing into lane
// 0's column. Element-wise operations only, so the results are the same values.
template <bool IS_G, class FLane>
__device__ __forceinline__ void sclf_fg_single(FLane& c, const SclFParams& p, int s, const float* row, int mode)
{
    const int h = 1 << (s - 1), G = c.G, half = 1 << (p.S - 1);
    const uint32_t o = __shfl_sync(0xFFFFFFFFu, c.own_llr.get(s), 0, G), ol = __shfl_sync(0xFFFFFFFFu, c.own_L.get(s - 1), 0, G);
    const int lb = p.bitL_off[s - 1];
    const typename FLane::View src = c.view(mode == 0 ? s : s - 1, mode == 0 ? p.llr_off[s] : p.llr_off[s - 1], o);
    const typename FLane::View dst = c.view(s - 1, p.llr_off[s - 1], 0u);
    for (int j0 = 0; j0 < h; j0 += 8 * G) {
        const int j = j0 + 8 * c.l;
        uint32_t bword = 0;
        if (IS_G) { // the decision words of this round, fetched by all lanes (tensor-memory loads are warp-collective)
            const int nq = (h - j0 + 31) >> 5 < ((8 * G + 31) >> 5) ? (h - j0 + 31) >> 5 : ((8 * G + 31) >> 5);
            for (int q = 0; q < nq; ++q) {
                const uint32_t w = c.ld(lb + (j0 >> 5) + q, ol);
                if (q == ((j - j0) >> 5)) bword = w;
            }
        }
        if (j < h) {
            float a[8], b[8], r[8];
            if (mode == 0 && src.quad) {
#pragma unroll
                for (int u = 0; u < 8; u += 4) {
                    const float4 va = src.ld4(j + u), vb = src.ld4(h + j + u);
                    a[u] = va.x, a[u + 1] = va.y, a[u + 2] = va.z, a[u + 3] = va.w;
                    b[u] = vb.x, b[u + 1] = vb.y, b[u + 2] = vb.z, b[u + 3] = vb.w;
                }
            } else
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (mode == 0) {
                    a[u] = src[j + u];
                    b[u] = src[h + j + u];
                } else if (mode == 1) {
                    a[u] = __ldg(row + j + u);
                    b[u] = __ldg(row + h + j + u);
                } else { // virtual stage S-1 before the root's g step
                    a[u] = f_f32(__ldg(row + j + u), __ldg(row + j + u + half));
                    b[u] = f_f32(__ldg(row + h + j + u), __ldg(row + h + j + u + half));
                }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) r[u] = IS_G ? g_f32(a[u], b[u], (bword >> ((j & 31) + u)) & 1u) : f_f32(a[u], b[u]);
            if (dst.quad) {
                dst.st4(j, make_float4(r[0], r[1], r[2], r[3]));
                dst.st4(j + 4, make_float4(r[4], r[5], r[6], r[7]));
            } else {
#pragma unroll
                for (int u = 0; u < 8; ++u) dst.set(j + u, r[u]);
            }
        }
    }
    c.own_llr.set(s - 1, (uint32_t)c.l);
}

// f / g step of a node of size 2^s (RateRNode::decode scl_avx_float.cpp:238-262). Every lane runs the step (decision
// words may sit in tensor memory, whose loads are warp-collective); dead lanes only skip their stores.
template <bool IS_G, class FLane>
__device__ __forceinline__ void sclf_fg(FLane& c, const SclFParams& p, int s, const float* row)
{
    const bool active = c.l < c.P;
    const int h = 1 << (s - 1);
    const bool from_row = (s == p.S) && !p.root_leaf;
    if (from_row && p.virt_top) { // the root's f / g results are never stored: their consumers recompute them
        c.root_right = IS_G ? 1 : 0;
        c.own_llr.set(s - 1, (uint32_t)c.l);
        return;
    }
    const bool virt = p.virt_top && s == p.S - 1;
    if (c.P == 1 && c.G >= 2 && h >= 16 && !(virt && c.root_right)) {
        sclf_fg_single<IS_G>(c, p, s, row, from_row ? 1 : (virt ? 2 : 0));
        return;
    }
    const uint32_t o = c.own_llr.get(s), ol = c.own_L.get(s - 1);
    const int lb = p.bitL_off[s - 1];
    FgSrc<FLane> src{ c, (from_row || virt) ? c.view(s - 1, p.llr_off[s - 1], (uint32_t)c.l) : c.view(s, p.llr_off[s], o), row,
                      from_row ? 1 : (virt ? 2 + c.root_right : 0), 1 << (p.S - 1), p.bitL_off[p.S - 1], c.own_L.get(p.S - 1),
                      { 0u, 0u }, { -1, -1 } };
    const typename FLane::View dst = c.view(s - 1, p.llr_off[s - 1], (uint32_t)c.l);
    if (FLane::kTT && h >= 8 && src.mode == 0 && s <= c.smem_top) {
        // both stages in shared memory and the CTA size known at compile time: 32-bit shared addresses whose element
        // offsets are immediates (the generic path below spends a 64-bit multiply-add and a generic load per element:
        // a third of all executed instructions in ncu r02_scl_f32_v1)
        constexpr uint32_t ES = 4u * (uint32_t)FLane::kTT;
        const uint32_t sa = (uint32_t)__cvta_generic_to_shared(c.sm) + ES * (uint32_t)p.llr_off[s] + 4u * (uint32_t)((int)o - c.l);
        const uint32_t da = (uint32_t)__cvta_generic_to_shared(c.sm) + ES * (uint32_t)p.llr_off[s - 1];
        uint32_t bw = 0;
        for (int j = 0; j < h; j += 8) {
            float a[8], b[8], r[8];
            const uint32_t ja = sa + ES * (uint32_t)j, jb = ja + ES * (uint32_t)h;
#pragma unroll
            for (int u = 0; u < 8; ++u) a[u] = lds_f32(ja + ES * u), b[u] = lds_f32(jb + ES * u);
            if (IS_G) {
                if ((j & 31) == 0) bw = c.ld(lb + (j >> 5), ol);
#pragma unroll
                for (int u = 0; u < 8; ++u) r[u] = g_f32(a[u], b[u], (bw >> ((j & 31) + u)) & 1u);
            } else {
#pragma unroll
                for (int u = 0; u < 8; ++u) r[u] = f_f32(a[u], b[u]);
            }
            if (active) {
                const uint32_t jd = da + ES * (uint32_t)j;
#pragma unroll
                for (int u = 0; u < 8; ++u) sts_f32(jd + ES * u, r[u]);
            }
        }
    } else if (h >= 4) {
        // UNR elements per step, all loads first (source and destination may alias as far as the compiler knows, so
        // a rolled loop serialises load -> f/g -> store; the big stages live in the L2 arena and want many loads in
        // flight); the channel row as 16-byte loads
        if (h >= 16 && src.mode == 0 && src.v.quad)
            fg_block<IS_G, 16>(c, src, dst, active, h, lb, ol); // a stored arena stage: eight 16-byte loads in flight per lane
        else if (h >= 8)
            fg_block<IS_G, 8>(c, src, dst, active, h, lb, ol);
        else
            fg_block<IS_G, 4>(c, src, dst, active, h, lb, ol);
    } else {
        uint32_t bw = 0;
        if (IS_G) bw = c.ld(lb, ol);
        for (int j = 0; j < h; ++j) {
            src.bits_for(j, 0);
            src.bits_for(h + j, 1);
            const float a = src.get(j, 0), b = src.get(h + j, 1);
            const float r = IS_G ? g_f32(a, b, (bw >> (j & 31)) & 1u) : f_f32(a, b);
            if (active) dst.set(j, r);
        }
    }
    c.own_llr.set(s - 1, (uint32_t)c.l);
}

// the three sums of a leaf (own stage buffer), each in the reference's order: lane k of 8 accumulates elements
// k, k+8, ... starting from +0, pads of a short node are 0, then x[0]+x[1]+...+x[7] left to right
// (RateZeroDecoder :316-337, RepetitionDecoder :428-455, reduce_add_ps avxconvenience.h:256-272)
template <bool WANT_ALL, class FLane>
__device__ __forceinline__ void sclf_sums(const FLane& c, const SclFParams& p, int s, float& neg, float& pos, float& all)
{
    const int n = 1 << s;
    // (no lane sharing here even while one path is alive: the order of the float additions is part of the result)
    c.with_own_view(s, p.llr_off[s], [&](const auto& v) {
    float an[8], ap[8], aa[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) an[k] = ap[k] = aa[k] = 0.0f;
    for (int i = 0; i < n; i += 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float x = (i + k < n) ? v[i + k] : 0.0f;
            an[k] = fadd(an[k], fmin_ps(x, 0.0f));
            if (WANT_ALL) {
                ap[k] = fadd(ap[k], fmax_ps(x, 0.0f));
                aa[k] = fadd(aa[k], x);
            }
        }
    }
    neg = reduce_add8(an);
    pos = WANT_ALL ? reduce_add8(ap) : 0.0f;
    all = WANT_ALL ? reduce_add8(aa) : 0.0f;
    });
}

// RateZeroDecoder::decode scl_avx_float.cpp:316-337: metric += sum(min(llr, 0)), decisions 0, no fork, no re-sort
template <class FLane>
__device__ __forceinline__ void sclf_r0(FLane& c, const SclFParams& p, int s, int side)
{
    float neg, pos, all; // every lane runs (dead lanes: garbage into their own state, replaced at the next fork)
    sclf_sums<false>(c, p, s, neg, pos, all);
    c.metric = fadd(c.metric, neg);
    const int dst = side ? p.bitR_off[s] : p.bitL_off[s], nb = stage_bwords(s);
    for (int j = 0; j < nb; ++j) c.st(dst + j, 0u);
    set_side(c, s, side);
}

// findWeakLlrs (arrayfuncs.h:209-231) on the |LLR| floats stored in place in the own stage buffer (the reference
// works on a scratch copy mTemp; the LLRs of a leaf are dead once their signs are taken): `want` passes of a
// selection sort that takes the FIRST minimum (strict '<') of positions [i, n) and swaps it to position i.
template <class FLane>
__device__ __forceinline__ void sclf_find_weak(const FLane& c, int s, int off, int want, int (&widx)[4], float (&wval)[4])
{
    const int n = 1 << s;
    WeakIdx ix;
    ix.init();
    const int lim = n - 1 < want ? n - 1 : want;
    c.with_own_view(s, off, [&](const auto& v) {
    for (int i = 0; i < lim; ++i) {
        const float vi = v[i];
        float bv = vi;
        int bp = i;
#pragma unroll 4
        for (int j = i + 1; j < n; ++j) {
            const float x = v[j];
            if (x < bv) {
                bv = x;
                bp = j;
            }
        }
        if (bp != i) {
            v.set(i, bv);
            v.set(bp, vi);
        }
        ix.swap(i, bp);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        widx[k] = ix.I[k];
        wval[k] = (k < n) ? v[k] : INFINITY; // pads of the 8-float block hold +inf (:363-365,:517-519)
    }
    });
}

// first phase of RateOne / SPC leaves: decisions = sign bits into the own side buffer, |llr| in place; returns the
// parity of the signs
template <class FLane>
__device__ __forceinline__ uint32_t sclf_signs_and_abs(const FLane& c, const SclFParams& p, int s, int side)
{
    const int n = 1 << s, dst = side ? p.bitR_off[s] : p.bitL_off[s];
    uint32_t result = 0;
    c.with_own_view(s, p.llr_off[s], [&](const auto& v) {
        uint32_t par = 0, acc = 0;
        if (n >= 4) { // four loads in flight, then the in-place stores
            for (int j = 0; j < n; j += 4) {
                const float x0 = v[j], x1 = v[j + 1], x2 = v[j + 2], x3 = v[j + 3];
                const uint32_t s4 = fsign(x0) | (fsign(x1) << 1) | (fsign(x2) << 2) | (fsign(x3) << 3);
                acc |= s4 << (j & 31);
                v.set(j, fabs_(x0));
                v.set(j + 1, fabs_(x1));
                v.set(j + 2, fabs_(x2));
                v.set(j + 3, fabs_(x3));
                if ((j & 31) == 28 || j == n - 4) {
                    par ^= acc;
                    c.st(dst + (j >> 5), acc);
                    acc = 0;
                }
            }
            result = __popc(par) & 1u;
            return;
        }
        for (int j = 0; j < n; ++j) {
            const float x = v[j];
            const uint32_t sg = fsign(x);
            par ^= sg;
            acc |= sg << (j & 31);
            v.set(j, fabs_(x));
            if ((j & 31) == 31 || j == n - 1) {
                c.st(dst + (j >> 5), acc);
                acc = 0;
            }
        }
        result = par;
    });
    return result;
}

// RepetitionDecoder::decode scl_avx_float.cpp:428-481 (only built for n < 8)
template <class FLane>
__device__ __forceinline__ void sclf_rep(FLane& c, const SclFParams& p, int s, int side)
{
    float v[2] = { -INFINITY, -INFINITY };
    float neg, pos, all;
    sclf_sums<true>(c, p, s, neg, pos, all);
    if (c.l < c.P) {
        v[0] = fadd(c.metric, neg);
        v[1] = fadd(c.metric, -pos);
    }
    const int newP = 2 * c.P < p.L ? 2 * c.P : p.L;
    float val;
    int idx;
    scl_sort_key64<2>(c, v, newP, val, idx);
    const bool live = c.l < newP;
    const int sp = live ? idx >> 1 : 0;
    scl_fork(c, sp, val, newP);
    // decisions: +|sum| for candidate 0, -|sum| for candidate 1 -> the sign is the candidate's parity (all lanes store)
    const int dst = side ? p.bitR_off[s] : p.bitL_off[s], nb = stage_bwords(s);
    const uint32_t fill = (idx & 1) ? 0xFFFFFFFFu : 0u;
    for (int j = 0; j < nb; ++j) c.st(dst + j, fill);
    set_side(c, s, side);
}

// RateOneDecoder::decode scl_avx_float.cpp:353-413
template <class FLane>
__device__ __forceinline__ void sclf_r1(FLane& c, const SclFParams& p, int s, int side)
{
    float v[4] = { -INFINITY, -INFINITY, -INFINITY, -INFINITY };
    int wi[4] = { 0, 0, 0, 0 };
    float wv[4];
    sclf_signs_and_abs(c, p, s, side);
    sclf_find_weak(c, s, p.llr_off[s], 2, wi, wv);
    if (c.l < c.P) {
        v[0] = c.metric;
        v[1] = fadd(c.metric, -wv[0]);
        v[2] = fadd(c.metric, -wv[1]);
        v[3] = fadd(v[1], -wv[1]);
    }
    __syncwarp();
    const int newP = 4 * c.P < p.L ? 4 * c.P : p.L;
    float val;
    int idx;
    scl_sort_key64<4>(c, v, newP, val, idx);
    const bool live = c.l < newP;
    const int sp = live ? idx >> 2 : 0;
    int fi[4];
    fi[0] = __shfl_sync(0xFFFFFFFFu, wi[0], sp, c.G);
    fi[1] = __shfl_sync(0xFFFFFFFFu, wi[1], sp, c.G);
    fi[2] = 0;
    fi[3] = 0;
    scl_fork(c, sp, val, newP);
    scl_take_bits(c, p, s, side, sp, live, idx & 3, fi);
    set_side(c, s, side);
}

// SpcDecoder::decode scl_avx_float.cpp:498-621
template <class FLane>
__device__ __forceinline__ void sclf_spc(FLane& c, const SclFParams& p, int s, int side)
{
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = -INFINITY;
    int wi[4] = { 0, 0, 0, 0 };
    float wv[4];
    const int parity = (int)sclf_signs_and_abs(c, p, s, side);
    sclf_find_weak(c, s, p.llr_off[s], 4, wi, wv);
    if (c.l < c.P) {
        // odd parity: the weakest bit is flipped in every candidate's base and its cost is already in `m`
        const float m = parity ? fadd(c.metric, -wv[0]) : c.metric;
        const float w = parity ? 0.0f : wv[0];
        const float mw = fadd(m, -w);
        v[0] = m;
        v[1] = fadd(mw, -wv[1]);
        v[2] = fadd(mw, -wv[2]);
        v[3] = fadd(mw, -wv[3]);
        v[4] = fadd(fadd(m, -wv[1]), -wv[2]);
        v[5] = fadd(fadd(m, -wv[1]), -wv[3]);
        v[6] = fadd(fadd(m, -wv[2]), -wv[3]);
        v[7] = fadd(fadd(v[1], -wv[2]), -wv[3]);
    }
    __syncwarp();
    const int newP = 8 * c.P < p.L ? 8 * c.P : p.L;
    float val;
    int idx;
    scl_sort_key64<8>(c, v, newP, val, idx);
    const bool live = c.l < newP;
    const int sp = live ? idx >> 3 : 0;
    int fi[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) fi[q] = __shfl_sync(0xFFFFFFFFu, wi[q], sp, c.G);
    const int spar = __shfl_sync(0xFFFFFFFFu, parity, sp, c.G);
    // flip sets per candidate as 4-bit masks over (idx0..idx3), scl_avx_float.cpp:535-593
    const uint32_t tab = spar ? 0xEDB78421u : 0xFCA69530u;
    scl_fork(c, sp, val, newP);
    scl_take_bits(c, p, s, side, sp, live, (int)((tab >> (4 * (idx & 7))) & 15u), fi);
    set_side(c, s, side);
}

// GG: lanes per group as a compile-time constant (0 = run-time p.G), as in k_scl_i8
template <int BM, int TT, int GG = 0>
// Launch bounds: the tensor-memory plans run ONE CTA of TT (or, for an odd PC_SCLF_T, up to 384) lanes per SM and may use
// the registers that leaves; the shared-memory plans (short codes: CTAs of <= 128 lanes, up to eight per SM) are held to
// the 96-100 registers of five resident CTAs -- without the second bound they compiled to 138 registers, three CTAs per
// SM, and N = 64 L = 4 fell from 10.9 to 7.9 Gbit/s (profiles/r02_c15_bisect.txt).
__global__ void __launch_bounds__(TT ? TT : ((BM & BM_TMEM) ? 384 : 128), (TT || (BM & BM_TMEM)) ? 1 : 5)
    k_scl_f32(const __grid_constant__ SclFParams p)
{
    using FLane = FLaneT<BM, TT>;
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ uint32_t tm_base_s;
    const int T = blockDim.x, t = threadIdx.x, warp = t >> 5;
    uint32_t* crc_table = smem; // 256 words
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.crc_table[i];
    if (BM & BM_TMEM) {
        if (warp == 0) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             (uint32_t)__cvta_generic_to_shared(&tm_base_s)),
                         "r"((uint32_t)p.tm_alloc_cols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
    }
    __syncthreads();
    if (BM & BM_TMEM) asm volatile("tcgen05.fence::after_thread_sync;");
    FLane c;
    c.sm = smem + 256 + t;
    c.ar = p.arena + ((size_t)blockIdx.x * T + t);
    c.AS = p.arena_stride;
    c.gl3 = 3 * ((size_t)blockIdx.x * T + t);
    c.aq = p.arena_quad;
    c.tm = (BM & BM_TMEM) ? tm_base_s + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * p.tm_cols_thread) : 0u;
    c.T = T;
    const int G = GG ? GG : p.G;
    c.G = G;
    c.l = t & (G - 1);
    c.smem_top = p.smem_top;
    const int gpc = T / G; // codewords per CTA
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    const unsigned gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << ((t & 31) & ~(G - 1));
    const unsigned long long B = p.count ? *p.count : p.B;
    const unsigned long long nbatch = (B + gpc - 1) / gpc; // 0 when there is nothing to do: straight to the dealloc
    for (unsigned long long bi = blockIdx.x; bi < nbatch; bi += gridDim.x) {
        const unsigned long long fl = bi * gpc + (unsigned long long)(t / G);
        const bool live_frame = fl < B;
        unsigned long long f = live_frame ? fl : B - 1;
        if (p.frame_idx) f = p.frame_idx[f];
        const float* row = p.llr + f * N;
        c.P = 1;
        c.metric = 0.0f;
        c.root_right = 0;
        c.own_llr = Tbl{ { 0u, 0u, 0u, 0u } };
        c.own_L = c.own_llr;
        c.own_R = c.own_llr;
        if (p.root_leaf && c.l == 0) // stage the channel row (the root is a leaf), setFirstPath :103-109
            for (int j = 0; j < N; ++j) c.lst(p.S, p.llr_off[p.S], j, row[j]);
        __syncwarp();
        for (int pc = 0;; ++pc) {
            const uint32_t ins = __ldg(p.code + pc);
            const uint32_t op = ins & 63u;
            const int s = (int)((ins >> 6) & 31u), side = (int)((ins >> 11) & 1u);
            if (op == SCL_END) break;
            switch (op) {
            case SCL_F: sclf_fg<false>(c, p, s, row); break;
            case SCL_G: sclf_fg<true>(c, p, s, row); break;
            case SCL_COMB: scl_comb(c, p, s, side); break;
            case SCL_R0: sclf_r0(c, p, s, side); break;
            case SCL_R1: sclf_r1(c, p, s, side); break;
            case SCL_REP: sclf_rep(c, p, s, side); break;
            default: sclf_spc(c, p, s, side); break;
            }
            // other lanes' buffers are read by the next step; `lockstep` keeps the CTA's warps on the same program step
            // (the kernel is ~90 KB of SASS; free-running warps thrash the instruction caches, see pc_scl_i8.cu)
            c.st_fence();
            if ((GG && (BM & BM_TMEM)) || p.lockstep)
                __syncthreads();
            else
                __syncwarp();
        }
        // extractBestPath (scl_avx_float.cpp:706-748)
        const bool active = c.l < c.P;
        const int top = p.bitR_off[p.S], nbw = stage_bwords(p.S);
        uint32_t owner = c.own_R.get(p.S);
        int off = top;
        if (!p.systematic || (top & kTmem)) {
            // private copy of the path's decisions in the lane's shared-memory scratch (every lane: the reads may be
            // warp-collective); non-systematic: u = x F^(x)n on that copy
            const int scr = p.scr_off;
            const uint32_t lim = lowmask(1 << (p.S < 5 ? p.S : 5));
            for (int w = 0; w < nbw; ++w) {
                uint32_t x = c.ld(top + w, owner);
                if (!p.systematic) {
                    x ^= (x >> 1) & 0x55555555u & lim;
                    if (p.S > 1) x ^= (x >> 2) & 0x33333333u & lim;
                    if (p.S > 2) x ^= (x >> 4) & 0x0F0F0F0Fu & lim;
                    if (p.S > 3) x ^= (x >> 8) & 0x00FF00FFu & lim;
                    if (p.S > 4) x ^= (x >> 16) & 0x0000FFFFu;
                }
                c.st(scr + w, x);
            }
            if (!p.systematic)
                for (int d = 1; d < nbw; d <<= 1)
                    for (int g = 0; g < nbw; g += 2 * d)
                        for (int i = g; i < g + d; ++i)
                            c.st(scr + i, c.ld(scr + i, (uint32_t)c.l) ^ c.ld(scr + i + d, (uint32_t)c.l));
            owner = (uint32_t)c.l;
            off = scr;
        }
        bool okp = false;
        uint8_t* out_row = p.info + f * kb;
        const SclBitsReader<FLane> rd{ c, off, owner };
        if (active) okp = extract_info_rd(rd, p.runs, p.nruns, p.K, out_row, p.crc, crc_table, false);
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, active && okp) & gmask;
        const int first_lane = bal ? (__ffs(bal) - 1) : (__ffs(gmask) - 1); // absolute lane in the warp
        if ((t & 31) == first_lane && live_frame) {
            extract_info_rd(rd, p.runs, p.nruns, p.K, out_row, p.crc, crc_table, true);
            if (p.ok) p.ok[f] = bal ? 1 : 0;
        }
        __syncwarp();
    }
    if (BM & BM_TMEM) {
        asm volatile("tcgen05.fence::before_thread_sync;");
        __syncthreads();
        if (warp == 0)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm_base_s), "r"((uint32_t)p.tm_alloc_cols));
    }
}

cudaError_t launch_scl_f32(const SclFLaunch& L, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                           const unsigned* d_frame_idx, const unsigned long long* d_count, cudaStream_t st)
{
    SclFParams p = L.p;
    p.llr = d_llr;
    p.info = d_info;
    p.ok = d_ok;
    p.B = B;
    p.frame_idx = d_frame_idx;
    p.count = d_count;
    // decision words in the arena (long codes) or in tensor memory -> the instantiation whose ld / st test that flag
    const bool ba = (p.bitR_off[p.S] & kArena) != 0;
    void (*kern)(SclFParams) = ba ? k_scl_f32<BM_ARENA, 0> : k_scl_f32<BM_SMEM, 0>;
    if (p.tm_alloc_cols) {
        switch (L.threads) { // the CTA sizes the tensor-memory planner picks get their own instantiation
        case 512: kern = k_scl_f32<BM_TMEM, 512>; break;
        case 384: kern = k_scl_f32<BM_TMEM, 384>; break;
        case 256: kern = k_scl_f32<BM_TMEM, 256>; break;
        default: kern = k_scl_f32<BM_TMEM, 0>; break;
        }
        // the metric's L = 8 (5 ... 8) at the default CTA size: group size and lockstep fold at compile time
        if (L.threads == 512 && p.G == 8 && p.lockstep == 1 && !getenv("PC_SCL_GRT")) kern = k_scl_f32<BM_TMEM, 512, 8>;
        if (L.threads == 384 && p.lockstep == 1 && !getenv("PC_SCL_GRT")) {
            if (p.G == 8) kern = k_scl_f32<BM_TMEM, 384, 8>;
            if (p.G == 4) kern = k_scl_f32<BM_TMEM, 384, 4>;
            if (p.G == 32) kern = k_scl_f32<BM_TMEM, 384, 32>; // BASELINE config 4
        }
    } else if (!ba && p.G == 4 && !getenv("PC_SCL_GRT")) {
        kern = k_scl_f32<BM_SMEM, 0, 4>; // BASELINE config 1: N = 64, L = 4 ('mixed' second pass)
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
    // several CTAs per SM only fit when the L1/shared split is at its shared-memory maximum; a single CTA keeps the
    // default split (the rest stays L1 for the channel rows and the arena)
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 L.per_sm > 1 ? (int)cudaSharedmemCarveoutMaxShared : (int)cudaSharedmemCarveoutDefault);
    if (e != cudaSuccess) return e;
    static const bool debug_occ = getenv("PC_DEBUG_OCC") != nullptr;
    if (debug_occ) {
        int occ = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L.threads, L.smem_bytes);
        fprintf(stderr, "k_scl_f32: %d threads, %zu B smem, grid %d, tmem cols %d, virt_top %d, smem_top %d, arena elems %d, occupancy %d\n",
                L.threads, L.smem_bytes, L.grid, p.tm_alloc_cols, p.virt_top, p.smem_top, p.arena_elems, occ);
    }
    const int gpc = L.threads / p.G;
    size_t nb = (B + gpc - 1) / gpc;
    const int grid = (int)(nb < (size_t)L.grid ? nb : (size_t)L.grid);
    kern<<<grid, L.threads, L.smem_bytes, st>>>(p);
    return cudaGetLastError();
}

// host planner: which LLR stages fit shared memory, CTA size and CTAs per SM
static int plan_scl_f32_smem(int S, int K, int L, bool root_leaf, size_t smem_max, int sms, SclFLaunch& out, const char** msg)
{
    if (L < 1 || L > 32) {
        *msg = "list size must be in [1, 32]";
        return 1;
    }
    if (S > 17) {
        *msg = "block length too large for the SCL kernel";
        return 1;
    }
    int G = 1;
    while (G < L) G <<= 1;
    SclFParams& p = out.p;
    p.S = S;
    p.K = K;
    p.L = L;
    p.G = G;
    p.root_leaf = root_leaf ? 1 : 0;
    const int top_llr = root_leaf ? S : S - 1;
    const size_t sm_total = 233472 < smem_max + 1024 ? 233472 : smem_max + 1024; // per SM, 1 KiB reserved per CTA
    // candidates in order of preference: (threads, CTAs per SM); the first whose decision words fit wins, then as
    // many of the small LLR stages as fit go to shared memory
    static const int cand[][2] = { { 128, 8 }, { 128, 6 }, { 128, 4 }, { 128, 3 }, { 128, 2 }, { 128, 1 }, { 64, 1 }, { 32, 1 } };
    int T = 0, per_sm = 0, smem_top = -1, bit_top = S;
    int force_t = -1, force_n = -1; // experiment overrides (tools/sweep_scl.sh)
    if (const char* e = getenv("PC_SCLF_THREADS")) force_t = atoi(e);
    if (const char* e = getenv("PC_SCLF_CTAS")) force_n = atoi(e);
    // decision words of the stages > bit_top go to the arena: only when nothing fits otherwise (long codes)
    const int bit_lo = S < 5 ? S : 5;
    for (bit_top = S; bit_top >= bit_lo && !T; --bit_top) {
        int bw_smem = 0;
        for (int s = 0; s <= S; ++s)
            if (s <= bit_top) bw_smem += 2 * (s >= 5 ? (1 << (s - 5)) : 1);
        for (int pass = 0; pass < 2 && !T; ++pass)
        for (size_t ci = 0; ci < sizeof(cand) / sizeof(cand[0]); ++ci) {
            const int t = cand[ci][0], n = cand[ci][1];
            if (t < G) continue;
            if ((force_t > 0 && t != force_t) || (force_n > 0 && n != force_n)) continue;
            if (bit_top < S && bit_top > 5 && t * n < 128) continue; // rather move another stage out than run 32 lanes
            size_t per_cta = sm_total / n - 1024;
            if (per_cta > smem_max) per_cta = smem_max;
            if (per_cta < 1024) continue;
            const long budget = (long)((per_cta - 1024) / (4 * (size_t)t)); // words per lane
            if (budget < bw_smem) continue;
            int w = bw_smem, top = -1;
            for (int s = 0; s <= top_llr && w + (1 << s) <= budget; ++s) {
                w += 1 << s;
                top = s;
            }
            // first pass: the whole tree on chip with as many CTAs per SM as fit (at least 256 lanes); second pass: the most lanes whose
            // decision words fit (measured, N=1024 L=8: 3 x 128 lanes 1.48 Gbit/s, 2 x 128 1.40, 1 x 128 0.82; the large
            // stages then live in the arena). More than 3 CTAs would not be resident anyway (96 registers per thread).
            if (force_n <= 0 && pass == 0 && (top < top_llr || t * n < 256)) continue;
            if (force_n <= 0 && pass == 1 && n > 3) continue;
            T = t;
            per_sm = n;
            smem_top = top;
            break;
        }
        if (T) break;
    }
    if (!T) {
        *msg = "SCL path state does not fit shared memory (block length too large for this list size)";
        return 1;
    }
    int w = 0, e = 0;
    for (int s = 0; s < 20; ++s) p.llr_off[s] = p.bitL_off[s] = p.bitR_off[s] = 0;
    for (int s = 0; s <= top_llr; ++s) {
        if (s <= smem_top) {
            p.llr_off[s] = w;
            w += 1 << s;
        } else {
            p.llr_off[s] = e;
            e += 1 << s;
        }
    }
    for (int s = 0; s <= S; ++s) {
        const int bw = s >= 5 ? (1 << (s - 5)) : 1;
        if (s <= bit_top) {
            p.bitL_off[s] = w;
            w += bw;
            p.bitR_off[s] = w;
            w += bw;
        } else { // 4-byte arena elements hold floats or decision words alike
            p.bitL_off[s] = kArena | e;
            e += bw;
            p.bitR_off[s] = kArena | e;
            e += bw;
        }
    }
    p.words = w;
    p.smem_top = smem_top;
    p.arena_elems = e;
    p.arena_quad = (smem_top >= 1 && !getenv("PC_SCLF_NOQUAD")) ? 1 : 0; // every arena LLR stage then holds whole quads at a quad offset
    p.tm_alloc_cols = 0;
    p.tm_cols_thread = 0;
    p.scr_off = p.bitL_off[S]; // non-systematic extraction works on the (free) left buffer of the top stage
    p.lockstep = getenv("PC_SCL_LOCKSTEP") ? atoi(getenv("PC_SCL_LOCKSTEP")) != 0 : 0;
    p.virt_top = 0;
    out.threads = T;
    out.smem_bytes = 1024 + 4 * (size_t)w * T;
    out.grid = sms * per_sm;
    out.per_sm = per_sm;
    // keep the arena within ~1 GiB
    while (out.grid > sms && (size_t)4 * e * out.grid * T > ((size_t)1 << 30)) out.grid -= sms;
    p.arena_stride = (size_t)out.grid * T;
    return 0;
}

// Planner, first choice: ONE CTA of T lanes per SM whose warps stay in step, all decision words in TENSOR MEMORY (which
// frees 136 shared-memory words per lane at N = 1024 for LLR stages), LLR stage S-1 virtual (recomputed from the channel
// row) and only the stage(s) that still do not fit in an arena small enough to stay in L2:
//   N = 1024, T = 384 (default): stages <= 6 in shared memory, stages 7 and 8 in an 87 MB arena (was: stages >= 4, 233 MB);
//   T = 256: the same stages, 58 MB arena; T = 224: stages <= 7 in shared memory, 34 MB arena.
// PC_SCLF_T picks T (multiple of 32 and of the group size), PC_SCLF_VIRT=0 stores stage S-1, PC_SCL_NOTMEM=1 or
// PC_SCLF_THREADS / PC_SCLF_CTAS fall back to the shared-memory planner above.
int plan_scl_f32(int S, int K, int L, bool root_leaf, bool virt_ok, size_t smem_max, int sms, SclFLaunch& out, const char** msg)
{
    if (L < 1 || L > 32) {
        *msg = "list size must be in [1, 32]";
        return 1;
    }
    const bool no_tmem = getenv("PC_SCL_NOTMEM") != nullptr || getenv("PC_SCLF_THREADS") != nullptr || getenv("PC_SCLF_CTAS") != nullptr;
    int G = 1;
    while (G < L) G <<= 1;
    // measured at N = 1024, L = 8 (profiles/r02_c4_perf.txt): T = 256 2.79 Gbit/s (58 MB arena, DRAM traffic 4.3x the
    // algorithmic bytes), 320 2.55, 384 3.32 (87 MB arena, 17x, still only 0.47 TB/s), 512 3.34 (stage 6 leaves shared memory)
    // with the quad arena layout and the group size at compile time (profiles/r02_c40_perf.txt): 384 lanes 4.35 Gbit/s, 512 lanes
    // 4.38 (N = 512 K = 128: 2.81 / 3.02); the 512-lane form has a G = 8 instantiation only, other list sizes keep 384
    int T = G == 8 ? 512 : 384;
    if (const char* e = getenv("PC_SCLF_T")) T = atoi(e);
    if (T < 32 || T > 512 || (T % 32) != 0 || (T % G) != 0) T = 384;
    if (T > 384 && T != 512) T = 384; // the run-time-size instantiation is bounded at 384 threads
    const bool lockstep = getenv("PC_SCL_LOCKSTEP") ? atoi(getenv("PC_SCL_LOCKSTEP")) != 0 : true;
    bool virt = virt_ok && !root_leaf && S >= 7;
    if (const char* e = getenv("PC_SCLF_VIRT")) virt = virt && atoi(e) != 0;
    const int top_llr = root_leaf ? S : S - 1;
    // (short codes keep their whole tree in shared memory with up to 1024 lanes per SM: the planner above)
    if (!no_tmem && S >= 9 && S <= 13) {
        const int tm_alloc = 512, tm_cols = tm_alloc / ((T / 32 + 3) / 4);
        size_t per_cta = 233472 - 1024;
        if (per_cta > smem_max) per_cta = smem_max;
        const long budget = ((long)per_cta - 1024 - 64) / (4L * T); // shared-memory words per lane
        SclFParams& p = out.p;
        for (int s = 0; s < 20; ++s) p.llr_off[s] = p.bitL_off[s] = p.bitR_off[s] = 0;
        int tcol = 0, tm_lo = S + 1;
        for (int s = S; s >= 0; --s) { // decision buffers, top stage first, into tensor memory while they fit
            const int bw = s >= 5 ? (1 << (s - 5)) : 1;
            if (tcol + 2 * bw > tm_cols) break;
            p.bitL_off[s] = kTmem | tcol;
            tcol += bw;
            p.bitR_off[s] = kTmem | tcol;
            tcol += bw;
            tm_lo = s;
        }
        int w = 0;
        for (int s = 0; s < tm_lo; ++s) {
            const int bw = s >= 5 ? (1 << (s - 5)) : 1;
            p.bitL_off[s] = w;
            w += bw;
            p.bitR_off[s] = w;
            w += bw;
        }
        int e = 0, smem_top = -1;
        bool open = true;
        for (int s = 0; s <= top_llr; ++s) {
            if (virt && s == S - 1) continue; // never stored
            if (open && w + (1 << s) <= budget) {
                p.llr_off[s] = w;
                w += 1 << s;
                smem_top = s;
            } else {
                open = false;
                p.llr_off[s] = e;
                e += 1 << s;
            }
        }
        // scratch of N/32 words for extractBestPath: LLR stage S-5 has exactly that size and is dead by then
        const int nbw = S >= 5 ? (1 << (S - 5)) : 1;
        int scr = -1;
        if (S - 5 >= 0 && S - 5 <= smem_top) {
            scr = p.llr_off[S - 5];
        } else if (w + nbw <= budget) {
            scr = w;
            w += nbw;
        }
        if (w <= budget && tm_lo <= S && scr >= 0 && smem_top >= (top_llr < 4 ? top_llr : 4)) {
            p.S = S;
            p.K = K;
            p.L = L;
            p.G = G;
            p.root_leaf = root_leaf ? 1 : 0;
            p.words = w;
            p.smem_top = smem_top;
            p.arena_elems = e;
            p.arena_quad = (smem_top >= 1 && !getenv("PC_SCLF_NOQUAD")) ? 1 : 0;
            p.tm_alloc_cols = tm_alloc;
            p.tm_cols_thread = tm_cols;
            p.scr_off = scr;
            p.lockstep = lockstep ? 1 : 0;
            p.virt_top = virt ? 1 : 0;
            out.threads = T;
            out.smem_bytes = 1024 + 4 * (size_t)w * T;
            out.grid = sms;
            out.per_sm = 1;
            p.arena_stride = (size_t)out.grid * T;
            return 0;
        }
    }
    return plan_scl_f32_smem(S, K, L, root_leaf, smem_max, sms, out, msg);
}

} // namespace pc

namespace pc {
// Adaptive decoder, between the passes: list of the frames whose first-pass verdict was false
// (AdaptiveFloat::decode adaptive_float.cpp:33-45: `if (!success && mListSize > 1)`), warp-aggregated append.
__global__ void __launch_bounds__(256) k_compact_failed(const uint8_t* __restrict__ ok, unsigned long long B,
                                                        unsigned* __restrict__ idx, unsigned long long* __restrict__ count)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < B; base += stride) {
        const unsigned long long i = base + threadIdx.x;
        const bool failed = i < B && ok[i] == 0;
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, failed);
        if (!bal) continue;
        const int lane = threadIdx.x & 31, leader = __ffs(bal) - 1;
        unsigned long long pos = 0;
        if (lane == leader) pos = atomicAdd(count, (unsigned long long)__popc(bal));
        pos = __shfl_sync(0xFFFFFFFFu, pos, leader);
        if (failed) idx[pos + __popc(bal & ((1u << lane) - 1u))] = (unsigned)i;
    }
}

cudaError_t launch_compact_failed(const uint8_t* d_ok, size_t B, unsigned* d_idx, unsigned long long* d_count, cudaStream_t st)
{
    cudaError_t e = cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    size_t g = (B + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    k_compact_failed<<<(int)g, 256, 0, st>>>(d_ok, B, d_idx, d_count);
    return cudaGetLastError();
}
} // namespace pc
