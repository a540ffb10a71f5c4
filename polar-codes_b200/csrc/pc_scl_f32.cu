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
#include <cstdlib>

#include "pc_fastssc_f32_core.h"
#include "pc_kernels.cuh"
#include "pc_scl_common.cuh"

namespace pc {

template <bool BA> // BA: some decision words live in the arena (long codes); off = no flag tests
struct FLaneT {
    uint32_t* sm;   // smem base + tid (word w at sm[w * T])
    float* ar;      // arena base + global lane id (element e at ar[e * AS])
    size_t AS;
    int T, G, l;    // threads per CTA, lanes per group, lane index inside the group
    int P;          // live paths (uniform)
    int smem_top;
    float metric;
    Tbl own_llr, own_L, own_R;
    // decision words: shared memory, or the arena for the largest stages of long codes (flag in the offset)
    __device__ __forceinline__ uint32_t ld(int word, uint32_t owner) const
    {
        if (BA && (word & kArena)) return reinterpret_cast<const uint32_t*>(ar)[(size_t)(word & (kArena - 1)) * AS + ((int)owner - l)];
        return sm[(size_t)word * T + ((int)owner - l)];
    }
    __device__ __forceinline__ void st(int word, uint32_t v) const
    {
        if (BA && (word & kArena))
            reinterpret_cast<uint32_t*>(ar)[(size_t)(word & (kArena - 1)) * AS] = v;
        else
            sm[(size_t)word * T] = v;
    }
    // LLR stage s (off = p.llr_off[s]) in lane `owner`'s space, resolved once per node step
    struct View {
        float* p;
        size_t stride;
        __device__ __forceinline__ float operator[](int i) const { return p[(size_t)i * stride]; }
        __device__ __forceinline__ void set(int i, float v) const { p[(size_t)i * stride] = v; }
    };
    __device__ __forceinline__ View view(int s, int off, uint32_t owner) const
    {
        const ptrdiff_t d = (int)owner - l;
        if (s <= smem_top) return View{ reinterpret_cast<float*>(sm) + (size_t)off * T + d, (size_t)T };
        return View{ ar + (size_t)off * AS + d, AS };
    }
    // LLR element i of stage s (off = p.llr_off[s]) in lane `owner`'s space
    __device__ __forceinline__ float lld(int s, int off, int i, uint32_t owner) const
    {
        const ptrdiff_t d = (int)owner - l;
        if (s <= smem_top) return __uint_as_float(sm[(size_t)(off + i) * T + d]);
        return ar[(size_t)(off + i) * AS + d];
    }
    __device__ __forceinline__ void lst(int s, int off, int i, float v) const
    {
        if (s <= smem_top)
            sm[(size_t)(off + i) * T] = __float_as_uint(v);
        else
            ar[(size_t)(off + i) * AS] = v;
    }
};

// _mm256_max_ps lane: returns the second operand unless a > b
__device__ __forceinline__ float fmax_ps(float a, float b) { return a > b ? a : b; }

template <bool IS_G, int UNR, class FLane>
__device__ __forceinline__ void fg_block(const FLane& c, const float* row, const typename FLane::View& src, const typename FLane::View& dst, bool from_row,
                                         int h, int lb, uint32_t ol)
{
    const bool row16 = from_row && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0);
    uint32_t bw = 0;
    for (int j = 0; j < h; j += UNR) {
        float a[UNR], b[UNR], r[UNR];
        if (row16) {
#pragma unroll
            for (int u = 0; u < UNR; u += 4) {
                const float4 va = __ldg(reinterpret_cast<const float4*>(row + j + u)), vb = __ldg(reinterpret_cast<const float4*>(row + h + j + u));
                a[u] = va.x, a[u + 1] = va.y, a[u + 2] = va.z, a[u + 3] = va.w;
                b[u] = vb.x, b[u + 1] = vb.y, b[u + 2] = vb.z, b[u + 3] = vb.w;
            }
        } else {
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                a[u] = from_row ? __ldg(row + j + u) : src[j + u];
                b[u] = from_row ? __ldg(row + h + j + u) : src[h + j + u];
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
#pragma unroll
        for (int u = 0; u < UNR; ++u) dst.set(j + u, r[u]);
    }
}

// f / g step of a node of size 2^s (RateRNode::decode scl_avx_float.cpp:238-262)
template <bool IS_G, class FLane>
__device__ __forceinline__ void sclf_fg(FLane& c, const SclFParams& p, int s, const float* row)
{
    const bool active = c.l < c.P;
    const int h = 1 << (s - 1);
    const bool from_row = (s == p.S) && !p.root_leaf;
    const uint32_t o = c.own_llr.get(s), ol = c.own_L.get(s - 1);
    const int lb = p.bitL_off[s - 1];
    const typename FLane::View src = from_row ? c.view(s - 1, p.llr_off[s - 1], (uint32_t)c.l) : c.view(s, p.llr_off[s], o);
    const typename FLane::View dst = c.view(s - 1, p.llr_off[s - 1], (uint32_t)c.l);
    if (active) {
        uint32_t bw = 0;
        if (h >= 4) {
            // UNR elements per step, all loads first (source and destination may alias as far as the compiler knows, so
            // a rolled loop serialises load -> f/g -> store; the big stages live in the L2 arena and want many loads in
            // flight); the channel row as 16-byte loads
            if (h >= 8)
                fg_block<IS_G, 8>(c, row, src, dst, from_row, h, lb, ol);
            else
                fg_block<IS_G, 4>(c, row, src, dst, from_row, h, lb, ol);
        } else
        for (int j = 0; j < h; ++j) {
            const float a = from_row ? __ldg(row + j) : src[j];
            const float b = from_row ? __ldg(row + h + j) : src[h + j];
            float r;
            if (IS_G) {
                if ((j & 31) == 0) bw = c.ld(lb + (j >> 5), ol);
                r = g_f32(a, b, (bw >> (j & 31)) & 1u);
            } else {
                r = f_f32(a, b);
            }
            dst.set(j, r);
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
    const typename FLane::View v = c.view(s, p.llr_off[s], (uint32_t)c.l);
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
}

// RateZeroDecoder::decode scl_avx_float.cpp:316-337: metric += sum(min(llr, 0)), decisions 0, no fork, no re-sort
template <class FLane>
__device__ __forceinline__ void sclf_r0(FLane& c, const SclFParams& p, int s, int side)
{
    if (c.l < c.P) {
        float neg, pos, all;
        sclf_sums<false>(c, p, s, neg, pos, all);
        c.metric = fadd(c.metric, neg);
        const int dst = side ? p.bitR_off[s] : p.bitL_off[s], nb = stage_bwords(s);
        for (int j = 0; j < nb; ++j) c.st(dst + j, 0u);
    }
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
    const typename FLane::View v = c.view(s, off, (uint32_t)c.l);
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
}

// first phase of RateOne / SPC leaves: decisions = sign bits into the own side buffer, |llr| in place; returns the
// parity of the signs
template <class FLane>
__device__ __forceinline__ uint32_t sclf_signs_and_abs(const FLane& c, const SclFParams& p, int s, int side)
{
    const int n = 1 << s, dst = side ? p.bitR_off[s] : p.bitL_off[s];
    const typename FLane::View v = c.view(s, p.llr_off[s], (uint32_t)c.l);
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
        return __popc(par) & 1u;
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
    return par;
}

// RepetitionDecoder::decode scl_avx_float.cpp:428-481 (only built for n < 8)
template <class FLane>
__device__ __forceinline__ void sclf_rep(FLane& c, const SclFParams& p, int s, int side)
{
    float v[2] = { -INFINITY, -INFINITY };
    if (c.l < c.P) {
        float neg, pos, all;
        sclf_sums<true>(c, p, s, neg, pos, all);
        v[0] = fadd(c.metric, neg);
        v[1] = fadd(c.metric, -pos);
    }
    const int newP = 2 * c.P < p.L ? 2 * c.P : p.L;
    float val;
    int idx;
    scl_sort<2>(c, v, newP, val, idx);
    const bool live = c.l < newP;
    const int sp = live ? idx >> 1 : 0;
    scl_fork(c, sp, val, newP);
    if (live) {
        // decisions: +|sum| for candidate 0, -|sum| for candidate 1 -> the sign is the candidate's parity
        const int dst = side ? p.bitR_off[s] : p.bitL_off[s], nb = stage_bwords(s);
        const uint32_t fill = (idx & 1) ? 0xFFFFFFFFu : 0u;
        for (int j = 0; j < nb; ++j) c.st(dst + j, fill);
    }
    set_side(c, s, side);
}

// RateOneDecoder::decode scl_avx_float.cpp:353-413
template <class FLane>
__device__ __forceinline__ void sclf_r1(FLane& c, const SclFParams& p, int s, int side)
{
    float v[4] = { -INFINITY, -INFINITY, -INFINITY, -INFINITY };
    int wi[4] = { 0, 0, 0, 0 };
    float wv[4];
    if (c.l < c.P) {
        sclf_signs_and_abs(c, p, s, side);
        sclf_find_weak(c, s, p.llr_off[s], 2, wi, wv);
        v[0] = c.metric;
        v[1] = fadd(c.metric, -wv[0]);
        v[2] = fadd(c.metric, -wv[1]);
        v[3] = fadd(v[1], -wv[1]);
    }
    __syncwarp();
    const int newP = 4 * c.P < p.L ? 4 * c.P : p.L;
    float val;
    int idx;
    scl_sort<4>(c, v, newP, val, idx);
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
    int parity = 0;
    if (c.l < c.P) {
        parity = (int)sclf_signs_and_abs(c, p, s, side);
        sclf_find_weak(c, s, p.llr_off[s], 4, wi, wv);
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
    scl_sort<8>(c, v, newP, val, idx);
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

template <bool BA>
__global__ void __launch_bounds__(128) k_scl_f32(const __grid_constant__ SclFParams p)
{
    using FLane = FLaneT<BA>;
    extern __shared__ __align__(16) uint32_t smem[];
    const int T = blockDim.x, t = threadIdx.x;
    uint32_t* crc_table = smem; // 256 words
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.crc_table[i];
    __syncthreads();
    FLane c;
    c.sm = smem + 256 + t;
    c.ar = p.arena + ((size_t)blockIdx.x * T + t);
    c.AS = p.arena_stride;
    c.T = T;
    c.G = p.G;
    c.l = t & (p.G - 1);
    c.smem_top = p.smem_top;
    const int gpc = T / p.G; // codewords per CTA
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    const unsigned gmask = (p.G == 32 ? 0xFFFFFFFFu : ((1u << p.G) - 1u)) << ((t & 31) & ~(p.G - 1));
    const unsigned long long B = p.count ? *p.count : p.B;
    if (B == 0) return;
    const unsigned long long nbatch = (B + gpc - 1) / gpc;
    for (unsigned long long bi = blockIdx.x; bi < nbatch; bi += gridDim.x) {
        const unsigned long long fl = bi * gpc + (unsigned long long)(t / p.G);
        const bool live_frame = fl < B;
        unsigned long long f = live_frame ? fl : B - 1;
        if (p.frame_idx) f = p.frame_idx[f];
        const float* row = p.llr + f * N;
        c.P = 1;
        c.metric = 0.0f;
        c.own_llr = Tbl{ { 0u, 0u, 0u } };
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
            __syncwarp(); // other lanes' buffers are read by the next step
        }
        // extractBestPath (scl_avx_float.cpp:706-748)
        const bool active = c.l < c.P;
        const int top = p.bitR_off[p.S], nbw = stage_bwords(p.S);
        uint32_t owner = c.own_R.get(p.S);
        int off = top;
        if (!p.systematic) {
            // u = x F^(x)n on a private copy (the left buffer of the top stage is free at this point)
            const int scr = p.bitL_off[p.S];
            if (active) {
                const uint32_t lim = lowmask(1 << (p.S < 5 ? p.S : 5));
                for (int w = 0; w < nbw; ++w) {
                    uint32_t x = c.ld(top + w, owner);
                    x ^= (x >> 1) & 0x55555555u & lim;
                    if (p.S > 1) x ^= (x >> 2) & 0x33333333u & lim;
                    if (p.S > 2) x ^= (x >> 4) & 0x0F0F0F0Fu & lim;
                    if (p.S > 3) x ^= (x >> 8) & 0x00FF00FFu & lim;
                    if (p.S > 4) x ^= (x >> 16) & 0x0000FFFFu;
                    c.st(scr + w, x);
                }
                for (int d = 1; d < nbw; d <<= 1)
                    for (int g = 0; g < nbw; g += 2 * d)
                        for (int i = g; i < g + d; ++i)
                            c.st(scr + i, c.ld(scr + i, (uint32_t)c.l) ^ c.ld(scr + i + d, (uint32_t)c.l));
            }
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
    // decision words in the arena (long codes) -> the instantiation whose ld / st test the offset flag
    const bool ba = (p.bitR_off[p.S] & kArena) != 0;
    void (*kern)(SclFParams) = ba ? k_scl_f32<true> : k_scl_f32<false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
    // several CTAs per SM only fit when the L1/shared split is at its shared-memory maximum; a single CTA keeps the
    // default split (the rest stays L1 for the channel rows and the arena)
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 L.per_sm > 1 ? (int)cudaSharedmemCarveoutMaxShared : (int)cudaSharedmemCarveoutDefault);
    if (e != cudaSuccess) return e;
    const int gpc = L.threads / p.G;
    size_t nb = (B + gpc - 1) / gpc;
    const int grid = (int)(nb < (size_t)L.grid ? nb : (size_t)L.grid);
    kern<<<grid, L.threads, L.smem_bytes, st>>>(p);
    return cudaGetLastError();
}

// host planner: which LLR stages fit shared memory, CTA size and CTAs per SM
int plan_scl_f32(int S, int K, int L, bool root_leaf, size_t smem_max, int sms, SclFLaunch& out, const char** msg)
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
    out.threads = T;
    out.smem_bytes = 1024 + 4 * (size_t)w * T;
    out.grid = sms * per_sm;
    out.per_sm = per_sm;
    // keep the arena within ~1 GiB
    while (out.grid > sms && (size_t)4 * e * out.grid * T > ((size_t)1 << 30)) out.grid -= sms;
    p.arena_stride = (size_t)out.grid * T;
    return 0;
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
