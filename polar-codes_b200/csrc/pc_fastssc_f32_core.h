// fp32 Fast-SSC decoder core: one THREAD decodes one codeword by interpreting a node program. Semantics follow
// FastSscAvxFloat (reference src/polarcode/decoding/fastssc_avx_float.cpp, include/polarcode/decoding/avx_float.h):
// the reference works on 8-lane AVX vectors, and every place where the lane structure decides a result -- the order
// of float additions, which lane wins a tie -- is restated on explicit 8-element arrays here, so the same IEEE
// operations happen in the same order (compile with -fmad=false / no contraction).
// Hard decisions are kept as one bit per element: the reference's "bit floats" are only ever consumed through their
// sign bit (g's sign select avx_float.h:68-80, Combine's xor :193-199, FloatContainer::getPackedInformationBits).
//
// __host__ __device__ throughout: tests/host_emul.cpp runs this exact code on the CPU against the reference.
#ifndef PC_FASTSSC_F32_CORE_H
#define PC_FASTSSC_F32_CORE_H
#include <math.h>
#include <string.h>

#include "pc_fastssc_i8_core.h" // lowmask, msb_first_word, extract_info_rd

namespace pc {

// node program of the float decoder: op[5:0] | s[10:6] | off[31:11]  (pc_program.h mk_instr)
enum FOp : uint32_t {
    FOP_F = 0, FOP_G = 1, FOP_G0 = 2, FOP_G1 = 3, FOP_COMB = 4, FOP_COMB0 = 5,
    FOP_R0 = 6, FOP_R1 = 7, FOP_REP = 8, FOP_SPC = 9, FOP_DREP = 10, FOP_DSPC = 11, FOP_DSPC8 = 12, FOP_ZSPC = 13,
    FOP_ZSPC8 = 14, FOP_TREP = 15, FOP_T5 = 16, FOP_REPONE8 = 17, FOP_END = 18, FOP_COUNT
};

PC_HD uint32_t f2u(float x)
{
#ifdef __CUDA_ARCH__
    return __float_as_uint(x);
#else
    uint32_t u;
    memcpy(&u, &x, 4);
    return u;
#endif
}
PC_HD float u2f(uint32_t u)
{
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float x;
    memcpy(&x, &u, 4);
    return x;
#endif
}
PC_HD float fadd(float a, float b)
{
#ifdef __CUDA_ARCH__
    return __fadd_rn(a, b); // never contracted into an FMA
#else
    return a + b;
#endif
}
PC_HD uint32_t fsign(float x) { return f2u(x) >> 31; }
PC_HD float fabs_(float x) { return u2f(f2u(x) & 0x7FFFFFFFu); }
// _mm256_min_ps lane: returns the second operand unless a < b (so NaN -> b)
PC_HD float fmin_ps(float a, float b) { return a < b ? a : b; }
// _mm256_polarf_ps / _mm256_polarg_ps (avx_float.h:55-81)
// (on the device the minimum of the two magnitudes is one FMNMX with |.| operand modifiers; fminf and fmin_ps differ
// only for NaN, which a magnitude of a finite or infinite LLR is not)
#ifdef __CUDA_ARCH__
PC_HD float f_f32(float a, float b) { return u2f(((f2u(a) ^ f2u(b)) & 0x80000000u) | f2u(fminf(fabsf(a), fabsf(b)))); }
#else
PC_HD float f_f32(float a, float b) { return u2f(((f2u(a) ^ f2u(b)) & 0x80000000u) | f2u(fmin_ps(fabs_(a), fabs_(b)))); }
#endif
PC_HD float g_f32(float a, float b, uint32_t bit) { return fadd(u2f(f2u(a) ^ (bit << 31)), b); }

// strided view of a float array (thread-interleaved on the device, stride 1 on the host and for the channel row)
// A view type V offers operator[](i), ld8(i, a) = eight consecutive elements (issued, not yet usable), wait() = the ld8
// results are usable, and kBatch = element pairs an f / g loop keeps in flight (a multiple of 8).
struct FView {
    static constexpr int kBatch = 8;
    const float* p;
    size_t stride;
    PC_HD float operator[](int i) const { return p[(size_t)i * stride]; }
    PC_HD void ld8(int i, float* a) const
    {
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
        for (int k = 0; k < 8; ++k) a[k] = p[(size_t)(i + k) * stride];
    }
    PC_HD void wait() const {}
};

struct FCtx {
    static constexpr bool kSoft = false; // true: the context also keeps the reference's float bit container (FSoftCtx)
    int S;
    const int* llr_off; // element offset of stage s inside the thread's LLR space (s < S, or s <= S when staged)
    float* llr;         // element 0 of the thread's LLR space
    size_t ls;          // element stride
    uint32_t* bits;     // hard-decision words, element i -> word i/32 bit i%32
    size_t bs;
    PC_HD FView stage(int s) const { return FView{ llr + (size_t)llr_off[s] * ls, ls }; }
    PC_HD void st(int s, int i, float v) const { llr[((size_t)llr_off[s] + i) * ls] = v; }
    PC_HD void st8(int s, int i, const float* v) const
    {
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
        for (int k = 0; k < 8; ++k) llr[((size_t)llr_off[s] + i + k) * ls] = v[k];
    }
    PC_HD uint32_t bld(int w) const { return bits[(size_t)w * bs]; }
    PC_HD void bst(int w, uint32_t v) const { bits[(size_t)w * bs] = v; }
    PC_HD uint32_t bit(int i) const { return (bld(i >> 5) >> (i & 31)) & 1u; }
    // writes n decisions (n <= 32 or a multiple of 32; off is a multiple of min(n, 32)) given by a generator
    template <class Gen>
    PC_HD void put_bits(int off, int n, const Gen& gen) const
    {
        if (n >= 32) {
            for (int w = 0; w < (n >> 5); ++w) {
                uint32_t v = 0;
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
    // a leaf's result as the FLOAT PATTERNS the reference stores in its bit container (the decision is the sign bit);
    // the hard-decision contexts keep the sign only
    template <class Gen>
    PC_HD void put_vals(int off, int n, const Gen& gen) const
    {
        put_bits(off, n, [&gen](int i) { return gen(i) >> 31; });
    }
};

// ---- soft-output context: the reference's float bit container next to the packed decisions (k_fastssc_f32_soft) ---
struct FSoftCtx : FCtx {
    static constexpr bool kSoft = true;
    uint32_t* soft; // element i of the thread's float bit container at soft[i * ss]
    size_t ss;
    template <class Gen>
    PC_HD void put_vals(int off, int n, const Gen& gen) const
    {
        for (int i0 = 0; i0 < n; i0 += 32) {
            const int m = n - i0 < 32 ? n - i0 : 32;
            uint32_t v = 0;
            for (int k = 0; k < m; ++k) {
                const uint32_t x = gen(i0 + k);
                soft[(size_t)(off + i0 + k) * ss] = x;
                v |= (x >> 31) << k;
            }
            if (m == 32) {
                bst((off + i0) >> 5, v);
            } else {
                const int sh = off & 31;
                const uint32_t msk = lowmask(m) << sh;
                bst(off >> 5, (bld(off >> 5) & ~msk) | (v << sh));
            }
        }
    }
    // ROneNode::rightDecode :204-219: right = the g value, left ^= right (patterns)
    PC_HD void soft_g1(int off, int h, int i, uint32_t gv) const
    {
        soft[(size_t)(off + h + i) * ss] = gv;
        soft[(size_t)(off + i) * ss] ^= gv;
    }
    // Combine / CombineBitsShort / Combine_0R
    PC_HD void soft_comb(int off, int h, bool zero_left) const
    {
        for (int i = 0; i < h; ++i) {
            const uint32_t r = soft[(size_t)(off + h + i) * ss];
            uint32_t* l = soft + (size_t)(off + i) * ss;
            *l = zero_left ? r : (*l ^ r);
        }
    }
};

// ---- 8-lane helpers ------------------------------------------------------------------------------------------
// lane-wise accumulation of the n/8 vectors of a node, starting from +0 (RepetitionDecoder fastssc_avx_float.cpp:273-281;
// lanes >= n of a short node are zeroed first, RepetitionPrepare avx_float.h:238-243)
template <class V>
PC_HD void accumulate8(const V& in, int n, float (&s)[8])
{
    for (int k = 0; k < 8; ++k) s[k] = 0.0f;
    for (int i = 0; i < n; i += 8)
        for (int k = 0; k < 8; ++k) s[k] = fadd(s[k], (i + k < n) ? in[i + k] : 0.0f);
}
// reduce_add_ps (avxconvenience.h:256-272): x[0] + x[1] + ... + x[7], left to right
PC_HD float reduce_add8(const float (&s)[8])
{
    float r = s[0];
    for (int k = 1; k < 8; ++k) r = fadd(r, s[k]);
    return r;
}
// _mm256_minidx_ps (avxconvenience.h:371-388): minimum of 8 and the FIRST lane holding it
PC_HD int minidx8(const float (&a)[8], float& mn)
{
    float m = a[0];
    for (int k = 1; k < 8; ++k) m = fmin_ps(a[k], m); // min tree; the value of a min is order independent (no NaN here)
    int idx = 0;
    for (int k = 7; k >= 0; --k)
        if (a[k] == m) idx = k;
    mn = a[idx];
    return idx;
}
// _mm256_spc_right4_ps (avx_float.h:294-308) on 4 values: parity over the 4 signs; EVERY lane whose magnitude equals
// the minimum is flipped (ties flip several). Returns the 4 stored values.
PC_HD void spc4_vals(const float (&y)[4], uint32_t (&out)[4])
{
    float m = fabs_(y[0]);
    uint32_t par = 0;
    for (int k = 0; k < 4; ++k) {
        m = fmin_ps(fabs_(y[k]), m);
        par ^= fsign(y[k]);
    }
    for (int k = 0; k < 4; ++k) out[k] = f2u(y[k]) ^ ((par && fabs_(y[k]) == m) ? 0x80000000u : 0u);
}

// ---- internal nodes --------------------------------------------------------------------------------------------
// RateRNode / ShortRateRNode / ROneNode / ZeroRNode (fastssc_avx_float.cpp:148-238): element-wise, lane structure
// does not matter
// Eight elements per step with all loads before the stores: `in` and the destination stage may alias as far as the
// compiler can tell, so a rolled loop serialises load -> op -> store (one memory latency per element on the device,
// where the stages live in a DRAM/L2 arena).
template <class C, class V, class Op>
PC_HD void fop_map(const C& c, const V& in, int s, const Op& op)
{
    const int h = 1 << (s - 1);
    constexpr int U = V::kBatch;
    if (U > 8 && h >= U) {
        for (int i = 0; i < h; i += U) {
            float a[U], b[U];
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < U; k += 8) in.ld8(i + k, a + k), in.ld8(i + k + h, b + k);
            in.wait();
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < U; ++k) a[k] = op(a[k], b[k], i + k);
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < U; k += 8) c.st8(s - 1, i + k, a + k);
        }
    } else if (h >= 8) {
        for (int i = 0; i < h; i += 8) {
            float a[8], b[8];
            in.ld8(i, a), in.ld8(i + h, b);
            in.wait();
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < 8; ++k) a[k] = op(a[k], b[k], i + k);
            c.st8(s - 1, i, a);
        }
    } else {
        for (int i = 0; i < h; ++i) c.st(s - 1, i, op(in[i], in[i + h], i));
    }
}
template <class C, class V>
PC_HD void fop_f(const C& c, const V& in, int s)
{
    fop_map(c, in, s, [](float a, float b, int) { return f_f32(a, b); });
}
template <class C, class V>
PC_HD void fop_g(const C& c, const V& in, int s, int off)
{
    const int h = 1 << (s - 1);
    constexpr int U = V::kBatch; // <= 32: the decisions of a batch sit in one word (off is a multiple of h)
    if (U > 8 && h >= U) {
        for (int i = 0; i < h; i += U) {
            float a[U], b[U];
            const uint32_t bw = c.bld((off + i) >> 5) >> ((off + i) & 31);
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < U; k += 8) in.ld8(i + k, a + k), in.ld8(i + k + h, b + k);
            in.wait();
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < U; ++k) a[k] = g_f32(a[k], b[k], (bw >> k) & 1u);
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < U; k += 8) c.st8(s - 1, i + k, a + k);
        }
    } else if (h >= 8) { // the 8 decisions of a step sit in one word
        for (int i = 0; i < h; i += 8) {
            float a[8], b[8];
            const uint32_t bw = c.bld((off + i) >> 5) >> ((off + i) & 31);
            in.ld8(i, a), in.ld8(i + h, b);
            in.wait();
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
            for (int k = 0; k < 8; ++k) a[k] = g_f32(a[k], b[k], (bw >> k) & 1u);
            c.st8(s - 1, i, a);
        }
    } else {
        for (int i = 0; i < h; ++i) c.st(s - 1, i, g_f32(in[i], in[i + h], c.bit(off + i)));
    }
}
template <class C, class V>
PC_HD void fop_g0(const C& c, const V& in, int s)
{
    fop_map(c, in, s, [](float a, float b, int) { return fadd(a, b); });
}
// right child rate-1: decisions = sign(g); parent = [left ^ right, right]   (ROneNode::rightDecode :204-219)
template <class C, class V>
PC_HD void fop_g1(const C& c, const V& in, int s, int off)
{
    const int h = 1 << (s - 1);
    if (h >= 32) {
        for (int w = 0; w < (h >> 5); ++w) {
            const uint32_t left = c.bld((off >> 5) + w);
            uint32_t r = 0;
            for (int k0 = 0; k0 < 32; k0 += 8) {
                float a[8], b[8];
                in.ld8(32 * w + k0, a), in.ld8(32 * w + k0 + h, b);
                in.wait();
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
                for (int k = 0; k < 8; ++k) {
                    const float gv = g_f32(a[k], b[k], (left >> (k0 + k)) & 1u);
                    r |= fsign(gv) << (k0 + k);
                    if constexpr (C::kSoft) c.soft_g1(off, h, 32 * w + k0 + k, f2u(gv));
                }
            }
            c.bst((off >> 5) + w, left ^ r);
            c.bst(((off + h) >> 5) + w, r);
        }
    } else {
        const int sh = off & 31;
        uint32_t word = c.bld(off >> 5);
        uint32_t r = 0;
        for (int k = 0; k < h; ++k) {
            const float gv = g_f32(in[k], in[k + h], (word >> (sh + k)) & 1u);
            r |= fsign(gv) << k;
            if constexpr (C::kSoft) c.soft_g1(off, h, k, f2u(gv));
        }
        word ^= r << sh;
        word = (word & ~(lowmask(h) << (sh + h))) | (r << (sh + h));
        c.bst(off >> 5, word);
    }
}
// Combine / CombineBitsShort / Combine_0R (avx_float.h:193-236)
template <class C>
PC_HD void fop_comb(const C& c, int s, int off, bool zero_left)
{
    const int h = 1 << (s - 1);
    if constexpr (C::kSoft) c.soft_comb(off, h, zero_left);
    if (h >= 32) {
        for (int w = 0; w < (h >> 5); ++w) {
            const uint32_t r = c.bld(((off + h) >> 5) + w);
            c.bst((off >> 5) + w, zero_left ? r : (c.bld((off >> 5) + w) ^ r));
        }
    } else {
        const int sh = off & 31;
        uint32_t word = c.bld(off >> 5);
        const uint32_t r = (word >> (sh + h)) & lowmask(h);
        if (zero_left)
            word = (word & ~(lowmask(h) << sh)) | (r << sh);
        else
            word ^= r << sh;
        c.bst(off >> 5, word);
    }
}

// ---- leaves ----------------------------------------------------------------------------------------------------
// Every leaf hands its result to the context as the float patterns the reference writes into its bit container (the
// decision is the sign bit): the hard-decision contexts keep the signs, the soft-output context the whole words.
// RateZeroDecoder :247 fills +inf; RepetitionDecoder :273-287 the sum
template <class C>
PC_HD void fleaf_fill(const C& c, int s, int off, uint32_t pattern)
{
    c.put_vals(off, 1 << s, [pattern](int) { return pattern; });
}
// RateOneDecoder :259-265: the LLRs themselves (-0.0 decodes to 1)
template <class C, class V>
PC_HD void fleaf_r1(const C& c, const V& in, int s, int off)
{
    c.put_vals(off, 1 << s, [&in](int i) { return f2u(in[i]); });
}
// RepetitionDecoder :273-287
template <class C, class V>
PC_HD void fleaf_rep(const C& c, const V& in, int s, int off)
{
    float acc[8];
    accumulate8(in, 1 << s, acc);
    fleaf_fill(c, s, off, f2u(reduce_add8(acc)));
}
// DoubleRepetitionDecoder :303-332 (n >= 8: every lane of the stored vector is its class sum -- the lanes add the same
// four numbers in orders that differ only by commutation)
template <class C, class V>
PC_HD void fleaf_drep(const C& c, const V& in, int s, int off)
{
    const int n = 1 << s;
    float a[8];
    accumulate8(in, n, a);
    float even, odd;
    if (n >= 8) {
        even = fadd(fadd(a[0], a[4]), fadd(a[2], a[6]));
        odd = fadd(fadd(a[1], a[5]), fadd(a[3], a[7]));
    } else {
        even = fadd(fadd(fadd(a[0], a[2]), a[4]), a[6]);
        odd = fadd(fadd(fadd(a[1], a[3]), a[5]), a[7]);
    }
    const uint32_t ve = f2u(even), vo = f2u(odd);
    c.put_vals(off, n, [ve, vo](int i) { return (i & 1) ? vo : ve; });
}
// SpcDecoder :342-373 (lanes >= n of a short node hold +inf, SpcPrepare avx_float.h:245-250): per vector the first
// lane holding the vector minimum, across vectors strict '<'; the stored value is the LLR, sign flipped at the minimum
template <class C, class V>
PC_HD void fleaf_spc(const C& c, const V& in, int s, int off)
{
    const int n = 1 << s;
    uint32_t par = 0;
    int minIdx = 0;
    float minAbs = INFINITY;
    for (int i = 0; i < n; i += 8) {
        float a[8];
        for (int k = 0; k < 8; ++k) {
            const float x = (i + k < n) ? in[i + k] : INFINITY;
            par ^= fsign(x);
            a[k] = fabs_(x);
        }
        float t;
        const int vm = minidx8(a, t);
        if (t < minAbs) {
            minIdx = vm + i;
            minAbs = t;
        }
    }
    c.put_vals(off, n, [&in, par, minIdx](int i) { return f2u(in[i]) ^ ((i == minIdx) ? (par << 31) : 0u); });
}
// DoubleSpcDecoder :425-466 (n >= 16): per lane the LAST occurrence of the lane minimum (the "lt" compare of
// avx_float.h:252-255 is a GT, so ties move the index), then the first even / odd lane holding the class minimum
template <class C, class V>
PC_HD void fleaf_dspc(const C& c, const V& in, int s, int off)
{
    const int n = 1 << s;
    float mv[8];
    int mi[8];
    uint32_t par[2] = { 0u, 0u };
    for (int k = 0; k < 8; ++k) {
        mv[k] = 3.402823466e+38f;
        mi[k] = 0;
    }
    for (int i = 0; i < n; i += 8)
        for (int k = 0; k < 8; ++k) {
            const float x = in[i + k], a = fabs_(x);
            par[k & 1] ^= fsign(x);
            if (!(a > mv[k])) {
                mv[k] = a;
                mi[k] = i + k;
            }
        }
    int idx[2];
    for (int q = 0; q < 2; ++q) {
        float m = mv[q];
        for (int k = q; k < 8; k += 2) m = fmin_ps(mv[k], m);
        int lane = q;
        for (int k = 6 + q; k >= q; k -= 2)
            if (mv[k] == m) lane = k;
        idx[q] = mi[lane];
    }
    const int ie = idx[0], io = idx[1];
    const uint32_t pe = par[0] << 31, po = par[1] << 31;
    c.put_vals(off, n, [&in, ie, io, pe, po](int i) {
        return f2u(in[i]) ^ ((i == ie) ? pe : 0u) ^ ((i == io) ? po : 0u);
    });
}
// DoubleSpcDecoderShort8 :473-488 (n == 8): every lane equal to its class minimum flips when the class parity is odd
template <class C, class V>
PC_HD void fleaf_dspc8(const C& c, const V& in, int off)
{
    float a[8];
    uint32_t par[2] = { 0u, 0u };
    for (int k = 0; k < 8; ++k) {
        a[k] = fabs_(in[k]);
        par[k & 1] ^= fsign(in[k]);
    }
    float m[2];
    for (int q = 0; q < 2; ++q) {
        m[q] = a[q];
        for (int k = q; k < 8; k += 2) m[q] = fmin_ps(a[k], m[q]);
    }
    const float m0 = m[0], m1 = m[1];
    const uint32_t p0 = par[0] << 31, p1 = par[1] << 31;
    c.put_vals(off, 8, [&in, m0, m1, p0, p1](int i) {
        const float mm = (i & 1) ? m1 : m0;
        return f2u(in[i]) ^ ((fabs_(in[i]) == mm) ? ((i & 1) ? p1 : p0) : 0u);
    });
}
// ZeroSpcDecoder :503-546 (n >= 16): the parity and the minimum come from left + right, but the stored value whose
// sign becomes the decision is `right` (:519-520) -- reproduced as is
template <class C, class V>
PC_HD void fleaf_zspc(const C& c, const V& in, int s, int off)
{
    const int n = 1 << s, h = n >> 1;
    uint32_t par = 0;
    int minIdx = 0;
    float minAbs = INFINITY;
    for (int i = 0; i < h; i += 8) {
        float a[8];
        for (int k = 0; k < 8; ++k) {
            const float llr = fadd(in[i + k], in[i + k + h]);
            par ^= fsign(llr);
            a[k] = fabs_(llr);
        }
        if (minAbs > 0) {
            float t;
            const int vm = minidx8(a, t);
            if (t < minAbs) {
                minIdx = vm + i;
                minAbs = t;
            }
        }
    }
    c.put_vals(off, n, [&in, h, par, minIdx](int i) {
        const int j = i >= h ? i - h : i;
        return f2u(in[j + h]) ^ ((j == minIdx) ? (par << 31) : 0u);
    });
}
// ZeroSpcDecoderShort8 :556-565 and TripleRepetitionDecoder :572-589: SPC over y[j] = v[j] + v[j+4]; the stored vector is
// the SPC result of the upper four lanes in both halves
template <class C>
PC_HD void fleaf_sum4_spc(const C& c, const float (&v)[8], int s, int off)
{
    float y[4];
    for (int j = 0; j < 4; ++j) y[j] = fadd(v[j + 4], v[j]); // upper lanes of input + swaplane are what is stored
    uint32_t o[4];
    spc4_vals(y, o);
    const uint32_t o0 = o[0], o1 = o[1], o2 = o[2], o3 = o[3];
    c.put_vals(off, 1 << s, [o0, o1, o2, o3](int i) { return (i & 2) ? ((i & 1) ? o3 : o2) : ((i & 1) ? o1 : o0); });
}
// TypeFiveDecoder :762-792 (rep_then_spc == true) and RepetitionRateOneDecoderShort8 :718-739 (false): on the lane sums
// v, a[j] = f(v[j], v[j+4]), rep = (a0 + a1) + (a2 + a3), y[j] = g(v[j], v[j+4], sign(rep)) (through the SPC for type
// five); the stored vector of 8, repeated over the node, is [+-1.0 with sign(rep) ^ sign(y[j]), y[j]]
template <class C>
PC_HD void fleaf_rep_then(const C& c, const float (&v)[8], int s, int off, bool spc)
{
    float a[4], y[4];
    for (int j = 0; j < 4; ++j) a[j] = f_f32(v[j + 4], v[j]);
    const float rep = fadd(fadd(a[0], a[1]), fadd(a[2], a[3]));
    const uint32_t rb = fsign(rep);
    for (int j = 0; j < 4; ++j) y[j] = g_f32(v[j], v[j + 4], rb);
    uint32_t o[4];
    if (spc)
        spc4_vals(y, o);
    else
        for (int j = 0; j < 4; ++j) o[j] = f2u(y[j]);
    const uint32_t rs = rb << 31, o0 = o[0], o1 = o[1], o2 = o[2], o3 = o[3];
    c.put_vals(off, 1 << s, [rs, o0, o1, o2, o3](int i) {
        const uint32_t yv = (i & 2) ? ((i & 1) ? o3 : o2) : ((i & 1) ? o1 : o0);
        return (i & 4) ? yv : (0x3F800000u | ((yv ^ rs) & 0x80000000u));
    });
}

// ---- interpreter -------------------------------------------------------------------------------------------------
// one node-program step on the input view `in` (C: where results go, V: where the node's LLRs are read)
template <class C, class V>
PC_HD void fastssc_f32_step(const C& c, const V& in, uint32_t op, int s, int off)
{
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

PC_HD void fastssc_f32_run(const FCtx& c, const uint32_t* code, const float* row, bool staged_root)
{
    const int S = c.S;
    if (staged_root)
        for (int i = 0; i < (1 << S); ++i) c.st(S, i, row[i]);
    for (int pc = 0;; ++pc) {
        const uint32_t ins = code[pc];
        const uint32_t op = instr_op(ins);
        const int s = (int)instr_s(ins), off = (int)instr_off(ins);
        if (op == FOP_END) break;
        const FView in = (s == S && !staged_root) ? FView{ row, 1 } : c.stage(s);
        fastssc_f32_step(c, in, op, s, off);
    }
}

} // namespace pc
#endif
