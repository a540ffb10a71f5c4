// Pieces shared by the two SCL kernels (pc_scl_i8.cu: SclFipChar, pc_scl_f32.cu: SclAvxFloat): the per-lane owner
// tables that replace the reference's ref-counted block tree, the emulated selection sort over candidate metrics,
// and the decision-bit steps (combine, take-with-flips, fork), which are identical for both LLR types because hard
// decisions are one bit per element in both kernels.
//
// A "lane type" LN provides: sm/T/G/l/P, metric, own_llr/own_L/own_R, ld(word, owner) and st(word, value) on the
// lane's shared-memory words. A "params type" PR provides bitL_off[] / bitR_off[] / L.
#ifndef PC_SCL_COMMON_CUH
#define PC_SCL_COMMON_CUH
#include <limits.h>
#include <stdint.h>

namespace pc {

// word offsets with this bit set address the lane's rows of the L2/HBM arena instead of shared memory: the largest LLR
// stages (so that more codewords fit one SM) and, for long codes, the decision words of the largest stages
constexpr int kArena = 1 << 30;
// word offsets with this bit set are COLUMNS of the lane's tensor-memory slice (decision words of the large stages):
// tcgen05.ld/st 32x32b can only reach the executing thread's own lane, so a read through the owner table is a load of
// the same column by every lane followed by a warp shuffle from the owner. Both are warp-collective: every function
// that touches decision words is executed by ALL lanes (dead lanes compute into their own, unreferenced buffers).
constexpr int kTmem = 1 << 29;
// bit modes of a kernel instantiation: which flags the decision-word accessors test
enum : int { BM_SMEM = 0, BM_ARENA = 1, BM_TMEM = 2 };

struct Tbl { // owner lane per stage, one byte each (stages 0 .. 15)
    uint32_t w[4];
    __device__ __forceinline__ uint32_t get(int s) const
    {
        const int q = s >> 2;
        const uint32_t v = q == 0 ? w[0] : (q == 1 ? w[1] : (q == 2 ? w[2] : w[3]));
        return __byte_perm(v, 0u, 0x4440u | (uint32_t)(s & 3)); // byte s % 4 into the low byte, zeros above
    }
    __device__ __forceinline__ void set(int s, uint32_t lane)
    {
        const int q = s >> 2;
        const uint32_t sh = 8u * (uint32_t)(s & 3), m = ~(0xFFu << sh), b = lane << sh;
        if (q == 0)
            w[0] = (w[0] & m) | b;
        else if (q == 1)
            w[1] = (w[1] & m) | b;
        else if (q == 2)
            w[2] = (w[2] & m) | b;
        else
            w[3] = (w[3] & m) | b;
    }
    __device__ __forceinline__ void copy_from(const Tbl& o, int src, int width)
    {
        w[0] = __shfl_sync(0xFFFFFFFFu, o.w[0], src, width);
        w[1] = __shfl_sync(0xFFFFFFFFu, o.w[1], src, width);
        w[2] = __shfl_sync(0xFFFFFFFFu, o.w[2], src, width);
        w[3] = __shfl_sync(0xFFFFFFFFu, o.w[3], src, width);
    }
};

// maximum over the G (power of two) lanes of a group, levels unrolled (the rolled loop spent as many instructions on
// its own control as on the shuffles)
__device__ __forceinline__ int group_max(int m, int G)
{
    switch (G) {
    case 32: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 16)); // fall through
    case 16: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 8));  // fall through
    case 8: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 4));   // fall through
    case 4: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 2));   // fall through
    case 2: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 1));   // fall through
    default: break;
    }
    return m;
}

__device__ __forceinline__ int stage_bwords(int s) { return s >= 5 ? (1 << (s - 5)) : 1; } // decision words

// ---- register-array helpers (dynamic index -> predicated selects) ---------------------------------
template <class V, int C>
__device__ __forceinline__ V sel(const V (&v)[C], int k)
{
    V r = v[0];
#pragma unroll
    for (int i = 1; i < C; ++i) r = (k == i) ? v[i] : r;
    return r;
}
template <class V, int C>
__device__ __forceinline__ void put(V (&v)[C], int k, V x)
{
#pragma unroll
    for (int i = 0; i < C; ++i) v[i] = (k == i) ? x : v[i];
}

// simplePartialSortDescending (reference include/polarcode/arrayfuncs.h:161-183) over the C * P candidate metrics
// held C per lane (position = lane * C + k), emulated pass by pass: pass i takes the FIRST maximum of positions
// [i, size) (strict '>'), then swaps values and indices of positions i and max. Returns Values[j] / Indices[j] for
// the caller's new path j = lane. V is int (SclFipChar's long metrics fit easily) or float (SclAvxFloat).
template <int C, class V, class LN>
__device__ __forceinline__ void scl_sort(const LN& c, V (&v)[C], int newP, V& out_val, int& out_idx)
{
    const int size = C * c.P, G = c.G;
    int idx[C];
#pragma unroll
    for (int k = 0; k < C; ++k) idx[k] = c.l * C + k;
    const int lim = size - 1 < newP ? size - 1 : newP;
    for (int i = 0; i < lim; ++i) {
        V bv = v[0];
        int bp = INT_MAX;
#pragma unroll
        for (int k = 0; k < C; ++k) {
            const int pos = c.l * C + k;
            if (pos >= i && pos < size && (bp == INT_MAX || v[k] > bv)) {
                bv = v[k];
                bp = pos;
            }
        }
        for (int off = G >> 1; off; off >>= 1) {
            const V ov = __shfl_xor_sync(0xFFFFFFFFu, bv, off);
            const int op = __shfl_xor_sync(0xFFFFFFFFu, bp, off);
            if (op != INT_MAX && (bp == INT_MAX || ov > bv || (ov == bv && op < bp))) {
                bv = ov;
                bp = op;
            }
        }
        const int li = i / C, ki = i % C, lm = bp / C, km = bp % C;
        const V vi = __shfl_sync(0xFFFFFFFFu, sel(v, ki), li, G);
        const int ii = __shfl_sync(0xFFFFFFFFu, sel(idx, ki), li, G);
        const int im = __shfl_sync(0xFFFFFFFFu, sel(idx, km), lm, G);
        if (c.l == lm) {
            put(v, km, vi);
            put(idx, km, ii);
        }
        if (c.l == li) {
            put(v, ki, bv);
            put(idx, ki, im);
        }
    }
    const int lj = (c.l / C) & (G - 1), kj = c.l % C;
    out_val = v[0];
    out_idx = 0;
#pragma unroll
    for (int k = 0; k < C; ++k) {
        const V tv = __shfl_sync(0xFFFFFFFFu, v[k], lj, G);
        const int ti = __shfl_sync(0xFFFFFFFFu, idx[k], lj, G);
        if (kj == k) {
            out_val = tv;
            out_idx = ti;
        }
    }
}

// The same selection sort for INTEGER metrics with |metric| < 2^22 (int8 SCL, N <= 32768): value and current position
// are packed into one key, (value << 8) | (255 - position), so "first maximum" is a plain integer max (a larger
// value wins, equal values go to the lower position) and one pass costs one max-reduction instead of a compare
// network over (value, position) pairs. The swap semantics are kept: the pass-i maximum is FINAL for new path i (its
// value and candidate index go straight to lane i), the former content of slot i moves to the maximum's slot and
// takes that slot's position bits; a finalised slot is marked consumed. The pass loop is unrolled over the slot index
// inside a lane so that slot i is a compile-time register.
template <int C, class LN>
__device__ __forceinline__ void scl_sort_packed(const LN& c, const int (&v)[C], int newP, int& out_val, int& out_idx)
{
    static_assert((C & (C - 1)) == 0, "C must be a power of two");
    constexpr int LOGC = C == 2 ? 1 : (C == 4 ? 2 : 3);
    const int size = C * c.P, G = c.G;
    int kv[C], idx[C];
#pragma unroll
    for (int k = 0; k < C; ++k) {
        const int pos = c.l * C + k;
        idx[k] = pos;
        kv[k] = pos < size ? (int)(((unsigned)v[k] << 8) | (unsigned)(255 - pos)) : INT_MIN;
    }
    const int lim = size - 1 < newP ? size - 1 : newP;
    out_val = 0;
    out_idx = 0;
    for (int li = 0; li * C < lim; ++li) {
#pragma unroll
        for (int ki = 0; ki < C; ++ki) {
            const int i = li * C + ki;
            if (i < lim) { // uniform
                int m = kv[0];
#pragma unroll
                for (int k = 1; k < C; ++k) m = max(m, kv[k]);
                // a full-warp group reduces with one redux.sync; sub-warp masks serialise per group in hardware and
                // measured slower than the shuffle tree (L=8: 4.9 vs 5.9 Gbit/s), L=32: 0.80 -> 1.00
                if (G == 32)
                    m = __reduce_max_sync(0xFFFFFFFFu, m);
                else
                    m = group_max(m, G);
                const int bp = 255 - (m & 255), lm = bp >> LOGC, km = bp & (C - 1);
                const int im = __shfl_sync(0xFFFFFFFFu, sel(idx, km), lm, G);
                const int vi = __shfl_sync(0xFFFFFFFFu, kv[ki], li, G);
                const int ii = __shfl_sync(0xFFFFFFFFu, idx[ki], li, G);
                if (c.l == lm) {
                    put(kv, km, (int)(((unsigned)vi & ~255u) | (unsigned)(255 - bp)));
                    put(idx, km, ii);
                }
                if (c.l == li) kv[ki] = INT_MIN; // consumed (also when the maximum already sat in slot i)
                if (c.l == i) {
                    out_val = m >> 8;
                    out_idx = im;
                }
            }
        }
    }
    if (lim < newP) { // size == newP: the last slot was never the target of a pass and keeps its content
        const int lj = lim >> LOGC, kj = lim & (C - 1);
        const int vj = __shfl_sync(0xFFFFFFFFu, sel(kv, kj), lj, G);
        const int ij = __shfl_sync(0xFFFFFFFFu, sel(idx, kj), lj, G);
        if (c.l == lim) {
            out_val = vj >> 8;
            out_idx = ij;
        }
    }
}

// The same emulation for groups of at most 8 lanes (C * G <= 64 candidates) and |metric| < 2^19 (N <= 4096): the candidate's
// ORIGINAL index rides in the key as well, key = (value << 12) | ((63 - position) << 6) | index, so the separate index
// array, its dynamic-index selects and two of the three shuffles per pass go away (the sort was 28 % of the executed
// instructions of k_scl_i8, profiles/r02_scl_i8_full_v2.txt). A pass: maximum of the keys (larger value wins, equal values
// go to the lower position), the content of slot i moves to the winner's slot and takes its position bits, slot i is
// consumed, new path i gets the winner's value and index.
template <int C, class LN>
__device__ __forceinline__ void scl_sort_packed2(const LN& c, const int (&v)[C], int newP, int& out_val, int& out_idx)
{
    static_assert((C & (C - 1)) == 0, "C must be a power of two");
    constexpr int LOGC = C == 2 ? 1 : (C == 4 ? 2 : 3);
    const int size = C * c.P, G = c.G;
    int kv[C];
#pragma unroll
    for (int k = 0; k < C; ++k) {
        const int pos = c.l * C + k;
        kv[k] = pos < size ? (int)(((unsigned)v[k] << 12) | ((unsigned)(63 - pos) << 6) | (unsigned)pos) : INT_MIN;
    }
    const int lim = size - 1 < newP ? size - 1 : newP;
    out_val = 0;
    out_idx = 0;
    for (int li = 0; li * C < lim; ++li) {
#pragma unroll
        for (int ki = 0; ki < C; ++ki) {
            const int i = li * C + ki;
            if (i < lim) { // uniform
                int m = kv[0];
#pragma unroll
                for (int k = 1; k < C; ++k) m = max(m, kv[k]);
                m = group_max(m, G);
                const int bp = 63 - ((m >> 6) & 63), lm = bp >> LOGC, km = bp & (C - 1);
                const int vi = __shfl_sync(0xFFFFFFFFu, kv[ki], li, G);
                if (c.l == lm) put(kv, km, (int)(((unsigned)vi & ~(63u << 6)) | ((unsigned)(63 - bp) << 6)));
                if (c.l == li) kv[ki] = INT_MIN; // consumed (also when the maximum already sat in slot i)
                if (c.l == i) {
                    out_val = m >> 12;
                    out_idx = m & 63;
                }
            }
        }
    }
    if (lim < newP) { // size == newP: the last slot was never the target of a pass and keeps its content
        const int lj = lim >> LOGC, kj = lim & (C - 1);
        const int vj = __shfl_sync(0xFFFFFFFFu, sel(kv, kj), lj, G);
        if (c.l == lim) {
            out_val = vj >> 12;
            out_idx = vj & 63;
        }
    }
}

// The emulation for FLOAT metrics: a 64-bit key, high word = the metric mapped to an unsigned integer of the same order
// (-0.0 is folded into +0.0 first, as the float compare treats them as equal), low word = ((0xFFFF - position) << 16) |
// candidate index. One unsigned 64-bit maximum per pass replaces the (value, position) compare network and the separate
// index array of scl_sort (which was 37 % of the executed instructions of k_scl_f32, profiles/r02_scl_f32_full_v3_T384.txt).
__device__ __forceinline__ uint32_t f32_order_key(float x)
{
    const uint32_t u = __float_as_uint(__fadd_rn(x, 0.0f));
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float f32_from_order_key(uint32_t k) { return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k); }
__device__ __forceinline__ unsigned long long group_max64(unsigned long long m, int G)
{
    switch (G) {
    case 32: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 16)); // fall through
    case 16: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 8));  // fall through
    case 8: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 4));   // fall through
    case 4: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 2));   // fall through
    case 2: m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, 1));   // fall through
    default: break;
    }
    return m;
}
template <int C, class LN>
__device__ __forceinline__ void scl_sort_key64(const LN& c, const float (&v)[C], int newP, float& out_val, int& out_idx)
{
    static_assert((C & (C - 1)) == 0, "C must be a power of two");
    constexpr int LOGC = C == 2 ? 1 : (C == 4 ? 2 : 3);
    typedef unsigned long long u64;
    const int size = C * c.P, G = c.G;
    u64 kv[C];
#pragma unroll
    for (int k = 0; k < C; ++k) {
        const int pos = c.l * C + k;
        kv[k] = pos < size ? (((u64)f32_order_key(v[k]) << 32) | ((u64)(0xFFFFu - (unsigned)pos) << 16) | (u64)pos) : 0ull;
    }
    const int lim = size - 1 < newP ? size - 1 : newP;
    out_val = 0.0f;
    out_idx = 0;
    for (int li = 0; li * C < lim; ++li) {
#pragma unroll
        for (int ki = 0; ki < C; ++ki) {
            const int i = li * C + ki;
            if (i < lim) { // uniform
                u64 m = kv[0];
#pragma unroll
                for (int k = 1; k < C; ++k) m = max(m, kv[k]);
                m = group_max64(m, G);
                const uint32_t lo = (uint32_t)m;
                const int bp = (int)(0xFFFFu - (lo >> 16)), lm = bp >> LOGC, km = bp & (C - 1);
                const u64 vi = __shfl_sync(0xFFFFFFFFu, kv[ki], li, G);
                if (c.l == lm) put(kv, km, (vi & ~(0xFFFFull << 16)) | ((u64)(0xFFFFu - (unsigned)bp) << 16));
                if (c.l == li) kv[ki] = 0ull; // consumed (also when the maximum already sat in slot i)
                if (c.l == i) {
                    out_val = f32_from_order_key((uint32_t)(m >> 32));
                    out_idx = (int)(lo & 0xFFFFu);
                }
            }
        }
    }
    if (lim < newP) { // size == newP: the last slot was never the target of a pass and keeps its content
        const int lj = lim >> LOGC, kj = lim & (C - 1);
        const u64 vj = __shfl_sync(0xFFFFFFFFu, sel(kv, kj), lj, G);
        if (c.l == lim) {
            out_val = f32_from_order_key((uint32_t)(vj >> 32));
            out_idx = (int)((uint32_t)vj & 0xFFFFu);
        }
    }
}

// Indices[] bookkeeping of findWeakLlrs (arrayfuncs.h:209-231): the first four entries plus a patch list for
// positions >= 4 (at most four swaps happen). swap(i, bp) is called once per pass with i < 4.
struct WeakIdx {
    int I[4], ppos[4], pidx[4], np;
    __device__ __forceinline__ void init()
    {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            I[k] = k;
            ppos[k] = -1;
            pidx[k] = 0;
        }
        np = 0;
    }
    __device__ __forceinline__ void swap(int i, int bp)
    {
        int im = bp;
        if (bp < 4) {
            im = sel(I, bp);
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < np && ppos[q] == bp) im = pidx[q]; // later patches override earlier ones
        }
        const int ii = sel(I, i);
        put(I, i, im);
        if (bp < 4) {
            if (bp != i) put(I, bp, ii);
        } else {
            put(ppos, np, bp);
            put(pidx, np, ii);
            np++;
        }
    }
};

template <class LN>
__device__ __forceinline__ void set_side(LN& c, int s, int side)
{
    if (side)
        c.own_R.set(s, (uint32_t)c.l);
    else
        c.own_L.set(s, (uint32_t)c.l);
}

// CombineBits (fip_char.h:173-207, avx_float.h:203-236): decisions(s, side) = [left ^ right, right] of stage s-1
template <class LN, class PR>
__device__ __forceinline__ void scl_comb(LN& c, const PR& p, int s, int side)
{
    // executed by every lane, live or not (see kTmem): a dead lane combines whatever its tables point at into its own
    // buffers, which no path references
    const int h = 1 << (s - 1);
    const uint32_t ol = c.own_L.get(s - 1), orr = c.own_R.get(s - 1);
    const int lo = p.bitL_off[s - 1], ro = p.bitR_off[s - 1], dst = side ? p.bitR_off[s] : p.bitL_off[s];
    if (h >= 32) {
        const int hw = h >> 5;
        if (hw >= 4) {
            for (int j = 0; j < hw; j += 4) { // loads first: the stores may alias them as far as the compiler knows
                uint32_t a[4], b[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) a[u] = c.ld(lo + j + u, ol), b[u] = c.ld(ro + j + u, orr);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    c.st(dst + j + u, a[u] ^ b[u]);
                    c.st(dst + hw + j + u, b[u]);
                }
            }
        } else {
            for (int j = 0; j < hw; ++j) {
                const uint32_t a = c.ld(lo + j, ol), b = c.ld(ro + j, orr);
                c.st(dst + j, a ^ b);
                c.st(dst + hw + j, b);
            }
        }
    } else {
        const uint32_t m = (1u << h) - 1u;
        const uint32_t a = c.ld(lo, ol) & m, b = c.ld(ro, orr) & m;
        c.st(dst, (a ^ b) | (b << h));
    }
    set_side(c, s, side);
}

// the new path in this lane takes the decisions of old path `sp` (stored in lane sp's side buffer) with the
// elements listed in fmask (bit q -> index fi[q]) flipped
template <class LN, class PR>
__device__ __forceinline__ void scl_take_bits(const LN& c, const PR& p, int s, int side, int sp, bool live, int fmask,
                                              const int (&fi)[4])
{
    // every lane stores (a dead lane copies path 0's words into its own unreferenced buffer): the tensor-memory store
    // is warp-collective
    (void)live;
    const int nb = stage_bwords(s), dst = side ? p.bitR_off[s] : p.bitL_off[s];
    c.st_fence(); // the leaf's own decisions were stored into these words a moment ago
    for (int w = 0; w < nb; ++w) {
        uint32_t t = c.ld(dst + w, (uint32_t)sp);
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (((fmask >> q) & 1) && (fi[q] >> 5) == w) t ^= 1u << (fi[q] & 31);
        __syncwarp();
        c.st(dst + w, t);
    }
    __syncwarp();
}

// duplicatePath + switchToNext (scl_fip_char.cpp:55-63,91-98): 9 shuffles instead of ref-count updates
template <class LN, class V>
__device__ __forceinline__ void scl_fork(LN& c, int sp, V val, int newP)
{
    Tbl a = c.own_llr, b = c.own_L, d = c.own_R;
    c.own_llr.copy_from(a, sp, c.G);
    c.own_L.copy_from(b, sp, c.G);
    c.own_R.copy_from(d, sp, c.G);
    c.metric = val;
    c.P = newP;
}

template <class LN>
struct SclBitsReader {
    const LN& c;
    int off;
    uint32_t owner;
    __device__ __forceinline__ uint32_t operator()(int word) const { return c.ld(off + word, owner); }
};

} // namespace pc
#endif
