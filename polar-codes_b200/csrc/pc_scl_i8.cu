// CRC-aided SCL int8 batch decoder for sm_100a: a group of G = 2^ceil(log2 L) lanes owns one codeword, ONE LANE PER
// LIST PATH. Replaces SclFipChar::decode of the reference (src/polarcode/decoding/scl_fip_char.cpp:801-855).
//
// * Per-path work (f, g, combine, candidate metrics of a leaf) runs in the path's lane with the same SWAR int8
//   arithmetic as the Fast-SSC kernel (pc_swar.h); all groups of a CTA execute the same node program in lockstep,
//   so the live path count P is uniform.
// * Lazy list copy: the reference's ref-counted block tree (PathList / DataPool, scl_fip_char.cpp:21-163,
//   datapool.txx:100-130) becomes three small owner tables per lane (LLR, left-decisions, right-decisions: which
//   lane's buffer holds stage s of my path), packed 5 bits per stage in registers. A fork is 9 warp shuffles; a
//   path always writes into its own lane's buffers, and a stage is only ever rewritten by all lanes at the same
//   program point, when its old contents are dead, so no copy-on-write is needed.
// * Path selection: the reference's swap-based selection sort (simplePartialSortDescending, arrayfuncs.h:161-183) is
//   emulated pass by pass with warp-shuffle arg-max (ties -> lowest position, then the swap), because the survivor
//   ORDER and tie handling are part of the bit-exact semantics. findWeakLlrs (arrayfuncs.h:209-231) likewise.
// * Hard decisions are one bit per element (only the MSB of the reference's bit bytes is ever consumed).
// * extractBestPath: every lane gathers its path's info bits and runs the detector; the first passing lane (ballot)
//   stores its block, else lane 0.
// * Optional indirection (frame_idx / count) lets the adaptive decoders (AdaptiveChar) run this kernel over the
//   frames whose Fast-SSC pass failed the detector only.
// Storage: word-granular, thread-interleaved (word w of a lane at smem[w * T + tid]); other lanes' buffers are read
// at their tid; the largest LLR stages live in an L2-resident arena (see kArena).
#include <cstdio>
#include <cstdlib>

#include "pc_fastssc_i8_core.h"
#include "pc_kernels.cuh"
#include "pc_scl_common.cuh"

namespace pc {

__device__ __forceinline__ uint32_t lds_u32(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

// BM: BM_ARENA some decision words live in the arena (long codes), BM_TMEM in tensor memory; 0 = no flag tests.
// TT: threads per CTA as a compile-time constant (0 = run-time): word w of a lane's shared-memory column sits at byte
// offset 4 TT w, an immediate of the load when w and TT are known
template <int BM, int TT, int GG = 0>
struct LaneT {
    static constexpr int kTT = TT;
    static constexpr int kGG = GG; // lanes per group when known at compile time (launch_scl_i8 picks it only with sort2 and lockstep on)
    uint32_t* sm;   // smem base + tid (word w at sm[w * T])
    uint32_t sm32;  // the same as a 32-bit shared address
    uint32_t* ar;   // arena base + global lane id (word w at ar[w * AS])
    size_t AS;
    uint32_t tm;    // tensor-memory address of this thread's column 0 (lane quarter of its warp + column base)
    int T, G, l;    // threads per CTA, lanes per group, lane index inside the group
    int P;          // live paths (uniform)
    int metric;     // path metric (the reference uses long; |metric| <= 128 N fits easily)
    Tbl own_llr, own_L, own_R;
    // decision words: shared memory, tensor memory (warp-collective: called by all 32 lanes with the same word), or
    // the arena for the largest stages of long codes (flag in the offset)
    __device__ __forceinline__ uint32_t ld(int word, uint32_t owner) const
    {
        if ((BM & BM_TMEM) && (word & kTmem)) {
            uint32_t v;
            tm_ld1_nowait(tm + (uint32_t)(word & 0xFFFF), v);
            tm_wait_ld();
            return __shfl_sync(0xFFFFFFFFu, v, (int)owner, G);
        }
        if ((BM & BM_ARENA) && (word & kArena)) return ar[(size_t)(word & (kArena - 1)) * AS + ((int)owner - l)];
        if (TT) return lds_u32(sm32 + 4u * ((uint32_t)word * (uint32_t)TT + (uint32_t)((int)owner - l)));
        return sm[(size_t)word * T + ((int)owner - l)];
    }
    __device__ __forceinline__ void st(int word, uint32_t v) const
    {
        if ((BM & BM_TMEM) && (word & kTmem)) {
            tm_st1(tm + (uint32_t)(word & 0xFFFF), v); // completion is awaited by st_fence(), once per step
        } else if ((BM & BM_ARENA) && (word & kArena))
            ar[(size_t)(word & (kArena - 1)) * AS] = v;
        else if (TT)
            sts_u32(sm32 + 4u * (uint32_t)word * (uint32_t)TT, v);
        else
            sm[(size_t)word * T] = v;
    }
    // tensor-memory stores complete: called before a step re-reads columns it wrote (scl_take_bits after the leaf's own
    // decisions) and at the end of every node-program step
    __device__ __forceinline__ void st_fence() const
    {
        if (BM & BM_TMEM) tm_wait_st();
    }
    // LLR stage starting at word offset `off` (p.llr_off[s]) in lane `owner`'s space, resolved once per node step
    struct View {
        uint32_t* p;
        uint32_t stride; // word index * stride stays below 2^32 (the planner bounds arena words * lanes)
        __device__ __forceinline__ uint32_t operator[](int w) const { return p[(uint32_t)w * stride]; }
        __device__ __forceinline__ void set(int w, uint32_t v) const { p[(uint32_t)w * stride] = v; }
    };
    // (Measured and not adopted: shared-address views for every access when all LLR stages sit in shared memory -- the
    // same 9.20 Gbit/s as these generic views, profiles/r02_c33_perf.txt; the decision-word ld / st above do use them.)
    __device__ __forceinline__ View view(int off, uint32_t owner) const
    {
        const ptrdiff_t d = (int)owner - l;
        if (off & kArena) return View{ ar + (size_t)(off & (kArena - 1)) * AS + d, (uint32_t)AS };
        return View{ sm + (size_t)off * T + d, (uint32_t)T };
    }
    // the lane's OWN stage in shared memory with the compile-time CTA size: 32-bit shared address, word w at + 4 TT w
    struct SView {
        uint32_t a;
        __device__ __forceinline__ uint32_t operator[](int w) const { return lds_u32(a + 4u * (uint32_t)(TT ? TT : 1) * (uint32_t)w); }
        __device__ __forceinline__ void set(int w, uint32_t v) const { sts_u32(a + 4u * (uint32_t)(TT ? TT : 1) * (uint32_t)w, v); }
    };
    // runs `f` on the own stage at `off`: the shared-memory form when the stage is there and TT is known (a shared load
    // with a multiply-free address instead of a generic load behind a 64-bit multiply-add), else the general view
    template <class F>
    __device__ __forceinline__ void with_own_view(int off, F&& f) const
    {
        if (TT != 0 && !(off & kArena))
            f(SView{ sm32 + 4u * (uint32_t)(TT ? TT : 1) * (uint32_t)off });
        else
            f(view(off, (uint32_t)l));
    }
};

__device__ __forceinline__ int stage_words(int s) { return s >= 2 ? (1 << (s - 2)) : 1; }   // LLR words
__device__ __forceinline__ uint32_t bytemask(int nbytes) { return nbytes >= 4 ? 0xFFFFFFFFu : ((1u << (8 * nbytes)) - 1u); }

// The same step while ONE path is alive (the whole frozen prefix of the tree, 15 % of a (1024, 512) frame's word steps
// in the lane-per-path form, where G - 1 lanes of a group only watch): all G lanes of the group share path 0's step,
// four words per lane and round, reading through lane 0's tables and writing into lane 0's column.
template <bool IS_G, class Lane>
__device__ __forceinline__ void scl_fg_single(Lane& c, const SclParams& p, int s, const int8_t* row)
{
    const int h = 1 << (s - 1), hw = h >> 2, G = c.G;
    const uint32_t* grow = reinterpret_cast<const uint32_t*>(row);
    const bool from_row = (s == p.S) && !p.root_leaf;
    const uint32_t o = __shfl_sync(0xFFFFFFFFu, c.own_llr.get(s), 0, G), ol = __shfl_sync(0xFFFFFFFFu, c.own_L.get(s - 1), 0, G);
    const int lb = p.bitL_off[s - 1];
    const typename Lane::View src = c.view(from_row ? p.llr_off[s - 1] : p.llr_off[s], o), dst = c.view(p.llr_off[s - 1], 0u);
    const bool row16 = from_row && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0);
    for (int j0 = 0; j0 < hw; j0 += 4 * G) {
        const int j = j0 + 4 * c.l;
        uint32_t bword = 0;
        if (IS_G) { // the decision words of this round, fetched by all lanes (tensor-memory loads are warp-collective)
            const int nq = (hw - j0 + 7) >> 3 < (G >> 1) ? (hw - j0 + 7) >> 3 : (G >> 1);
            for (int q = 0; q < nq; ++q) {
                const uint32_t w = c.ld(lb + (j0 >> 3) + q, ol);
                if (q == (c.l >> 1)) bword = w;
            }
        }
        if (j < hw) {
            uint32_t a[4], b[4], r[4];
            if (row16) {
                const uint4 va = __ldg(reinterpret_cast<const uint4*>(grow + j)), vb = __ldg(reinterpret_cast<const uint4*>(grow + hw + j));
                a[0] = va.x, a[1] = va.y, a[2] = va.z, a[3] = va.w;
                b[0] = vb.x, b[1] = vb.y, b[2] = vb.z, b[3] = vb.w;
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    a[u] = from_row ? __ldg(grow + j + u) : src[j + u];
                    b[u] = from_row ? __ldg(grow + hw + j + u) : src[hw + j + u];
                }
            }
            if (IS_G) {
                const uint32_t bw = bword >> (4 * (j & 7));
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = sw_g(a[u], b[u], sw_expand4(bw >> (4 * u)));
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = sw_f(a[u], b[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) dst.set(j + u, r[u]);
        }
    }
    c.own_llr.set(s - 1, (uint32_t)c.l);
}

// f / g step of a node of size 2^s: parent LLRs (stage s, through the owner table or the channel row) -> child
// LLRs (stage s-1, own buffer). RateRNode::decode scl_fip_char.cpp:315-349.
template <bool IS_G, class Lane>
__device__ __forceinline__ void scl_fg(Lane& c, const SclParams& p, int s, const int8_t* row)
{
    if (c.P == 1 && c.G >= 2 && (1 << (s - 1)) >= 32) { // at least eight words: two or more lanes of the group have work
        scl_fg_single<IS_G>(c, p, s, row);
        return;
    }
    // Every lane runs the step (the decision-word reads may be warp-collective tensor-memory loads): a dead lane
    // reads through its (valid, copied) tables -- the same addresses as the path it was copied from, so no extra
    // traffic -- and only its stores are suppressed.
    const bool active = c.l < c.P;
    const int h = 1 << (s - 1);
    const uint32_t* grow = reinterpret_cast<const uint32_t*>(row);
    const bool from_row = (s == p.S) && !p.root_leaf;
    const uint32_t o = c.own_llr.get(s), ol = c.own_L.get(s - 1);
    const int lb = p.bitL_off[s - 1];
    const typename Lane::View src = c.view(from_row ? p.llr_off[s - 1] : p.llr_off[s], o), dst = c.view(p.llr_off[s - 1], (uint32_t)c.l);
    if (Lane::kTT && h >= 16 && !from_row && !((p.llr_off[s] | p.llr_off[s - 1]) & kArena)) {
        // both stages in shared memory and the CTA size known at compile time: 32-bit shared addresses whose word
        // offsets are immediates (the view path below pays a multiply-add per access, 12 % of the executed
        // instructions in ncu r02_scl_i8_v1)
        constexpr uint32_t ES = 4u * (uint32_t)Lane::kTT;
        const uint32_t smb = c.sm32;
        const uint32_t sa = smb + ES * (uint32_t)p.llr_off[s] + 4u * (uint32_t)((int)o - c.l), da = smb + ES * (uint32_t)p.llr_off[s - 1];
        const int hw = h >> 2;
        uint32_t bword = 0;
        for (int j = 0; j < hw; j += 4) {
            uint32_t a[4], b[4], r[4];
            const uint32_t ja = sa + ES * (uint32_t)j, jb = ja + ES * (uint32_t)hw;
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] = lds_u32(ja + ES * u), b[u] = lds_u32(jb + ES * u);
            if (IS_G) {
                if ((j & 7) == 0) bword = c.ld(lb + (j >> 3), ol); // one decision word covers eight LLR words
                const uint32_t bw = bword >> (4 * (j & 7));
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = sw_g(a[u], b[u], sw_expand4(bw >> (4 * u)));
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = sw_f(a[u], b[u]);
            }
            if (active) {
                const uint32_t jd = da + ES * (uint32_t)j;
#pragma unroll
                for (int u = 0; u < 4; ++u) sts_u32(jd + ES * u, r[u]);
            }
        }
    } else if (h >= 16) {
        // four words per step: all loads first (source and destination may alias as far as the compiler knows, so
        // a rolled loop serialises load -> f/g -> store), the channel row as 16-byte loads
        const int hw = h >> 2;
        const bool row16 = from_row && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0);
        uint32_t bword = 0;
        for (int j = 0; j < hw; j += 4) {
            uint32_t a[4], b[4], r[4];
            if (row16) {
                const uint4 va = __ldg(reinterpret_cast<const uint4*>(grow + j)), vb = __ldg(reinterpret_cast<const uint4*>(grow + hw + j));
                a[0] = va.x, a[1] = va.y, a[2] = va.z, a[3] = va.w;
                b[0] = vb.x, b[1] = vb.y, b[2] = vb.z, b[3] = vb.w;
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    a[u] = from_row ? __ldg(grow + j + u) : src[j + u];
                    b[u] = from_row ? __ldg(grow + hw + j + u) : src[hw + j + u];
                }
            }
            if (IS_G) {
                if ((j & 7) == 0) bword = c.ld(lb + (j >> 3), ol); // one decision word covers eight LLR words
                const uint32_t bw = bword >> (4 * (j & 7));
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = sw_g(a[u], b[u], sw_expand4(bw >> (4 * u)));
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) r[u] = sw_f(a[u], b[u]);
            }
            if (active) {
#pragma unroll
                for (int u = 0; u < 4; ++u) dst.set(j + u, r[u]);
            }
        }
    } else if (h >= 4) {
        const int hw = h >> 2;
        uint32_t bword = 0;
        if (IS_G) bword = c.ld(lb, ol); // h <= 8: at most two LLR words, one decision word
        for (int j = 0; j < hw; ++j) {
            const uint32_t a = from_row ? __ldg(grow + j) : src[j];
            const uint32_t b = from_row ? __ldg(grow + hw + j) : src[hw + j];
            uint32_t r;
            if (IS_G)
                r = sw_g(a, b, sw_expand4(bword >> (4 * (j & 7))));
            else
                r = sw_f(a, b);
            if (active) dst.set(j, r);
        }
    } else {
        const uint32_t w = from_row ? __ldg(grow) : src[0];
        const uint32_t hi = w >> (8 * h);
        uint32_t r;
        if (IS_G)
            r = sw_g(w, hi, sw_expand4(c.ld(lb, ol)));
        else
            r = sw_f(w, hi);
        if (active) dst.set(0, r & bytemask(h));
    }
    c.own_llr.set(s - 1, (uint32_t)c.l);
}

// sums of the negative / positive LLRs of the own stage-s buffer (exact integer sums, dp4a)
template <class Lane>
__device__ __forceinline__ void llr_sums(const Lane& c, const SclParams& p, int s, int& neg, int& pos)
{
    const int n = 1 << s, nw = stage_words(s);
    neg = 0;
    pos = 0;
    if (c.P == 1 && c.G >= 2 && nw >= 8) {
        // one live path: the lanes of the group sum slices of ITS buffer (lane 0's column) and add up (exact integers)
        const typename Lane::View v0 = c.view(p.llr_off[s], 0u);
        for (int j = c.l; j < nw; j += c.G) {
            const uint32_t w = v0[j], f = sw_signfill(w);
            neg = __dp4a((int)(w & f), 0x01010101, neg);
            pos = __dp4a((int)(w & ~f), 0x01010101, pos);
        }
        for (int off = c.G >> 1; off; off >>= 1) {
            neg += __shfl_xor_sync(0xFFFFFFFFu, neg, off);
            pos += __shfl_xor_sync(0xFFFFFFFFu, pos, off);
        }
        return;
    }
    c.with_own_view(p.llr_off[s], [&](const auto& v) {
    if (nw >= 4) {
        int n1 = 0, p1 = 0; // two accumulator pairs, four loads in flight
        for (int j = 0; j < nw; j += 4) {
            const uint32_t w0 = v[j], w1 = v[j + 1], w2 = v[j + 2], w3 = v[j + 3];
            const uint32_t f0 = sw_signfill(w0), f1 = sw_signfill(w1), f2 = sw_signfill(w2), f3 = sw_signfill(w3);
            neg = __dp4a((int)(w0 & f0), 0x01010101, neg);
            pos = __dp4a((int)(w0 & ~f0), 0x01010101, pos);
            n1 = __dp4a((int)(w1 & f1), 0x01010101, n1);
            p1 = __dp4a((int)(w1 & ~f1), 0x01010101, p1);
            neg = __dp4a((int)(w2 & f2), 0x01010101, neg);
            pos = __dp4a((int)(w2 & ~f2), 0x01010101, pos);
            n1 = __dp4a((int)(w3 & f3), 0x01010101, n1);
            p1 = __dp4a((int)(w3 & ~f3), 0x01010101, p1);
        }
        neg += n1;
        pos += p1;
        return;
    }
    for (int j = 0; j < nw; ++j) {
        uint32_t w = v[j];
        if (n < 4) w &= bytemask(n);
        const uint32_t f = sw_signfill(w);
        neg = __dp4a((int)(w & f), 0x01010101, neg);
        pos = __dp4a((int)(w & ~f), 0x01010101, pos);
    }
    });
}

// RateZeroDecoder::decode scl_fip_char.cpp:387-421: metric += sum(min(llr, 0)), decisions 0, no fork, no re-sort
template <class Lane>
__device__ __forceinline__ void scl_r0(Lane& c, const SclParams& p, int s, int side)
{
    int neg, pos; // every lane runs (dead lanes: garbage into their own state, replaced at the next fork)
    llr_sums(c, p, s, neg, pos);
    c.metric += neg;
    const int dst = side ? p.bitR_off[s] : p.bitL_off[s], nb = stage_bwords(s);
    for (int j = 0; j < nb; ++j) c.st(dst + j, 0u);
    set_side(c, s, side);
}

// findWeakLlrs (arrayfuncs.h:209-231) on the |LLR| bytes stored in place in the own stage buffer: `want` passes of
// a selection sort that takes the FIRST minimum (strict '<') of positions [i, n) and swaps it to position i. The
// swaps mutate the array, which decides later ties, so they are performed for real. The Indices array of the
// reference is tracked as its first four entries plus a patch list for positions >= 4.
template <class Lane>
__device__ __forceinline__ void scl_find_weak(const Lane& c, int off, int n, int want, int (&widx)[4], int (&wval)[4])
{
    int I[4] = { 0, 1, 2, 3 }, ppos[4] = { -1, -1, -1, -1 }, pidx[4] = { 0, 0, 0, 0 };
    int np = 0;
    c.with_own_view(off, [&](const auto& v) {
    const int lim = n - 1 < want ? n - 1 : want;
    const int nw = (n + 3) >> 2;
    for (int i = 0; i < lim; ++i) {
        int bv = 256, bp = i;
        if (nw <= 256) {
            // first minimum of [i, n) as the minimum of 16-bit keys (|llr| << 8 | word index), one running minimum per
            // byte lane (VIMNMX.U16x2); excluded bytes are forced to 0xFF, above every |llr| <= 127
            // word 0 (the only one with excluded bytes) peeled, the rest four words per round (nw is a power of two)
            const uint32_t x0 = v[0] | bytemask(i) | (n < 4 ? ~bytemask(n) : 0u);
            uint32_t me = __byte_perm(x0, 0u, 0x2404), mo = __byte_perm(x0, 0u, 0x3414);
            if (nw >= 4) {
#pragma unroll
                for (int u = 1; u < 4; ++u) {
                    const uint32_t x = v[u];
                    me = __vminu2(me, __byte_perm(x, (uint32_t)u, 0x2404));
                    mo = __vminu2(mo, __byte_perm(x, (uint32_t)u, 0x3414));
                }
                for (int w = 4; w < nw; w += 4) {
                    uint32_t x[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) x[u] = v[w + u];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        me = __vminu2(me, __byte_perm(x[u], (uint32_t)(w + u), 0x2404));
                        mo = __vminu2(mo, __byte_perm(x[u], (uint32_t)(w + u), 0x3414));
                    }
                }
            } else if (nw == 2) {
                const uint32_t x = v[1];
                me = __vminu2(me, __byte_perm(x, 1u, 0x2404));
                mo = __vminu2(mo, __byte_perm(x, 1u, 0x3414));
            }
            uint32_t best = 0xFFFFFFFFu;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const uint32_t k = ((b & 1) ? mo : me) >> ((b & 2) ? 16 : 0) & 0xFFFFu;
                best = min(best, ((k >> 8) << 10) | ((k & 0xFFu) << 2) | (uint32_t)b);
            }
            bv = (int)(best >> 10);
            bp = (int)(best & 1023u);
        } else {
            for (int w = i >> 2; w < nw; ++w) {
                const uint32_t x = v[w];
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int pos = 4 * w + b, val = (int)((x >> (8 * b)) & 0xFFu);
                    if (pos >= i && pos < n && val < bv) {
                        bv = val;
                        bp = pos;
                    }
                }
            }
        }
        // swap bytes i <-> bp
        const uint32_t wi = v[i >> 2];
        const int vi = (int)((wi >> (8 * (i & 3))) & 0xFFu);
        if (bp != i) {
            v.set(i >> 2, (wi & ~(0xFFu << (8 * (i & 3)))) | ((uint32_t)bv << (8 * (i & 3))));
            const uint32_t wm = v[bp >> 2];
            v.set(bp >> 2, (wm & ~(0xFFu << (8 * (bp & 3)))) | ((uint32_t)vi << (8 * (bp & 3))));
        }
        // Indices[i] <-> Indices[bp]
        int im = bp;
        if (bp < 4) {
            im = sel(I, bp);
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < np && ppos[q] == bp) im = pidx[q]; // later patches override earlier ones
        }
        const int ii = sel(I, i); // i < 4 always (want <= 4)
        put(I, i, im);
        if (bp < 4) {
            if (bp != i) put(I, bp, ii);
        } else {
            put(ppos, np, bp);
            put(pidx, np, ii);
            np++;
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        widx[k] = I[k];
        const uint32_t x = v[k >> 2];
        wval[k] = (k < n) ? (int)((x >> (8 * (k & 3))) & 0xFFu) : 0;
    }
    });
}

// first phase of RateOne / SPC leaves: decisions = signs into the own side buffer, |max(llr, -127)| in place,
// returns the XOR of all LLR words (its byte MSBs give the parity)
template <class Lane>
__device__ __forceinline__ uint32_t scl_signs_and_abs(const Lane& c, const SclParams& p, int s, int side)
{
    const int n = 1 << s, nw = stage_words(s), dst = side ? p.bitR_off[s] : p.bitL_off[s];
    uint32_t par = 0;
    c.with_own_view(p.llr_off[s], [&](const auto& v) {
        uint32_t acc = 0;
        if (nw >= 8) { // eight words (one decision word) per step, loads first
            for (int j = 0; j < nw; j += 8) {
                uint32_t w[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) w[u] = v[j + u];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    par ^= w[u];
                    acc |= sw_movemask(w[u]) << (4 * u);
                    v.set(j + u, sw_abs_sat(w[u]));
                }
                c.st(dst + (j >> 3), acc);
                acc = 0;
            }
            return;
        }
        for (int j = 0; j < nw; ++j) {
            uint32_t w = v[j];
            if (n < 4) w &= bytemask(n);
            par ^= w;
            acc |= sw_movemask(w) << (4 * (j & 7));
            v.set(j, sw_abs_sat(w));
            if ((j & 7) == 7 || j == nw - 1) {
                c.st(dst + (j >> 3), acc);
                acc = 0;
            }
        }
    });
    return par;
}

// RepetitionDecoder::decode scl_fip_char.cpp:508-581
template <class Lane>
__device__ __forceinline__ void scl_rep(Lane& c, const SclParams& p, int s, int side)
{
    int v[2] = { INT_MIN, INT_MIN };
    int neg, pos;
    llr_sums(c, p, s, neg, pos);
    if (c.l < c.P) {
        v[0] = c.metric + neg;
        v[1] = c.metric - pos;
    }
    const int newP = 2 * c.P < p.L ? 2 * c.P : p.L;
    int val, idx;
    if ((Lane::kGG && Lane::kGG <= 8) || p.sort2)
        scl_sort_packed2<2>(c, v, newP, val, idx);
    else
        scl_sort_packed<2>(c, v, newP, val, idx);
    const bool live = c.l < newP;
    const int sp = live ? idx >> 1 : 0;
    scl_fork(c, sp, val, newP);
    const int dst = side ? p.bitR_off[s] : p.bitL_off[s], nb = stage_bwords(s);
    const uint32_t fill = (idx & 1) ? 0xFFFFFFFFu : 0u;
    for (int j = 0; j < nb; ++j) c.st(dst + j, fill); // all lanes (tensor-memory stores are warp-collective)
    set_side(c, s, side);
}

// RateOneDecoder::decode scl_fip_char.cpp:423-505
template <class Lane>
__device__ __forceinline__ void scl_r1(Lane& c, const SclParams& p, int s, int side)
{
    int v[4] = { INT_MIN, INT_MIN, INT_MIN, INT_MIN };
    int wi[4] = { 0, 0, 0, 0 }, wv[4];
    scl_signs_and_abs(c, p, s, side);
    scl_find_weak(c, p.llr_off[s], 1 << s, 2, wi, wv);
    if (c.l < c.P) {
        v[0] = c.metric;
        v[1] = c.metric - wv[0];
        v[2] = c.metric - wv[1];
        v[3] = c.metric - wv[0] - wv[1];
    }
    __syncwarp();
    const int newP = 4 * c.P < p.L ? 4 * c.P : p.L;
    int val, idx;
    if ((Lane::kGG && Lane::kGG <= 8) || p.sort2)
        scl_sort_packed2<4>(c, v, newP, val, idx);
    else
        scl_sort_packed<4>(c, v, newP, val, idx);
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

// SpcDecoder::decode scl_fip_char.cpp:583-726
template <class Lane>
__device__ __forceinline__ void scl_spc(Lane& c, const SclParams& p, int s, int side)
{
    int v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = INT_MIN;
    int wi[4] = { 0, 0, 0, 0 }, wv[4];
    const uint32_t par = scl_signs_and_abs(c, p, s, side);
    const int parity = (int)parity_of_msbs(par);
    scl_find_weak(c, p.llr_off[s], 1 << s, 4, wi, wv);
    if (c.l < c.P) {
        int m = c.metric, weakest = 0;
        if (parity)
            m -= wv[0];
        else
            weakest = wv[0];
        v[0] = m;
        v[1] = m - weakest - wv[1];
        v[2] = m - weakest - wv[2];
        v[3] = m - weakest - wv[3];
        v[4] = m - wv[1] - wv[2];
        v[5] = m - wv[1] - wv[3];
        v[6] = m - wv[2] - wv[3];
        v[7] = m - weakest - wv[1] - wv[2] - wv[3];
    }
    __syncwarp();
    const int newP = 8 * c.P < p.L ? 8 * c.P : p.L;
    int val, idx;
    if ((Lane::kGG && Lane::kGG <= 8) || p.sort2)
        scl_sort_packed2<8>(c, v, newP, val, idx);
    else
        scl_sort_packed<8>(c, v, newP, val, idx);
    const bool live = c.l < newP;
    const int sp = live ? idx >> 3 : 0;
    int fi[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) fi[q] = __shfl_sync(0xFFFFFFFFu, wi[q], sp, c.G);
    const int spar = __shfl_sync(0xFFFFFFFFu, parity, sp, c.G);
    // flip sets per candidate as 4-bit masks over (idx0..idx3), scl_fip_char.cpp:641-692
    const uint32_t tab = spar ? 0xEDB78421u : 0xFCA69530u;
    scl_fork(c, sp, val, newP);
    scl_take_bits(c, p, s, side, sp, live, (int)((tab >> (4 * (idx & 7))) & 15u), fi);
    set_side(c, s, side);
}

// GG: lanes per group as a compile-time constant (0 = run-time p.G): the shuffle widths, the group-maximum tree of the
// sorts and the lane arithmetic then fold
template <int BM, int TT, int GG = 0>
__global__ void __launch_bounds__(512, 1) k_scl_i8(const __grid_constant__ SclParams p)
{
    using Lane = LaneT<BM, TT, GG>;
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
    Lane c;
    c.sm = smem + 256 + t;
    c.sm32 = (uint32_t)__cvta_generic_to_shared(c.sm);
    c.ar = p.arena + ((size_t)blockIdx.x * T + t);
    c.AS = p.arena_stride;
    // a warp reaches the 32 tensor-memory lanes of its quarter (warp % 4); warps sharing a quarter are stacked along columns
    c.tm = (BM & BM_TMEM) ? tm_base_s + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * p.tm_cols_thread) : 0u;
    c.T = T;
    const int G = GG ? GG : p.G;
    c.G = G;
    c.l = t & (G - 1);
    const int gpc = T / G; // codewords per CTA
    const int N = 1 << p.S, kb = (p.K + 7) >> 3;
    const unsigned gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << ((t & 31) & ~(G - 1));
    const unsigned long long nfr = p.count ? *p.count : p.B;
    const unsigned long long nbatch = (nfr + gpc - 1) / gpc; // 0 when there is nothing to do: straight to the dealloc
    for (unsigned long long bi = blockIdx.x; bi < nbatch; bi += gridDim.x) {
        const unsigned long long fl = bi * gpc + (unsigned long long)(t / G);
        const bool live_frame = fl < nfr;
        unsigned long long f = live_frame ? fl : nfr - 1;
        if (p.frame_idx) f = p.frame_idx[f];
        const int8_t* row = p.llr + f * N;
        c.P = 1;
        c.metric = 0;
        c.own_llr = Tbl{ { 0u, 0u, 0u, 0u } };
        c.own_L = c.own_llr;
        c.own_R = c.own_llr;
        if (p.root_leaf && c.l == 0) { // stage the channel row (the root is a leaf)
            const int nw = stage_words(p.S);
            if (N >= 4) {
                const typename Lane::View v = c.view(p.llr_off[p.S], (uint32_t)c.l);
                for (int j = 0; j < nw; ++j) v.set(j, reinterpret_cast<const uint32_t*>(row)[j]);
            } else {
                c.view(p.llr_off[p.S], (uint32_t)c.l).set(0, (uint32_t)(uint8_t)row[0] | ((uint32_t)(uint8_t)row[1] << 8));
            }
        }
        __syncwarp();
        for (int pc = 0;; ++pc) {
            const uint32_t ins = __ldg(p.code + pc);
            const uint32_t op = ins & 63u;
            const int s = (int)((ins >> 6) & 31u), side = (int)((ins >> 11) & 1u);
            if (op == SCL_END) break;
            switch (op) {
            case SCL_F: scl_fg<false>(c, p, s, row); break;
            case SCL_G: scl_fg<true>(c, p, s, row); break;
            case SCL_COMB: scl_comb(c, p, s, side); break;
            case SCL_R0: scl_r0(c, p, s, side); break;
            case SCL_R1: scl_r1(c, p, s, side); break;
            case SCL_REP: scl_rep(c, p, s, side); break;
            default: scl_spc(c, p, s, side); break;
            }
            // other lanes' buffers are read by the next step. With `lockstep` the whole CTA meets here: the kernel is
            // 130 KB of SASS, and warps spread over the program (ncu r02_scl_i8_v0: `no_instruction` 5.3 cycles per
            // issue with 16 free-running warps) thrash the instruction caches; in step, a fetched line serves all warps.
            c.st_fence();
            if ((GG && (BM & BM_TMEM)) || p.lockstep)
                __syncthreads();
            else
                __syncwarp();
        }
        // extractBestPath (scl_fip_char.cpp:816-855)
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
        const SclBitsReader<Lane> rd{ c, off, owner };
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

cudaError_t launch_scl_i8(const SclLaunch& L, const int8_t* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                          const unsigned* d_frame_idx, const unsigned long long* d_count, cudaStream_t st)
{
    SclParams p = L.p;
    p.llr = d_llr;
    p.info = d_info;
    p.ok = d_ok;
    p.B = B;
    p.frame_idx = d_frame_idx;
    p.count = d_count;
    // decision words in the arena (long codes) or in tensor memory -> the instantiation whose ld / st test that flag
    const bool ba = (p.bitR_off[p.S] & kArena) != 0;
    void (*kern)(SclParams) = ba ? k_scl_i8<BM_ARENA, 0> : k_scl_i8<BM_SMEM, 0>;
    if (p.tm_alloc_cols) kern = L.threads == 512 ? k_scl_i8<BM_TMEM, 512> : (L.threads == 256 ? k_scl_i8<BM_TMEM, 256> : k_scl_i8<BM_TMEM, 0>);
    // group size (and with it sort form and lockstep) at compile time for the list sizes of the metric (L = 5 ... 8) and of
    // BASELINE config 4 (L = 17 ... 32)
    if (p.tm_alloc_cols && L.threads == 512 && p.lockstep == 1 && !getenv("PC_SCL_GRT")) {
        if (p.G == 8 && p.sort2) kern = k_scl_i8<BM_TMEM, 512, 8>;
        if (p.G == 4 && p.sort2) kern = k_scl_i8<BM_TMEM, 512, 4>; // pcsim's listlength sweep: L = 3, 4
        if (p.G == 32 && !p.sort2) kern = k_scl_i8<BM_TMEM, 512, 32>;
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes);
    // several CTAs per SM only fit when the L1/shared split is at its shared-memory maximum; a single CTA keeps the
    // default split (the rest stays L1 for the channel rows and the arena)
    // (the tensor-memory plan needs less shared memory: ask for what its CTAs use, the rest stays L1)
    int carve = L.per_sm > 1 ? (int)cudaSharedmemCarveoutMaxShared : (int)cudaSharedmemCarveoutDefault;
    if (p.tm_alloc_cols) {
        carve = (int)(((size_t)L.per_sm * (L.smem_bytes + 1024) * 100 + 233471) / 233472) + 3;
        if (carve > 100) carve = 100;
    }
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    if (e != cudaSuccess) return e;
    static const bool debug_occ = getenv("PC_DEBUG_OCC") != nullptr;
    if (debug_occ) {
        int occ = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L.threads, L.smem_bytes);
        fprintf(stderr, "k_scl_i8: %d threads, %zu B smem, grid %d, tmem cols %d, occupancy %d CTAs/SM\n", L.threads, L.smem_bytes,
                L.grid, p.tm_alloc_cols, occ);
    }
    const int gpc = L.threads / p.G;
    size_t nb = (B + gpc - 1) / gpc;
    const int grid = (int)(nb < (size_t)L.grid ? nb : (size_t)L.grid);
    kern<<<grid, L.threads, L.smem_bytes, st>>>(p);
    return cudaGetLastError();
}

// host planner without tensor memory: per-lane word layout and CTA size
static int plan_scl_i8_smem(int S, int K, int L, bool root_leaf, size_t smem_max, int sms, SclLaunch& out, const char** msg)
{
    if (L < 1 || L > 32) {
        *msg = "list size must be in [1, 32]";
        return 1;
    }
    if (S > 15) { // path metrics are packed into 24-bit sort keys (scl_sort_packed): |metric| <= 127 N < 2^22
        *msg = "block length too large for the SCL kernel";
        return 1;
    }
    int G = 1;
    while (G < L) G <<= 1;
    SclParams& p = out.p;
    p.S = S;
    p.K = K;
    p.L = L;
    p.G = G;
    p.root_leaf = root_leaf ? 1 : 0;
    const int top_llr = root_leaf ? S : S - 1;
    // Lanes (= list paths) resident per SM decide throughput: the kernel is latency bound (one lane per path, long
    // dependent chains), so the planner trades shared-memory residency of the largest LLR stages for more lanes.
    // Candidates in order of preference; the first one that keeps stages <= `pref` in shared memory wins. Measured
    // on B200 (N=1024, L=8, tools/perf_decoder.py): see DESIGN.md section 3.4.
    int pref = 7, force_lanes = -1, force_top = -2;
    if (const char* e = getenv("PC_SCL_PREF")) pref = atoi(e);
    if (const char* e = getenv("PC_SCL_LANES")) force_lanes = atoi(e);
    if (const char* e = getenv("PC_SCL_SMEMTOP")) force_top = atoi(e);
    static const int cand[] = { 512, 384, 256, 192, 128, 96, 64, 32 };
    int T = 0, per_sm = 0, smem_top = -1, bit_top = S;
    // decision words of the stages > bit_top go to the arena: only when nothing fits otherwise (N >= 16384 at L = 8)
    const int bit_lo = S < 5 ? S : 5;
    for (bit_top = S; bit_top >= bit_lo && !T; --bit_top) {
        int bw_smem = 0;
        for (int s = 0; s <= S; ++s)
            if (s <= bit_top) bw_smem += 2 * (s >= 5 ? (1 << (s - 5)) : 1);
        for (int pass = 0; pass < 2 && !T; ++pass) {
            for (size_t ci = 0; ci < sizeof(cand) / sizeof(cand[0]); ++ci) {
                const int lanes = cand[ci];
                if (force_lanes > 0 && lanes != force_lanes) continue;
                if (bit_top < S && bit_top > 5 && lanes < 128) continue; // rather move another stage out than run 32 lanes
                const int t = lanes <= 256 ? lanes : 128, n = lanes / t;
                if (t < G || (t % G) != 0) continue;
                size_t per_cta = 233472 / (size_t)n - 1024;
                if (per_cta > smem_max) per_cta = smem_max;
                const long budget = ((long)per_cta - 1024) / (4L * t); // words per lane
                if (budget < bw_smem) continue;
                int w = bw_smem, top = -1;
                for (int st = 0; st <= top_llr; ++st) {
                    const int sw = st >= 2 ? (1 << (st - 2)) : 1;
                    if (w + sw > budget || (force_top > -2 && st > force_top)) break;
                    w += sw;
                    top = st;
                }
                // first pass: insist on the preferred residency; second pass: anything that fits
                if (pass == 0 && force_lanes <= 0 && top < (top_llr < pref ? top_llr : pref)) continue;
                T = t;
                per_sm = n;
                smem_top = top;
                break;
            }
        }
        if (T) break;
    }
    if (!T) {
        *msg = "SCL path state does not fit shared memory (block length too large for this list size)";
        return 1;
    }
    int w = 0, aw = 0;
    for (int s = 0; s < 20; ++s) p.llr_off[s] = p.bitL_off[s] = p.bitR_off[s] = 0;
    for (int s = 0; s <= top_llr; ++s) {
        const int sw = s >= 2 ? (1 << (s - 2)) : 1;
        if (s <= smem_top) {
            p.llr_off[s] = w;
            w += sw;
        } else {
            p.llr_off[s] = kArena | aw;
            aw += sw;
        }
    }
    for (int s = 0; s <= S; ++s) {
        const int bw = s >= 5 ? (1 << (s - 5)) : 1;
        if (s <= bit_top) {
            p.bitL_off[s] = w;
            w += bw;
            p.bitR_off[s] = w;
            w += bw;
        } else {
            p.bitL_off[s] = kArena | aw;
            aw += bw;
            p.bitR_off[s] = kArena | aw;
            aw += bw;
        }
    }
    p.words = w;
    p.arena_words = aw;
    p.tm_alloc_cols = 0;
    p.tm_cols_thread = 0;
    p.lockstep = getenv("PC_SCL_LOCKSTEP") ? atoi(getenv("PC_SCL_LOCKSTEP")) != 0 : 0;
    p.sort2 = (G <= 8 && S <= 12 && !getenv("PC_SCL_SORT1")) ? 1 : 0;
    p.scr_off = p.bitL_off[S]; // non-systematic extraction works on the (free) left buffer of the top stage
    out.threads = T;
    out.smem_bytes = 1024 + 4 * (size_t)w * T;
    out.grid = sms * per_sm;
    out.per_sm = per_sm;
    p.arena_stride = (size_t)out.grid * T;
    if ((unsigned long long)aw * p.arena_stride >= (1ull << 32)) {
        *msg = "SCL arena too large for 32-bit word indexing";
        return 1;
    }
    return 0;
}

// Planner, first choice: 2 CTAs of 256 lanes per SM (16 warps; 128 registers per thread is the register file's limit)
// with the decision words of the large stages in TENSOR MEMORY. The kernel is bound by dependent integer chains
// (ncu r01/r02: issue slots 37 % busy with 2 warps per scheduler, top stall `wait`), so resident warps decide its
// throughput, and shared memory alone holds only 256 lanes of path state (about 200 words per lane at N = 1024).
// Tensor memory gives every lane 128 more columns: decision buffers descending from the top stage go there while
// they fit, the rest (and the LLR stages, ascending, while they fit) to shared memory, the largest LLR stages to the
// L2-resident arena as before. Falls back to plan_scl_i8_smem when the code is too long for this shape (decision
// words needed in the arena) or on request (PC_SCL_NOTMEM=1, PC_SCL_LANES).
int plan_scl_i8(int S, int K, int L, bool root_leaf, size_t smem_max, int sms, SclLaunch& out, const char** msg)
{
    if (L < 1 || L > 32) {
        *msg = "list size must be in [1, 32]";
        return 1;
    }
    const bool no_tmem = getenv("PC_SCL_NOTMEM") != nullptr || getenv("PC_SCL_LANES") != nullptr;
    int G = 1;
    while (G < L) G <<= 1;
    // shape: ONE CTA of 512 lanes whose warps stay in step (default), or PC_SCL_T=256: two CTAs of 256
    int T = 512;
    if (const char* e = getenv("PC_SCL_T")) T = atoi(e) == 256 ? 256 : (atoi(e) == 128 ? 128 : 512);
    const int per_sm = 512 / T, tm_alloc = 512 / per_sm, tm_cols = tm_alloc / ((T / 32 + 3) / 4); // 128 columns per thread
    const bool lockstep = getenv("PC_SCL_LOCKSTEP") ? atoi(getenv("PC_SCL_LOCKSTEP")) != 0 : true;
    const int top_llr = root_leaf ? S : S - 1;
    if (!no_tmem && S >= 8 && S <= 15) { // shorter codes fit shared memory whole (planner above)
        size_t per_cta = 233472 / (size_t)per_sm - 1024;
        if (per_cta > smem_max) per_cta = smem_max;
        const long budget = ((long)per_cta - 1024 - 64) / (4L * T); // shared-memory words per lane
        SclParams& p = out.p;
        for (int s = 0; s < 20; ++s) p.llr_off[s] = p.bitL_off[s] = p.bitR_off[s] = 0;
        // decision buffers, top stage first, into tensor memory while they fit
        int tcol = 0, tm_lo = S + 1;
        for (int s = S; s >= 6; --s) {
            const int bw = 1 << (s - 5);
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
        int aw = 0, smem_top = -1;
        bool open = true;
        for (int s = 0; s <= top_llr; ++s) {
            const int sw = s >= 2 ? (1 << (s - 2)) : 1;
            if (open && w + sw <= budget) {
                p.llr_off[s] = w;
                w += sw;
                smem_top = s;
            } else {
                open = false;
                p.llr_off[s] = kArena | aw;
                aw += sw;
            }
        }
        // scratch of N/32 words for extractBestPath: the (dead by then) LLR stage S-3 has exactly that size
        const int nbw = 1 << (S - 5);
        int scr = -1;
        if (S - 3 >= 2 && S - 3 <= smem_top) {
            scr = p.llr_off[S - 3];
        } else if (w + nbw <= budget) {
            scr = w;
            w += nbw;
        }
        const bool fits = w <= budget && tm_lo <= S && scr >= 0 && smem_top >= (top_llr < 6 ? top_llr : 6) &&
                          (unsigned long long)aw * ((unsigned long long)sms * per_sm * T) < (1ull << 32);
        if (fits) {
            p.S = S;
            p.K = K;
            p.L = L;
            p.G = G;
            p.root_leaf = root_leaf ? 1 : 0;
            p.words = w;
            p.arena_words = aw;
            p.tm_alloc_cols = tm_alloc;
            p.tm_cols_thread = tm_cols;
            p.scr_off = scr;
            p.lockstep = lockstep ? 1 : 0;
            p.sort2 = (G <= 8 && S <= 12 && !getenv("PC_SCL_SORT1")) ? 1 : 0; // 8 G <= 64 candidates, |metric| < 2^19
            out.threads = T;
            out.smem_bytes = 1024 + 4 * (size_t)w * T;
            out.grid = sms * per_sm;
            out.per_sm = per_sm;
            p.arena_stride = (size_t)out.grid * T;
            return 0;
        }
    }
    return plan_scl_i8_smem(S, K, L, root_leaf, smem_max, sms, out, msg);
}

} // namespace pc
