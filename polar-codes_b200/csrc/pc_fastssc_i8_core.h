// Fast-SSC int8 decoder core: one THREAD decodes one codeword by interpreting the node program of
// pc_program.h. Four int8 LLRs per 32-bit word (pc_swar.h); hard decisions are one bit per element.
//
// Storage per thread (see "where things live" below): LLR stage s (n = 2^s int8) is a run of 16-byte chunks in
// shared memory, tensor memory or the L2-resident arena, chosen per stage by the host planner; stages 4 and 5 are
// always shared-memory chunks 0 and 1..2; stages n = 2, 4, 8 are the four "small" words; hard decisions are
// bit-packed words (element i -> word i/32, bit i%32). The channel row is read straight from HBM.
// All threads of a warp execute the same node program with the same addresses (tcgen05.ld/st are warp-collective).
//
// Semantics follow reference src/polarcode/decoding/fastssc_fip_char.cpp (node decode functions
// :202-492) and include/polarcode/decoding/fip_char.h; the hard decision of a node is the MSB of the
// reference's "bit byte", which is all any consumer reads (g's blendv, CombineBits, info extraction).
//
// __host__ __device__ throughout: tests/host_emul.cpp runs this exact code for one emulated thread.
#ifndef PC_FASTSSC_I8_CORE_H
#define PC_FASTSSC_I8_CORE_H
#include "pc_crc.h"
#include "pc_program.h"
#include "pc_swar.h"

namespace pc {

struct alignas(16) u4 {
    uint32_t x, y, z, w;
};

// ---- where things live -------------------------------------------------------------------------------
// Every LLR stage s >= 4 (n = 2^s int8 values = 2^(s-4) chunks of 16 bytes) is placed by the host planner
// (pc_api.cu plan_fastssc) in one of three memories, all private to the thread:
//   PK_SMEM   shared memory, thread-interleaved chunks (one conflict-free LDS.128/STS.128 per warp access)
//   PK_TMEM   tensor memory used as plain per-thread storage: lane = thread, 4 columns per chunk, moved with
//             tcgen05.ld/st 32x32b.x4 (measured 459 B/clk/SM vs 128 B/clk/SM for shared memory, tools/tmem_bench.cu)
//   PK_ARENA  L2-resident HBM arena, thread-interleaved like shared memory (long codes only)
// The channel row itself is read straight from HBM. Stages s <= 3 and the hard-decision scratch word are in
// the "small" shared-memory words.
enum : uint32_t { PK_SMEM = 0, PK_TMEM = 1, PK_ARENA = 2 };
PC_HD uint32_t mk_place(uint32_t kind, uint32_t base) { return kind | (base << 2); }

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void tm_st4(uint32_t taddr, const u4& v)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void tm_ld4_nowait(uint32_t taddr, u4& v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tm_st1(uint32_t taddr, uint32_t v)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(v) : "memory");
}
__device__ __forceinline__ void tm_ld1_nowait(uint32_t taddr, uint32_t& v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
#else
// host emulation (tests/host_emul.cpp): "tensor memory" is a plain array of columns, the address its index
extern thread_local uint32_t* g_tm_emul;
inline void tm_st4(uint32_t taddr, const u4& v)
{
    g_tm_emul[taddr] = v.x; g_tm_emul[taddr + 1] = v.y; g_tm_emul[taddr + 2] = v.z; g_tm_emul[taddr + 3] = v.w;
}
inline void tm_ld4_nowait(uint32_t taddr, u4& v)
{
    v = u4{ g_tm_emul[taddr], g_tm_emul[taddr + 1], g_tm_emul[taddr + 2], g_tm_emul[taddr + 3] };
}
inline void tm_st1(uint32_t taddr, uint32_t v) { g_tm_emul[taddr] = v; }
inline void tm_ld1_nowait(uint32_t taddr, uint32_t& v) { v = g_tm_emul[taddr]; }
inline void tm_wait_ld() {}
inline void tm_wait_st() {}
#endif

// One stage (or the channel row) as seen by one thread
struct Mem {
    uint32_t* p;   // generic address of chunk 0 (shared, arena or channel row); unused for tensor memory
    int stride;    // words between consecutive chunks
    uint32_t tcol; // tensor-memory address of chunk 0
    bool tm;
    PC_HD u4 ld4(int i) const
    { // general accessor (leaves); the hot loops below use the templated forms
        u4 v;
        if (tm) {
            tm_ld4_nowait(tcol + 4u * (uint32_t)i, v);
            tm_wait_ld();
        } else {
            v = *reinterpret_cast<const u4*>(p + (size_t)i * stride);
        }
        return v;
    }
    PC_HD void st4(int i, const u4& v) const
    {
        if (tm) {
            tm_st4(tcol + 4u * (uint32_t)i, v);
            tm_wait_st();
        } else {
            *reinterpret_cast<u4*>(p + (size_t)i * stride) = v;
        }
    }
};

// per-thread state
struct Ctx {
    int S;             // log2 N
    uint32_t cur;      // hard decisions of the 32-element block being decoded (nodes with n <= 32)
    const uint32_t* place; // place[s], s = 4 .. S (mk_place)
    uint32_t* small;   // word j (0..3) at small[j * ws]: stages n = 2, 4, 8
    int ws;
    uint32_t* sm;      // this thread's slot of shared-memory chunk 0, chunk stride scs words
    int scs;
    uint32_t* ar;      // same for the arena
    size_t acs;
    uint32_t tm;       // tensor-memory address of this thread's column 0 (lane field of its warp + column base)
    // hard-decision words, 32 per word, element i -> word i/32 bit i%32
    uint32_t bits_kind;
    uint32_t* bits;    // word 0 (shared or arena), stride bws
    size_t bws;
    uint32_t tbits;    // tensor-memory address of word 0

    PC_HD Mem stage(int s) const
    {
        const uint32_t pl = place[s];
        const uint32_t kind = pl & 3u, base = pl >> 2;
        Mem m;
        m.tm = kind == PK_TMEM;
        m.tcol = tm + base;
        if (kind == PK_ARENA) {
            m.p = ar + (size_t)base * acs;
            m.stride = (int)acs;
        } else {
            m.p = sm + (size_t)base * scs;
            m.stride = scs;
        }
        return m;
    }
    PC_HD uint32_t sld(int j) const { return small[j * ws]; }
    PC_HD void sst(int j, uint32_t v) const { small[j * ws] = v; }
    PC_HD uint32_t bld(int word) const
    {
        if (bits_kind == PK_TMEM) {
            uint32_t v;
            tm_ld1_nowait(tbits + (uint32_t)word, v);
            tm_wait_ld();
            return v;
        }
        return bits[(size_t)word * bws];
    }
    PC_HD void bst(int word, uint32_t v) const
    {
        if (bits_kind == PK_TMEM) {
            tm_st1(tbits + (uint32_t)word, v);
            tm_wait_st();
        } else {
            bits[(size_t)word * bws] = v;
        }
    }
};

PC_HD uint32_t lowmask(int nbits) { return nbits >= 32 ? 0xFFFFFFFFu : ((1u << nbits) - 1u); }

PC_HD u4 f4(const u4& a, const u4& b) { return u4{ sw_f(a.x, b.x), sw_f(a.y, b.y), sw_f(a.z, b.z), sw_f(a.w, b.w) }; }
PC_HD u4 adds4(const u4& a, const u4& b)
{
    return u4{ sw_adds(a.x, b.x), sw_adds(a.y, b.y), sw_adds(a.z, b.z), sw_adds(a.w, b.w) };
}
// 16 hard decisions select add/sub per lane
PC_HD u4 g4(const u4& a, const u4& b, uint32_t bits16)
{
    return u4{ sw_g(a.x, b.x, sw_expand4(bits16)), sw_g(a.y, b.y, sw_expand4(bits16 >> 4)),
               sw_g(a.z, b.z, sw_expand4(bits16 >> 8)), sw_g(a.w, b.w, sw_expand4(bits16 >> 12)) };
}
PC_HD uint32_t movemask16(const u4& v)
{
    return sw_movemask(v.x) | (sw_movemask(v.y) << 4) | (sw_movemask(v.z) << 8) | (sw_movemask(v.w) << 12);
}

// stages 4 and 5 (always shared memory, chunks 0 and 1..2)
PC_HD u4 ld_s4(const Ctx& c) { return *reinterpret_cast<const u4*>(c.sm); }
PC_HD void st_s4(const Ctx& c, const u4& v) { *reinterpret_cast<u4*>(c.sm) = v; }
PC_HD u4 ld_s5(const Ctx& c, int i) { return *reinterpret_cast<const u4*>(c.sm + (size_t)(1 + i) * c.scs); }

// ------------------------------------------------------------------ internal nodes -----------------
// One step of a node with n = 2^s >= 64 elements: child = op(lo half, hi half) in batches of two 16-byte
// chunks (32 child LLRs = one hard-decision word). All loads of a batch are issued before the arithmetic.
//   BIG_F   child = f(lo, hi)                     RateRNode::decode   fastssc_fip_char.cpp:401-404
//   BIG_G   child = g(lo, hi, left decisions)     :406-408
//   BIG_G0  child = sat(lo + hi)                  ZeroRNode::decode   :476-483
//   BIG_G1  right child rate-1: decisions = sign(g), combine fused    ROneNode :427-449
enum { BIG_F = 0, BIG_G = 1, BIG_G0 = 2, BIG_G1 = 3 };

// Measured and NOT adopted (round 2, profiles/r02_fastssc_hints.txt): reading the two adjacent chunks of a channel-row
// batch as one 32-byte LDG.E.256 (212 Gbit/s against 213 with the extra code path compiled in), the same with
// L1::no_allocate + L2::evict_first (179: DRAM reads grow from 3.9 to 6.1 GB per 2^20 frames, because the second pass
// over a row -- g after the left half -- otherwise still finds most of it in L2), and an L2 evict-last policy on the arena
// chunks (212); merely compiling the three variants into this loop cost 8 % (231 -> 213) through instruction fetch.
// ST / DT (source / destination in tensor memory) are warp-uniform RUN-TIME flags on purpose: one copy of the
// arithmetic per OP keeps the hot code small. With the 16 specialised bodies (60-70 KB of hot SASS) the kernel's top
// stall was instruction fetch (profiles/r01_fastssc_i8_full_v3.txt); this alone took N=1024 from 201 to 233 Gbit/s.
// (Making OP a run-time value as well shrinks the code further but measured no faster: 230.)
template <int OP>
PC_HD void big_loop(const Ctx& c, const Mem& src, const Mem& dst, int s, int off)
{
    const bool ST = src.tm, DT = (OP != BIG_G1) && dst.tm;
    const int half = 1 << (s - 5), nb = half >> 1, wo = off >> 5;
    for (int j = 0; j < nb; ++j) {
        uint32_t bl = 0;
        if (OP == BIG_G || OP == BIG_G1) bl = c.bld(wo + j);
        u4 a0, a1, b0, b1;
        if (ST) {
            const uint32_t t = src.tcol + 8u * (uint32_t)j;
            tm_ld4_nowait(t, a0);
            tm_ld4_nowait(t + 4u, a1);
            tm_ld4_nowait(t + 4u * (uint32_t)half, b0);
            tm_ld4_nowait(t + 4u * (uint32_t)half + 4u, b1);
            tm_wait_ld();
        } else {
            const uint32_t* p = src.p + (size_t)(2 * j) * src.stride;
            a0 = *reinterpret_cast<const u4*>(p);
            a1 = *reinterpret_cast<const u4*>(p + src.stride);
            b0 = *reinterpret_cast<const u4*>(p + (size_t)half * src.stride);
            b1 = *reinterpret_cast<const u4*>(p + (size_t)(half + 1) * src.stride);
        }
        u4 r0, r1;
        if (OP == BIG_F) {
            r0 = f4(a0, b0);
            r1 = f4(a1, b1);
        } else if (OP == BIG_G0) {
            r0 = adds4(a0, b0);
            r1 = adds4(a1, b1);
        } else {
            r0 = g4(a0, b0, bl & 0xFFFFu);
            r1 = g4(a1, b1, bl >> 16);
        }
        if (OP == BIG_G1) {
            const uint32_t r = movemask16(r0) | (movemask16(r1) << 16);
            c.bst(wo + nb + j, r);
            c.bst(wo + j, bl ^ r);
        } else if (DT) {
            const uint32_t t = dst.tcol + 8u * (uint32_t)j;
            tm_st4(t, r0);
            tm_st4(t + 4u, r1);
        } else {
            uint32_t* q = dst.p + (size_t)(2 * j) * dst.stride;
            *reinterpret_cast<u4*>(q) = r0;
            *reinterpret_cast<u4*>(q + dst.stride) = r1;
        }
    }
    if (DT && OP != BIG_G1) tm_wait_st();
}

template <int OP>
PC_HD void big_op(Ctx& c, const Mem& src, int s, int off)
{
    if (s == 5) { // one chunk per half; the left decisions are still in the block register
        const u4 a = src.ld4(0), b = src.ld4(1);
        const uint32_t left = c.cur & 0xFFFFu;
        if (OP == BIG_F) {
            st_s4(c, f4(a, b));
        } else if (OP == BIG_G0) {
            st_s4(c, adds4(a, b));
        } else if (OP == BIG_G) {
            st_s4(c, g4(a, b, left));
        } else {
            const uint32_t r = movemask16(g4(a, b, left));
            c.cur = (left ^ r) | (r << 16);
        }
        return;
    }
    const Mem dst = c.stage(s - 1);
    big_loop<OP>(c, src, dst, s, off);
}

// the same steps for nodes with n <= 16 (sources: stage 4 chunk or the small words)
PC_HD void op_f_small(const Ctx& c, int s)
{
    if (s == 4) {
        u4 v = ld_s4(c);
        c.sst(2, sw_f(v.x, v.z));
        c.sst(3, sw_f(v.y, v.w));
    } else if (s == 3) {
        c.sst(1, sw_f(c.sld(2), c.sld(3)));
    } else { // s == 2
        uint32_t w = c.sld(1);
        c.sst(0, sw_f(w, w >> 16));
    }
}
PC_HD void op_g0_small(const Ctx& c, int s)
{
    if (s == 4) {
        u4 v = ld_s4(c);
        c.sst(2, sw_adds(v.x, v.z));
        c.sst(3, sw_adds(v.y, v.w));
    } else if (s == 3) {
        c.sst(1, sw_adds(c.sld(2), c.sld(3)));
    } else {
        uint32_t w = c.sld(1);
        c.sst(0, sw_adds(w, w >> 16));
    }
}
PC_HD void op_g_small(const Ctx& c, int s, int off)
{
    const uint32_t lb = c.cur >> (off & 31);
    if (s == 4) {
        u4 v = ld_s4(c);
        c.sst(2, sw_g(v.x, v.z, sw_expand4(lb)));
        c.sst(3, sw_g(v.y, v.w, sw_expand4(lb >> 4)));
    } else if (s == 3) {
        c.sst(1, sw_g(c.sld(2), c.sld(3), sw_expand4(lb)));
    } else {
        uint32_t w = c.sld(1);
        c.sst(0, sw_g(w, w >> 16, sw_expand4(lb & 3u)));
    }
}
PC_HD void op_g1_small(Ctx& c, int s, int off)
{
    const int pos = off & 31, h = 1 << (s - 1);
    const uint32_t lb = c.cur >> pos;
    uint32_t r;
    if (s == 4) {
        u4 v = ld_s4(c);
        r = sw_movemask(sw_g(v.x, v.z, sw_expand4(lb))) | (sw_movemask(sw_g(v.y, v.w, sw_expand4(lb >> 4))) << 4);
    } else if (s == 3) {
        r = sw_movemask(sw_g(c.sld(2), c.sld(3), sw_expand4(lb)));
    } else {
        uint32_t w = c.sld(1);
        r = sw_movemask(sw_g(w, w >> 16, sw_expand4(lb & 3u))) & 3u;
    }
    c.cur ^= r << pos;       // left ^= right
    c.cur |= r << (pos + h); // right (those positions were still zero)
}

// CombineInPlace / CombineBitsShort (fip_char.h:164-207) on one bit per element
PC_HD void op_comb(Ctx& c, int s, int off, bool zero_left)
{
    if (s >= 6) {
        const int wo = off >> 5, hw = 1 << (s - 6);
        for (int j = 0; j < hw; ++j) {
            uint32_t r = c.bld(wo + hw + j);
            c.bst(wo + j, zero_left ? r : (c.bld(wo + j) ^ r));
        }
    } else {
        const int pos = off & 31, h = 1 << (s - 1);
        uint32_t r = (c.cur >> (pos + h)) & lowmask(h);
        c.cur ^= r << pos; // left bits are zero when zero_left, so xor == copy
    }
}

// ------------------------------------------------------------------ leaves --------------------------

// LLRs of a node with n = 2^S <= 32 elements into x[0 .. max(1, n/4)); unused high bytes zeroed for n = 2
template <int S>
PC_HD void load_small(const Ctx& c, uint32_t (&x)[8])
{
    if (S == 5) {
        u4 a = ld_s5(c, 0), b = ld_s5(c, 1);
        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
        x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
    } else if (S == 4) {
        u4 a = ld_s4(c);
        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
    } else if (S == 3) {
        x[0] = c.sld(2);
        x[1] = c.sld(3);
    } else if (S == 2) {
        x[0] = c.sld(1);
    } else {
        x[0] = c.sld(0) & 0xFFFFu;
    }
}
// x = sat(lo + hi) of a node with n = 2^S <= 32 elements: n/2 lanes in x[0 .. max(1, n/8))
template <int S>
PC_HD void load_small_sum(const Ctx& c, uint32_t (&x)[8])
{
    if (S == 5) {
        u4 r = adds4(ld_s5(c, 0), ld_s5(c, 1));
        x[0] = r.x; x[1] = r.y; x[2] = r.z; x[3] = r.w;
    } else if (S == 4) {
        u4 a = ld_s4(c);
        x[0] = sw_adds(a.x, a.z);
        x[1] = sw_adds(a.y, a.w);
    } else if (S == 3) {
        x[0] = sw_adds(c.sld(2), c.sld(3));
    } else { // S == 2: two lanes
        uint32_t w = c.sld(1);
        x[0] = sw_adds(w, w >> 16) & 0xFFFFu;
    }
}

// reduce_adds_epi8 (avxconvenience.h:92-101) over NW words = 4*NW lanes, zero lanes above: the pairing
// (i,i+16) (i,i+8) (i,i+4) (i,i+2) (0,1) is normative because saturation is not associative.
// Returns the word after the stride-4 level (lanes 0..3 live); callers finish with sw_hreduce2/1.
template <int NW>
PC_HD uint32_t reduce_words(uint32_t (&x)[8])
{
    if (NW == 8) {
        x[0] = sw_adds(x[0], x[4]); x[1] = sw_adds(x[1], x[5]);
        x[2] = sw_adds(x[2], x[6]); x[3] = sw_adds(x[3], x[7]);
    }
    if (NW >= 4) {
        x[0] = sw_adds(x[0], x[2]);
        x[1] = sw_adds(x[1], x[3]);
    }
    if (NW >= 2) x[0] = sw_adds(x[0], x[1]);
    return x[0];
}

// per-lane saturating accumulation of the n/32 vectors of a long node, in order
// (RepetitionDecoder::decode fastssc_fip_char.cpp:225-236)
PC_HD void accumulate_vectors(const Ctx& c, int s, uint32_t (&acc)[8])
{
    const int nv = 1 << (s - 5);
    const Mem m = c.stage(s);
    u4 a = m.ld4(0), b = m.ld4(1);
    acc[0] = a.x; acc[1] = a.y; acc[2] = a.z; acc[3] = a.w;
    acc[4] = b.x; acc[5] = b.y; acc[6] = b.z; acc[7] = b.w; // sat(0 + x) == x
    for (int v = 1; v < nv; ++v) {
        a = m.ld4(2 * v);
        b = m.ld4(2 * v + 1);
        acc[0] = sw_adds(acc[0], a.x); acc[1] = sw_adds(acc[1], a.y);
        acc[2] = sw_adds(acc[2], a.z); acc[3] = sw_adds(acc[3], a.w);
        acc[4] = sw_adds(acc[4], b.x); acc[5] = sw_adds(acc[5], b.y);
        acc[6] = sw_adds(acc[6], b.z); acc[7] = sw_adds(acc[7], b.w);
    }
}

PC_HD void store_block_bits(Ctx& c, int s, int off, uint32_t word)
{ // same 32-bit pattern for every word of a node with n >= 64
    const int wo = off >> 5, nw = 1 << (s - 5);
    for (int j = 0; j < nw; ++j) c.bst(wo + j, word);
}

// RepetitionDecoder / ShortRepetitionDecoder (fastssc_fip_char.cpp:225-240,265-272)
template <int S>
PC_HD void leaf_rep_small(Ctx& c, int off)
{
    uint32_t x[8];
    load_small<S>(c, x);
    uint32_t t = reduce_words<(S >= 2 ? (1 << (S - 2)) : 1)>(x);
    t = sw_hreduce1(sw_hreduce2(t));
    if (sw_byte0_negative(t)) c.cur |= lowmask(1 << S) << (off & 31);
}
PC_HD void leaf_rep_big(Ctx& c, int s, int off)
{
    uint32_t acc[8];
    accumulate_vectors(c, s, acc);
    uint32_t t = sw_hreduce1(sw_hreduce2(reduce_words<8>(acc)));
    store_block_bits(c, s, off, sw_byte0_negative(t) ? 0xFFFFFFFFu : 0u);
}

// DoubleRepetitionDecoder (fastssc_fip_char.cpp:249-263), n >= 32: half_reduce_adds_epi8 stops at stride 2,
// even elements take the sign of the even-lane sum, odd elements of the odd-lane sum
PC_HD void leaf_drep(Ctx& c, int s, int off)
{
    uint32_t acc[8];
    accumulate_vectors(c, s, acc);
    uint32_t t = sw_hreduce2(reduce_words<8>(acc));
    uint32_t word = (((t >> 7) & 1u) ? 0x55555555u : 0u) | (((t >> 15) & 1u) ? 0xAAAAAAAAu : 0u);
    if (s == 5)
        c.cur = word;
    else
        store_block_bits(c, s, off, word);
}

// RateOneDecoder (fastssc_fip_char.cpp:210-215): hard decision = sign (LLR 0 -> bit 0)
template <int S>
PC_HD void leaf_r1_small(Ctx& c, int off)
{
    uint32_t x[8];
    load_small<S>(c, x);
    constexpr int NW = S >= 2 ? (1 << (S - 2)) : 1;
    uint32_t bits = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) bits |= sw_movemask(x[w]) << (4 * w);
    c.cur |= (bits & lowmask(1 << S)) << (off & 31);
}
PC_HD void leaf_r1_big(Ctx& c, int s, int off)
{
    const int wo = off >> 5, nw = 1 << (s - 5);
    const Mem m = c.stage(s);
    for (int j = 0; j < nw; ++j) c.bst(wo + j, movemask16(m.ld4(2 * j)) | (movemask16(m.ld4(2 * j + 1)) << 16));
}

// minpos_epu8 (src/polarcode/avxconvenience.cpp:13-56) restated on keys: value*32 + lane in 16-bit
// fields, so an unsigned 16-bit minimum returns the FIRST lane holding the smallest |LLR|.
PC_HD uint32_t minu16x2(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    return __vminu2(a, b);
#else
    uint32_t lo = (a & 0xFFFFu) < (b & 0xFFFFu) ? (a & 0xFFFFu) : (b & 0xFFFFu);
    uint32_t hi = (a >> 16) < (b >> 16) ? (a >> 16) : (b >> 16);
    return lo | (hi << 16);
#endif
}
// folds word w (lanes 4w..4w+3) of |LLR| bytes into the running key minimum
PC_HD uint32_t minkey_word(uint32_t kmin, uint32_t absw, int w)
{
    const uint32_t ce = (uint32_t)(4 * w) | ((uint32_t)(4 * w + 2) << 16);
    const uint32_t co = (uint32_t)(4 * w + 1) | ((uint32_t)(4 * w + 3) << 16);
    kmin = minu16x2(kmin, ((absw << 5) & 0x1FE01FE0u) | ce);
    kmin = minu16x2(kmin, ((absw >> 3) & 0x1FE01FE0u) | co);
    return kmin;
}
PC_HD uint32_t minkey_finish(uint32_t kmin)
{
    uint32_t lo = kmin & 0xFFFFu, hi = kmin >> 16;
    return lo < hi ? lo : hi; // (min |LLR|) * 32 + first lane
}
PC_HD uint32_t parity_of_msbs(uint32_t p)
{
    p ^= p >> 16;
    p ^= p >> 8;
    return (p >> 7) & 1u;
}

// ShortSpcDecoder (fastssc_fip_char.cpp:305-319) on NW words of real lanes (pad lanes hold 127 in the
// reference: never the unique minimum, never change the parity). Returns the hard decisions.
// The flip is a two's complement negate of the stored byte (:316), so 0 and -128 keep their sign.
template <int NW>
PC_HD uint32_t spc_short_bits(const uint32_t (&x)[8])
{
    uint32_t bits = 0, par = 0, kmin = 0xFFFFFFFFu;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        bits |= sw_movemask(x[w]) << (4 * w);
        par ^= x[w];
        kmin = minkey_word(kmin, sw_abs(x[w]), w);
    }
    uint32_t k = minkey_finish(kmin), m = k >> 5, idx = k & 31u;
    uint32_t flip = parity_of_msbs(par) & (m != 0u) & (m != 128u);
    return bits ^ (flip << idx);
}
template <int S>
PC_HD void leaf_spc_small(Ctx& c, int off)
{
    uint32_t x[8];
    load_small<S>(c, x);
    c.cur |= spc_short_bits<(1 << (S - 2))>(x) << (off & 31);
}

// SpcDecoder (n > 32, fastssc_fip_char.cpp:274-303) and ZeroSpcDecoder (n >= 64, :321-359) share the
// cross-vector search: minAbs starts at 127, a 32-lane vector is only searched while minAbs > 0 and only
// a strictly smaller per-vector minimum (signed-char compare) moves minIdx; the flip is a negate.
struct SpcSearch {
    uint32_t par = 0;
    int minAbs = 127, minIdx = 0;
    uint32_t flipOk = 0;
    PC_HD void vector(const uint32_t (&x)[8], int v)
    {
        uint32_t kmin = 0xFFFFFFFFu;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            par ^= x[w];
            kmin = minkey_word(kmin, sw_abs(x[w]), w);
        }
        if (v == 0) { // minIdx stays 0 when no vector has a minimum below 127
            uint32_t b0 = x[0] & 0xFFu;
            flipOk = (b0 != 0u) & (b0 != 0x80u);
        }
        uint32_t k = minkey_finish(kmin), m = k >> 5;
        int testAbs = m == 128u ? -128 : (int)m; // value is returned through a char* (avxconvenience.cpp:51)
        if (minAbs > 0 && testAbs < minAbs) {
            minAbs = testAbs;
            minIdx = (int)(k & 31u) + 32 * v;
            flipOk = (m != 0u) & (m != 128u);
        }
    }
    PC_HD bool flips() const { return (parity_of_msbs(par) & flipOk) != 0u; }
};

// Applies the SPC flip to the stored decisions (and to the copy `dup` words further on for ZeroSpc). With the
// decision words in tensor memory every access must use a warp-uniform address, so all nv words are rewritten.
PC_HD void spc_flip(const Ctx& c, const SpcSearch& q, int wo, int nv, int dup)
{
    const bool fl = q.flips();
    if (c.bits_kind == PK_TMEM) {
        const uint32_t mask = fl ? (1u << (q.minIdx & 31)) : 0u;
        const int fw = q.minIdx >> 5;
        for (int v = 0; v < nv; ++v) {
            const uint32_t nb = c.bld(wo + v) ^ (v == fw ? mask : 0u);
            c.bst(wo + v, nb);
            if (dup) c.bst(wo + v + dup, nb);
        }
    } else if (fl) {
        const int w = wo + (q.minIdx >> 5);
        const uint32_t nb = c.bld(w) ^ (1u << (q.minIdx & 31));
        c.bst(w, nb);
        if (dup) c.bst(w + dup, nb);
    }
}

PC_HD void leaf_spc_big(Ctx& c, int s, int off)
{
    const int wo = off >> 5, nv = 1 << (s - 5);
    const Mem m = c.stage(s);
    SpcSearch q;
    for (int v = 0; v < nv; ++v) {
        u4 a = m.ld4(2 * v), b = m.ld4(2 * v + 1);
        uint32_t x[8] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
        c.bst(wo + v, movemask16(a) | (movemask16(b) << 16));
        q.vector(x, v);
    }
    spc_flip(c, q, wo, nv, 0);
}

// ShortZeroOneDecoder (fastssc_fip_char.cpp:390-399): x = sat(lo + hi), bits = [x, x]
template <int S>
PC_HD void leaf_s01(Ctx& c, int off)
{
    uint32_t x[8];
    load_small_sum<S>(c, x);
    constexpr int H = 1 << (S - 1);
    constexpr int NW = H >= 4 ? H / 4 : 1;
    uint32_t bits = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) bits |= sw_movemask(x[w]) << (4 * w);
    bits &= lowmask(H);
    c.cur |= (bits | (bits << H)) << (off & 31);
}

// ShortZeroSpcDecoder (fastssc_fip_char.cpp:361-388), 8 <= n <= 32
template <int S>
PC_HD void leaf_0spc_small(Ctx& c, int off)
{
    uint32_t x[8];
    load_small_sum<S>(c, x);
    constexpr int H = 1 << (S - 1);
    uint32_t bits = spc_short_bits<H / 4>(x);
    c.cur |= (bits | (bits << H)) << (off & 31);
}
// ZeroSpcDecoder (fastssc_fip_char.cpp:321-359), n >= 64
template <class Src>
PC_HD void leaf_0spc_big(Ctx& c, const Src& src, int s, int off)
{
    const int half = 1 << (s - 5), wo = off >> 5, nv = half >> 1; // n/2 elements = nv vectors of 32
    SpcSearch q;
    for (int v = 0; v < nv; ++v) {
        u4 a = adds4(src.ld4(2 * v), src.ld4(half + 2 * v));
        u4 b = adds4(src.ld4(2 * v + 1), src.ld4(half + 2 * v + 1));
        uint32_t x[8] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
        uint32_t bits = movemask16(a) | (movemask16(b) << 16);
        c.bst(wo + v, bits);
        c.bst(wo + nv + v, bits);
        q.vector(x, v);
    }
    spc_flip(c, q, wo, nv, nv);
}

// RateZeroDecoder (fastssc_fip_char.cpp:202-208): all hard decisions 0
PC_HD void leaf_r0(Ctx& c, int s, int off)
{
    if (s >= 6) store_block_bits(c, s, off, 0u); // n <= 32: the block word already holds zeros there
}

// ------------------------------------------------------------------ interpreter --------------------

// Copies the channel row into stage S (FastSscProgram::staged_root: the root is a leaf or N < 32). Rows of
// N < 16 bytes are not 16-byte aligned in HBM, so those go byte by byte.
PC_HD void stage_root(const Ctx& c, const int8_t* row)
{
    const int S = c.S;
    if (S >= 4) {
        const int nc = 1 << (S - 4);
        const Mem m = c.stage(S);
        const u4* g = reinterpret_cast<const u4*>(row);
        for (int i = 0; i < nc; ++i) m.st4(i, g[i]);
    } else {
        uint32_t w[2] = { 0u, 0u };
        for (int i = 0; i < (1 << S); ++i) w[i >> 2] |= (uint32_t)(uint8_t)row[i] << (8 * (i & 3));
        if (S == 3) {
            c.sst(2, w[0]);
            c.sst(3, w[1]);
        } else {
            c.sst(S - 1, w[0]); // S == 2 -> word 1, S == 1 -> word 0
        }
    }
}

template <int S>
PC_HD void leaf_small(Ctx& c, uint32_t op, int off)
{
    switch (op) {
    case OP_R1: leaf_r1_small<S>(c, off); break;
    case OP_REP: leaf_rep_small<S>(c, off); break;
    case OP_SPC:
        if (S >= 2) leaf_spc_small<(S >= 2 ? S : 2)>(c, off);
        break;
    case OP_S01:
        if (S >= 2) leaf_s01<(S >= 2 ? S : 2)>(c, off);
        break;
    case OP_0SPC:
        if (S >= 3) leaf_0spc_small<(S >= 3 ? S : 3)>(c, off);
        break;
    default: break; // OP_R0: nothing to do
    }
}

// Runs the node program for one codeword. `row` is the channel LLR row (N int8) in HBM.
PC_HD void fastssc_i8_run(Ctx& c, const uint32_t* code, const int8_t* row, bool staged_root)
{
    const int S = c.S;
    c.cur = 0u;
    if (staged_root) stage_root(c, row);
    Mem groot;
    groot.p = reinterpret_cast<uint32_t*>(const_cast<int8_t*>(row));
    groot.stride = 4;
    groot.tcol = 0;
    groot.tm = false;
    for (int pc = 0;; ++pc) {
        const uint32_t ins = code[pc];
        const uint32_t op = instr_op(ins);
        const int s = (int)instr_s(ins), off = (int)instr_off(ins);
        if (op == OP_END) break;
        if (op <= OP_G1) { // internal steps
            if (s >= 5) {
                const Mem src = (s == S && !staged_root) ? groot : c.stage(s);
                switch (op) {
                case OP_F: big_op<BIG_F>(c, src, s, off); break;
                case OP_G: big_op<BIG_G>(c, src, s, off); break;
                case OP_G0: big_op<BIG_G0>(c, src, s, off); break;
                default: big_op<BIG_G1>(c, src, s, off); break;
                }
            } else {
                switch (op) {
                case OP_F: op_f_small(c, s); break;
                case OP_G: op_g_small(c, s, off); break;
                case OP_G0: op_g0_small(c, s); break;
                default: op_g1_small(c, s, off); break;
                }
            }
            continue;
        }
        switch (op) {
        case OP_COMB: op_comb(c, s, off, false); break;
        case OP_COMB0: op_comb(c, s, off, true); break;
        case OP_FLUSH:
            c.bst(off >> 5, c.cur);
            c.cur = 0u;
            break;
        default: // leaves
            if (s >= 6 || (s == 5 && op == OP_DREP)) {
                switch (op) {
                case OP_R0: leaf_r0(c, s, off); break;
                case OP_R1: leaf_r1_big(c, s, off); break;
                case OP_REP: leaf_rep_big(c, s, off); break;
                case OP_DREP: leaf_drep(c, s, off); break;
                case OP_SPC: leaf_spc_big(c, s, off); break;
                case OP_0SPC: leaf_0spc_big(c, (s == S && !staged_root) ? groot : c.stage(s), s, off); break;
                default: break;
                }
            } else {
                switch (s) {
                case 5: leaf_small<5>(c, op, off); break;
                case 4: leaf_small<4>(c, op, off); break;
                case 3: leaf_small<3>(c, op, off); break;
                case 2: leaf_small<2>(c, op, off); break;
                default: leaf_small<1>(c, op, off); break;
                }
            }
            break;
        }
    }
}

// ------------------------------------------------------------------ output --------------------------

// Non-systematic codes: the information is u = x F^{(x)n} (fastssc_fip_char.cpp:619-622 re-encodes the
// hard decisions with the non-systematic butterfly). In-place on the bit words: stage d xors element
// i + 2^d into element i (reference src/polarcode/encoding/butterfly_fip.cpp:15-63).
PC_HD void transform_bits_inplace(const Ctx& c)
{
    const int nw = c.S > 5 ? (1 << (c.S - 5)) : 1;
    const uint32_t lim = lowmask(1 << (c.S < 5 ? c.S : 5));
    for (int w = 0; w < nw; ++w) {
        uint32_t x = c.bld(w);
        x ^= (x >> 1) & 0x55555555u & lim;
        if (c.S > 1) x ^= (x >> 2) & 0x33333333u & lim;
        if (c.S > 2) x ^= (x >> 4) & 0x0F0F0F0Fu & lim;
        if (c.S > 3) x ^= (x >> 8) & 0x00FF00FFu & lim;
        if (c.S > 4) x ^= (x >> 16) & 0x0000FFFFu;
        c.bst(w, x);
    }
    for (int d = 1; d < nw; d <<= 1)
        for (int g = 0; g < nw; g += 2 * d)
            for (int i = g; i < g + d; ++i) c.bst(i, c.bld(i) ^ c.bld(i + d));
}

// element i%32 in bit i%32 (LSB first) -> the reference's MSB-first bytes as a little-endian word
PC_HD uint32_t msb_first_word(uint32_t w)
{
#ifdef __CUDA_ARCH__
    return __byte_perm(__brev(w), 0, 0x0123);
#else
    uint32_t r = 0;
    for (int i = 0; i < 32; ++i)
        if (w & (1u << i)) r |= 1u << (8 * (i >> 3) + 7 - (i & 7));
    return r;
#endif
}

// CharContainer::getPackedInformationBits (src/polarcode/bitcontainer.cpp:536-564) through the run list,
// then Detector::check over the ceil(K/8) output bytes (fastssc_fip_char.cpp:627-629).
// Returns the detector verdict. `crc_table` may live in shared memory.
template <class Rd>
PC_HD bool extract_info_rd(const Rd& rd, const uint32_t* runs, int nruns, int K, uint8_t* out_row, const CrcSpec& crc,
                           const uint32_t* crc_table, bool store)
{
    const int kb = (K + 7) >> 3;
    const bool word_stores = (kb & 3) == 0;
    const int payload = crc.reflected ? 4 * ((kb >> 2) - 1) : kb - crc.bytes;
    CrcStream cs;
    cs.begin(crc);
    uint32_t acc = 0u;
    int ob = 0; // output byte index
    for (int r = 0; r < nruns; ++r) {
        const uint32_t run = runs[r];
        const uint32_t sword = run & 0x7FFu, sshift = (run >> 11) & 31u, len = ((run >> 16) & 31u) + 1u,
                       dshift = (run >> 21) & 31u, last = (run >> 26) & 1u;
        acc |= ((rd((int)sword) >> sshift) & lowmask((int)len)) << dshift;
        if (last) {
            const uint32_t w = msb_first_word(acc);
            const int nb = (kb - ob) < 4 ? (kb - ob) : 4;
            if (!store) {
            } else if (word_stores) {
                *reinterpret_cast<uint32_t*>(out_row + ob) = w;
            } else {
                for (int j = 0; j < nb; ++j) out_row[ob + j] = (uint8_t)(w >> (8 * j));
            }
            if (crc.width) {
                for (int j = 0; j < nb; ++j) {
                    const uint32_t byte = (w >> (8 * j)) & 0xFFu;
                    const int idx = ob + j;
                    if (idx < payload)
                        cs.payload_byte(crc, crc_table, byte);
                    else if (idx < payload + crc.bytes)
                        cs.stored_byte(byte);
                }
            }
            ob += nb;
            acc = 0u;
        }
    }
    return cs.ok(crc);
}

struct CtxBitsReader {
    const Ctx& c;
    PC_HD uint32_t operator()(int word) const { return c.bld(word); }
};
PC_HD bool extract_info(const Ctx& c, const uint32_t* runs, int nruns, int K, uint8_t* out_row, const CrcSpec& crc,
                        const uint32_t* crc_table, bool store = true)
{
    return extract_info_rd(CtxBitsReader{ c }, runs, nruns, K, out_row, crc, crc_table, store);
}

} // namespace pc
#endif
