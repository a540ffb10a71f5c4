// Packed polar encoder, register-block form, for sm_100a (32 <= N <= 32768). Replaces Encoding::ButterflyFipPacked::encode
// of the reference (src/polarcode/encoding/butterfly_fip_packed.cpp:45-59) for B codewords per launch: scatter the K
// info bits to the non-frozen positions (PackedContainer::insertPackedInformationBits, src/polarcode/bitcontainer.cpp:
// 886-898), transform x = u F^(x)n (butterfly_fip.cpp:15-63) and, for systematic codes, zero the frozen positions
// (resetFrozenBits, bitcontainer.cpp:1004-1010) and transform again.
//
// Why a second kernel: k_encode (pc_encode.cu) gives every lane ONE code word, so five of the ten stages of an
// N = 1024 transform are warp shuffles (three instructions per word and stage) and the bit deposit is a per-lane loop
// over the runs of the lane's own mask; it measured 8 % of HBM bandwidth for an op whose algorithmic traffic is just
// K/8 + N/8 bytes per codeword. Here a THREAD owns R = min(N/32, 32) consecutive code words in registers:
//   * stages 0..4 inside a word: x ^= (x << 2^d) & mask_d on MSB-first words (element i = bit 31 - i%32, i.e. the
//     byte-swapped memory word), so the shift is a left shift (IMAD.SHL, FMA pipe) next to one LOP3 (ALU pipe);
//   * stages 5..9 across the thread's registers: one XOR per word pair, no shuffles;
//   * stages >= 10 (N > 1024): the 32 lanes of a warp hold 32 DIFFERENT codewords and the warps of a group hold the
//     word blocks, so the exchange is a shared-memory transpose to a second register layout (words r * TPC + t)
//     in which the top five stages are register XORs again; the butterfly stages commute, so the systematic form
//     runs  A: 0..9 -> B: 10.. -> mask -> B: 10.. -> A: 0..9  with two transposes;
//   * the deposit of a word is Hacker's-Delight "expand" with five precomputed move masks per word (five shift +
//     bit-select pairs, no data-dependent loop); masks and offsets sit in __constant__ memory and are the same for
//     all lanes of a warp (the word index depends on the register index and the warp only), so they are constant-
//     bank operands, not loads;
//   * info rows are staged coalesced into shared memory (byte-swapped on the way), code rows are stored by their
//     owner thread as whole 32-byte sectors.
#include <cstring>
#include <mutex>

#include "pc_fastssc_i8_core.h"
#include "pc_internal.h"

namespace pc {

constexpr int kEncMaxWords = 512;                  // N <= 16384
__constant__ uint32_t c_enc[8 * kEncMaxWords];     // per code word: mask, q1, pos, mv[0..4] (see build_table)

struct Enc2Params {
    const uint8_t* info;  // [B][kb] (check bytes already in place)
    uint8_t* code;        // [B][N/8]
    unsigned long long B;
    int S, kb, kw;        // kw: words per info row, ceil(kb / 4)
    int tpc;              // threads (= warps of a group) per codeword, N / (32 R)
    int systematic;
    int vec_in;           // info rows are 16-byte aligned and every thread's share is whole 16-byte pieces
    int v8;               // code rows are 32-byte aligned: 256-bit stores
    const uint32_t* tab;  // the code's table in HBM (multi-warp codewords read it through L1: 16 warps walking 16 KB of
                          // __constant__ memory with different offsets miss the constant cache on every access)
    CrcSpec crc;          // fused Detector::generate (single-thread codewords only; width 0 = none / done before)
    const uint32_t* crc_table;
};

__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

// in-word stages on an MSB-first word (element i at bit 31 - i): element i takes element i + 2^d
__device__ __forceinline__ uint32_t inword_msb(uint32_t x)
{
    x ^= (x << 1) & 0xAAAAAAAAu;
    x ^= (x << 2) & 0xCCCCCCCCu;
    x ^= (x << 4) & 0xF0F0F0F0u;
    x ^= (x << 8) & 0xFF00FF00u;
    x ^= (x << 16);
    return x;
}

template <int R>
__device__ __forceinline__ void stages_inreg(uint32_t (&x)[R], int first_bit)
{
    // register-index bit b pairs registers r and r | (1 << b); stages below first_bit were done in the other layout
#pragma unroll
    for (int b = 0; (1 << b) < R; ++b) {
        if (b >= first_bit) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (!(r & (1 << b))) x[r] ^= x[r | (1 << b)];
        }
    }
}

// bit j of src goes to the j-th lowest set bit of the mask (Hacker's Delight 7-5, "expand"), move masks precomputed
__device__ __forceinline__ uint32_t expand32(uint32_t x, const uint32_t* mv, uint32_t m)
{
#pragma unroll
    for (int i = 4; i >= 0; --i) {
        const uint32_t t = x << (1 << i);
        x = (x & ~mv[i]) | (t & mv[i]);
    }
    return x & m;
}

template <int R, bool MULTI> // MULTI: several warps per codeword (N > 1024); otherwise the word block index is 0
__global__ void __launch_bounds__(512, 2) k_encode2(const __grid_constant__ Enc2Params p)
{
    extern __shared__ __align__(16) uint32_t sm[];
    const int T = blockDim.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tpc = MULTI ? p.tpc : 1, grp = warp / tpc, t = MULTI ? warp - grp * tpc : 0; // group of tpc warps = 32 codewords
    const int gpc = (T >> 5) / tpc, cpc = 32 * gpc;                // codewords per CTA
    const int W = R * tpc, Wp = W + 1;                             // words per codeword (+1: conflict-free transposes)
    // shared memory: the staged info rows [kw][cpc] (word q of codeword slot at q * cpc + slot); MULTI: the transpose
    // buffer [cpc][Wp] ALIASES them (the rows are dead once deposited) and the code's table (8 W words) follows;
    // otherwise the detector table (256 words) follows the rows
    uint32_t* s_info = sm;
    uint32_t* s_x = sm;
    const size_t head = MULTI ? ((size_t)cpc * Wp > (size_t)cpc * p.kw ? (size_t)cpc * Wp : (size_t)cpc * p.kw) : (size_t)cpc * p.kw;
    uint32_t* s_tab = sm + head;                                   // MULTI: 8 * W words
    uint32_t* s_crc = sm + head;                                   // !MULTI: 256 words (fused detector only)
    if (MULTI) {
        for (int i = threadIdx.x; i < 8 * W; i += T) s_tab[i] = __ldg(p.tab + i);
    } else if (p.crc.width) {
        for (int i = threadIdx.x; i < 256; i += T) s_crc[i] = p.crc_table[i];
    }
    const int slot = grp * 32 + lane;
    const int nbytes = 4 * W;
    int ltpc = 0;
    while ((1 << ltpc) < tpc) ++ltpc;
    const int share = (p.kw + tpc - 1) / tpc; // info words of its row that one thread stages
    for (unsigned long long base = (unsigned long long)blockIdx.x * cpc; base < p.B; base += (unsigned long long)gridDim.x * cpc) {
        const unsigned long long left = p.B - base;
        const int ncw = left < (unsigned long long)cpc ? (int)left : cpc;
        const bool live = slot < ncw;
        // ---- stage the info rows: thread (slot, t) loads words [t share, (t+1) share) of ITS row -- a warp's 16-byte
        // loads cover whole 32-byte sectors between them -- byte-swaps them so that a word is MSB-first (info bit k =
        // bit 31 - k % 32) and stores them word-interleaved (conflict-free for the stores and for the reads below)
        __syncthreads();
        {
            const uint8_t* rowp = p.info + (base + (live ? slot : 0)) * (unsigned long long)p.kb;
            uint32_t* dst = s_info + slot;
            const int q0 = t * share, q1e = q0 + share < p.kw ? q0 + share : p.kw;
            if (p.vec_in) {
                const uint4* r4 = reinterpret_cast<const uint4*>(rowp) + (q0 >> 2);
                for (int q = q0; q < q1e; q += 16) { // up to four 16-byte loads in flight
                    uint4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (q + 4 * u < q1e) v[u] = __ldg(r4 + ((q - q0) >> 2) + u);
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (q + 4 * u < q1e) {
                            dst[(q + 4 * u) * cpc] = bswap32(v[u].x);
                            dst[(q + 4 * u + 1) * cpc] = bswap32(v[u].y);
                            dst[(q + 4 * u + 2) * cpc] = bswap32(v[u].z);
                            dst[(q + 4 * u + 3) * cpc] = bswap32(v[u].w);
                        }
                }
            } else {
                for (int q = q0; q < q1e; ++q) {
                    uint32_t w = 0;
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                        if (4 * q + b < p.kb) w |= (uint32_t)rowp[4 * q + b] << (24 - 8 * b);
                    dst[q * cpc] = w;
                }
            }
        }
        __syncthreads();
        uint32_t* myinfo = s_info + slot;
        if (!MULTI && p.crc.width) {
            // Detector::generate fused (ButterflyFipPacked::encode runs it on the info block first,
            // butterfly_fip_packed.cpp:47-49): the row's own thread runs the table-driven CRC over its staged words
            // (MSB-first, so byte b of the row is bits 31-8(b%4).. of word b/4) and writes the check bytes into them
            CrcStream cs;
            cs.begin(p.crc);
            const int payload = p.crc.reflected ? 4 * ((p.kb >> 2) - 1) : p.kb - p.crc.bytes;
            for (int q = 0; 4 * q < payload; ++q) {
                const uint32_t w = myinfo[q * cpc];
#pragma unroll
                for (int b = 0; b < 4; ++b)
                    if (4 * q + b < payload) cs.payload_byte(p.crc, s_crc, (w >> (24 - 8 * b)) & 0xFFu);
            }
            const uint32_t val = cs.value(p.crc);
            for (int i = 0; i < p.crc.bytes; ++i) {
                const uint32_t byte = p.crc.reflected ? (val >> (8 * i)) & 0xFFu : (val >> (8 * (p.crc.bytes - 1 - i))) & 0xFFu;
                const int pos = payload + i, sh = 24 - 8 * (pos & 3);
                uint32_t* wp = myinfo + (pos >> 2) * cpc;
                *wp = (*wp & ~(0xFFu << sh)) | (byte << sh);
            }
        }
        // ---- deposit: code word w takes the info bits ending at stream bit e - 1, right-aligned, expanded onto its mask
        uint32_t x[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            // table entry of code word t R + r: __constant__ operands for single-thread codewords, the shared-memory copy
            // otherwise (the same address for all lanes of a warp: one broadcast read)
            const uint32_t* Eg = s_tab + 8 * (t * R + r);
            const uint32_t* Ec = c_enc + 8 * r;
            const uint32_t m = MULTI ? Eg[0] : Ec[0];
            uint32_t v = 0;
            if (m) { // uniform
                const int q1 = (int)(MULTI ? Eg[1] : Ec[1]), pos = (int)(MULTI ? Eg[2] : Ec[2]);
                const uint32_t lo = myinfo[q1 * cpc], hi = q1 ? myinfo[(q1 - 1) * cpc] : 0u;
                v = __funnelshift_r(lo, hi, pos);
                if (m != 0xFFFFFFFFu) {
#pragma unroll
                    for (int i = 4; i >= 0; --i) {
                        const uint32_t mv = MULTI ? Eg[3 + i] : Ec[3 + i];
                        const uint32_t sh = v << (1 << i);
                        v = (v & ~mv) | (sh & mv);
                    }
                    v &= m;
                }
            }
            x[r] = v;
        }
        // ---- transform (the stages commute): layout A does stages 0..9, layout B (tpc > 1) stages >= 10
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] = inword_msb(x[r]);
        stages_inreg<R>(x, 0);
        uint32_t* mine = s_x + (size_t)slot * Wp;
        if (tpc > 1) {
            // layout A (words t R + r) -> layout B (words r tpc + t). R == 32 here; register-index bit b in layout B is
            // word-index bit ltpc + b = stage 5 + ltpc + b, of which stages < 10 (b < 5 - ltpc) were done above.
            __syncthreads();
#pragma unroll
            for (int r = 0; r < R; ++r) mine[t * R + r] = x[r];
            __syncthreads();
#pragma unroll
            for (int r = 0; r < R; ++r) x[r] = mine[r * tpc + t];
            stages_inreg<R>(x, 5 - ltpc);
        }
        if (p.systematic) { // resetFrozenBits, then the second transform starting in the layout we are in
            if (tpc > 1) {
#pragma unroll
                for (int r = 0; r < R; ++r) x[r] &= s_tab[8 * (r * tpc + t)];
                stages_inreg<R>(x, 5 - ltpc);
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) x[r] &= c_enc[8 * r];
            }
        }
        if (tpc > 1) { // back to layout A (the owner of 128 contiguous output bytes)
            __syncthreads();
#pragma unroll
            for (int r = 0; r < R; ++r) mine[r * tpc + t] = x[r];
            __syncthreads();
#pragma unroll
            for (int r = 0; r < R; ++r) x[r] = mine[t * R + r];
        }
        if (p.systematic) {
#pragma unroll
            for (int r = 0; r < R; ++r) x[r] = inword_msb(x[r]);
            stages_inreg<R>(x, 0);
        }
        // ---- store: the thread's R consecutive code words, memory order = byte-swapped MSB-first words
        if (live) {
            uint32_t* out = reinterpret_cast<uint32_t*>(p.code + (base + slot) * (unsigned long long)nbytes) + t * R;
            if (R >= 8 && p.v8) {
#pragma unroll
                for (int r = 0; r < R; r += 8)
                    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(out + r), "r"(bswap32(x[r])),
                                 "r"(bswap32(x[r + 1])), "r"(bswap32(x[r + 2])), "r"(bswap32(x[r + 3])), "r"(bswap32(x[r + 4])),
                                 "r"(bswap32(x[r + 5])), "r"(bswap32(x[r + 6])), "r"(bswap32(x[r + 7]))
                                 : "memory");
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) out[r] = bswap32(x[r]);
            }
        }
    }
}

// ---- host side ---------------------------------------------------------------------------------------------------
// Table of one code (8 words per code word), in the kernel's MSB-first register convention: mask bit 31 - i is set when
// element 32 w + i is an information position; q1 / pos locate the last info bit of the word in the byte-swapped info
// stream (word (e-1) / 32, bit 31 - (e-1) % 32); mv[0..4] are expand's move masks for that mask.
static void build_table(const pc_code* c, std::vector<uint32_t>& tab)
{
    const int W = c->N / 32;
    tab.assign((size_t)8 * W, 0u);
    uint32_t e = 0; // info bits up to and including this word
    for (int w = 0; w < W; ++w) {
        const uint32_t lsb = c->prog.info_mask[w]; // bit i%32 set for info element i
        uint32_t m = 0;
        for (int i = 0; i < 32; ++i)
            if (lsb & (1u << i)) m |= 1u << (31 - i);
        e += (uint32_t)__builtin_popcount(m);
        uint32_t* E = &tab[(size_t)8 * w];
        E[0] = m;
        if (!m) continue;
        E[1] = (e - 1) >> 5;
        E[2] = 31 - ((e - 1) & 31);
        uint32_t mm = m, mk = ~m << 1;
        for (int i = 0; i < 5; ++i) {
            uint32_t mp = mk ^ (mk << 1);
            mp ^= mp << 2;
            mp ^= mp << 4;
            mp ^= mp << 8;
            mp ^= mp << 16;
            const uint32_t mv = mp & mm;
            E[3 + i] = mv;
            mm = (mm ^ mv) | (mv >> (1 << i));
            mk &= ~mp;
        }
    }
}

// the __constant__ table holds ONE code at a time (per device); it is re-uploaded in stream order when another code
// object (or the same one after a change of its frozen set -- never happens -- ) encodes
static std::mutex g_enc_mu;
static const pc_code* g_enc_owner[64] = { nullptr };
static unsigned long long g_enc_serial[64] = { 0 };

// tpc <= 16 (a group of 16 warps = 512 threads): N <= 16384; longer codes keep k_encode
bool encode2_supported(const pc_code* c) { return c->N >= 32 && c->N <= 16384; }

void encode2_forget(const pc_code* c)
{
    std::lock_guard<std::mutex> lk(g_enc_mu);
    for (auto& o : g_enc_owner)
        if (o == c) o = nullptr;
}

template <int R, bool MULTI>
static cudaError_t launch2(const Enc2Params& p, int grid, int threads, size_t smem, cudaStream_t st)
{
    auto kern = k_encode2<R, MULTI>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    kern<<<grid, threads, smem, st>>>(p);
    return cudaGetLastError();
}

bool encode2_fuses_crc(const pc_code* c) { return encode2_supported(c) && c->N <= 1024; }

// fuse_crc: run the code's detector on the staged info rows inside the kernel (only where encode2_fuses_crc)
int encode2_device_raw(pc_code* c, const uint8_t* src, uint8_t* d_code, size_t B, cudaStream_t st, bool fuse_crc)
{
    const int W = c->N / 32, R = W < 32 ? W : 32, tpc = W / R;
    const int kb = (c->K + 7) / 8;
    {
        std::lock_guard<std::mutex> lk(g_enc_mu);
        int dev = c->device;
        if (dev < 0 || dev >= 64) return fail(PC_ERR_UNSUPPORTED, "device index out of range for the encoder table");
        if (g_enc_owner[dev] != c) {
            std::vector<uint32_t> tab;
            build_table(c, tab);
            // another code's encode kernels (on any stream) may still be reading the table: drain the device first.
            // Switches are rare (a process usually encodes with one code); a code that keeps the table pays nothing.
            cudaError_t e = cudaDeviceSynchronize();
            if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_enc, tab.data(), tab.size() * 4, 0, cudaMemcpyHostToDevice);
            if (e != cudaSuccess) return cuda_fail(e, "encoder table upload");
            g_enc_owner[dev] = c;
            g_enc_serial[dev]++;
        }
        if (tpc > 1 && !c->d_enc_tab) { // the same table in HBM for the multi-warp kernel
            std::vector<uint32_t> tab;
            build_table(c, tab);
            cudaError_t e = cudaMalloc(&c->d_enc_tab, tab.size() * 4);
            if (e == cudaSuccess) e = cudaMemcpy(c->d_enc_tab, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice);
            if (e != cudaSuccess) return cuda_fail(e, "encoder table (HBM)");
        }
    }
    Enc2Params p;
    p.info = src;
    p.code = d_code;
    p.B = B;
    p.S = c->S;
    p.kb = kb;
    p.kw = (kb + 3) / 4;
    p.tpc = tpc;
    p.systematic = c->systematic;
    // every thread stages kw / tpc words of its own row: 16-byte loads when rows and shares are whole 16-byte pieces
    p.vec_in = (kb % 16 == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0 && ((kb / 4) % (4 * tpc)) == 0) ? 1 : 0;
    p.v8 = ((reinterpret_cast<uintptr_t>(d_code) & 31u) == 0 && (4 * R) % 32 == 0) ? 1 : 0;
    p.tab = c->d_enc_tab;
    crc_spec(fuse_crc && tpc == 1 ? c->crc_kind : 0, p.crc);
    p.crc_table = c->d_crc_table;
    // CTA: whole groups of tpc warps (32 codewords each); 256 threads for short codes, 512 when a group needs them
    int threads = tpc <= 8 ? 256 : 32 * tpc;
    const int gpc = (threads / 32) / tpc, cpc = 32 * gpc;
    const size_t rows_w = (size_t)cpc * p.kw, xpose_w = (size_t)cpc * (W + 1);
    const size_t smem = 4 * (tpc > 1 ? (rows_w > xpose_w ? rows_w : xpose_w) + 8 * (size_t)W : rows_w + 256);
    size_t ctas = (B + cpc - 1) / cpc;
    const int grid = (int)(ctas < 148 * 8 ? ctas : 148 * 8);
    cudaError_t e;
    switch (R) {
    case 1: e = launch2<1, false>(p, grid, threads, smem, st); break;
    case 2: e = launch2<2, false>(p, grid, threads, smem, st); break;
    case 4: e = launch2<4, false>(p, grid, threads, smem, st); break;
    case 8: e = launch2<8, false>(p, grid, threads, smem, st); break;
    case 16: e = launch2<16, false>(p, grid, threads, smem, st); break;
    default: e = tpc > 1 ? launch2<32, true>(p, grid, threads, smem, st) : launch2<32, false>(p, grid, threads, smem, st); break;
    }
    if (e != cudaSuccess) return cuda_fail(e, "encode2 launch");
    count_launch();
    return PC_OK;
}

} // namespace pc
