// Soft outputs of the int8 Fast-SSC decoder for sm_100a: Decoder::getSoftCodeword / getSoftInformation of the reference
// (src/polarcode/decoding/decoder.cpp:147-152) return the decoder's BIT CONTAINER as bytes -- every node writes whole
// int8 values (R1: the LLRs, REP: the saturated sum, SPC: the LLRs with the flipped one negated, ...) and the combine
// steps XOR whole bytes (CombineInPlace fip_char.h:164-171), of which the hard-decision kernel (pc_fastssc_i8.cu) keeps
// only the MSB. This kernel keeps the bytes: one thread per codeword interprets the SAME node program with the decision
// store widened from 1 bit to 1 byte per element (N bytes per thread next to the LLR stages, word-interleaved in an
// HBM arena so that every warp access is coalesced). Four elements per 32-bit word with the SWAR primitives of
// pc_swar.h wherever a step covers whole words; the leaves whose result depends on the 32-lane order of the AVX2 code
// (REP / DoubleREP reductions, SPC minimum search) are restated byte by byte in that order
// (fastssc_fip_char.cpp:202-492, avxconvenience.h:92-101,202-210, avxconvenience.cpp:13-56).
// It is the soft-output path, not the throughput path: pc_decode_batch stays the bit-packed kernel.
#include "pc_fastssc_i8_core.h"
#include "pc_internal.h"

namespace pc {

struct SoftParams {
    const int8_t* llr;      // [B][N]
    int8_t* soft_code;      // [B][N] bit container after decode() (may be null)
    int8_t* soft_info;      // [B][K] container bytes at the information positions (may be null)
    unsigned long long B;
    int S, K, staged_root;
    const uint32_t* code;   // Fast-SSC int8 node program (HBM copy)
    const uint32_t* lut;    // K information positions
    uint32_t* arena;        // [words][stride]
    size_t stride;
    int llr_off[24];        // word offset of LLR stage s
    int bits_off;           // word offset of the N decision bytes
};

struct SoftCtx {
    uint32_t* ar;
    size_t st;
    const int* llr_off;
    int bits_off;
    __device__ __forceinline__ uint32_t& L(int s, int w) const { return ar[(size_t)(llr_off[s] + w) * st]; }
    __device__ __forceinline__ uint32_t& Bw(int w) const { return ar[(size_t)(bits_off + w) * st]; }
    __device__ __forceinline__ int8_t lget(int s, int i) const { return (int8_t)(L(s, i >> 2) >> (8 * (i & 3))); }
    __device__ __forceinline__ int8_t bget(int i) const { return (int8_t)(Bw(i >> 2) >> (8 * (i & 3))); }
    __device__ __forceinline__ void bset(int i, int8_t v) const
    {
        uint32_t& w = Bw(i >> 2);
        const int sh = 8 * (i & 3);
        w = (w & ~(0xFFu << sh)) | ((uint32_t)(uint8_t)v << sh);
    }
};

__device__ __forceinline__ int8_t s_adds8(int a, int b)
{
    const int x = a + b;
    return (int8_t)(x > 127 ? 127 : (x < -128 ? -128 : x));
}
__device__ __forceinline__ int8_t s_abs8(int8_t a) { return (int8_t)(a < 0 ? (uint8_t)(-(int)a) : a); } // |-128| = -128 (0x80)
__device__ __forceinline__ uint32_t bytes_mask(int nbytes) { return nbytes >= 4 ? 0xFFFFFFFFu : ((1u << (8 * nbytes)) - 1u); }

// reduce_adds_epi8 (avxconvenience.h:92-101): x[i] + x[i+16], then strides 8, 4, 2, 1, every add saturating
__device__ __forceinline__ int8_t reduce_adds32(int8_t (&x)[32])
{
    for (int st = 16; st; st >>= 1)
        for (int i = 0; i < st; ++i) x[i] = s_adds8(x[i], x[i + st]);
    return x[0];
}
// minpos_epu8 (avxconvenience.cpp:13-56): first minimum of 32 UNSIGNED bytes; the value is returned as a char
__device__ __forceinline__ int minpos32(const int8_t (&x)[32], int8_t& val)
{
    int best = 0;
    unsigned bv = (uint8_t)x[0];
    for (int i = 1; i < 32; ++i)
        if ((uint8_t)x[i] < bv) {
            bv = (uint8_t)x[i];
            best = i;
        }
    val = (int8_t)bv;
    return best;
}
// max(x, -127) per byte (CombineBitsShort fip_char.h:187-207 clamps both operands): 0x80 -> 0x81, exact per byte
__device__ __forceinline__ uint32_t clamp127(uint32_t x)
{
    const uint32_t t = x ^ kH;
    const uint32_t z = ~(((t & kL) + kL) | t | kL); // 0x80 in every byte of t that is zero
    return x + (z >> 7);
}

// SpcDecoder (n > 32, :274-303) / ZeroSpcDecoder (:321-359) core on the decision bytes [off, off + n) that already
// hold the values: per 32-vector minimum with the cross-vector rules of the reference, then the negate
__device__ __forceinline__ int spc_search_long(const SoftCtx& c, int off, int n, uint8_t& par)
{
    int minIdx = 0;
    int8_t testAbs, minAbs = 127;
    par = 0;
    for (int v = 0; v < n / 32; ++v) {
        int8_t ab[32];
        for (int i = 0; i < 32; ++i) {
            const int8_t x = c.bget(off + 32 * v + i);
            par ^= (uint8_t)x;
            ab[i] = s_abs8(x);
        }
        if (minAbs > 0) {
            const int vm = minpos32(ab, testAbs);
            if (testAbs < minAbs) {
                minIdx = vm + 32 * v;
                minAbs = testAbs;
            }
        }
    }
    return minIdx;
}

__global__ void __launch_bounds__(128) k_fastssc_i8_soft(const __grid_constant__ SoftParams p)
{
    const size_t GT = (size_t)gridDim.x * blockDim.x, gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    SoftCtx c{ p.arena + gt, p.stride, p.llr_off, p.bits_off };
    const int S = p.S, N = 1 << S;
    for (unsigned long long f = gt; f < p.B; f += GT) {
        const int8_t* row = p.llr + f * (unsigned long long)N;
        const uint32_t* grow = reinterpret_cast<const uint32_t*>(row);
        if (p.staged_root) {
            if (N >= 4) {
                for (int w = 0; w < N / 4; ++w) c.L(S, w) = grow[w];
            } else {
                uint32_t w = 0;
                for (int i = 0; i < N; ++i) w |= (uint32_t)(uint8_t)row[i] << (8 * i);
                c.L(S, 0) = w;
            }
        }
        for (int pc = 0;; ++pc) {
            const uint32_t ins = p.code[pc];
            const uint32_t op = instr_op(ins);
            const int s = (int)instr_s(ins), off = (int)instr_off(ins);
            if (op == OP_END) break;
            const int n = 1 << s, h = n >> 1;
            const bool from_row = (s == S) && !p.staged_root;
            // source word w of this node's LLRs
            auto SRC = [&](int w) -> uint32_t { return from_row ? __ldg(grow + w) : c.L(s, w); };
            switch (op) {
            case OP_F:
            case OP_G:
            case OP_G0:
            case OP_G1: {
                if (h >= 4) {
                    const int hw = h >> 2;
                    for (int w = 0; w < hw; ++w) {
                        const uint32_t a = SRC(w), b = SRC(w + hw);
                        if (op == OP_F) {
                            c.L(s - 1, w) = sw_f(a, b);
                        } else if (op == OP_G0) {
                            c.L(s - 1, w) = sw_adds(a, b);
                        } else {
                            const uint32_t left = c.Bw((off >> 2) + w);
                            const uint32_t r = sw_g(a, b, sw_signfill(left));
                            if (op == OP_G) {
                                c.L(s - 1, w) = r;
                            } else { // ROneNode :427-449: right decisions = g as bytes, parent = [left ^ right, right]
                                c.Bw((off >> 2) + w) = left ^ r;
                                c.Bw(((off + h) >> 2) + w) = r;
                            }
                        }
                    }
                } else { // n = 4 or 2: both halves inside one word
                    const uint32_t w = SRC(0), hi = w >> (8 * h), m = bytes_mask(h);
                    if (op == OP_F) {
                        c.L(s - 1, 0) = sw_f(w, hi) & m;
                    } else if (op == OP_G0) {
                        c.L(s - 1, 0) = sw_adds(w, hi) & m;
                    } else {
                        const int sh = 8 * (off & 3);
                        uint32_t bw = c.Bw(off >> 2);
                        const uint32_t left = (bw >> sh) & m;
                        const uint32_t r = sw_g(w, hi, sw_signfill(left)) & m;
                        if (op == OP_G) {
                            c.L(s - 1, 0) = r;
                        } else {
                            // ShortROneNode :451-474 (n <= 32 always here): [left ^ right, right]
                            bw = (bw & ~((m | (m << (8 * h))) << sh)) | (((left ^ r) | (r << (8 * h))) << sh);
                            c.Bw(off >> 2) = bw;
                        }
                    }
                }
                break;
            }
            case OP_COMB:
            case OP_COMB0: {
                const bool zero_left = op == OP_COMB0, shortn = n <= 32;
                if (h >= 4) {
                    const int hw = h >> 2;
                    for (int w = 0; w < hw; ++w) {
                        uint32_t& lw = c.Bw((off >> 2) + w);
                        uint32_t& rw = c.Bw(((off + h) >> 2) + w);
                        if (zero_left) { // Combine_0R / Combine_0RShort fip_char.h:209-220: left = right
                            lw = rw;
                        } else if (shortn) { // CombineBitsShort :187-207: both operands clamped to >= -127
                            const uint32_t r = clamp127(rw);
                            lw = clamp127(lw) ^ r;
                            rw = r;
                        } else { // CombineInPlace :164-171
                            lw ^= rw;
                        }
                    }
                } else {
                    const int sh = 8 * (off & 3);
                    const uint32_t m = bytes_mask(h);
                    uint32_t bw = c.Bw(off >> 2);
                    uint32_t l = (bw >> sh) & m, r = (bw >> (sh + 8 * h)) & m;
                    if (zero_left) {
                        l = r;
                    } else {
                        r = clamp127(r) & m;
                        l = (clamp127(l) & m) ^ r;
                    }
                    bw = (bw & ~((m | (m << (8 * h))) << sh)) | ((l | (r << (8 * h))) << sh);
                    c.Bw(off >> 2) = bw;
                }
                break;
            }
            case OP_R0: // RateZeroDecoder :202-208
                for (int i = 0; i < n; ++i) c.bset(off + i, 127);
                break;
            case OP_R1: // RateOneDecoder :210-215
                for (int i = 0; i < n; ++i) c.bset(off + i, c.lget(s, i));
                break;
            case OP_REP: { // RepetitionDecoder :225-240 / ShortRepetitionDecoder :265-272
                int8_t acc[32];
                for (int i = 0; i < 32; ++i) acc[i] = 0;
                if (n > 32) {
                    for (int v = 0; v < n / 32; ++v)
                        for (int i = 0; i < 32; ++i) acc[i] = s_adds8(acc[i], c.lget(s, 32 * v + i));
                } else {
                    for (int i = 0; i < n; ++i) acc[i] = c.lget(s, i);
                }
                const int8_t r = reduce_adds32(acc);
                for (int i = 0; i < n; ++i) c.bset(off + i, r);
                break;
            }
            case OP_DREP: { // DoubleRepetitionDecoder :249-263, half_reduce_adds_epi8 avxconvenience.h:202-210
                int8_t acc[32];
                for (int i = 0; i < 32; ++i) acc[i] = 0;
                for (int v = 0; v < n / 32; ++v)
                    for (int i = 0; i < 32; ++i) acc[i] = s_adds8(acc[i], c.lget(s, 32 * v + i));
                for (int st = 16; st >= 2; st >>= 1)
                    for (int i = 0; i < st; ++i) acc[i] = s_adds8(acc[i], acc[i + st]);
                for (int i = 0; i < n; ++i) c.bset(off + i, acc[i & 1]);
                break;
            }
            case OP_SPC: {
                if (n > 32) { // SpcDecoder :274-303
                    for (int i = 0; i < n; ++i) c.bset(off + i, c.lget(s, i));
                    uint8_t par;
                    const int mi = spc_search_long(c, off, n, par);
                    if (par & 0x80) c.bset(off + mi, (int8_t)(-c.bget(off + mi)));
                } else { // ShortSpcDecoder :305-319: lanes >= n padded with 127
                    int8_t v[32], ab[32];
                    uint8_t par = 0;
                    for (int i = 0; i < 32; ++i) {
                        v[i] = i < n ? c.lget(s, i) : (int8_t)127;
                        par ^= (uint8_t)v[i];
                        ab[i] = s_abs8(v[i]);
                    }
                    if (par & 0x80) {
                        int8_t dummy;
                        const int vm = minpos32(ab, dummy);
                        v[vm] = (int8_t)(-v[vm]);
                    }
                    for (int i = 0; i < n; ++i) c.bset(off + i, v[i]);
                }
                break;
            }
            case OP_S01: // ShortZeroOneDecoder :390-399
                for (int i = 0; i < h; ++i) {
                    const int8_t x = s_adds8(c.lget(s, i), c.lget(s, i + h));
                    c.bset(off + i, x);
                    c.bset(off + h + i, x);
                }
                break;
            case OP_0SPC: {
                // the root itself can be a 0SPC leaf only when staged (compile_fastssc_i8: staged_root for leaf roots)
                if (n > 32) { // ZeroSpcDecoder :321-359
                    for (int i = 0; i < h; ++i) {
                        const int8_t x = s_adds8(c.lget(s, i), c.lget(s, i + h));
                        c.bset(off + i, x);
                        c.bset(off + h + i, x);
                    }
                    uint8_t par;
                    const int mi = spc_search_long(c, off, h, par);
                    if (par & 0x80) {
                        c.bset(off + mi, (int8_t)(-c.bget(off + mi)));
                        c.bset(off + h + mi, (int8_t)(-c.bget(off + h + mi)));
                    }
                } else { // ShortZeroSpcDecoder :361-388
                    int8_t v[32], ab[32];
                    uint8_t par = 0;
                    for (int i = 0; i < 32; ++i) {
                        v[i] = i < h ? s_adds8(c.lget(s, i), c.lget(s, i + h)) : (int8_t)127;
                        par ^= (uint8_t)v[i];
                        ab[i] = s_abs8(v[i]);
                    }
                    if (par & 0x80) {
                        int8_t dummy;
                        const int vm = minpos32(ab, dummy);
                        v[vm] = (int8_t)(-v[vm]);
                    }
                    for (int i = 0; i < h; ++i) {
                        c.bset(off + i, v[i]);
                        c.bset(off + h + i, v[i]);
                    }
                }
                break;
            }
            default: break; // OP_FLUSH: only the bit-packed kernel keeps a block register
            }
        }
        if (p.soft_code) {
            int8_t* out = p.soft_code + f * (unsigned long long)N;
            if (N >= 4 && (reinterpret_cast<uintptr_t>(out) & 3u) == 0) {
                for (int w = 0; w < N / 4; ++w) reinterpret_cast<uint32_t*>(out)[w] = c.Bw(w);
            } else {
                for (int i = 0; i < N; ++i) out[i] = c.bget(i);
            }
        }
        if (p.soft_info) { // CharContainer::getSoftInformation bitcontainer.cpp:582-590
            int8_t* out = p.soft_info + f * (unsigned long long)p.K;
            for (int k = 0; k < p.K; ++k) out[k] = c.bget((int)__ldg(p.lut + k));
        }
    }
}

int soft_decode_device(pc_code* c, const int8_t* d_llr, int8_t* d_soft_code, int8_t* d_soft_info, size_t B, cudaStream_t st)
{
    SoftParams p;
    p.llr = d_llr;
    p.soft_code = d_soft_code;
    p.soft_info = d_soft_info;
    p.B = B;
    p.S = c->S;
    p.K = c->K;
    p.staged_root = c->prog.staged_root ? 1 : 0;
    p.code = c->d_code;
    p.lut = c->d_lut;
    const int top = c->prog.staged_root ? c->S : c->S - 1;
    int w = 0;
    for (int s = 0; s < 24; ++s) p.llr_off[s] = 0;
    for (int s = 0; s <= top; ++s) {
        p.llr_off[s] = w;
        w += s >= 2 ? (1 << (s - 2)) : 1;
    }
    p.bits_off = w;
    w += c->N >= 4 ? c->N / 4 : 1;
    const int threads = 128;
    size_t ctas = (B + threads - 1) / threads;
    const size_t max_ctas = 148 * 8;
    // keep the arena within 1 GiB for long codes
    size_t cap = ((size_t)1 << 30) / ((size_t)4 * w * threads);
    if (cap < 1) cap = 1;
    if (ctas > max_ctas) ctas = max_ctas;
    if (ctas > cap) ctas = cap;
    p.stride = ctas * threads;
    const size_t need = (size_t)4 * w * p.stride;
    int s = ensure_bytes(reinterpret_cast<void**>(&c->soft_arena), &c->soft_arena_bytes, need);
    if (s != PC_OK) return s;
    p.arena = c->soft_arena;
    k_fastssc_i8_soft<<<(int)ctas, threads, 0, st>>>(p);
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "soft decode launch");
}

} // namespace pc
