// SIMD-in-word int8 arithmetic: four int8 LLRs per 32-bit register, exact lane-for-lane restatement of
// the saturating AVX2 byte ops the reference's `char` decoders are built from
// (reference include/polarcode/decoding/fip_char.h:44-73, include/polarcode/avxconvenience.h:92-101).
//
// Every function is __host__ __device__: tests/host_emul.cpp compiles this header with g++ and runs the
// same interpreter on the CPU for ONE emulated thread (debug harness; not a product path).
// On the device the byte ops map to LOP3/IADD3/PRMT/VABSDIFF4 sequences (sm_100a has no native
// saturating byte add; see DESIGN.md "SWAR cost table").
#ifndef PC_SWAR_H
#define PC_SWAR_H
#include <stdint.h>

#if defined(__CUDACC__)
#define PC_HD __host__ __device__ __forceinline__
#else
#define PC_HD inline
#endif

namespace pc {

static constexpr uint32_t kH = 0x80808080u; // byte MSBs
static constexpr uint32_t kL = 0x7f7f7f7fu;
static constexpr uint32_t k1 = 0x01010101u;

// 0xFF in every byte whose MSB is set, else 0x00
PC_HD uint32_t sw_signfill(uint32_t x)
{
#ifdef __CUDA_ARCH__
    // PRMT with the sign-replicate selector bit (the __byte_perm intrinsic masks it off, so inline PTX)
    uint32_t r;
    asm("prmt.b32 %0, %1, %1, 0xba98;" : "=r"(r) : "r"(x));
    return r;
#else
    return ((x >> 7) & k1) * 0xFFu;
#endif
}

// per-byte |a - b|, bytes read as SIGNED (result 0..255) -- the one packed-byte ALU op sm_100a has natively
// (SASS VABSDIFF4; vadd4/vmin4/vmax4 are all multi-instruction emulations, see DESIGN.md "SWAR cost table")
PC_HD uint32_t sw_absdiff_s(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("vabsdiff4.u32.s32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(0u));
    return r;
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        int x = (int8_t)(a >> (8 * i)), y = (int8_t)(b >> (8 * i));
        r |= (uint32_t)(x > y ? x - y : y - x) << (8 * i);
    }
    return r;
#endif
}
// per-byte |a - b|, bytes read as UNSIGNED
PC_HD uint32_t sw_absdiff_u(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    uint32_t r;
    asm("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(0u));
    return r;
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        int x = (a >> (8 * i)) & 0xFF, y = (b >> (8 * i)) & 0xFF;
        r |= (uint32_t)(x > y ? x - y : y - x) << (8 * i);
    }
    return r;
#endif
}

// _mm256_adds_epi8 on four lanes
PC_HD uint32_t sw_adds(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    return __vaddss4(a, b);
#else
    uint32_t s = ((a & kL) + (b & kL)) ^ ((a ^ b) & kH);
    uint32_t ovf = (~(a ^ b)) & (a ^ s) & kH;
    uint32_t fill = sw_signfill(ovf);
    uint32_t satv = kL + ((a >> 7) & k1); // 0x7f (a >= 0) or 0x80 (a < 0)
    return (s & ~fill) | (satv & fill);
#endif
}

// per-byte wrap-around a - b
PC_HD uint32_t sw_sub(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    return __vsub4(a, b);
#else
    return ((a | kH) - (b & kL)) ^ ((a ^ ~b) & kH);
#endif
}

// |x| with |-128| = 127  (== |max(x,-127)|, fip_char.h:52-55)
PC_HD uint32_t sw_abs_sat(uint32_t a)
{
#ifdef __CUDA_ARCH__
    return __vabsss4(a);
#else
    uint32_t f = sw_signfill(a);
    uint32_t r = sw_sub(a ^ f, f); // two's complement negate of negative lanes; -128 stays 0x80
    return r ^ sw_signfill(r);     // 0x80 -> 0x7f
#endif
}

// _mm256_abs_epi8: |-128| = 0x80 (128 when read unsigned)
PC_HD uint32_t sw_abs(uint32_t a)
{
#ifdef __CUDA_ARCH__
    return __vabs4(a);
#else
    uint32_t f = sw_signfill(a);
    return sw_sub(a ^ f, f);
#endif
}

// unsigned byte min / max
PC_HD uint32_t sw_minu(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    return __vminu4(a, b);
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        uint32_t x = (a >> (8 * i)) & 0xFF, y = (b >> (8 * i)) & 0xFF;
        r |= (x < y ? x : y) << (8 * i);
    }
    return r;
#endif
}
PC_HD uint32_t sw_maxu(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    return __vmaxu4(a, b);
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        uint32_t x = (a >> (8 * i)) & 0xFF, y = (b >> (8 * i)) & 0xFF;
        r |= (x > y ? x : y) << (8 * i);
    }
    return r;
#endif
}

// F_function_calc (fip_char.h:44-65): sign(a)sign(b) * max(min(|a|',|b|'),1), |.|' clamps -128 to 127;
// a zero input counts as +1 (its sign bit is clear), the result is never 0 and never -128.
PC_HD uint32_t sw_f(uint32_t a, uint32_t b)
{
    // 15 ALU instructions per 4 lanes, no inter-byte carries anywhere:
    uint32_t x = sw_absdiff_s(a, 0u);         // |a| in [0,128]
    x ^= sw_signfill(x);                      // 128 -> 127, so x + y <= 255 below
    const uint32_t y = sw_absdiff_s(b, 0u);   // |b| in [0,128]; min(x, y) <= 127 without clamping y
    const uint32_t m = (x + y - sw_absdiff_u(x, y)) >> 1;      // min(x, y) = (x + y - |x - y|) / 2
    const uint32_t mag = (m + k1 + sw_absdiff_u(m, k1)) >> 1;  // max(m, 1) = (m + 1 + |m - 1|) / 2
    const uint32_t fill = sw_signfill(a ^ b);
    return (mag ^ fill) + (fill & k1);        // negate lanes whose sign product is negative: ~mag + 1 <= 0xFF
}

// G_function_calc (fip_char.h:67-73): bit ? sat(b - a) : sat(b + a). `fill` = 0xFF in lanes whose bit is 1.
// sat(b - a) == ~sat(~b + a), so one saturating add serves both signs.
PC_HD uint32_t sw_g(uint32_t a, uint32_t b, uint32_t fill) { return sw_adds(b ^ fill, a) ^ fill; }

// 4-bit mask -> 0xFF/0x00 per byte (bit j -> byte j): bit j times 2^(7j+7) lands on the MSB of byte j, the
// sign-replicating PRMT turns MSBs into fills and ignores every other bit
PC_HD uint32_t sw_expand4(uint32_t nib) { return sw_signfill((nib & 0xFu) * 0x10204080u); }

// MSBs of the four bytes -> 4-bit mask (byte j -> bit j)
PC_HD uint32_t sw_movemask(uint32_t w) { return (((w >> 7) & k1) * 0x01020408u) >> 24; }

// reduce_adds_epi8's last two levels inside one word: lanes (0,2),(1,3) then (0,1); result in byte 0
PC_HD uint32_t sw_hreduce2(uint32_t x) { return sw_adds(x, x >> 16); } // byte0 = l0+l2, byte1 = l1+l3
PC_HD uint32_t sw_hreduce1(uint32_t x) { return sw_adds(x, x >> 8); }

PC_HD int sw_byte0_negative(uint32_t x) { return (x >> 7) & 1; }

} // namespace pc
#endif
