// Error detectors of the reference's ErrorDetection::Detector family as one table-driven byte-wise
// engine (reference src/polarcode/errordetection/{crc8,crc16,crc16nr,crc24nrc,crc11nr,crc6nr,crc32}.cpp,
// parameter sets in src/polarcode/errordetection/CRC.h:1307-1313,1513-1519,1704-1710,1986-1992,2155-2161).
//
// Non-reflected CRCs of any width w <= 32 run on a LEFT-ALIGNED 32-bit register: reg = init << (32-w),
// per byte reg = (reg << 8) ^ T[(reg >> 24) ^ byte] with T built for poly << (32-w); crc = reg >> (32-w).
// CRC-32C (crc32.cpp:33-41 uses the x86 crc32 instruction on little-endian words, init 0, no final
// xor) is the reflected byte-wise form reg = (reg >> 8) ^ T[(reg ^ byte) & 0xFF], poly 0x82F63B78.
// The check bytes sit at the tail of the block: big-endian for the non-reflected family
// (crc24nrc.cpp:44-53 etc.), native little-endian word for CRC-32C (crc32.cpp:58-69).
#ifndef PC_CRC_H
#define PC_CRC_H
#include <stdint.h>

#include "pc_swar.h"

namespace pc {

struct CrcSpec {
    int kind;      // PC_CRC_* value
    int width;     // bits of the polynomial
    int bytes;     // whole bytes occupied at the tail of the block
    int reflected; // 1 only for CRC-32C
    uint32_t poly, init;
};

// kind: 0 none, 6, 8, 11, 16 (CCITT-FALSE), 17 (16 "NR" = XMODEM), 24, 32
PC_HD bool crc_spec(int kind, CrcSpec& s)
{
    switch (kind) {
    case 0: s = CrcSpec{ 0, 0, 0, 0, 0, 0 }; return true;
    case 6: s = CrcSpec{ 6, 6, 1, 0, 0x21u, 0 }; return true;
    case 8: s = CrcSpec{ 8, 8, 1, 0, 0x07u, 0 }; return true;
    case 11: s = CrcSpec{ 11, 11, 2, 0, 0x621u, 0 }; return true;
    case 16: s = CrcSpec{ 16, 16, 2, 0, 0x1021u, 0xFFFFu }; return true;
    case 17: s = CrcSpec{ 17, 16, 2, 0, 0x1021u, 0 }; return true;
    case 24: s = CrcSpec{ 24, 24, 3, 0, 0xB2B117u, 0 }; return true;
    case 32: s = CrcSpec{ 32, 32, 4, 1, 0x82F63B78u, 0 }; return true;
    }
    return false;
}

inline void crc_build_table(const CrcSpec& s, uint32_t* table /*256*/)
{
    if (s.width == 0) {
        for (int i = 0; i < 256; ++i) table[i] = 0;
        return;
    }
    if (s.reflected) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t r = i;
            for (int k = 0; k < 8; ++k) r = (r & 1u) ? ((r >> 1) ^ s.poly) : (r >> 1);
            table[i] = r;
        }
        return;
    }
    const uint32_t p = s.poly << (32 - s.width);
    for (uint32_t i = 0; i < 256; ++i) {
        uint32_t r = i << 24;
        for (int k = 0; k < 8; ++k) r = (r & 0x80000000u) ? ((r << 1) ^ p) : (r << 1);
        table[i] = r;
    }
}

// streaming state: feed the block's bytes in order; the first `payload` bytes go through the CRC, the
// remaining `s.bytes` are collected as the stored check value
struct CrcStream {
    uint32_t reg, stored;
    PC_HD void begin(const CrcSpec& s)
    {
        reg = s.reflected ? s.init : (s.width ? (s.init << (32 - s.width)) : 0u);
        stored = 0;
    }
    PC_HD void payload_byte(const CrcSpec& s, const uint32_t* table, uint32_t byte)
    {
        if (s.reflected)
            reg = (reg >> 8) ^ table[(reg ^ byte) & 0xFFu];
        else
            reg = (reg << 8) ^ table[(reg >> 24) ^ byte];
    }
    PC_HD void stored_byte(uint32_t byte) { stored = (stored << 8) | byte; }
    PC_HD uint32_t value(const CrcSpec& s) const { return s.reflected ? reg : (s.width ? reg >> (32 - s.width) : 0u); }
    // check value as the big-endian number formed by the tail bytes
    PC_HD uint32_t tail_value(const CrcSpec& s) const
    {
        uint32_t v = value(s);
        if (s.reflected) // stored as a native little-endian word
            v = (v >> 24) | ((v >> 8) & 0xFF00u) | ((v << 8) & 0xFF0000u) | (v << 24);
        return v;
    }
    PC_HD bool ok(const CrcSpec& s) const { return s.width == 0 || tail_value(s) == stored; }
};

// Detector::calculate(data, bits): CRC::CalculateBits for the non-reflected parameter sets (reference
// src/polarcode/errordetection/CRC.h:613-645, 870-988, 1059-1150) with the reference's register widths (8-bit register
// for widths <= 8, 16-bit for <= 16, else 32) and its handling of a partial last byte: that byte is XORed in whole and
// the register shifted `bits % 8` times. CRC-8 / CRC-16 round `bits` down to whole bytes (crc8.cpp:36-39,
// crc16.cpp:20-23); CRC-32C takes bits/8 as the number of little-endian WORDS (crc32.cpp:28-31).
PC_HD uint32_t crc_calculate_bits(const CrcSpec& s, const uint8_t* data, size_t bits)
{
    if (s.width == 0) return 0u;
    if (s.reflected) {
        uint32_t crc = 0;
        for (size_t i = 0; i < (bits / 8) * 4; ++i) {
            crc ^= data[i];
            for (int k = 0; k < 8; ++k) crc = (crc & 1u) ? ((crc >> 1) ^ s.poly) : (crc >> 1);
        }
        return crc;
    }
    if (s.kind == 8 || s.kind == 16) bits = (bits / 8) * 8;
    const int regbits = s.width <= 8 ? 8 : (s.width <= 16 ? 16 : 32);
    const uint32_t regmask = regbits == 32 ? 0xFFFFFFFFu : ((1u << regbits) - 1u);
    uint32_t rem = s.init;
    const size_t whole = bits / 8, tail = bits % 8;
    if (s.width >= 8) {
        const int sh = s.width - 8;
        const uint32_t top = 1u << (s.width - 1);
        for (size_t i = 0; i < whole; ++i) {
            rem = (rem ^ ((uint32_t)data[i] << sh)) & regmask;
            for (int k = 0; k < 8; ++k) rem = ((rem & top) ? ((rem << 1) ^ s.poly) : (rem << 1)) & regmask;
        }
        if (tail) {
            rem = (rem ^ ((uint32_t)data[whole] << sh)) & regmask;
            for (size_t k = 0; k < tail; ++k) rem = ((rem & top) ? ((rem << 1) ^ s.poly) : (rem << 1)) & regmask;
        }
    } else {
        const int sh = 8 - s.width;
        const uint32_t poly = (s.poly << sh) & regmask;
        if (whole) {
            rem = (rem << sh) & regmask;
            for (size_t i = 0; i < whole; ++i) {
                rem = (rem ^ data[i]) & regmask;
                for (int k = 0; k < 8; ++k) rem = ((rem & 0x80u) ? ((rem << 1) ^ poly) : (rem << 1)) & regmask;
            }
            rem = (rem >> sh) & regmask;
        }
        if (tail) {
            rem = ((rem << sh) ^ data[whole]) & regmask;
            for (size_t k = 0; k < tail; ++k) rem = ((rem & 0x80u) ? ((rem << 1) ^ poly) : (rem << 1)) & regmask;
            rem = (rem >> sh) & regmask;
        }
    }
    const uint32_t wmask = s.width == 32 ? 0xFFFFFFFFu : ((1u << s.width) - 1u);
    return rem & wmask;
}

} // namespace pc
#endif
