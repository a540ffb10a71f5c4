/* TEST ORACLE -- plain scalar C restatement of the reference's hot path. Test infrastructure only:
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it; the product never does.
 * Pinned against oracle/_ref (the unmodified reference library) and the reference's own KATs by
 * tests/test_oracle_*.py. Every function cites the reference file:line it follows. */
#ifndef PC_ORACLE_H
#define PC_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* construction (src/polarcode/construction/{bhattacharrya,betaexpansion}.cpp) -> sorted frozen set, returns N-K */
int pco_frozen_bits_bb(int N, int K, float design_snr, unsigned* frozen_out);
int pco_frozen_bits_be(int N, int K, unsigned* frozen_out);

/* int8 primitives (include/polarcode/decoding/fip_char.h:44-73) on n lanes */
void pco_f_i8(const int8_t* a, const int8_t* b, int8_t* out, int n);
void pco_g_i8(const int8_t* a, const int8_t* b, const int8_t* bits, int8_t* out, int n);

/* float -> int8 quantiser (src/polarcode/bitcontainer.cpp:34-39,449-516) */
void pco_quantize(int N, const float* in, int8_t* out);

/* Fast-SSC int8 decoder (src/polarcode/decoding/fastssc_fip_char.cpp). soft_bits (N int8, may be NULL)
 * receives the root bit bytes (= Decoder::getSoftCodeword). info_out gets ceil(K/8) bytes. */
void pco_fastssc_i8_decode(int N, const unsigned* frozen, int nf, int systematic, const int8_t* llr,
                           uint8_t* info_out, int8_t* soft_bits);
/* node census of the tree createDecoder builds (fastssc_fip_char.cpp:496-569); counts[20] */
void pco_fastssc_i8_census(int N, const unsigned* frozen, int nf, int* counts);

/* packed butterfly encoder (src/polarcode/encoding/butterfly_fip_packed.cpp:45-59); info holds
 * ceil(K/8) bytes with the CRC already in place; code_out N/8 bytes */
void pco_encode(int N, const unsigned* frozen, int nf, int systematic, const uint8_t* info, uint8_t* code_out);

/* CRC family (src/polarcode/errordetection/). kind: 6,8,11,16 (CCITT-FALSE),17 (=16 "NR"/XMODEM),24,32 */
uint64_t pco_crc_calculate(int kind, const uint8_t* data, size_t bits);
void pco_crc_generate(int kind, uint8_t* data, int bytes);
int pco_crc_check(int kind, const uint8_t* data, int bytes);

/* int8 CRC-aided SCL decoder SclFipChar (src/polarcode/decoding/scl_fip_char.cpp), pc_oracle_scl.c.
 * Returns the detector verdict; info_out gets ceil(K/8) bytes. */
int pco_scl_i8_decode(int N, const unsigned* frozen, int nf, int L, int crc_kind, int systematic, const int8_t* llr,
                      uint8_t* info_out);

#ifdef __cplusplus
}
#endif
#endif
