/* polar_b200 -- C-ABI of the B200-native batched polar codec (libpolar_b200.so).
 *
 * Drop-in boundary for the hot path of ant-uni-bremen/polar-codes: every entry point below is what a
 * binding of the reference's per-codeword classes would call once the loop over codewords moves onto
 * the GPU. Reference interface each one replaces (paths relative to the reference tree):
 *
 *   pc_frozen_bits          PolarCode::Construction::frozen_bits            include/polarcode/construction/constructor.h:58-62
 *   pc_code_create          PolarCode::Decoding::create / makeDecoder       src/polarcode/decoding/decoder.cpp:26-87
 *                           + Encoding::ButterflyFipPacked ctor             include/polarcode/encoding/butterfly_fip_packed.h:25-49
 *   pc_code_set_crc         Decoder/Encoder::setErrorDetection              decoder.cpp:133-136, src/polarcode/encoding/encoder.cpp
 *   pc_code_set_systematic  Decoder/Encoder::setSystematic                  decoder.h:118, encoder.h
 *   pc_decode_batch         Decoder::decode_vector(const char* | const float*, void*)   decoder.cpp:154-181   (B codewords per call)
 *   pc_decode_soft_batch    Decoder::getSoftCodeword / getSoftInformation   decoder.cpp:147-152 (int8 Fast-SSC; _f32: fp32 Fast-SSC)
 *   pc_encode_batch         Encoder::encode_vector(void*, void*)            src/polarcode/encoding/encoder.cpp:79-90
 *   pc_crc_generate_batch   ErrorDetection::Detector::generate              include/polarcode/errordetection/errordetector.h:57-63
 *   pc_crc_check_batch      Detector::check                                 errordetector.h:65-70
 *   pc_crc_calculate        Detector::calculate                             errordetector.h:48
 *   pc_quantize_batch       CharContainer::insertLlr(const float*)          src/polarcode/bitcontainer.cpp:505-516
 *   pc_awgn_generate        pcsim SimulationWorker::generateData/encode/modulate/transmit  src/simulation/simulator.cpp:632-672,832-867
 *   pc_count_errors         SimulationWorker::countErrors                   src/simulation/simulator.cpp:939-965
 *
 * Conventions: plain pointers and sizes only; no exceptions cross this boundary (int status, 0 = OK,
 * message via pc_last_error()). Pointers named d_* are DEVICE pointers on the current CUDA device,
 * h_* are host pointers. `stream` is a cudaStream_t passed as void* (NULL = default stream).
 * Bit order is the reference's: code/info bit i lives in byte i/8, bit 7-(i%8) (MSB first).
 * LLR >= 0 decodes to bit 0. Frames are rows: llr[B][N], info[B][ceil(K/8)], code[B][N/8].
 * There is no CPU fallback: every compute entry point fails with PC_ERR_CUDA when no GPU is usable.
 *
 * Devices, streams, alignment:
 *  - A pc_code belongs to the CUDA device that was current at pc_code_create (tables, constant-pool slice, scratch
 *    arenas and streams live there). Every entry point taking a pc_code returns PC_ERR_ARG when another device is
 *    current; a host process that drives several GPUs creates one code object per device. The detector batch calls
 *    without a code object keep one table per (device, kind).
 *  - One pc_code may be in flight on ONE stream at a time: the calls share per-code scratch (quantiser output,
 *    int8 -> float copy, failed-frame list, decoder arenas). Calls on the same stream serialise by stream order;
 *    for concurrency create one code object per stream.
 *  - d_llr / d_llr_out must be 16-byte aligned (rows of 16 bytes or more), d_info 4-byte aligned when ceil(K/8) is a multiple of 4, d_code
 *    4-byte aligned for N >= 32 (PC_ERR_ARG otherwise; cudaMalloc'd bases and whole rows of them always qualify). */
#ifndef POLAR_B200_H
#define POLAR_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct pc_code pc_code;

/* decoder kinds -- the implementations Decoding::makeDecoder selects (decoder.cpp:54-87) */
enum {
    PC_DEC_FASTSSC_I8 = 0, /* 'char',  L == 1 : FastSscFipChar  */
    PC_DEC_FASTSSC_F32 = 1, /* 'float', L == 1 : FastSscAvxFloat */
    PC_DEC_SCL_I8 = 2,     /* 'char',  L  > 1 : SclFipChar      */
    PC_DEC_SCL_F32 = 3,    /* 'float', L  > 1 : SclAvxFloat     */
    PC_DEC_ADAPTIVE_F32 = 4, /* 'mixed', L  > 1 : AdaptiveFloat   */
    /* the two wrappers pcsim builds directly (src/simulation/simulator.cpp:721-733), not reachable through
     * Decoding::create's type strings: */
    PC_DEC_ADAPTIVE_I8 = 5,   /* AdaptiveChar  (adaptive_char.cpp:35-46):  FastSscFipChar, then SclFipChar  */
    PC_DEC_ADAPTIVE_MIXED = 6, /* AdaptiveMixed (adaptive_mixed.cpp:36-47): FastSscFipChar, then SclAvxFloat */
    /* pcsim's 'depthfirst' decoder (simulator.cpp:715-716): DepthFirst(N, trialLimit = L, frozen), SC-Flip on float LLRs
     * (src/polarcode/decoding/depth_first.cpp:556-640); L is the trial limit (1 .. 32) */
    PC_DEC_SCFLIP_F32 = 7
};

/* error detectors -- ErrorDetection::create(size, type) (errordetector.cpp:27-88) */
enum {
    PC_CRC_NONE = 0,  /* Dummy: check() is always true */
    PC_CRC_6NR = 6,   /* poly 0x21,     init 0, stored in 1 byte  */
    PC_CRC_8 = 8,     /* poly 0x07,     init 0                    */
    PC_CRC_11NR = 11, /* poly 0x621,    init 0, stored in 2 bytes */
    PC_CRC_16 = 16,   /* CCITT-FALSE 0x1021 init 0xFFFF           */
    PC_CRC_16NR = 17, /* XMODEM      0x1021 init 0 ("nr" type)    */
    PC_CRC_24NRC = 24, /* poly 0xB2B117, init 0                   */
    PC_CRC_32C = 32   /* CRC-32C over little-endian words         */
};

enum {
    PC_OK = 0,
    PC_ERR_ARG = 1,    /* invalid argument (std::invalid_argument / std::logic_error in the reference) */
    PC_ERR_CUDA = 2,   /* CUDA runtime failure or no device */
    PC_ERR_UNSUPPORTED = 3
};

const char* pc_last_error(void);
const char* pc_version(void);

/* host-only: frozen set for (N, K) by "BB" (Bhattacharyya, design SNR in dB) or "BE" (beta expansion), the
 * substring match of Construction::frozen_bits (constructor.cpp:41-63); writes N-K sorted indices to out
 * (capacity N), returns N-K or a negative status. "5G" (FiveGList, fiveGList.cpp:30-43): the N-K least reliable
 * positions of the 3GPP TS 38.212 polar sequence restricted to indices < N -- identical to the reference for N = 1024;
 * for N < 1024 the reference's unfiltered table prefix returns indices >= N, here the standard's filtered sequence
 * (what the reference's python/polar_5g_parameters.py computes) is used. N > 1024 is PC_ERR_ARG as in the reference. */
int pc_frozen_bits(int N, int K, float design_snr_db, const char* type, unsigned* out);

/* error detector id from the reference's (size, type) pair; negative on the sizes create() rejects */
int pc_crc_kind(unsigned size, const char* type);
int pc_crc_bytes(int crc_kind); /* bytes the detector occupies at the tail of the info block */

/* Builds the code object: host-side tree compilation (node program, info-bit runs, masks) and upload to
 * the current device. kind is one of PC_DEC_*; L is the list size (1 for Fast-SSC). Defaults mirror
 * the reference: systematic = true, detector = CRC-8 (makeDecoder installs one, decoder.cpp:85). */
pc_code* pc_code_create(int N, const unsigned* frozen, int n_frozen, int kind, int L);
void pc_code_destroy(pc_code* c);
int pc_code_set_crc(pc_code* c, int crc_kind);
int pc_code_set_systematic(pc_code* c, int systematic);
int pc_code_N(const pc_code* c);
int pc_code_K(const pc_code* c);
int pc_code_info_bytes(const pc_code* c); /* ceil(K/8) */
/* introspection for tests and reports: number of node-program instructions, frames per CTA, smem bytes */
int pc_code_program_length(const pc_code* c);
int pc_code_program(const pc_code* c, uint32_t* out, int capacity);
int pc_code_launch_info(const pc_code* c, int* threads_per_cta, int* smem_bytes, int* frames_per_cta);

/* Decode B codewords. llr_is_float selects the float32 overload of decode_vector (quantised on device
 * with the reference's quantiser for the int8 kinds), otherwise int8. d_ok (may be NULL) receives the
 * detector verdict per frame (decode()'s return value). Asynchronous on `stream`. */
int pc_decode_batch(pc_code* c, const void* d_llr, int llr_is_float, uint8_t* d_info, uint8_t* d_ok, size_t B,
                    void* stream);
/* Same call with HOST buffers: pinned staging, H2D, decode, D2H, synchronised on return. This is the
 * path the per-codeword decode_vector() wrappers use (B = 1) and bench.py's e2e leg. */
int pc_decode_batch_host(pc_code* c, const void* h_llr, int llr_is_float, uint8_t* h_info, uint8_t* h_ok,
                         size_t B);

/* Soft outputs of the int8 Fast-SSC decoder ('char', L = 1), Decoder::getSoftCodeword / getSoftInformation
 * (src/polarcode/decoding/decoder.cpp:147-152) after decode(): the decoder's bit container as BYTES, byte for byte what
 * the reference holds (leaves write LLR-valued bytes, combine steps XOR whole bytes; only the sign is a decision).
 * d_soft_code[B][N] int8 and / or d_soft_info[B][K] int8 (the container bytes at the information positions,
 * CharContainer::getSoftInformation bitcontainer.cpp:582-590); either may be NULL. Other decoder kinds: PC_ERR_UNSUPPORTED. */
int pc_decode_soft_batch(pc_code* c, const void* d_llr, int llr_is_float, int8_t* d_soft_code, int8_t* d_soft_info, size_t B,
                         void* stream);
int pc_decode_soft_batch_host(pc_code* c, const void* h_llr, int llr_is_float, int8_t* h_soft_code, int8_t* h_soft_info,
                              size_t B);
/* The same getters for the fp32 Fast-SSC decoder ('float', L = 1): FloatContainer::getSoftBits / getSoftInformation
 * (src/polarcode/bitcontainer.cpp:294-297, 331-339) -- the bit container as 32-bit FLOAT PATTERNS, bit for bit what the
 * reference holds: leaves write LLR-valued floats (+inf for frozen blocks), combine steps XOR the patterns
 * (include/polarcode/decoding/avx_float.h:190-236), so only the sign bit of an inner word is meaningful and a word may be a
 * NaN pattern -- compare as uint32. d_soft_code[B][N] and / or d_soft_info[B][K] float32; either may be NULL. */
int pc_decode_soft_batch_f32(pc_code* c, const void* d_llr, int llr_is_float, float* d_soft_code, float* d_soft_info, size_t B,
                             void* stream);
int pc_decode_soft_batch_f32_host(pc_code* c, const void* h_llr, int llr_is_float, float* h_soft_code, float* h_soft_info,
                                  size_t B);

/* Encode B info blocks of ceil(K/8) bytes. The detector's check bytes are generated into the tail of a
 * device-side copy of each block (the reference writes them into the caller's buffer,
 * butterfly_fip_packed.cpp:47-49); d_info_crc (may be NULL) receives that copy. */
int pc_encode_batch(pc_code* c, const uint8_t* d_info, uint8_t* d_code, uint8_t* d_info_crc, size_t B, void* stream);
int pc_encode_batch_host(pc_code* c, const uint8_t* h_info, uint8_t* h_code, uint8_t* h_info_crc, size_t B);

/* Detector batch operations on rows of `bytes` bytes (device pointers) */
int pc_crc_generate_batch(int crc_kind, uint8_t* d_data, int bytes, size_t B, void* stream);
int pc_crc_check_batch(int crc_kind, const uint8_t* d_data, int bytes, uint8_t* d_ok, size_t B, void* stream);
/* host-only single call: Detector::calculate(data, bits) */
uint64_t pc_crc_calculate(int crc_kind, const uint8_t* h_data, size_t bits);

/* float32 -> int8 LLR with the reference quantiser for block length N (rounding mode depends on N) */
int pc_quantize_batch(int N, const float* d_in, int8_t* d_out, size_t B, void* stream);

/* Device-side channel simulation for BLER sweeps (Philox4x32-10, counter = first_frame + row index):
 * random info bits -> detector generate -> encode -> BPSK (0 -> +1) -> AWGN at Eb/N0 -> LLR.
 * llr_is_float == 0: int8 LLR = quantise(amplification * y); == 1: float LLR = amplification * y.
 * d_info_out[B][ceil(K/8)] receives the transmitted info blocks (incl. check bytes). */
int pc_awgn_generate(pc_code* c, uint64_t seed, uint64_t first_frame, float ebn0_db, float amplification,
                     int llr_is_float, void* d_llr_out, uint8_t* d_info_out, size_t B, void* stream);
/* counts[0] += frames, [1] += frame errors, [2] += bit errors over the K info bits, [3] += frames
 * whose detector verdict was false. d_counts is uint64[4] on the device. */
int pc_count_errors(const uint8_t* d_info_ref, const uint8_t* d_info_dec, const uint8_t* d_ok, int K, size_t B,
                    unsigned long long* d_counts, void* stream);

/* Puncturer for block lengths that are not a power of two (PolarCode::Puncturer, include/polarcode/puncturer.h:37-101,
 * src/polarcode/puncturer.cpp:52-89; pybind python/bindings/puncturer_python.cc). The parent length is the next power
 * of two; the first (parent - block_length) entries of `frozen` are punctured. NULL + pc_last_error() when there are
 * not enough frozen positions (std::out_of_range in the reference). */
typedef struct pc_puncturer pc_puncturer;
pc_puncturer* pc_puncturer_create(size_t block_length, const unsigned* frozen, int n_frozen);
void pc_puncturer_destroy(pc_puncturer* p);
size_t pc_puncturer_block_length(const pc_puncturer* p);        /* Puncturer::blockLength()       */
size_t pc_puncturer_parent_block_length(const pc_puncturer* p); /* Puncturer::parentBlockLength() */
int pc_puncturer_positions(const pc_puncturer* p, unsigned* out); /* blockOutputPositions(); returns the count */
/* puncture<T>: d_out[B][M] = gather of d_in[B][Np]; depuncture<T>: d_out[B][Np] = scatter of d_in[B][M], punctured
 * positions 0. elem_bytes 1 (bits as bytes, int8 LLRs), 4 (float) or 8 (double); asynchronous on `stream`. */
int pc_puncture_batch(const pc_puncturer* p, const void* d_in, void* d_out, int elem_bytes, size_t B, void* stream);
int pc_depuncture_batch(const pc_puncturer* p, const void* d_in, void* d_out, int elem_bytes, size_t B, void* stream);
/* puncturePacked: MSB-first packed bits, d_in[B][Np/8] -> d_out[B][M/8] (both lengths multiples of 8) */
int pc_puncture_packed_batch(const pc_puncturer* p, const uint8_t* d_in, uint8_t* d_out, size_t B, void* stream);

/* Allocation-free decode. No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The fp32 Fast-SSC decode of pc_decode_batch (kind PC_DEC_FASTSSC_F32, float32 LLRs on the device, same outputs bit for
 * bit) with its per-thread scratch in a file-scope device array of 32 MiB that every GPU loading the library holds until
 * exit; the call itself performs no cudaMalloc / cudaFree (pc_code_create / pc_code_set_crc, called earlier, do).
 * Limits: block length N <= 4096 (PC_ERR_UNSUPPORTED beyond, and for other decoder kinds); any B (the grid strides).
 * All calls on a device share the array, which keeps the previous call's contents (nothing in it needs to start at
 * zero): calls that overlap on the GPU -- two streams, two host threads -- race; serialise them. Asynchronous on `stream`. */
int pc_decode_batch_f32_static(pc_code* c, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, void* stream);

/* device memory helpers so that non-CUDA hosts (ctypes, cgo, JNI) can drive the library alone */
void* pc_device_alloc(size_t bytes);
void pc_device_free(void* d);
void* pc_host_alloc(size_t bytes); /* pinned host memory for the *_host entry points */
void pc_host_free(void* h);
int pc_memcpy_h2d(void* d, const void* h, size_t bytes);
int pc_memcpy_d2h(void* h, const void* d, size_t bytes);
int pc_device_synchronize(void);
/* number of kernel launches issued by this library since load (bench.py's gpu_launches) */
unsigned long long pc_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
