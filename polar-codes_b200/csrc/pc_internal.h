// Internal definition of the opaque pc_code handle of include/polar_b200.h.
#ifndef PC_INTERNAL_H
#define PC_INTERNAL_H
#include <cuda_runtime.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/polar_b200.h"
#include "pc_crc.h"
#include "pc_kernels.cuh"
#include "pc_program.h"

struct pc_code {
    int N = 0, K = 0, S = 0, kind = 0, L = 1;
    int systematic = 1;
    int crc_kind = PC_CRC_8;
    int device = 0;
    std::vector<unsigned> frozen;
    pc::FastSscProgram prog;      // Fast-SSC int8 node program (also used for its lut / info mask by the encoder)
    // device copies
    uint32_t* d_code = nullptr;
    uint32_t* d_runs = nullptr;
    uint32_t* d_crc_table = nullptr;
    uint32_t* d_lut = nullptr;        // K info positions
    uint32_t* d_info_mask = nullptr;  // N/32 words
    // encoder tables: info-bit scatter runs grouped by code word (pc_encode.cu)
    uint2* d_enc_runs = nullptr;      // unused since the encoder deposits bits from (mask, prefix count)
    uint32_t* d_enc_first = nullptr;  // number of info positions below each code word, N/32 + 1 entries
    uint32_t* d_enc_tab = nullptr;    // pc_encode2.cu table (8 words per code word) for multi-warp codewords
    uint8_t* d_enc_info = nullptr;    // scratch copy of the info blocks when the detector writes check bytes
    size_t enc_info_bytes = 0;
    void* d_quant = nullptr;          // int8 LLRs quantised from a float batch
    size_t quant_bytes = 0;
    void* d_chan_code = nullptr;      // code words of the channel generator (pc_channel.cu)
    size_t chan_code_bytes = 0;
    int const_code = -1, const_runs = -1, const_words = 0;
    // SCL int8 (kind PC_DEC_SCL_I8)
    pc::SclProgram scl_prog;
    uint32_t* d_scl_code = nullptr;
    pc::SclLaunch scl_plan;
    bool scl_plan_valid = false;
    uint32_t* scl_arena = nullptr;
    size_t scl_arena_bytes = 0;
    // SCL fp32 (kind PC_DEC_SCL_F32, second pass of PC_DEC_ADAPTIVE_F32)
    pc::SclProgram sclf_prog;
    uint32_t* d_sclf_code = nullptr;
    pc::SclFLaunch sclf_plan;
    bool sclf_plan_valid = false;
    float* sclf_arena = nullptr;
    size_t sclf_arena_bytes = 0;
    // adaptive: frames whose first pass failed (index list + count on the device)
    void* d_fail_idx = nullptr;
    size_t fail_idx_bytes = 0;
    unsigned long long* d_fail_count = nullptr;
    void* d_ok_tmp = nullptr;
    size_t ok_tmp_bytes = 0;
    // fp32 Fast-SSC (kind PC_DEC_FASTSSC_F32, first pass of PC_DEC_ADAPTIVE_F32)
    pc::FloatProgram f32_prog;
    uint32_t* d_f32_code = nullptr;
    pc::F32Launch f32_plan;
    bool f32_plan_valid = false;
    float* f32_arena = nullptr;
    size_t f32_arena_bytes = 0;
    void* d_f32_in = nullptr;         // float LLRs converted from an int8 batch
    size_t f32_in_bytes = 0;
    // Fast-SSC launch plan
    pc::FastSscLaunch plan;
    bool plan_valid = false;
    pc::FastSscLaunch plan_small; // the same code planned with smaller CTAs, for batches that would leave SMs idle
    int plan_small_cap = 0;       // CTA size cap plan_small was made for (0 = none)
    uint32_t* arena = nullptr;
    size_t arena_bytes = 0;
    float* scf_arena = nullptr;       // pc_scflip_f32.cu: LLR stages, decision words, leaf reliabilities, flip words
    size_t scf_arena_bytes = 0;
    uint32_t* soft_arena = nullptr;   // pc_fastssc_i8_soft.cu: LLR stages + N decision BYTES per thread
    size_t soft_arena_bytes = 0;
    // staging for the *_host entry points
    void* d_stage_in = nullptr;
    void* d_stage_out = nullptr;
    uint8_t* d_stage_ok = nullptr;
    size_t stage_in_bytes = 0, stage_out_bytes = 0, stage_ok_bytes = 0;
    cudaStream_t streams[2] = { nullptr, nullptr };
    cudaEvent_t ev_decode = nullptr; // chains the decode kernels of consecutive host-batch slices
};

namespace pc {
void set_error(const std::string& s);
int fail(int status, const std::string& s);
int cuda_fail(cudaError_t e, const char* what);
void count_launch(unsigned long long n = 1);
int ensure_bytes(void** d, size_t* have, size_t need);
// A pc_code lives on the device that was current at pc_code_create (its tables, constant-pool slice, arenas and
// streams belong to it): every entry point that takes one checks that this device is current.
int check_code_device(const pc_code* c);
// device pointers the kernels access with vector loads / word stores: PC_ERR_ARG instead of a misaligned-address fault
int check_aligned(const void* p, size_t align, const char* what);
int crc_generate_rows(int crc_kind, uint8_t* d_data, int bytes, size_t B, cudaStream_t st);
const uint32_t* crc_table_device(int kind);
int quantize_device(int N, const float* d_in, int8_t* d_out, size_t B, cudaStream_t st);
int encode_device(pc_code* c, const uint8_t* d_info, uint8_t* d_code, uint8_t* d_info_crc, size_t B, cudaStream_t st);
int build_encoder_tables(pc_code* c);
int encode_device_raw(pc_code* c, const uint8_t* d_info, uint8_t* d_code, size_t B, cudaStream_t st);
int scflip_decode_device(pc_code* c, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, const unsigned* d_frame_idx,
                         const unsigned long long* d_count, int trials, cudaStream_t st);
int soft_decode_device(pc_code* c, const int8_t* d_llr, int8_t* d_soft_code, int8_t* d_soft_info, size_t B, cudaStream_t st);
} // namespace pc
#endif
