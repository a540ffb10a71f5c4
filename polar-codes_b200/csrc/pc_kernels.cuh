// Kernel-side declarations shared by the C-ABI translation unit (pc_api.cu) and the kernel files.
#ifndef PC_KERNELS_CUH
#define PC_KERNELS_CUH
#include <cuda_runtime.h>
#include <stdint.h>

#include "pc_crc.h"

namespace pc {

// words of __constant__ memory shared by all live code objects (node programs + gather runs)
constexpr int kConstWords = 12288; // 48 KiB of the 64 KiB bank

struct FastSscLaunch {
    // code description
    int S, K, nruns;
    int staged_root, systematic;
    CrcSpec crc;
    int const_code, const_runs;   // offsets into the constant pool, or -1 when the program lives in HBM
    const uint32_t* g_code;       // program in HBM (long codes)
    const uint32_t* g_runs;
    const uint32_t* g_crc_table;  // 256 words
    // geometry
    int threads;        // per CTA
    int grid;           // CTAs (grid-stride over frames)
    int smem_chunks;    // 16-byte chunks per thread held in shared memory
    int gsplit;         // first chunk index held in the HBM arena (INT_MAX: none)
    int g_chunks;       // chunks per thread in the HBM arena
    int bits_in_smem;   // hard-decision words in shared memory (else HBM arena)
    int nbw;            // hard-decision words per frame
    size_t smem_bytes;
    uint32_t* arena;    // HBM arena: [g_chunks][grid*threads] chunks, then [nbw][grid*threads] words
};

cudaError_t const_pool_upload(int offset_words, const uint32_t* h_src, int nwords, cudaStream_t st);
cudaError_t launch_fastssc_i8(const FastSscLaunch& L, const int8_t* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                              cudaStream_t st);
cudaError_t fastssc_i8_set_smem_limit(size_t bytes);

} // namespace pc
#endif
