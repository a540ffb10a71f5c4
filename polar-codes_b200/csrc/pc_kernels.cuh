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
    int threads;        // per CTA (multiple of 32)
    int grid;           // CTAs (grid-stride over frames)
    int ctas_per_sm;
    size_t smem_bytes;
    // placement (pc_fastssc_i8_core.h "where things live")
    uint32_t place[24];  // per stage s: mk_place(kind, base)
    int smem_chunks;     // 16-byte chunk slots per thread in shared memory
    int tmem_alloc_cols; // tcgen05.alloc size of the CTA (0: tensor memory unused)
    int tmem_cols_thread; // columns owned by each thread (warps sharing a lane quarter are stacked along columns)
    int arena_chunks;    // chunk slots per thread in the HBM arena
    int bits_kind, bits_base; // decision words: PK_*; base = first column (tensor memory)
    int nbw;             // decision words per frame
    uint32_t* arena;     // [arena_chunks][grid*threads] chunks, then [nbw][grid*threads] words when bits_kind == PK_ARENA
};

// ---- SCL int8 (pc_scl_i8.cu) ----
struct SclParams {
    const int8_t* llr;
    uint8_t* info;
    uint8_t* ok;
    unsigned long long B;
    const unsigned* frame_idx;         // optional indirection (adaptive second pass): frame = frame_idx[i]
    const unsigned long long* count;   // optional device-side frame count (overrides B when non-null)
    int S, K, L, G, nruns, systematic, root_leaf;
    CrcSpec crc;
    const uint32_t* code;
    const uint32_t* runs;
    const uint32_t* crc_table;
    int llr_off[20];  // word offset of LLR stage s in a lane's space (s = 0 .. S)
    int bitL_off[20]; // word offset of the left / right decision buffer of stage s
    int bitR_off[20];
    int words;        // shared-memory words per lane
    int arena_words;  // arena words per lane (LLR stages whose llr_off carries the arena flag)
    size_t arena_stride; // lanes the arena rows are laid out for (grid * threads)
    uint32_t* arena;
    // tensor memory as private storage for the decision words of the large stages (offsets carry kTmem)
    int tm_alloc_cols;   // tcgen05.alloc size of one CTA (0: unused)
    int tm_cols_thread;  // columns per thread (warps sharing a lane quarter are stacked along columns)
    int scr_off;         // smem word offset of a per-lane scratch of N/32 words for extractBestPath
    int lockstep;        // CTA barrier after every node-program step (keeps the warps on the same instruction lines)
    int sort2;           // scl_sort_packed2 (candidate index inside the sort key): groups of <= 8 lanes, N <= 4096
};

struct SclLaunch {
    SclParams p;
    int threads, grid, per_sm;
    size_t smem_bytes;
};

cudaError_t launch_scl_i8(const SclLaunch& L, const int8_t* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                          const unsigned* d_frame_idx, const unsigned long long* d_count, cudaStream_t st);
int plan_scl_i8(int S, int K, int L, bool root_leaf, size_t smem_max, int sms, SclLaunch& out, const char** msg);

// ---- SCL fp32 (pc_scl_f32.cu) ----
struct SclFParams {
    const float* llr;
    uint8_t* info;
    uint8_t* ok;
    unsigned long long B;
    const unsigned* frame_idx;         // optional indirection (adaptive second pass): frame = frame_idx[i]
    const unsigned long long* count;   // optional device-side frame count (overrides B when non-null)
    int S, K, L, G, nruns, systematic, root_leaf;
    CrcSpec crc;
    const uint32_t* code;
    const uint32_t* runs;
    const uint32_t* crc_table;
    int smem_top;     // LLR stages s <= smem_top live in shared memory, larger ones in the arena
    int llr_off[20];  // word offset (shared memory) or element offset (arena) of LLR stage s in a lane's space
    int bitL_off[20]; // word offset of the left / right decision buffer of stage s (shared memory)
    int bitR_off[20];
    int words;        // shared-memory words per lane
    int arena_elems;  // arena floats per lane
    size_t arena_stride; // lanes the arena rows are laid out for (grid * threads)
    float* arena;     // [arena_elems][arena_stride]
    // tensor memory for the decision words (offsets carry kTmem), see SclParams
    int tm_alloc_cols, tm_cols_thread, scr_off, lockstep;
    int virt_top;     // LLR stage S-1 is not stored: its consumers recompute it from the channel row (pc_scl_f32.cu FgSrc)
    int arena_quad;   // arena LLR stages in the quad layout (pc_scl_f32.cu FLaneT::View)
};
struct SclFLaunch {
    SclFParams p;
    int threads, grid, per_sm;
    size_t smem_bytes;
};
cudaError_t launch_scl_f32(const SclFLaunch& L, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                           const unsigned* d_frame_idx, const unsigned long long* d_count, cudaStream_t st);
cudaError_t launch_compact_failed(const uint8_t* d_ok, size_t B, unsigned* d_idx, unsigned long long* d_count, cudaStream_t st);
// virt_ok: no leaf of the node program sits at stage S-1, so that stage may be virtual
int plan_scl_f32(int S, int K, int L, bool root_leaf, bool virt_ok, size_t smem_max, int sms, SclFLaunch& out, const char** msg);

// ---- fp32 Fast-SSC (pc_fastssc_f32.cu) ----
struct F32Params {
    const float* llr;
    uint8_t* info;
    uint8_t* ok;
    unsigned long long B;
    int S, K, nruns, staged_root, systematic;
    CrcSpec crc;
    const uint32_t* code;
    const uint32_t* runs;
    const uint32_t* crc_table;
    int llr_off[20];
    int elems;     // floats per thread in the arena
    int nbw;       // decision words per frame
    size_t arena_stride; // threads the arena rows are laid out for
    float* arena;  // [elems][arena_stride]
    uint32_t* bits_arena; // [nbw][grid*threads]
    int g8_cw_words; // group-of-8 kernel: shared-memory words per codeword (elems + nbw, padded to 8 mod 32)
    int sm_top;      // on-chip kernel: highest LLR stage kept in shared memory
    int sm_tm_cols;  // on-chip kernel: tensor-memory columns of a CTA holding stage S-1 (0: every stage in shared memory)
};

struct F32Launch {
    F32Params p;
    int threads, grid;
    int g8, g8_threads, g8_grid; // group-of-8 kernel (tree in shared memory) chosen for this code
    size_t g8_smem;
    int sm, sm_threads, sm_grid; // on-chip kernel (k_fastssc_f32_sm): threads per CTA (96 / 128), CTAs
    int sm_quad;                 // its shared-memory stages in the quad layout
    size_t sm_smem;
};
cudaError_t launch_fastssc_f32(const F32Launch& L, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, cudaStream_t st);
// resident CTAs per SM of the on-chip kernel for this plan (registers, shared memory), <= 0 on error
int fastssc_f32_sm_occupancy(const F32Launch& L);
// soft-output form (pc_decode_soft_batch_f32): p.arena / p.bits_arena laid out for grid * threads threads with N more words per thread
cudaError_t launch_fastssc_f32_soft(const F32Params& p, int grid, int threads, float* d_soft_code, float* d_soft_info, const uint32_t* d_lut,
                                    cudaStream_t st);
cudaError_t launch_i8_to_f32(const int8_t* d_in, float* d_out, size_t n, cudaStream_t st);

cudaError_t const_pool_upload(int offset_words, const uint32_t* h_src, int nwords, cudaStream_t st);
cudaError_t launch_fastssc_i8(const FastSscLaunch& L, const int8_t* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B,
                              cudaStream_t st);

} // namespace pc
#endif
