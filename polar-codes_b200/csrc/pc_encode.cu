// Packed polar encoder for sm_100a: a group of G = min(32, N/32) lanes owns one codeword, lane l holds code words
// l, l+32, ... (32 bits each) in registers. Replaces Encoding::ButterflyFipPacked::encode of the reference
// (src/polarcode/encoding/butterfly_fip_packed.cpp:45-59): scatter the K info bits to the non-frozen positions
// (PackedContainer::insertPackedInformationBits, src/polarcode/bitcontainer.cpp:886-898), transform
// x = u F^(x)n (butterfly_fip.cpp:15-63), and for systematic codes zero the frozen positions
// (resetFrozenBits, bitcontainer.cpp:1004-1010) and transform again.
//   stages 0..4   inside a word: x ^= (x >> 2^d) & mask_d          (element i takes element i + 2^d)
//   stages 5..9   across lanes: __shfl_xor_sync, lanes with the stage bit clear take their partner
//   stages >= 10  across the R registers of a lane
// I/O is coalesced: a group reads its info row through shared memory and writes N/8 contiguous bytes.
// Bit order (reference): element i is bit 7-(i%8) of byte i/8; in registers element i is bit i%32 of word i/32.
#include <cstdlib>

#include "pc_fastssc_i8_core.h"
#include "pc_internal.h"

namespace pc {

struct EncParams {
    const uint8_t* info;  // [B][kb] (check bytes already in place)
    uint8_t* code;        // [B][N/8]
    unsigned long long B;
    int S, K, kb, kbp;    // kbp = kb rounded up to a multiple of 4 (row stride in shared memory)
    int systematic;
    int direct;           // info rows are word aligned: read them without staging
    const uint32_t* cnt;       // cnt[w] = number of info positions below code word w (N/32 + 1 entries)
    const uint32_t* info_mask; // bit i%32 of word i/32 set for non-frozen positions
};

PC_HD uint32_t inword_transform(uint32_t x, int S)
{
    const uint32_t lim = lowmask(1 << (S < 5 ? S : 5));
    x ^= (x >> 1) & 0x55555555u & lim;
    if (S > 1) x ^= (x >> 2) & 0x33333333u & lim;
    if (S > 2) x ^= (x >> 4) & 0x0F0F0F0Fu & lim;
    if (S > 3) x ^= (x >> 8) & 0x00FF00FFu & lim;
    if (S > 4) x ^= (x >> 16) & 0x0000FFFFu;
    return x;
}

// software PDEP: the low bits of `src` go, in order, to the set bits of `mask`, one contiguous run of the mask at a
// time (polar information sets are made of few long runs per word)
__device__ __forceinline__ uint32_t deposit_bits(uint32_t src, uint32_t mask)
{
    uint32_t out = 0;
    while (mask) {
        const uint32_t low = mask & (0u - mask);   // lowest set bit
        const uint32_t t = mask + low;             // clears the lowest run (carry ripples through it)
        const uint32_t run = mask & ~t;            // the lowest run of ones
        const int pos = __ffs(low) - 1, len = __popc(run);
        out |= (src << pos) & run;
        src >>= len;
        mask &= t;
    }
    return out;
}

template <int R>
__device__ __forceinline__ void transform_regs(uint32_t (&x)[R], int S, int G, int lane_in_group)
{
#pragma unroll
    for (int r = 0; r < R; ++r) x[r] = inword_transform(x[r], S);
    // word-index bits 0..4 live in the lane index
    for (int b = 1; b < G; b <<= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t v = __shfl_xor_sync(0xFFFFFFFFu, x[r], b);
            if (!(lane_in_group & b)) x[r] ^= v;
        }
    }
    // word-index bits >= 5 live in the register index
#pragma unroll
    for (int b = 1; b < R; b <<= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (!(r & b)) x[r] ^= x[r | b];
    }
}

template <int R>
__global__ void __launch_bounds__(256) k_encode(const EncParams p)
{
    extern __shared__ __align__(16) uint8_t s_info[];
    const int W = p.S > 5 ? (1 << (p.S - 5)) : 1; // code words per codeword
    const int G = W < 32 ? W : 32;                // lanes per codeword
    const int per_warp = 32 / G, per_cta = (blockDim.x >> 5) * per_warp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lig = lane % G, slot = warp * per_warp + lane / G; // codeword slot inside the CTA
    const int nbytes = p.S >= 3 ? (1 << (p.S - 3)) : 1;
    // the lane's code words never change: keep their masks and info offsets in registers (short and medium codes)
    constexpr bool HOIST = R <= 16;
    constexpr int HR = HOIST ? R : 1;
    uint32_t mk[HR], cn_[HR];
#pragma unroll
    for (int r = 0; r < HR; ++r) {
        mk[r] = HOIST ? __ldg(p.info_mask + r * 32 + lig) : 0u;
        cn_[r] = HOIST ? __ldg(p.cnt + r * 32 + lig) : 0u;
    }
    for (unsigned long long base = (unsigned long long)blockIdx.x * per_cta; base < p.B;
         base += (unsigned long long)gridDim.x * per_cta) {
        const unsigned long long left = p.B - base;
        const int ncw = left < (unsigned long long)per_cta ? (int)left : per_cta;
        const bool live = slot < ncw;
        const uint32_t* row;
        if (p.direct) {
            // rows are 4-byte aligned (kb % 4 == 0): the lanes of a group read their info words straight from HBM
            // (one row is 1-2 cache lines shared by the whole group), no staging, no CTA barrier
            row = reinterpret_cast<const uint32_t*>(p.info + (base + (live ? slot : 0)) * (unsigned long long)p.kb);
        } else {
            // stage the contiguous info rows of this batch, padded to a 4-byte row stride
            __syncthreads();
            const uint8_t* src = p.info + base * (unsigned long long)p.kb;
            for (int i = threadIdx.x; i < ncw * p.kb; i += blockDim.x) s_info[(i / p.kb) * p.kbp + (i % p.kb)] = src[i];
            __syncthreads();
            row = reinterpret_cast<const uint32_t*>(s_info + (live ? slot : 0) * p.kbp);
        }
        // scatter (PackedContainer::insertPackedInformationBits): code word w takes the info bits cnt[w] .. cnt[w+1]-1,
        // a contiguous piece of the info stream (funnel shift of two row words), deposited at the set bits of its mask
        uint32_t x[R];
        const int nrw = p.kbp >> 2;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t m = HOIST ? mk[r < HR ? r : 0] : __ldg(p.info_mask + r * 32 + lig);
            const uint32_t cn = HOIST ? cn_[r < HR ? r : 0] : __ldg(p.cnt + r * 32 + lig);
            uint32_t acc = 0;
            if (m) {
                const int idx = (int)(cn >> 5), sh = (int)(cn & 31u);
                const uint32_t lo = msb_first_word(p.direct ? __ldg(row + idx) : row[idx]); // info bits, LSB first
                const uint32_t hi = (idx + 1 < nrw) ? msb_first_word(p.direct ? __ldg(row + idx + 1) : row[idx + 1]) : 0u;
                acc = deposit_bits(__funnelshift_r(lo, hi, sh), m);
            }
            x[r] = acc;
        }
        transform_regs<R>(x, p.S, G, lig);
        if (p.systematic) {
#pragma unroll
            for (int r = 0; r < R; ++r) x[r] &= p.info_mask[r * 32 + lig];
            transform_regs<R>(x, p.S, G, lig);
        }
        if (live) {
            uint8_t* out = p.code + (base + slot) * (unsigned long long)nbytes;
            if (p.S >= 5) {
#pragma unroll
                for (int r = 0; r < R; ++r) reinterpret_cast<uint32_t*>(out)[r * 32 + lig] = msb_first_word(x[r]);
            } else {
                const uint32_t w = msb_first_word(x[0]);
                for (int j = 0; j < nbytes; ++j) out[j] = (uint8_t)(w >> (8 * j));
            }
        }
    }
}

template <int R>
static cudaError_t launch_encode(const EncParams& p, int grid, size_t smem, cudaStream_t st)
{
    auto kern = k_encode<R>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    kern<<<grid, 256, smem, st>>>(p);
    return cudaGetLastError();
}

// host side of the scatter table: prefix count of information positions per code word
int build_encoder_tables(pc_code* c)
{
    const FastSscProgram& g = c->prog;
    const int W = (c->N + 31) / 32;
    std::vector<uint32_t> cnt(W + 1, 0u);
    for (int w = 0; w < W; ++w) cnt[w + 1] = cnt[w] + (uint32_t)__builtin_popcount(g.info_mask[w]);
    cudaFree(c->d_enc_first);
    c->d_enc_first = nullptr;
    cudaError_t e = cudaMalloc(&c->d_enc_first, cnt.size() * 4);
    if (e == cudaSuccess) e = cudaMemcpy(c->d_enc_first, cnt.data(), cnt.size() * 4, cudaMemcpyHostToDevice);
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "encoder tables");
}

bool encode2_supported(const pc_code* c);
bool encode2_fuses_crc(const pc_code* c);
int encode2_device_raw(pc_code* c, const uint8_t* src, uint8_t* d_code, size_t B, cudaStream_t st, bool fuse_crc);
static const bool g_force_old_encoder = getenv("PC_ENC_OLD") != nullptr; // measurements: keep this file's kernel

int encode_device_raw(pc_code* c, const uint8_t* src, uint8_t* d_code, size_t B, cudaStream_t st)
{
    // 32 <= N <= 16384: the register-block kernel (pc_encode2.cu); PC_ENC_OLD=1 keeps this file's kernel for measurements
    if (!g_force_old_encoder && encode2_supported(c)) return encode2_device_raw(c, src, d_code, B, st, false);
    const int kb = (c->K + 7) / 8;
    EncParams p;
    p.info = src;
    p.code = d_code;
    p.B = B;
    p.S = c->S;
    p.K = c->K;
    p.kb = kb;
    p.kbp = (kb + 3) & ~3;
    p.systematic = c->systematic;
    p.direct = (kb % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 3u) == 0) ? 1 : 0;
    p.cnt = c->d_enc_first;
    p.info_mask = c->d_info_mask;
    const int W = c->S > 5 ? (1 << (c->S - 5)) : 1, G = W < 32 ? W : 32, R = W / G;
    const int per_cta = 8 * (32 / G);
    const size_t smem = (size_t)per_cta * p.kbp;
    size_t ctas = (B + per_cta - 1) / per_cta;
    int grid = (int)(ctas < 148 * 8 ? ctas : 148 * 8);
    cudaError_t e;
    switch (R) {
    case 1: e = launch_encode<1>(p, grid, smem, st); break;
    case 2: e = launch_encode<2>(p, grid, smem, st); break;
    case 4: e = launch_encode<4>(p, grid, smem, st); break;
    case 8: e = launch_encode<8>(p, grid, smem, st); break;
    case 16: e = launch_encode<16>(p, grid, smem, st); break;
    case 32: e = launch_encode<32>(p, grid, smem, st); break;
    case 64: e = launch_encode<64>(p, grid, smem, st); break;
    default: return fail(PC_ERR_UNSUPPORTED, "block length too large for the encoder");
    }
    if (e != cudaSuccess) return cuda_fail(e, "encode launch");
    count_launch();
    return PC_OK;
}


int encode_device(pc_code* c, const uint8_t* d_info, uint8_t* d_code, uint8_t* d_info_crc, size_t B, cudaStream_t st)
{
    if (!c || !d_info || !d_code) return fail(PC_ERR_ARG, "null argument");
    if (B == 0) return PC_OK;
    if (c->N < 8) return fail(PC_ERR_UNSUPPORTED, "encoder needs N >= 8 (the reference's packed container asserts N % 8 == 0)");
    if (int ds = check_code_device(c)) return ds;
    if (c->N >= 32)
        if (int ds = check_aligned(d_code, 4, "d_code")) return ds; // code rows are stored as words
    const int kb = (c->K + 7) / 8;
    CrcSpec crc;
    crc_spec(c->crc_kind, crc);
    const uint8_t* src = d_info;
    // nobody wants the info blocks with their check bytes back and the code is short enough for one thread per
    // codeword: the detector runs inside the encoder kernel on the staged rows (no copy, no extra pass over HBM)
    if (crc.width && !d_info_crc && !g_force_old_encoder && encode2_fuses_crc(c) && !getenv("PC_ENC_NOFUSE"))
        return encode2_device_raw(c, d_info, d_code, B, st, true);
    if (crc.width || d_info_crc) {
        // the reference writes the check bytes into the caller's buffer (butterfly_fip_packed.cpp:47-49); here they go
        // into a device-side copy
        uint8_t* copy = d_info_crc;
        if (!copy) {
            int s = ensure_bytes(reinterpret_cast<void**>(&c->d_enc_info), &c->enc_info_bytes, (size_t)kb * B);
            if (s != PC_OK) return s;
            copy = c->d_enc_info;
        }
        cudaError_t e = cudaMemcpyAsync(copy, d_info, (size_t)kb * B, cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return cuda_fail(e, "info copy");
        int s = crc_generate_rows(c->crc_kind, copy, kb, B, st);
        if (s != PC_OK) return s;
        src = copy;
    }
    return encode_device_raw(c, src, d_code, B, st);
}

} // namespace pc

using namespace pc;

extern "C" {

int pc_encode_batch(pc_code* c, const uint8_t* d_info, uint8_t* d_code, uint8_t* d_info_crc, size_t B, void* stream)
{
    return encode_device(c, d_info, d_code, d_info_crc, B, static_cast<cudaStream_t>(stream));
}

int pc_encode_batch_host(pc_code* c, const uint8_t* h_info, uint8_t* h_code, uint8_t* h_info_crc, size_t B)
{
    if (!c || !h_info || !h_code) return fail(PC_ERR_ARG, "null argument");
    if (B == 0) return PC_OK;
    if (int ds = check_code_device(c)) return ds;
    const size_t kb = (size_t)(c->K + 7) / 8, nb = (size_t)(c->N >= 8 ? c->N / 8 : 1);
    // staging layout: [info B*kb | info+crc B*kb] in d_stage_in, code in d_stage_out
    int s = ensure_bytes(&c->d_stage_in, &c->stage_in_bytes, 2 * kb * B + 16);
    if (s == PC_OK) s = ensure_bytes(&c->d_stage_out, &c->stage_out_bytes, nb * B);
    if (s != PC_OK) return s;
    uint8_t* d_info = static_cast<uint8_t*>(c->d_stage_in);
    uint8_t* d_icrc = d_info + ((kb * B + 15) & ~(size_t)15);
    cudaError_t e = cudaMemcpy(d_info, h_info, kb * B, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return cuda_fail(e, "H2D");
    s = encode_device(c, d_info, static_cast<uint8_t*>(c->d_stage_out), d_icrc, B, nullptr);
    if (s != PC_OK) return s;
    e = cudaMemcpy(h_code, c->d_stage_out, nb * B, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && h_info_crc) e = cudaMemcpy(h_info_crc, d_icrc, kb * B, cudaMemcpyDeviceToHost);
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "encode (host batch)");
}

} // extern "C"
