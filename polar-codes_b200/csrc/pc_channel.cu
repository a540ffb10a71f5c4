// Device-side channel simulation for BLER sweeps: what pcsim's SimulationWorker does per block on the host
// (reference src/simulation/simulator.cpp:632-672: generateData :850-867, encode, Bpsk::modulate
// src/signalprocessing/modulation/bpsk.cpp:54-81, Awgn::transmit src/signalprocessing/transmission/awgn.cpp:30-43,
// Scale, countErrors :939-965) as kernels, so a sweep never touches the host.
// Random numbers: Philox4x32-10 keyed by the seed, counter = (global frame index, word index, stream id), so results
// do not depend on how frames are sharded over launches or GPUs. (The reference draws from an RDRAND-seeded LCG,
// random.cpp:14-35; parity runs replay the generated LLRs through the oracle, so the generator itself is free.)
#include "pc_internal.h"

namespace pc {

struct Philox {
    uint32_t k0, k1;
    __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const
    {
        uint32_t a = k0, b = k1;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            c0 = hi1 ^ c1 ^ a;
            c1 = lo1;
            c2 = hi0 ^ c3 ^ b;
            c3 = lo0;
            a += 0x9E3779B9u;
            b += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};

// uniform random info blocks; bits beyond K in the last byte are zero
__global__ void __launch_bounds__(256) k_random_info(uint8_t* info, unsigned long long B, int K, int kb,
                                                     unsigned long long seed, unsigned long long first_frame)
{
    const Philox ph{ (uint32_t)seed, (uint32_t)(seed >> 32) };
    const int groups = (kb + 15) / 16; // 16 bytes per Philox call
    const unsigned long long total = B * (unsigned long long)groups;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long row = i / groups;
        const int gidx = (int)(i % groups);
        const unsigned long long frame = first_frame + row;
        const uint4 r = ph((uint32_t)frame, (uint32_t)(frame >> 32), (uint32_t)gidx, 0x1F0u);
        const uint32_t w[4] = { r.x, r.y, r.z, r.w };
        uint8_t* dst = info + row * (unsigned long long)kb;
        for (int j = 0; j < 16; ++j) {
            const int b = gidx * 16 + j;
            if (b >= kb) break;
            uint32_t v = (w[j >> 2] >> (8 * (j & 3))) & 0xFFu;
            if (b == kb - 1 && (K & 7)) v &= (0xFF00u >> (K & 7)) & 0xFFu;
            dst[b] = (uint8_t)v;
        }
    }
}

__device__ __forceinline__ int quantize_one(float x, int mode)
{
    int v;
    if (mode == 0) {
        v = (x >= -2147483648.0f && x < 2147483648.0f) ? __float2int_rn(x) : INT_MIN;
        v = v > 127 ? 127 : (v < -128 ? -128 : v);
    } else if (mode == 1) {
        float y = (x > -128.0f) ? x : -128.0f;
        y = (y < 127.0f) ? y : 127.0f;
        v = __float2int_rn(y);
    } else {
        v = (int)roundf(fminf(fmaxf(x, -128.0f), 127.0f));
    }
    return v;
}

// BPSK (bit 0 -> +1, bit 1 -> -1), y = x + sigma * n, LLR = amplification * y; four code bits per thread
template <bool FLOAT_OUT>
__global__ void __launch_bounds__(256) k_awgn(const uint8_t* __restrict__ code, void* __restrict__ llr_out,
                                              unsigned long long B, int N, float sigma, float amp, int qmode,
                                              unsigned long long seed, unsigned long long first_frame)
{
    const Philox ph{ (uint32_t)seed, (uint32_t)(seed >> 32) };
    const int quads = N >> 2, nbytes = N >> 3;
    const int lq = 31 - __clz(quads); // N is a power of two: row / column by shift and mask, not a 64-bit division
    const unsigned long long total = B * (unsigned long long)quads;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long row = i >> lq;
        const int q = (int)(i & (unsigned long long)(quads - 1));
        const unsigned long long frame = first_frame + row;
        const uint4 r = ph((uint32_t)frame, (uint32_t)(frame >> 32), (uint32_t)q, 0xA36Eu);
        // Box-Muller on two pairs of uniforms in (0,1]
        const float u0 = ((float)r.x + 1.0f) * 2.3283064365386963e-10f, u1 = (float)r.y * 2.3283064365386963e-10f;
        const float u2 = ((float)r.z + 1.0f) * 2.3283064365386963e-10f, u3 = (float)r.w * 2.3283064365386963e-10f;
        const float m0 = sqrtf(-2.0f * __logf(u0)), m1 = sqrtf(-2.0f * __logf(u2));
        float s0, c0, s1, c1;
        __sincosf(6.283185307179586f * u1, &s0, &c0);
        __sincosf(6.283185307179586f * u3, &s1, &c1);
        const float n[4] = { m0 * c0, m0 * s0, m1 * c1, m1 * s1 };
        const uint32_t byte = code[row * (unsigned long long)nbytes + (q >> 1)];
        const uint32_t nib = (q & 1) ? (byte & 0xFu) : (byte >> 4); // elements 4q..4q+3, MSB first
        float y[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float x = ((nib >> (3 - k)) & 1u) ? -1.0f : 1.0f;
            y[k] = amp * (x + sigma * n[k]);
        }
        if (FLOAT_OUT) {
            reinterpret_cast<float4*>(llr_out)[i] = make_float4(y[0], y[1], y[2], y[3]);
        } else {
            uint32_t w = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) w |= (uint32_t)(quantize_one(y[k], qmode) & 0xFF) << (8 * k);
            reinterpret_cast<uint32_t*>(llr_out)[i] = w;
        }
    }
}

// SimulationWorker::countErrors (simulator.cpp:939-965): frames, frame errors, bit errors over the K info bits,
// frames whose detector verdict was false
__global__ void __launch_bounds__(256) k_count_errors(const uint8_t* __restrict__ ref, const uint8_t* __restrict__ dec,
                                                      const uint8_t* __restrict__ ok, int K, int kb, unsigned long long B,
                                                      unsigned long long* counts)
{
    unsigned long long fe = 0, be = 0, nf = 0, rep = 0;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < B;
         r += (unsigned long long)gridDim.x * blockDim.x) {
        const uint8_t* a = ref + r * (unsigned long long)kb;
        const uint8_t* b = dec + r * (unsigned long long)kb;
        unsigned e = 0;
        for (int j = 0; j < kb; ++j) {
            uint32_t d = (uint32_t)(a[j] ^ b[j]);
            if (j == kb - 1 && (K & 7)) d &= (0xFF00u >> (K & 7)) & 0xFFu;
            e += __popc(d);
        }
        nf++;
        fe += e ? 1 : 0;
        be += e;
        if (ok && !ok[r]) rep++;
    }
    for (int o = 16; o; o >>= 1) {
        fe += __shfl_xor_sync(0xFFFFFFFFu, fe, o);
        be += __shfl_xor_sync(0xFFFFFFFFu, be, o);
        nf += __shfl_xor_sync(0xFFFFFFFFu, nf, o);
        rep += __shfl_xor_sync(0xFFFFFFFFu, rep, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (nf) atomicAdd(&counts[0], nf);
        if (fe) atomicAdd(&counts[1], fe);
        if (be) atomicAdd(&counts[2], be);
        if (rep) atomicAdd(&counts[3], rep);
    }
}

} // namespace pc

using namespace pc;

extern "C" {

int pc_awgn_generate(pc_code* c, uint64_t seed, uint64_t first_frame, float ebn0_db, float amplification,
                     int llr_is_float, void* d_llr_out, uint8_t* d_info_out, size_t B, void* stream)
{
    if (!c || !d_llr_out || !d_info_out) return fail(PC_ERR_ARG, "null argument");
    if (c->N < 8) return fail(PC_ERR_UNSUPPORTED, "channel generator needs N >= 8");
    if (B == 0) return PC_OK;
    if (int ds = check_code_device(c)) return ds;
    if (int ds = check_aligned(d_llr_out, 16, "d_llr_out")) return ds;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int kb = (c->K + 7) / 8, nbytes = c->N / 8;
    int s = ensure_bytes(&c->d_chan_code, &c->chan_code_bytes, (size_t)nbytes * B);
    if (s != PC_OK) return s;
    uint8_t* d_code = static_cast<uint8_t*>(c->d_chan_code);
    const int groups = (kb + 15) / 16;
    size_t g = (B * groups + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    k_random_info<<<(int)g, 256, 0, st>>>(d_info_out, B, c->K, kb, seed, first_frame);
    count_launch();
    s = crc_generate_rows(c->crc_kind, d_info_out, kb, B, st);
    if (s != PC_OK) return s;
    s = encode_device_raw(c, d_info_out, d_code, B, st);
    if (s != PC_OK) return s;
    // sigma = 1 / sqrt(2 Es/N0), Es/N0 = Eb/N0 * K/N (simulator.cpp:832-838, awgn.cpp:38-43)
    const float ebn0 = powf(10.0f, ebn0_db / 10.0f);
    const float sigma = 1.0f / sqrtf(2.0f * ebn0 * (float)c->K / (float)c->N);
    const int qmode = c->N >= 32 ? 0 : 1;
    g = ((size_t)B * (c->N / 4) + 255) / 256;
    if (g > 148 * 32) g = 148 * 32;
    if (llr_is_float)
        k_awgn<true><<<(int)g, 256, 0, st>>>(d_code, d_llr_out, B, c->N, sigma, amplification, qmode, seed, first_frame);
    else
        k_awgn<false><<<(int)g, 256, 0, st>>>(d_code, d_llr_out, B, c->N, sigma, amplification, qmode, seed, first_frame);
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "awgn launch");
}

int pc_count_errors(const uint8_t* d_info_ref, const uint8_t* d_info_dec, const uint8_t* d_ok, int K, size_t B,
                    unsigned long long* d_counts, void* stream)
{
    if (!d_info_ref || !d_info_dec || !d_counts || K < 1) return fail(PC_ERR_ARG, "bad argument");
    if (B == 0) return PC_OK;
    size_t g = (B + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    k_count_errors<<<(int)g, 256, 0, static_cast<cudaStream_t>(stream)>>>(d_info_ref, d_info_dec, d_ok, K, (K + 7) / 8, B,
                                                                          d_counts);
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "count_errors launch");
}

} // extern "C"
