// C-ABI entry points other than decode: detectors, quantiser (this file), encoder (pc_encode.cu), channel
// simulation (pc_channel.cu). Kernels are hand-written for sm_100a; no CPU fallback.
#include <mutex>

#include "pc_internal.h"

using namespace pc;

namespace pc {

// ------------------------------------------------------------------ detectors ----------------------
// One thread per row (rows are the info blocks, <= a few KiB); the 256-entry table sits in shared memory.
// generate: CRC over the first bytes-c bytes, stored in the last c bytes (Detector::generate, reference
// src/polarcode/errordetection/crc24nrc.cpp:44-53 and siblings); check: recompute and compare (:34-42).
template <bool GENERATE>
__global__ void __launch_bounds__(256) k_crc_rows(uint8_t* data, int bytes, uint8_t* ok, unsigned long long B, CrcSpec crc,
                                                  const uint32_t* g_table)
{
    __shared__ uint32_t table[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) table[i] = g_table[i];
    __syncthreads();
    const int payload = crc.reflected ? 4 * ((bytes >> 2) - 1) : bytes - crc.bytes;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < B;
         r += (unsigned long long)gridDim.x * blockDim.x) {
        uint8_t* row = data + r * (unsigned long long)bytes;
        CrcStream cs;
        cs.begin(crc);
        int i = 0;
        if ((reinterpret_cast<uintptr_t>(row) & 15u) == 0) {
            // whole 16-byte pieces of the row through one load each (a byte load per CRC step left the kernel waiting on
            // the load unit: 62 dependent byte loads per 64-byte row)
            const uint4* r4 = reinterpret_cast<const uint4*>(row);
            for (; i + 16 <= payload; i += 16) {
                const uint4 v = r4[i >> 4];
                const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
                for (int k = 0; k < 16; ++k) cs.payload_byte(crc, table, (w[k >> 2] >> (8 * (k & 3))) & 0xFFu);
            }
        }
        for (; i < payload; ++i) cs.payload_byte(crc, table, row[i]);
        if (GENERATE) {
            const uint32_t v = cs.value(crc);
            if (crc.reflected) { // native little-endian word (crc32.cpp:58-69)
                for (int i = 0; i < 4; ++i) row[payload + i] = (uint8_t)(v >> (8 * i));
            } else {
                for (int i = 0; i < crc.bytes; ++i) row[payload + i] = (uint8_t)(v >> (8 * (crc.bytes - 1 - i)));
            }
        } else {
            for (int i = 0; i < crc.bytes; ++i) cs.stored_byte(row[payload + i]);
            ok[r] = cs.ok(crc) ? 1 : 0;
        }
    }
}

__global__ void k_crc_true(uint8_t* ok, unsigned long long B)
{
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < B;
         r += (unsigned long long)gridDim.x * blockDim.x)
        ok[r] = 1;
}

__global__ void k_crc_calculate(const uint8_t* data, unsigned long long bits, CrcSpec crc, unsigned long long* out)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = crc_calculate_bits(crc, data, (size_t)bits);
}

// shared table cache: one device copy per (device, detector kind)
static std::mutex g_tab_mu;
constexpr int kMaxDevices = 64;
static uint32_t* g_tables[kMaxDevices][40] = { { nullptr } };

const uint32_t* crc_table_device(int kind)
{
    std::lock_guard<std::mutex> lk(g_tab_mu);
    if (kind < 0 || kind >= 40) return nullptr;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
    uint32_t** g_tab = g_tables[dev];
    if (g_tab[kind]) return g_tab[kind];
    CrcSpec s;
    if (!crc_spec(kind, s)) return nullptr;
    uint32_t h[256];
    crc_build_table(s, h);
    uint32_t* d = nullptr;
    if (cudaMalloc(&d, sizeof(h)) != cudaSuccess) return nullptr;
    if (cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaFree(d);
        return nullptr;
    }
    g_tab[kind] = d;
    return d;
}

static int grid_for(size_t B, int threads)
{
    size_t g = (B + threads - 1) / threads;
    if (g > 148 * 16) g = 148 * 16;
    return (int)(g ? g : 1);
}

int crc_generate_rows(int crc_kind, uint8_t* d_data, int bytes, size_t B, cudaStream_t st)
{
    CrcSpec s;
    if (!crc_spec(crc_kind, s)) return fail(PC_ERR_ARG, "CRC INVALID SIZE!");
    if (s.width == 0 || B == 0) return PC_OK; // Dummy::generate does nothing
    if (bytes < s.bytes || (s.reflected && bytes < 4)) return fail(PC_ERR_ARG, "row shorter than the check value");
    const uint32_t* tab = crc_table_device(crc_kind);
    if (!tab) return fail(PC_ERR_CUDA, "CRC table upload failed");
    k_crc_rows<true><<<grid_for(B, 256), 256, 0, st>>>(d_data, bytes, nullptr, B, s, tab);
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "crc generate launch");
}

// ------------------------------------------------------------------ quantiser ----------------------
// CharContainer::insertLlr(const float*) (reference src/polarcode/bitcontainer.cpp:505-516): three code paths by
// block length. N >= 32: cvtps_epi32 (round to nearest even; NaN and |x| >= 2^31 give INT_MIN) then two
// saturating packs, so +huge and NaN become -128 (:468-503). 8 <= N < 32: clamp to [-128,127] with max_ps/min_ps
// (NaN -> -128), then round to nearest even (:449-466). N < 8: clamp, round half away from zero (:34-39).
__global__ void __launch_bounds__(256) k_quantize(const float* __restrict__ in, int8_t* __restrict__ out,
                                                  unsigned long long n, int mode)
{
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const float x = in[i];
        int v;
        if (mode == 0) {
            v = (x >= -2147483648.0f && x < 2147483648.0f) ? __float2int_rn(x) : INT_MIN;
            v = v > 127 ? 127 : (v < -128 ? -128 : v);
        } else if (mode == 1) {
            float y = (x > -128.0f) ? x : -128.0f;
            y = (y < 127.0f) ? y : 127.0f;
            v = __float2int_rn(y);
        } else {
            float y = fminf(fmaxf(x, -128.0f), 127.0f); // NaN -> -128 like the reference's std::max/std::min order
            if (x != x) y = -128.0f;
            v = (int)roundf(y);
        }
        out[i] = (int8_t)v;
    }
}

// four floats -> one packed word per thread for the N >= 32 path (coalesced 16-byte loads, 4-byte stores)
__global__ void __launch_bounds__(256) k_quantize4(const float4* __restrict__ in, uint32_t* __restrict__ out,
                                                   unsigned long long n4)
{
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const float4 x = in[i];
        const float f[4] = { x.x, x.y, x.z, x.w };
        uint32_t w = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int v = (f[k] >= -2147483648.0f && f[k] < 2147483648.0f) ? __float2int_rn(f[k]) : INT_MIN;
            v = v > 127 ? 127 : (v < -128 ? -128 : v);
            w |= (uint32_t)(v & 0xFF) << (8 * k);
        }
        out[i] = w;
    }
}

int quantize_device(int N, const float* d_in, int8_t* d_out, size_t B, cudaStream_t st)
{
    if (N < 1 || !d_in || !d_out) return fail(PC_ERR_ARG, "pc_quantize_batch: bad argument");
    if (B == 0) return PC_OK;
    const unsigned long long n = (unsigned long long)N * B;
    if (N >= 32 && (reinterpret_cast<uintptr_t>(d_in) & 15) == 0 && (reinterpret_cast<uintptr_t>(d_out) & 3) == 0) {
        size_t g = (n / 4 + 255) / 256;
        if (g > 148 * 32) g = 148 * 32;
        k_quantize4<<<(int)g, 256, 0, st>>>(reinterpret_cast<const float4*>(d_in), reinterpret_cast<uint32_t*>(d_out), n / 4);
    } else {
        size_t g = (n + 255) / 256;
        if (g > 148 * 32) g = 148 * 32;
        k_quantize<<<(int)g, 256, 0, st>>>(d_in, d_out, n, N >= 32 ? 0 : (N >= 8 ? 1 : 2));
    }
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "quantize launch");
}

} // namespace pc

extern "C" {

int pc_crc_generate_batch(int crc_kind, uint8_t* d_data, int bytes, size_t B, void* stream)
{
    if (!d_data && B) return fail(PC_ERR_ARG, "null argument");
    return crc_generate_rows(crc_kind, d_data, bytes, B, static_cast<cudaStream_t>(stream));
}

int pc_crc_check_batch(int crc_kind, const uint8_t* d_data, int bytes, uint8_t* d_ok, size_t B, void* stream)
{
    CrcSpec s;
    if (!crc_spec(crc_kind, s)) return fail(PC_ERR_ARG, "CRC INVALID SIZE!");
    if ((!d_data || !d_ok) && B) return fail(PC_ERR_ARG, "null argument");
    if (B == 0) return PC_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (s.width == 0) { // Dummy::check is always true (dummy.cpp:25-33)
        k_crc_true<<<grid_for(B, 256), 256, 0, st>>>(d_ok, B);
    } else {
        if (bytes < s.bytes || (s.reflected && bytes < 4)) return fail(PC_ERR_ARG, "row shorter than the check value");
        const uint32_t* tab = crc_table_device(crc_kind);
        if (!tab) return fail(PC_ERR_CUDA, "CRC table upload failed");
        k_crc_rows<false><<<grid_for(B, 256), 256, 0, st>>>(const_cast<uint8_t*>(d_data), bytes, d_ok, B, s, tab);
    }
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "crc check launch");
}

// Detector::calculate on the device (one thread: this is the per-call scalar API, kept off the CPU on purpose)
uint64_t pc_crc_calculate(int crc_kind, const uint8_t* h_data, size_t bits)
{
    CrcSpec s;
    if (!crc_spec(crc_kind, s)) {
        set_error("CRC INVALID SIZE!");
        return 0;
    }
    if (s.width == 0) return 0;
    const size_t bytes = s.reflected ? (bits / 8) * 4 : (bits + 7) / 8;
    uint8_t* d = nullptr;
    unsigned long long* d_out = nullptr;
    unsigned long long h_out = 0;
    cudaError_t e = cudaMalloc(&d, bytes + 8);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, 8);
    if (e == cudaSuccess && bytes) e = cudaMemcpy(d, h_data, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_crc_calculate<<<1, 32>>>(d, bits, s, d_out);
        count_launch();
        e = cudaMemcpy(&h_out, d_out, 8, cudaMemcpyDeviceToHost);
    }
    cudaFree(d);
    cudaFree(d_out);
    if (e != cudaSuccess) cuda_fail(e, "pc_crc_calculate");
    return h_out;
}

int pc_quantize_batch(int N, const float* d_in, int8_t* d_out, size_t B, void* stream)
{
    return quantize_device(N, d_in, d_out, B, static_cast<cudaStream_t>(stream));
}

} // extern "C"
