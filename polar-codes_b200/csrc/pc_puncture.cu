// Batch puncturer / depuncturer: the index gather / scatter of the reference's
// PolarCode::Puncturer (src/polarcode/puncturer.cpp:52-89, include/polarcode/puncturer.h:66-101) over a batch of
// code words. HBM-bound copies: one thread per OUTPUT element so that every store is coalesced; the position table
// (and its inverse for the scatter direction) stays in L1/L2.
//
//   puncture<T>:      out[b][j] = in[b][pos[j]]            j < M   (M = punctured length, pos sorted ascending)
//   puncturePacked:   the same on bits, MSB-first bytes     (puncturer.cpp:71-89)
//   depuncture<T>:    out[b][p] = p punctured ? 0 : in[b][rank(p)]   p < Np = 2^ceil(log2 M)
#include <algorithm>
#include <string>
#include <vector>

#include "pc_internal.h"

struct pc_puncturer {
    size_t M = 0, Np = 0;
    std::vector<unsigned> pos; // output positions, ascending (Puncturer::mOutputPositions)
    unsigned* d_pos = nullptr; // [M]
    int* d_inv = nullptr;      // [Np]: index into the punctured word or -1
};

namespace pc {

template <class T>
__global__ void __launch_bounds__(256) k_puncture(const T* __restrict__ in, T* __restrict__ out, const unsigned* __restrict__ pos,
                                                  unsigned M, unsigned Np, unsigned long long total)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned long long b = i / M;
        const unsigned j = (unsigned)(i - b * M);
        out[i] = in[b * Np + __ldg(pos + j)];
    }
}

template <class T>
__global__ void __launch_bounds__(256) k_depuncture(const T* __restrict__ in, T* __restrict__ out, const int* __restrict__ inv,
                                                    unsigned M, unsigned Np, unsigned long long total)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned long long b = i / Np;
        const int j = __ldg(inv + (unsigned)(i - b * Np));
        out[i] = j >= 0 ? in[b * M + (unsigned)j] : T(0); // std::fill(..., 0) then the scatter, puncturer.h:94-100
    }
}

// V outputs per thread, written with one V * sizeof(T)-byte store (the 1-byte element kernels are store-instruction bound
// otherwise); used when both row lengths are multiples of V and the output base is aligned
template <class T, int V>
struct alignas(sizeof(T) * V) Pack {
    T v[V];
};
template <class T, int V>
__global__ void __launch_bounds__(256) k_puncture_v(const T* __restrict__ in, Pack<T, V>* __restrict__ out, const unsigned* __restrict__ pos,
                                                    unsigned Mv, unsigned Np, unsigned long long total, bool row_aligned)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned long long b = i / Mv;
        const unsigned jv = (unsigned)(i - b * Mv);
        const T* row = in + b * Np;
        if constexpr (sizeof(T) == 1 && V == 16) {
            // positions ascend, so first/last 15 apart means 16 consecutive bytes: two index loads instead of sixteen
            // (the index table is 4x the payload for 1-byte elements), the bytes through aligned words + funnel shifts
            const unsigned p0 = __ldg(pos + (size_t)jv * V), p15 = __ldg(pos + (size_t)jv * V + 15);
            if (p15 - p0 == 15u && row_aligned) {
                const uint32_t* w = reinterpret_cast<const uint32_t*>(row + (p0 & ~3u));
                const unsigned sh = (p0 & 3u) * 8u;
                const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = sh ? w[4] : 0u;
                uint4 o4;
                o4.x = __funnelshift_r(w0, w1, sh);
                o4.y = __funnelshift_r(w1, w2, sh);
                o4.z = __funnelshift_r(w2, w3, sh);
                o4.w = __funnelshift_r(w3, w4, sh);
                *reinterpret_cast<uint4*>(out + i) = o4;
                continue;
            }
        }
        const Pack<unsigned, V> p = *reinterpret_cast<const Pack<unsigned, V>*>(pos + (size_t)jv * V);
        Pack<T, V> o;
#pragma unroll
        for (int k = 0; k < V; ++k) o.v[k] = row[p.v[k]];
        out[i] = o;
    }
}
template <class T, int V>
__global__ void __launch_bounds__(256) k_depuncture_v(const T* __restrict__ in, Pack<T, V>* __restrict__ out, const int* __restrict__ inv,
                                                      unsigned M, unsigned Npv, unsigned long long total, bool row_aligned)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned long long b = i / Npv;
        const unsigned pv = (unsigned)(i - b * Npv);
        const T* row = in + b * M;
        if constexpr (sizeof(T) == 1 && V == 16) {
            const int j0 = __ldg(inv + (size_t)pv * V), j15 = __ldg(inv + (size_t)pv * V + 15);
            if (j0 >= 0 && j15 - j0 == 15 && row_aligned) { // sixteen surviving positions in a row
                const uint32_t* w = reinterpret_cast<const uint32_t*>(row + ((unsigned)j0 & ~3u));
                const unsigned sh = ((unsigned)j0 & 3u) * 8u;
                const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = sh ? w[4] : 0u;
                uint4 o4;
                o4.x = __funnelshift_r(w0, w1, sh);
                o4.y = __funnelshift_r(w1, w2, sh);
                o4.z = __funnelshift_r(w2, w3, sh);
                o4.w = __funnelshift_r(w3, w4, sh);
                *reinterpret_cast<uint4*>(out + i) = o4;
                continue;
            }
        }
        const Pack<int, V> j = *reinterpret_cast<const Pack<int, V>*>(inv + (size_t)pv * V);
        Pack<T, V> o;
#pragma unroll
        for (int k = 0; k < V; ++k) o.v[k] = j.v[k] >= 0 ? row[j.v[k]] : T(0);
        out[i] = o;
    }
}

// one thread per FOUR output bytes: bit 7-i of a byte is input bit pos[8 byte + i] (MSB first on both sides)
__global__ void __launch_bounds__(256) k_puncture_packed(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                         const unsigned* __restrict__ pos, unsigned Mb, unsigned Npb, unsigned long long total)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned Mq = (Mb + 3) >> 2; // groups of four output bytes per row
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned long long b = i / Mq;
        const unsigned q = (unsigned)(i - b * Mq);
        const uint8_t* row = in + b * Npb;
        uint8_t* orow = out + b * Mb;
        const unsigned nbytes = Mb - 4 * q < 4 ? Mb - 4 * q : 4;
        unsigned word = 0;
        for (unsigned y = 0; y < nbytes; ++y) {
            const uint4 p0 = __ldg(reinterpret_cast<const uint4*>(pos + 8 * (4 * q + y)));
            const uint4 p1 = __ldg(reinterpret_cast<const uint4*>(pos + 8 * (4 * q + y) + 4));
            const unsigned pp[8] = { p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w };
            unsigned v = 0;
            if (pp[7] - pp[0] == 7u) { // the usual case: eight consecutive input bits (positions ascend) -> one shift
                const unsigned by = pp[0] >> 3, sh = pp[0] & 7u;
                v = row[by];
                if (sh) v = ((v << 8) | row[by + 1]) >> (8 - sh); // sh != 0: the run ends in byte by + 1 < Npb
                v &= 0xFFu;
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) v |= ((row[pp[k] >> 3] >> (7 - (pp[k] & 7))) & 1u) << (7 - k);
            }
            word |= v << (8 * y);
        }
        if (nbytes == 4 && ((reinterpret_cast<uintptr_t>(orow) | (4 * q)) & 3u) == 0) {
            *reinterpret_cast<unsigned*>(orow + 4 * q) = word;
        } else {
            for (unsigned y = 0; y < nbytes; ++y) orow[4 * q + y] = (uint8_t)(word >> (8 * y));
        }
    }
}

static int grid_for(unsigned long long total)
{
    unsigned long long g = (total + 255) / 256;
    const unsigned long long cap = 148ull * 16;
    return (int)(g < cap ? (g ? g : 1) : cap);
}

template <class T, int V>
static int run_gather(const pc_puncturer* p, const void* d_in, void* d_out, size_t B, cudaStream_t st, bool scatter)
{
    const bool vec = V > 1 && (p->M % V) == 0 && (p->Np % V) == 0 && (reinterpret_cast<uintptr_t>(d_out) % (sizeof(T) * V)) == 0;
    if (!vec && V > 4) return run_gather<T, 4>(p, d_in, d_out, B, st, scatter);
    if (vec) {
        const bool in_al = (reinterpret_cast<uintptr_t>(d_in) & 3u) == 0; // rows start 4-byte aligned (both lengths are multiples of V)
        const unsigned long long total = (unsigned long long)B * ((scatter ? p->Np : p->M) / V);
        if (scatter)
            k_depuncture_v<T, V><<<grid_for(total), 256, 0, st>>>(static_cast<const T*>(d_in), static_cast<Pack<T, V>*>(d_out), p->d_inv,
                                                                  (unsigned)p->M, (unsigned)(p->Np / V), total, in_al);
        else
            k_puncture_v<T, V><<<grid_for(total), 256, 0, st>>>(static_cast<const T*>(d_in), static_cast<Pack<T, V>*>(d_out), p->d_pos,
                                                                (unsigned)(p->M / V), (unsigned)p->Np, total, in_al);
    } else {
        const unsigned long long total = (unsigned long long)B * (scatter ? p->Np : p->M);
        if (scatter)
            k_depuncture<T><<<grid_for(total), 256, 0, st>>>(static_cast<const T*>(d_in), static_cast<T*>(d_out), p->d_inv, (unsigned)p->M,
                                                             (unsigned)p->Np, total);
        else
            k_puncture<T><<<grid_for(total), 256, 0, st>>>(static_cast<const T*>(d_in), static_cast<T*>(d_out), p->d_pos, (unsigned)p->M,
                                                           (unsigned)p->Np, total);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, scatter ? "depuncture launch" : "puncture launch");
    count_launch();
    return PC_OK;
}

static int dispatch(const pc_puncturer* p, const void* d_in, void* d_out, int elem_bytes, size_t B, void* stream, bool scatter)
{
    if (!p) return fail(PC_ERR_ARG, "null puncturer");
    if (B == 0) return PC_OK;
    if (!d_in || !d_out) return fail(PC_ERR_ARG, "null argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (elem_bytes) {
    case 1: return run_gather<uint8_t, 16>(p, d_in, d_out, B, st, scatter);
    case 4: return run_gather<uint32_t, 4>(p, d_in, d_out, B, st, scatter); // float / int32: moved as bits, the fill is +0.0f
    case 8: return run_gather<unsigned long long, 2>(p, d_in, d_out, B, st, scatter);
    default: return fail(PC_ERR_ARG, "element size must be 1, 4 or 8 bytes");
    }
}

} // namespace pc

extern "C" {

pc_puncturer* pc_puncturer_create(size_t block_length, const unsigned* frozen, int n_frozen)
{
    using namespace pc;
    if (block_length == 0 || (!frozen && n_frozen > 0) || n_frozen < 0) {
        set_error("bad puncturer arguments");
        return nullptr;
    }
    size_t np = 1;
    while (np < block_length) np <<= 1; // round_up_power_of_two, puncturer.cpp:23-33
    const size_t n_punct = np - block_length;
    if (n_punct > (size_t)n_frozen) { // std::out_of_range in the reference, puncturer.cpp:57-60
        set_error("Number of required puncturing positions exceeds frozen bit positions!");
        return nullptr;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no usable CUDA device (this library has no CPU path)");
        return nullptr;
    }
    // the first n_punct frozen positions (in the caller's order) are punctured; outputs are the remaining positions
    // in ascending order (inverse_set_difference, puncturer.cpp:35-49 -- std::set_difference expects the punctured
    // set sorted, which holds for the sorted frozen sets the constructors return)
    std::vector<char> gone(np, 0);
    for (size_t i = 0; i < n_punct; ++i) {
        if (frozen[i] >= np) {
            set_error("frozen bit position outside the parent block");
            return nullptr;
        }
        gone[frozen[i]] = 1;
    }
    pc_puncturer* p = new pc_puncturer();
    p->M = block_length;
    p->Np = np;
    std::vector<int> inv(np, -1);
    for (size_t q = 0; q < np; ++q)
        if (!gone[q]) {
            inv[q] = (int)p->pos.size();
            p->pos.push_back((unsigned)q);
        }
    if (p->pos.size() != block_length) { // duplicate entries among the punctured positions
        set_error("punctured positions are not distinct");
        delete p;
        return nullptr;
    }
    cudaError_t e = cudaMalloc(&p->d_pos, sizeof(unsigned) * block_length);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_inv, sizeof(int) * np);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_pos, p->pos.data(), sizeof(unsigned) * block_length, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_inv, inv.data(), sizeof(int) * np, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cuda_fail(e, "puncturer tables");
        cudaFree(p->d_pos);
        cudaFree(p->d_inv);
        delete p;
        return nullptr;
    }
    return p;
}

void pc_puncturer_destroy(pc_puncturer* p)
{
    if (!p) return;
    cudaFree(p->d_pos);
    cudaFree(p->d_inv);
    delete p;
}

size_t pc_puncturer_block_length(const pc_puncturer* p) { return p ? p->M : 0; }
size_t pc_puncturer_parent_block_length(const pc_puncturer* p) { return p ? p->Np : 0; }
int pc_puncturer_positions(const pc_puncturer* p, unsigned* out)
{
    if (!p || !out) return pc::fail(PC_ERR_ARG, "null argument");
    std::copy(p->pos.begin(), p->pos.end(), out);
    return (int)p->pos.size();
}

int pc_puncture_batch(const pc_puncturer* p, const void* d_in, void* d_out, int elem_bytes, size_t B, void* stream)
{
    return pc::dispatch(p, d_in, d_out, elem_bytes, B, stream, false);
}

int pc_depuncture_batch(const pc_puncturer* p, const void* d_in, void* d_out, int elem_bytes, size_t B, void* stream)
{
    return pc::dispatch(p, d_in, d_out, elem_bytes, B, stream, true);
}

int pc_puncture_packed_batch(const pc_puncturer* p, const uint8_t* d_in, uint8_t* d_out, size_t B, void* stream)
{
    using namespace pc;
    if (!p) return fail(PC_ERR_ARG, "null puncturer");
    if (B != 0 && (!d_in || !d_out)) return fail(PC_ERR_ARG, "null argument");
    if ((p->M % 8) != 0 || (p->Np % 8) != 0) return fail(PC_ERR_ARG, "packed puncturing needs block lengths that are multiples of 8");
    if (B == 0) return PC_OK;
    const unsigned long long total = (unsigned long long)B * ((p->M / 8 + 3) / 4);
    k_puncture_packed<<<grid_for(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_in, d_out, p->d_pos, (unsigned)(p->M / 8),
                                                                                       (unsigned)(p->Np / 8), total);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "puncture (packed) launch");
    count_launch();
    return PC_OK;
}
}
