// Micro-benchmark (not product): TMEM as per-thread private storage via tcgen05.st/ld 32x32b.x4.
#include <cstdio>
#include <cstdint>
#include <vector>
__device__ __forceinline__ void tm_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tm_ld4(uint32_t taddr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int MODE> // 0: TMEM, 1: smem
__global__ void k(uint32_t* out, long long* cycles, int cols_per_warp, int iters)
{
    extern __shared__ __align__(16) uint32_t sm[];
    __shared__ uint32_t tbase;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (MODE == 0) {
        if (w == 0) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbase)));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;");
    }
    const uint32_t my = MODE == 0 ? tbase + ((uint32_t)(32 * (w & 3)) << 16) + (uint32_t)((w >> 2) * cols_per_warp) : 0;
    uint32_t* smp = sm + 4 * threadIdx.x;
    const int cs = 4 * blockDim.x;
    const int nch = cols_per_warp / 4;
    uint32_t acc = 0;
    // fill
    for (int c = 0; c < nch; ++c) {
        uint32_t v = threadIdx.x * 1000 + c;
        if (MODE == 0) tm_st4(my + 4 * c, v, v + 1, v + 2, v + 3);
        else *reinterpret_cast<uint4*>(smp + c * cs) = make_uint4(v, v + 1, v + 2, v + 3);
    }
    if (MODE == 0) tm_wait_st();
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        // read two chunks, combine, write one (like f/g), stride through the region
        for (int c = 0; c < nch / 2; ++c) {
            uint32_t a0, a1, a2, a3, b0, b1, b2, b3;
            if (MODE == 0) {
                tm_ld4(my + 4 * c, a0, a1, a2, a3);
                tm_ld4(my + 4 * (c + nch / 2), b0, b1, b2, b3);
                tm_wait_ld();
                tm_st4(my + 4 * c, a0 ^ b0, a1 + b1, a2 ^ b2, a3 + b3);
            } else {
                uint4 a = *reinterpret_cast<uint4*>(smp + c * cs), b = *reinterpret_cast<uint4*>(smp + (c + nch / 2) * cs);
                *reinterpret_cast<uint4*>(smp + c * cs) = make_uint4(a.x ^ b.x, a.y + b.y, a.z ^ b.z, a.w + b.w);
                a0 = a.x;
            }
            acc += a0;
        }
        if (MODE == 0) tm_wait_st();
    }
    long long t1 = clock64();
    // verify readback of last chunk
    uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0;
    if (MODE == 0) { tm_ld4(my + 4 * (nch - 1), r0, r1, r2, r3); tm_wait_ld(); }
    else { uint4 a = *reinterpret_cast<uint4*>(smp + (nch - 1) * cs); r0 = a.x; r1 = a.y; r2 = a.z; r3 = a.w; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + r0 + r1 + r2 + r3;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    __syncthreads();
    if (MODE == 0 && w == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase));
}
int main()
{
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    for (int T : { 128, 256, 512 }) {
        int cpw = 512 / (T / 128);
        std::vector<uint32_t> h0(148 * T), h1(148 * T);
        for (int mode = 0; mode < 2; ++mode) {
            const int iters = 200;
            size_t smem = mode ? (size_t)T * cpw * 4 : 0;
            if (smem > 220000) { cpw = 220000 / (T * 4) / 8 * 8; smem = (size_t)T * cpw * 4; }
            if (mode) { cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); k<1><<<148, T, smem>>>(out, cyc, cpw, iters); }
            else k<0><<<148, T>>>(out, cyc, cpw, iters);
            cudaError_t e = cudaDeviceSynchronize();
            long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            cudaMemcpy(mode ? h1.data() : h0.data(), out, 148 * T * 4, cudaMemcpyDeviceToHost);
            double accesses = (double)iters * (cpw / 8) * 3; // per thread: 2 ld + 1 st of 16 B
            printf("T=%d mode=%s cols/warp=%d: %s, %lld cycles, %.2f cyc per 16B access per warp, %.1f B/clk/SM\n", T, mode ? "smem" : "tmem", cpw,
                   cudaGetErrorString(e), c, c / accesses, accesses * 16.0 * T / c);
        }
        if (cpw == 512 / (T / 128)) { int bad = 0; for (int i = 0; i < 148 * T; ++i) bad += h0[i] != h1[i]; printf("  tmem vs smem result mismatches: %d\n", bad); }
    }
    return 0;
}
