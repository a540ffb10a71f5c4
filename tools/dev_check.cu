// Debug harness (not product): compares the device paths of pc_swar.h primitives with their host paths.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../polar-codes_b200/csrc/pc_fastssc_i8_core.h"
using namespace pc;
__global__ void k(const uint32_t* a, const uint32_t* b, uint32_t* o, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t x = a[i], y = b[i];
    uint32_t* r = o + (size_t)i * 16;
    r[0] = sw_signfill(x); r[1] = sw_adds(x, y); r[2] = sw_sub(x, y); r[3] = sw_abs_sat(x); r[4] = sw_abs(x);
    r[5] = sw_minu(x, y); r[6] = sw_maxu(x, y); r[7] = sw_f(x, y); r[8] = sw_g(x, y, sw_expand4(x >> 3));
    r[9] = sw_expand4(x); r[10] = sw_movemask(x); r[11] = sw_hreduce1(sw_hreduce2(x)); r[12] = minu16x2(x, y);
    r[13] = msb_first_word(x); r[14] = minkey_word(0xFFFFFFFFu, sw_abs(x), 3); r[15] = parity_of_msbs(x);
}
static void href(uint32_t x, uint32_t y, uint32_t* r)
{
    r[0] = sw_signfill(x); r[1] = sw_adds(x, y); r[2] = sw_sub(x, y); r[3] = sw_abs_sat(x); r[4] = sw_abs(x);
    r[5] = sw_minu(x, y); r[6] = sw_maxu(x, y); r[7] = sw_f(x, y); r[8] = sw_g(x, y, sw_expand4(x >> 3));
    r[9] = sw_expand4(x); r[10] = sw_movemask(x); r[11] = sw_hreduce1(sw_hreduce2(x)); r[12] = minu16x2(x, y);
    r[13] = msb_first_word(x); r[14] = minkey_word(0xFFFFFFFFu, sw_abs(x), 3); r[15] = parity_of_msbs(x);
}
int main()
{
    const int n = 1 << 20;
    std::vector<uint32_t> a(n), b(n), o((size_t)n * 16);
    srand(1);
    const uint32_t special[] = { 0, 0x80808080u, 0x7f7f7f7fu, 0xffffffffu, 0x01010101u, 0x807f0001u, 0x81818181u };
    for (int i = 0; i < n; ++i) {
        a[i] = (uint32_t)rand() ^ ((uint32_t)rand() << 16);
        b[i] = (uint32_t)rand() ^ ((uint32_t)rand() << 16);
        if (i < 49) { a[i] = special[i % 7]; b[i] = special[i / 7]; }
    }
    uint32_t *da, *db, *dout;
    cudaMalloc(&da, n * 4); cudaMalloc(&db, n * 4); cudaMalloc(&dout, (size_t)n * 64);
    cudaMemcpy(da, a.data(), n * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(db, b.data(), n * 4, cudaMemcpyHostToDevice);
    k<<<n / 256, 256>>>(da, db, dout, n);
    cudaError_t e = cudaMemcpy(o.data(), dout, (size_t)n * 64, cudaMemcpyDeviceToHost);
    printf("cuda: %s\n", cudaGetErrorString(e));
    const char* names[16] = { "signfill", "adds", "sub", "abs_sat", "abs", "minu", "maxu", "f", "g", "expand4", "movemask",
                              "hreduce", "minu16x2", "msb_first", "minkey", "parity" };
    long bad[16] = { 0 };
    for (int i = 0; i < n; ++i) {
        uint32_t r[16];
        href(a[i], b[i], r);
        for (int j = 0; j < 16; ++j)
            if (r[j] != o[(size_t)i * 16 + j]) {
                if (bad[j]++ < 3) printf("%s a=%08x b=%08x host=%08x dev=%08x\n", names[j], a[i], b[i], r[j], o[(size_t)i * 16 + j]);
            }
    }
    for (int j = 0; j < 16; ++j) printf("%-10s mismatches %ld\n", names[j], bad[j]);
    return 0;
}
