/* TEST ORACLE -- plain scalar C restatement of the reference's hot path (see pc_oracle.h).
 *
 * Style: every SIMD register of the reference is restated as an explicit array of V=32 int8 lanes
 * (include/polarcode/avxconvenience.h:52-54), including the pad values the "short" node variants write,
 * so that lane order, saturation order and tie-breaking are reproduced literally instead of being
 * argued. Parity pinned by tests/test_oracle_*.py against oracle/_ref (the reference itself) and the
 * reference's KATs (test/polarcode/ sources, python/qa_pypolar_detector.py).
 *
 * Not reproduced (undefined behaviour in the reference): minpos_epu8 on a vector whose 32 lanes are all
 * -128 indexes p[4] out of bounds (src/polarcode/avxconvenience.cpp:41-49); the restatement returns
 * lane 0 / value -128 there and tests avoid that input.
 */
#include "pc_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define V 32

static inline int8_t sat8(int x) { return (int8_t)(x > 127 ? 127 : (x < -128 ? -128 : x)); }
/* _mm256_adds_epi8 / _mm256_subs_epi8 lane */
static inline int8_t adds8(int8_t a, int8_t b) { return sat8((int)a + (int)b); }
static inline int8_t subs8(int8_t a, int8_t b) { return sat8((int)a - (int)b); }
/* _mm256_abs_epi8 lane: |-128| wraps to -128 */
static inline int8_t abs8(int8_t a) { return (int8_t)(a < 0 ? (uint8_t)(-(int)a) : a); }

/* ---------------------------------------------------------------- construction -------------------- */

/* trackingSorter::stableSortDescending, src/polarcode/arrayfuncs.cpp:108-124 (insertion sort, strict '<') */
static void stable_sort_desc(double* data, int* perm, int size)
{
    for (int i = 1; i < size; ++i) {
        double x = data[i];
        int y = perm[i];
        int j = i - 1;
        while (j >= 0 && data[j] < x) {
            data[j + 1] = data[j];
            perm[j + 1] = perm[j];
            j--;
        }
        data[j + 1] = x;
        perm[j + 1] = y;
    }
}

static int cmp_u(const void* a, const void* b)
{
    unsigned x = *(const unsigned*)a, y = *(const unsigned*)b;
    return (x > y) - (x < y);
}

/* Bhattacharrya::{setParameterByDesignSNR,calculateChannelParameters,construct},
 * src/polarcode/construction/bhattacharrya.cpp:39-78. Types as in the reference: the linear SNR and the
 * initial parameter are float, the recursion is double. */
int pco_frozen_bits_bb(int N, int K, float design_snr, unsigned* frozen_out)
{
    float lin = (float)pow(10.0, design_snr / 10.0);
    float init = (float)exp(-2.0 * lin * (size_t)K / (size_t)N);
    double* z = (double*)malloc(sizeof(double) * N);
    int* perm = (int*)malloc(sizeof(int) * N);
    int n = 0;
    while ((1 << n) < N) n++;
    z[0] = init;
    for (int stage = n - 1; stage >= 0; --stage) {
        int Bq = 1 << stage;
        for (int j = 0; j < N; j += (Bq << 1)) {
            volatile double T = z[j];
            volatile double T2 = T * T; /* volatile: keep the product rounded (no FMA contraction) */
            z[j + Bq] = T2;
            z[j] = 2 * T - T2;
        }
    }
    for (int i = 0; i < N; ++i) perm[i] = i;
    stable_sort_desc(z, perm, N);
    for (int i = 0; i < N - K; ++i) frozen_out[i] = (unsigned)perm[i];
    qsort(frozen_out, N - K, sizeof(unsigned), cmp_u);
    free(z);
    free(perm);
    return N - K;
}

/* BetaExpansion::{calculateChannelParameters,construct}, src/polarcode/construction/betaexpansion.cpp:42-82.
 * The reference argsorts with std::sort (not stable); weights of distinct indices are distinct sums of
 * distinct powers of 2^(1/4) except for exact ties, which we break by index (tests compare with _ref). */
typedef struct {
    double w;
    int i;
} be_item;
static int cmp_be(const void* a, const void* b)
{
    const be_item *x = (const be_item*)a, *y = (const be_item*)b;
    if (x->w < y->w) return -1;
    if (x->w > y->w) return 1;
    return (x->i > y->i) - (x->i < y->i);
}
int pco_frozen_bits_be(int N, int K, unsigned* frozen_out)
{
    int n = 0;
    while ((1 << n) < N) n++;
    const double beta = pow(2.0, 1.0 / 4.0);
    double wts[32];
    for (int j = 0; j < n; ++j) wts[j] = pow(beta, (double)j);
    be_item* it = (be_item*)malloc(sizeof(be_item) * N);
    for (int i = 0; i < N; ++i) {
        double w = 0.0;
        for (int j = 0; j < n; ++j) w += wts[j] * ((i >> j) & 1);
        it[i].w = w;
        it[i].i = i;
    }
    qsort(it, N, sizeof(be_item), cmp_be);
    for (int i = 0; i < N - K; ++i) frozen_out[i] = (unsigned)it[i].i;
    qsort(frozen_out, N - K, sizeof(unsigned), cmp_u);
    free(it);
    return N - K;
}

/* ---------------------------------------------------------------- int8 primitives ----------------- */

/* FastSscFip::F_function_calc, include/polarcode/decoding/fip_char.h:44-65 */
static inline int8_t f8(int8_t l, int8_t r)
{
    int8_t x = (int8_t)((l ^ r) | 1);            /* xor signs, never zero */
    int8_t a = l < -127 ? -127 : l;              /* max(l,-127) */
    int8_t b = r < -127 ? -127 : r;
    a = abs8(a);
    b = abs8(b);
    if (a < 1) a = 1;
    if (b < 1) b = 1;
    int8_t m = a < b ? a : b;
    return x < 0 ? (int8_t)-m : m;               /* _mm256_sign_epi8(m, x), x != 0 */
}
/* FastSscFip::G_function_calc, fip_char.h:67-73: blendv selects on the MSB of the bit byte */
static inline int8_t g8(int8_t l, int8_t r, int8_t bit) { return (bit & 0x80) ? subs8(r, l) : adds8(r, l); }

void pco_f_i8(const int8_t* a, const int8_t* b, int8_t* out, int n)
{
    for (int i = 0; i < n; ++i) out[i] = f8(a[i], b[i]);
}
void pco_g_i8(const int8_t* a, const int8_t* b, const int8_t* bits, int8_t* out, int n)
{
    for (int i = 0; i < n; ++i) out[i] = g8(a[i], b[i], bits[i]);
}

/* reduce_adds_epi8, include/polarcode/avxconvenience.h:92-101: (i,i+16) -> (i,i+8) -> (i,i+4) -> (i,i+2) -> (0,1) */
static int8_t reduce_adds32(const int8_t* x)
{
    int8_t t[16];
    for (int i = 0; i < 16; ++i) t[i] = adds8(x[i], x[i + 16]);
    for (int w = 8; w >= 1; w >>= 1)
        for (int i = 0; i < w; ++i) t[i] = adds8(t[i], t[i + w]);
    return t[0];
}
/* half_reduce_adds_epi8, avxconvenience.h:202-210: same tree stopped at stride 2 */
static void half_reduce_adds32(const int8_t* x, int8_t* even, int8_t* odd)
{
    int8_t t[16];
    for (int i = 0; i < 16; ++i) t[i] = adds8(x[i], x[i + 16]);
    for (int w = 8; w >= 2; w >>= 1)
        for (int i = 0; i < w; ++i) t[i] = adds8(t[i], t[i + w]);
    *even = t[0];
    *odd = t[1];
}
/* minpos_epu8, src/polarcode/avxconvenience.cpp:13-56: four groups of 8 (first min each, unsigned), then first
 * group whose min is the smallest among {m0..m3,127,127,127,127}; value returned through a signed char */
static unsigned minpos32(const int8_t* x, int8_t* val)
{
    unsigned gm[4], gi[4];
    for (int g = 0; g < 4; ++g) {
        gm[g] = 256;
        gi[g] = 0;
        for (int i = 0; i < 8; ++i) {
            unsigned u = (uint8_t)x[8 * g + i];
            if (u < gm[g]) {
                gm[g] = u;
                gi[g] = i;
            }
        }
    }
    unsigned best = 256, lane = 0;
    for (int g = 0; g < 4; ++g)
        if (gm[g] < best) {
            best = gm[g];
            lane = g;
        }
    if (best > 127) { /* dummy 127 entries win: out-of-bounds in the reference (UB); see file header */
        if (val) *val = (int8_t)-128;
        return 0;
    }
    if (val) *val = (int8_t)gm[lane];
    return gi[lane] + lane * 8;
}

/* ---------------------------------------------------------------- quantiser ----------------------- */

/* _mm256_cvtps_epi32 lane under the default MXCSR: round-to-nearest-even, out of range / NaN -> INT_MIN */
static inline int32_t cvtps(float x)
{
    if (!(x >= -2147483648.0f && x < 2147483648.0f)) return INT32_MIN;
    return (int32_t)nearbyintf(x);
}
static inline int8_t packs(int32_t v)
{ /* packs_epi32 then packs_epi16: saturate twice == saturate once */
    return sat8(v > 32767 ? 32767 : (v < -32768 ? -32768 : v));
}
/* CharContainer::insertLlr(const float*), src/polarcode/bitcontainer.cpp:505-516 */
void pco_quantize(int N, const float* in, int8_t* out)
{
    if (N >= 32) { /* convert_f32_to_int8_large :468-503 */
        for (int i = 0; i < N; ++i) out[i] = packs(cvtps(in[i]));
    } else if (N >= 8) { /* vectorizedFtoC :449-466: max_ps/min_ps return the 2nd operand on NaN */
        for (int i = 0; i < N; ++i) {
            float x = in[i];
            x = (x > -128.0f) ? x : -128.0f; /* _mm256_max_ps(x, min) */
            x = (x < 127.0f) ? x : 127.0f;   /* _mm256_min_ps(x, max) */
            x = nearbyintf(x);
            out[i] = (int8_t)cvtps(x);
        }
    } else { /* convertFtoC :34-39: clamp, round() half away from zero */
        for (int i = 0; i < N; ++i) {
            float x = fminf(fmaxf(in[i], -128.0f), 127.0f);
            out[i] = (int8_t)roundf(x);
        }
    }
}

/* ---------------------------------------------------------------- Fast-SSC int8 ------------------- */

enum {
    C_R0, C_R1, C_REP, C_SREP, C_SPC, C_SSPC, C_DREP, C_S01, C_S0SPC, C_SR1N, C_S0R, C_SRR, C_0SPC, C_R1N, C_0R, C_RR,
    C_COUNT
};

typedef struct {
    const unsigned* fz; /* sorted frozen indices relative to this node */
    int nf;
} fset;

/* splitFrozenBits, src/polarcode/polarcode.cpp:14-34 */
static void split(const unsigned* fz, int nf, int half, unsigned* l, int* nl, unsigned* r, int* nr)
{
    *nl = *nr = 0;
    for (int i = 0; i < nf; ++i) {
        if ((int)fz[i] < half)
            l[(*nl)++] = fz[i];
        else
            r[(*nr)++] = fz[i] - half;
    }
}

/* classification order of FastSscFip::createDecoder, src/polarcode/decoding/fastssc_fip_char.cpp:496-569 */
static int classify(int n, const unsigned* fz, int nf, int nl, int nr)
{
    if (nf == n) return C_R0;
    if (nf == 0) return C_R1;
    if (nf == n - 1) return n <= V ? C_SREP : C_REP;
    if (nf == 1) return n <= V ? C_SSPC : C_SPC;
    if (nf == n - 2 && n >= V) return C_DREP;
    (void)fz;
    if (n <= V) {
        if (nl == n / 2 && nr == 0) return C_S01;
        if (nl == n / 2 && nr == 1) return C_S0SPC;
        if (nr == 0) return C_SR1N;
        if (nl == n / 2) return C_S0R;
        return C_SRR;
    }
    if (nl == n / 2 && nr == 1) return C_0SPC;
    if (nr == 0) return C_R1N;
    if (nl == n / 2) return C_0R;
    return C_RR;
}

/* SpcDecoder::decode (long, n>32), fastssc_fip_char.cpp:274-303: bits = llr copy; cross-vector search with
 * signed-char compare and "minAbs > 0" early exit; flip is a two's complement negate of the stored byte */
static void spc_long(const int8_t* x, int8_t* bits, int n)
{
    uint8_t par = 0;
    unsigned minIdx = 0;
    int8_t testAbs, minAbs = 127;
    for (int v = 0; v < n / V; ++v) {
        int8_t ab[V];
        for (int i = 0; i < V; ++i) {
            bits[v * V + i] = x[v * V + i];
            par ^= (uint8_t)x[v * V + i];
            ab[i] = abs8(x[v * V + i]);
        }
        if (minAbs > 0) {
            unsigned vm = minpos32(ab, &testAbs);
            if (testAbs < minAbs) {
                minIdx = vm + v * V;
                minAbs = testAbs;
            }
        }
    }
    if (par & 0x80) bits[minIdx] = (int8_t)(-bits[minIdx]);
}
/* ShortSpcDecoder::decode (n<=32), :305-319: pad lanes >= n with 127 (SpcPrepare, fip_char.h:230-235) */
static void spc_short(const int8_t* x, int8_t* bits, int n)
{
    int8_t v[V], ab[V];
    uint8_t par = 0;
    for (int i = 0; i < V; ++i) {
        v[i] = i < n ? x[i] : 127;
        par ^= (uint8_t)v[i];
        ab[i] = abs8(v[i]);
    }
    if (par & 0x80) {
        unsigned vm = minpos32(ab, NULL);
        v[vm] = (int8_t)(-v[vm]);
    }
    memcpy(bits, v, n); /* lanes >= n of BitsOut are never read by a parent */
}

static void dec_node(int n, const unsigned* fz, int nf, const int8_t* llr, int8_t* bits, int* census)
{
    unsigned* lf = (unsigned*)malloc(sizeof(unsigned) * (nf + 1) * 2);
    unsigned* rf = lf + nf + 1;
    int nl, nr, h = n / 2;
    split(fz, nf, h, lf, &nl, rf, &nr);
    int c = classify(n, fz, nf, nl, nr);
    if (census) census[c]++;
    if (!llr) { /* census only */
        if (c >= C_SR1N && c != C_0SPC) {
            if (c != C_S0R && c != C_0R) dec_node(h, lf, nl, NULL, NULL, census);
            if (c != C_SR1N && c != C_R1N) dec_node(h, rf, nr, NULL, NULL, census);
            /* note: R1N/0R parents still construct both children in the reference; the absorbed
             * R0/R1 child objects are never decoded and not counted here */
        }
        free(lf);
        return;
    }
    int8_t* child = (int8_t*)malloc(n); /* ChildLlr */
    switch (c) {
    case C_R0: /* RateZeroDecoder :202-208 */
        memset(bits, 127, n);
        break;
    case C_R1: /* RateOneDecoder :210-215 */
        memcpy(bits, llr, n);
        break;
    case C_REP: { /* RepetitionDecoder :225-240: per-lane accumulate over vectors in order, then tree */
        int8_t acc[V];
        memset(acc, 0, V);
        for (int v = 0; v < n / V; ++v)
            for (int i = 0; i < V; ++i) acc[i] = adds8(acc[i], llr[v * V + i]);
        memset(bits, reduce_adds32(acc), n);
        break;
    }
    case C_SREP: { /* ShortRepetitionDecoder :265-272 with RepetitionPrepare zero pad */
        int8_t v[V];
        for (int i = 0; i < V; ++i) v[i] = i < n ? llr[i] : 0;
        memset(bits, reduce_adds32(v), n);
        break;
    }
    case C_DREP: { /* DoubleRepetitionDecoder :249-263 */
        int8_t acc[V], e, o;
        memset(acc, 0, V);
        for (int v = 0; v < n / V; ++v)
            for (int i = 0; i < V; ++i) acc[i] = adds8(acc[i], llr[v * V + i]);
        half_reduce_adds32(acc, &e, &o);
        for (int i = 0; i < n; ++i) bits[i] = (i & 1) ? o : e;
        break;
    }
    case C_SPC:
        spc_long(llr, bits, n);
        break;
    case C_SSPC:
        spc_short(llr, bits, n);
        break;
    case C_S01: { /* ShortZeroOneDecoder :390-399 */
        for (int i = 0; i < h; ++i) bits[i] = bits[i + h] = adds8(llr[i], llr[i + h]);
        break;
    }
    case C_S0SPC: { /* ShortZeroSpcDecoder :361-388: x = sat(a+b), pad 127, negate first min, duplicate */
        int8_t v[V], ab[V];
        uint8_t par = 0;
        for (int i = 0; i < V; ++i) {
            v[i] = i < h ? adds8(llr[i], llr[i + h]) : 127;
            par ^= (uint8_t)v[i];
            ab[i] = abs8(v[i]);
        }
        if (par & 0x80) {
            unsigned vm = minpos32(ab, NULL);
            v[vm] = (int8_t)(-v[vm]);
        }
        memcpy(bits, v, h);
        memcpy(bits + h, v, h);
        break;
    }
    case C_0SPC: { /* ZeroSpcDecoder :321-359 (n>=64, sub vectors of 32) */
        uint8_t par = 0;
        unsigned minIdx = 0;
        int8_t testAbs, minAbs = 127;
        for (int v = 0; v < h / V; ++v) {
            int8_t ab[V];
            for (int i = 0; i < V; ++i) {
                int8_t s = adds8(llr[v * V + i], llr[h + v * V + i]);
                bits[v * V + i] = bits[h + v * V + i] = s;
                par ^= (uint8_t)s;
                ab[i] = abs8(s);
            }
            if (minAbs > 0) {
                unsigned vm = minpos32(ab, &testAbs);
                if (testAbs < minAbs) {
                    minIdx = vm + v * V;
                    minAbs = testAbs;
                }
            }
        }
        if (par & 0x80) {
            bits[minIdx] = (int8_t)(-bits[minIdx]);
            bits[minIdx + h] = (int8_t)(-bits[minIdx + h]);
        }
        break;
    }
    case C_SRR:
    case C_RR: { /* RateRNode :401-412 / ShortRateRNode :414-425 */
        for (int i = 0; i < h; ++i) child[i] = f8(llr[i], llr[i + h]);
        dec_node(h, lf, nl, child, bits, census);
        for (int i = 0; i < h; ++i) child[i] = g8(llr[i], llr[i + h], bits[i]);
        dec_node(h, rf, nr, child, bits + h, census);
        if (c == C_RR) { /* CombineInPlace fip_char.h:164-171 */
            for (int i = 0; i < h; ++i) bits[i] = (int8_t)(bits[i] ^ bits[i + h]);
        } else { /* CombineBitsShort fip_char.h:187-207: max(-127) on both, xor; right half = max(right,-127) */
            for (int i = 0; i < h; ++i) {
                int8_t l = bits[i] < -127 ? -127 : bits[i];
                int8_t r = bits[i + h] < -127 ? -127 : bits[i + h];
                bits[i] = (int8_t)(l ^ r);
                bits[i + h] = r;
            }
        }
        break;
    }
    case C_SR1N:
    case C_R1N: { /* ROneNode :427-449 / ShortROneNode :451-474 */
        for (int i = 0; i < h; ++i) child[i] = f8(llr[i], llr[i + h]);
        dec_node(h, lf, nl, child, bits, census);
        for (int i = 0; i < h; ++i) {
            int8_t o = g8(llr[i], llr[i + h], bits[i]);
            bits[i] = (int8_t)(bits[i] ^ o);
            bits[i + h] = o;
        }
        break;
    }
    case C_S0R:
    case C_0R: { /* ZeroRNode :476-483 / ShortZeroRNode :485-492 */
        for (int i = 0; i < h; ++i) child[i] = adds8(llr[i], llr[i + h]);
        dec_node(h, rf, nr, child, bits + h, census);
        memcpy(bits, bits + h, h);
        break;
    }
    }
    free(child);
    free(lf);
}

void pco_fastssc_i8_census(int N, const unsigned* frozen, int nf, int* counts)
{
    memset(counts, 0, sizeof(int) * 20);
    dec_node(N, frozen, nf, NULL, NULL, counts);
}

static void build_lut(int N, const unsigned* frozen, int nf, unsigned* lut)
{ /* BitContainer::calculateLUT, src/polarcode/bitcontainer.cpp:68-84 */
    int k = 0, fc = 0;
    for (int i = 0; i < N; ++i) {
        if (fc < nf && (int)frozen[fc] == i)
            fc++;
        else
            lut[k++] = i;
    }
}

/* natural-order polar transform x = u F^{(x)n} on one bit per byte: stage s xors element i+2^s into i
 * (src/polarcode/encoding/butterfly_fip.cpp:31-63, src/polarcode/avxconvenience.cpp:58-84) */
static void transform_bits(uint8_t* b, int N)
{
    for (int d = 1; d < N; d <<= 1)
        for (int g = 0; g < N; g += 2 * d)
            for (int i = g; i < g + d; ++i) b[i] ^= b[i + d];
}

/* FastSscFipChar::decode, fastssc_fip_char.cpp:616-630 (without the detector check) */
void pco_fastssc_i8_decode(int N, const unsigned* frozen, int nf, int systematic, const int8_t* llr,
                           uint8_t* info_out, int8_t* soft_bits)
{
    int K = N - nf;
    int8_t* bits = (int8_t*)malloc(N);
    unsigned* lut = (unsigned*)malloc(sizeof(unsigned) * (K > 8 ? K : 8));
    dec_node(N, frozen, nf, llr, bits, NULL);
    if (soft_bits) memcpy(soft_bits, bits, N);
    build_lut(N, frozen, nf, lut);
    uint8_t* hard = (uint8_t*)malloc(N);
    for (int i = 0; i < N; ++i) hard[i] = ((uint8_t)bits[i]) >> 7;
    if (!systematic) transform_bits(hard, N); /* :619-622 re-encode non-systematically, then extract */
    memset(info_out, 0, (K + 7) / 8);
    for (int k = 0; k < K; ++k) /* CharContainer::getPackedInformationBits, bitcontainer.cpp:536-564 */
        info_out[k / 8] |= (uint8_t)(hard[lut[k]] << (7 - (k % 8)));
    free(hard);
    free(lut);
    free(bits);
}

/* ---------------------------------------------------------------- encoder ------------------------- */

/* ButterflyFipPacked::encode, src/polarcode/encoding/butterfly_fip_packed.cpp:45-59 (CRC already in info) */
void pco_encode(int N, const unsigned* frozen, int nf, int systematic, const uint8_t* info, uint8_t* code_out)
{
    int K = N - nf;
    uint8_t* b = (uint8_t*)calloc(N, 1);
    unsigned* lut = (unsigned*)malloc(sizeof(unsigned) * (K > 8 ? K : 8));
    build_lut(N, frozen, nf, lut);
    for (int k = 0; k < K; ++k) b[lut[k]] = (info[k / 8] >> (7 - (k % 8))) & 1; /* insertPackedInformationBits */
    transform_bits(b, N);
    if (systematic) {
        for (int i = 0; i < nf; ++i) b[frozen[i]] = 0; /* resetFrozenBits :1004-1010 */
        transform_bits(b, N);
    }
    memset(code_out, 0, N / 8);
    for (int i = 0; i < N; ++i) code_out[i / 8] |= (uint8_t)(b[i] << (7 - (i % 8)));
    free(lut);
    free(b);
}

/* ---------------------------------------------------------------- CRC ----------------------------- */

typedef struct {
    int width;
    uint32_t poly, init;
    int regbits; /* width of CRC++'s CRCType register for this parameter set */
} crcp;

static int crc_params(int kind, crcp* p)
{
    switch (kind) {
    case 6: *p = (crcp){ 6, 0x21, 0, 8 }; return 1;        /* CRC_6_NR   CRC.h:1307-1313 */
    case 8: *p = (crcp){ 8, 0x07, 0, 8 }; return 1;        /* crc8.cpp:18-31 */
    case 11: *p = (crcp){ 11, 0x621, 0, 16 }; return 1;    /* CRC_11_NR  CRC.h:1513-1519 */
    case 16: *p = (crcp){ 16, 0x1021, 0xFFFF, 16 }; return 1; /* CCITT-FALSE CRC.h:1704-1710 */
    case 17: *p = (crcp){ 16, 0x1021, 0, 16 }; return 1;   /* XMODEM = "CRC16NR" CRC.h:1986-1992 */
    case 24: *p = (crcp){ 24, 0xB2B117, 0, 32 }; return 1; /* CRC_24_NRC CRC.h:2155-2161 */
    }
    return 0;
}

/* CRC::CalculateBits / CalculateRemainder / CalculateRemainderBits for non-reflected parameter sets,
 * src/polarcode/errordetection/CRC.h:613-645, 870-988, 1059-1150. The partial last byte is XORed in whole
 * and then shifted numBits times, exactly like the reference (low bits of that byte leak into the result). */
static uint32_t crcpp_bits(const crcp* p, const uint8_t* data, size_t bits)
{
    const uint32_t regmask = p->regbits == 32 ? 0xFFFFFFFFu : ((1u << p->regbits) - 1);
    uint32_t rem = p->init;
    size_t whole = bits / 8, tail = bits % 8;
    if (p->width >= 8) {
        const int sh = p->width - 8;
        const uint32_t top = 1u << (p->width - 1);
        for (size_t i = 0; i < whole; ++i) {
            rem = (rem ^ ((uint32_t)data[i] << sh)) & regmask;
            for (int k = 0; k < 8; ++k) rem = ((rem & top) ? ((rem << 1) ^ p->poly) : (rem << 1)) & regmask;
        }
        if (tail) {
            rem = (rem ^ ((uint32_t)data[whole] << sh)) & regmask;
            for (size_t k = 0; k < tail; ++k) rem = ((rem & top) ? ((rem << 1) ^ p->poly) : (rem << 1)) & regmask;
        }
    } else {
        const int sh = 8 - p->width;
        const uint32_t poly = (p->poly << sh) & regmask;
        if (whole) {
            rem = (rem << sh) & regmask;
            for (size_t i = 0; i < whole; ++i) {
                rem = (rem ^ data[i]) & regmask;
                for (int k = 0; k < 8; ++k) rem = ((rem & 0x80) ? ((rem << 1) ^ poly) : (rem << 1)) & regmask;
            }
            rem = (rem >> sh) & regmask;
        }
        if (tail) {
            rem = ((rem << sh) ^ data[whole]) & regmask;
            for (size_t k = 0; k < tail; ++k) rem = ((rem & 0x80) ? ((rem << 1) ^ poly) : (rem << 1)) & regmask;
            rem = (rem >> sh) & regmask;
        }
    }
    const uint32_t wmask = p->width == 32 ? 0xFFFFFFFFu : ((1u << p->width) - 1);
    return rem & wmask; /* Finalize: finalXOR 0, no reflection */
}

/* _mm_crc32_u32 (CRC-32C, reflected 0x82F63B78) over little-endian 32-bit words, init 0, no final xor:
 * CRC32::gen, src/polarcode/errordetection/crc32.cpp:33-41 */
static uint32_t crc32c_words(const uint8_t* data, int words)
{
    uint32_t crc = 0;
    for (int i = 0; i < words * 4; ++i) {
        crc ^= data[i];
        for (int k = 0; k < 8; ++k) crc = (crc & 1) ? ((crc >> 1) ^ 0x82F63B78u) : (crc >> 1);
    }
    return crc;
}

static int crc_bytes(int kind)
{
    switch (kind) {
    case 6: case 8: return 1;
    case 11: case 16: case 17: return 2;
    case 24: return 3;
    case 32: return 4;
    }
    return 0;
}

/* Detector::calculate of each class. Note the reference's unit quirks, restated as they are:
 * CRC8/CRC16 take bits/8 whole bytes (crc8.cpp:36-39, crc16.cpp:20-23); CRC32 passes bits/8 as the WORD
 * count (crc32.cpp:28-31), i.e. it reads 4x the bytes -- callers must own bits/2 bytes of data. */
uint64_t pco_crc_calculate(int kind, const uint8_t* data, size_t bits)
{
    crcp p;
    if (kind == 32) return crc32c_words(data, (int)(bits / 8));
    if (!crc_params(kind, &p)) return 0;
    if (kind == 8 || kind == 16) return crcpp_bits(&p, data, (bits / 8) * 8);
    return crcpp_bits(&p, data, bits);
}

/* Detector::generate: CRC over the first bytes-c bytes, stored big-endian in the last c bytes
 * (crc24nrc.cpp:44-53, crc16.cpp:34-42, crc11nr.cpp:42-50, crc6nr.cpp:41-48, crc8.cpp:48-53);
 * CRC32 stores the native little-endian word (crc32.cpp:58-69) */
void pco_crc_generate(int kind, uint8_t* data, int bytes)
{
    crcp p;
    int c = crc_bytes(kind);
    if (c == 0) return;
    if (kind == 32) {
        int words = bytes / 4 - 1;
        uint32_t v = crc32c_words(data, words);
        memcpy(data + 4 * words, &v, 4);
        return;
    }
    crc_params(kind, &p);
    uint32_t v = crcpp_bits(&p, data, (size_t)8 * (bytes - c));
    for (int i = 0; i < c; ++i) data[bytes - c + i] = (uint8_t)(v >> (8 * (c - 1 - i)));
}

int pco_crc_check(int kind, const uint8_t* data, int bytes)
{
    crcp p;
    int c = crc_bytes(kind);
    if (c == 0) return 1; /* Dummy::check, dummy.cpp:25-33 */
    if (kind == 32) {
        int words = (bytes >> 2) - 1;
        uint32_t v = crc32c_words(data, words), s;
        memcpy(&s, data + 4 * words, 4);
        return v == s;
    }
    crc_params(kind, &p);
    uint32_t v = crcpp_bits(&p, data, (size_t)8 * (bytes - c)), s = 0;
    for (int i = 0; i < c; ++i) s = (s << 8) | data[bytes - c + i];
    return v == s;
}
