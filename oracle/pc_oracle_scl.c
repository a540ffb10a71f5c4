/* TEST ORACLE -- scalar C restatement of the reference's int8 CRC-aided SCL decoder SclFipChar
 * (src/polarcode/decoding/scl_fip_char.cpp). Test infrastructure only (see pc_oracle.h).
 *
 * The reference keeps per-path trees of ref-counted blocks with lazy copies (PathList, DataPool); the restatement
 * copies whole path states on a fork instead -- same values, no sharing. Everything that decides a result is
 * restated literally: int64 metrics (larger = better), the candidate metrics of each leaf type, the swap-based
 * selection sorts (include/polarcode/arrayfuncs.h:161-183, 209-231) whose tie order depends on the swaps, the
 * `~` bit flips, and extractBestPath's first-CRC-pass rule (:816-855).
 * Pinned against oracle/_ref by tests/test_oracle_scl.py. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "pc_oracle.h"

static inline int8_t sat8(int x) { return (int8_t)(x > 127 ? 127 : (x < -128 ? -128 : x)); }
/* F_function_calc / G_function_calc, include/polarcode/decoding/fip_char.h:44-73 */
static inline int8_t f8(int8_t a, int8_t b)
{
    int x = a < -127 ? 127 : (a < 0 ? -a : a), y = b < -127 ? 127 : (b < 0 ? -b : b);
    int m = x < y ? x : y;
    if (m < 1) m = 1;
    return (int8_t)(((a < 0) != (b < 0)) ? -m : m);
}
static inline int8_t g8(int8_t a, int8_t b, int8_t bit) { return (bit & 0x80) ? sat8((int)b - a) : sat8((int)b + a); }

typedef struct {
    int N, S, L;
    int P;          /* live paths */
    /* per path: llr[s] (2^s int8), bit[s], lbit[s] for s = 0..S, laid out as arrays of size 2N per kind */
    int8_t *llr, *bit, *lbit;
    long long* metric;
    /* next generation */
    int8_t *nllr, *nbit, *nlbit;
    long long* nmetric;
    const unsigned* frozen;
    int nf;
} scl;

#define AT(base, p, s) ((base) + (size_t)(p) * 2 * d->N + ((size_t)1 << (s)))

/* simplePartialSortDescending, arrayfuncs.h:161-183 */
static void sort_desc(int* idx, long long* val, int n, int size)
{
    for (int i = 0; i < size; ++i) idx[i] = i;
    int lim = size - 1 < n ? size - 1 : n;
    for (int i = 0; i < lim; ++i) {
        int index = i;
        for (int j = i + 1; j < size; ++j)
            if (val[j] > val[index]) index = j;
        long long tv = val[i]; val[i] = val[index]; val[index] = tv;
        int ti = idx[i]; idx[i] = idx[index]; idx[index] = ti;
    }
}
/* findWeakLlrs, arrayfuncs.h:209-231 (mutates val) */
static void find_weak(int* idx, int8_t* val, int size, int n)
{
    for (int i = 0; i < size; ++i) idx[i] = i;
    int lim = size - 1 < n ? size - 1 : n;
    for (int i = 0; i < lim; ++i) {
        int index = i;
        for (int j = i + 1; j < size; ++j)
            if (val[j] < val[index]) index = j;
        int8_t tv = val[i]; val[i] = val[index]; val[index] = tv;
        int ti = idx[i]; idx[i] = idx[index]; idx[index] = ti;
    }
}

/* duplicatePath + switchToNext (scl_fip_char.cpp:55-63,91-98): the new path takes every stage >= `stage` of src */
static void fork_paths(scl* d, const int* idx, int c, int newP, int stage)
{
    for (int p = 0; p < newP; ++p) {
        int src = idx[p] / c;
        for (int s = stage; s <= d->S; ++s) {
            size_t n = (size_t)1 << s;
            memcpy(AT(d->nllr, p, s), AT(d->llr, src, s), n);
            memcpy(AT(d->nbit, p, s), AT(d->bit, src, s), n);
            memcpy(AT(d->nlbit, p, s), AT(d->lbit, src, s), n);
        }
    }
}
static void switch_next(scl* d, int newP)
{
    int8_t* t;
    t = d->llr; d->llr = d->nllr; d->nllr = t;
    t = d->bit; d->bit = d->nbit; d->nbit = t;
    t = d->lbit; d->lbit = d->nlbit; d->nlbit = t;
    long long* m = d->metric; d->metric = d->nmetric; d->nmetric = m;
    d->P = newP;
}

/* RateZeroDecoder::decode :387-421 */
static void leaf_r0(scl* d, int s)
{
    int n = 1 << s;
    for (int p = 0; p < d->P; ++p) {
        int8_t* l = AT(d->llr, p, s);
        long long pun = 0;
        for (int i = 0; i < n; ++i) pun += l[i] < 0 ? l[i] : 0;
        memset(AT(d->bit, p, s), 127, n);
        d->metric[p] += pun;
    }
}

/* RepetitionDecoder::decode :508-581 (the candidate bit bytes are result / ~result: MSB 0 for candidate 0, 1 for 1) */
static void leaf_rep(scl* d, int s)
{
    int n = 1 << s, P = d->P;
    long long* met = (long long*)malloc(sizeof(long long) * 2 * P);
    int* idx = (int*)malloc(sizeof(int) * 2 * P);
    for (int p = 0; p < P; ++p) {
        int8_t* l = AT(d->llr, p, s);
        long long neg = 0, pos = 0;
        for (int i = 0; i < n; ++i) {
            if (l[i] < 0) neg += l[i]; else pos += l[i];
        }
        met[2 * p] = d->metric[p] + neg;
        met[2 * p + 1] = d->metric[p] - pos;
    }
    int newP = 2 * P < d->L ? 2 * P : d->L;
    sort_desc(idx, met, newP, 2 * P);
    fork_paths(d, idx, 2, newP, s);
    for (int p = 0; p < newP; ++p) {
        d->nmetric[p] = met[p];
        memset(AT(d->nbit, p, s), (idx[p] & 1) ? 0x80 : 0x00, n);
    }
    switch_next(d, newP);
    free(met); free(idx);
}

/* RateOneDecoder::decode :423-505 */
static void leaf_r1(scl* d, int s)
{
    int n = 1 << s, P = d->P;
    long long* met = (long long*)malloc(sizeof(long long) * 4 * P);
    int* idx = (int*)malloc(sizeof(int) * 4 * P);
    int* hints = (int*)malloc(sizeof(int) * 4 * P * 2);
    int* cnt = (int*)malloc(sizeof(int) * 4 * P);
    int* widx = (int*)malloc(sizeof(int) * (n > 4 ? n : 4));
    int8_t* tmp = (int8_t*)malloc(n > 4 ? n : 4);
    for (int p = 0; p < P; ++p) {
        int8_t* l = AT(d->llr, p, s);
        for (int i = 0; i < n; ++i) {
            int v = l[i] < -127 ? -127 : l[i];
            tmp[i] = (int8_t)(v < 0 ? -v : v);
        }
        find_weak(widx, tmp, n, 2);
        long long m = d->metric[p];
        met[4 * p] = m;
        met[4 * p + 1] = m - tmp[0];
        if (n == 1) {
            met[4 * p + 2] = -0x100000000000LL;
            met[4 * p + 3] = -0x100000000000LL;
        } else {
            met[4 * p + 2] = m - tmp[1];
            met[4 * p + 3] = m - tmp[0] - tmp[1];
        }
        hints[(4 * p + 1) * 2] = widx[0];
        hints[(4 * p + 2) * 2] = n > 1 ? widx[1] : 0;
        hints[(4 * p + 3) * 2] = widx[0];
        hints[(4 * p + 3) * 2 + 1] = n > 1 ? widx[1] : 0;
        cnt[4 * p] = 0; cnt[4 * p + 1] = 1; cnt[4 * p + 2] = 1; cnt[4 * p + 3] = 2;
    }
    int newP = 4 * P < d->L ? 4 * P : d->L;
    sort_desc(idx, met, newP, 4 * P);
    fork_paths(d, idx, 4, newP, s);
    for (int p = 0; p < newP; ++p) {
        d->nmetric[p] = met[p];
        int8_t* b = AT(d->nbit, p, s);
        memcpy(b, AT(d->nllr, p, s), n);
        for (int i = 0; i < cnt[idx[p]]; ++i) {
            int k = hints[idx[p] * 2 + i];
            b[k] = (int8_t)~b[k];
        }
    }
    switch_next(d, newP);
    free(met); free(idx); free(hints); free(cnt); free(widx); free(tmp);
}

/* SpcDecoder::decode :583-726 */
static void leaf_spc(scl* d, int s)
{
    int n = 1 << s, P = d->P;
    long long* met = (long long*)malloc(sizeof(long long) * 8 * P);
    int* idx = (int*)malloc(sizeof(int) * 8 * P);
    int* hints = (int*)malloc(sizeof(int) * 8 * P * 4);
    int* cnt = (int*)malloc(sizeof(int) * 8 * P);
    int* widx = (int*)malloc(sizeof(int) * (n > 32 ? n : 32));
    int8_t* tmp = (int8_t*)malloc(n > 32 ? n : 32);
    for (int p = 0; p < P; ++p) {
        int8_t* l = AT(d->llr, p, s);
        int parity = 0;
        for (int i = 0; i < n; ++i) {
            parity ^= (l[i] < 0);
            int v = l[i] < -127 ? -127 : l[i];
            tmp[i] = (int8_t)(v < 0 ? -v : v);
        }
        find_weak(widx, tmp, n, 4);
        long long t0 = tmp[0], t1 = tmp[1], t2 = tmp[2], t3 = tmp[3];
        long long m = d->metric[p], weakest = 0;
        int* h = hints + (size_t)8 * p * 4;
        int* c = cnt + 8 * p;
        for (int k = 0; k < 8; ++k) c[k] = 0;
#define HINT(k, v) h[(k) * 4 + c[k]++] = (v)
        if (parity) {
            m -= t0;
            HINT(0, widx[0]); HINT(4, widx[0]); HINT(5, widx[0]); HINT(6, widx[0]);
        } else {
            HINT(1, widx[0]); HINT(2, widx[0]); HINT(3, widx[0]); HINT(7, widx[0]);
            weakest = t0;
        }
        long long* q = met + 8 * p;
        q[0] = m;
        q[1] = m - weakest - t1;
        q[2] = m - weakest - t2;
        q[3] = m - weakest - t3;
        q[4] = m - t1 - t2;
        q[5] = m - t1 - t3;
        q[6] = m - t2 - t3;
        q[7] = m - weakest - t1 - t2 - t3;
        HINT(1, widx[1]); HINT(2, widx[2]); HINT(3, widx[3]);
        HINT(4, widx[1]); HINT(4, widx[2]);
        HINT(5, widx[1]); HINT(5, widx[3]);
        HINT(6, widx[2]); HINT(6, widx[3]);
        HINT(7, widx[1]); HINT(7, widx[2]); HINT(7, widx[3]);
#undef HINT
    }
    int newP = 8 * P < d->L ? 8 * P : d->L;
    sort_desc(idx, met, newP, 8 * P);
    fork_paths(d, idx, 8, newP, s);
    for (int p = 0; p < newP; ++p) {
        d->nmetric[p] = met[p];
        int8_t* b = AT(d->nbit, p, s);
        memcpy(b, AT(d->nllr, p, s), n);
        for (int i = 0; i < cnt[idx[p]]; ++i) {
            int k = hints[(size_t)idx[p] * 4 + i];
            b[k] = (int8_t)~b[k];
        }
    }
    switch_next(d, newP);
    free(met); free(idx); free(hints); free(cnt); free(widx); free(tmp);
}

/* createDecoder :729-755 and RateRNode::decode :315-349; `fz` = sorted frozen indices relative to the node */
static void node(scl* d, int s, const unsigned* fz, int nf)
{
    int n = 1 << s;
    if (nf == n) { leaf_r0(d, s); return; }
    if (nf == 0) { leaf_r1(d, s); return; }
    if (nf == n - 1) { leaf_rep(d, s); return; }
    if (nf == 1) { leaf_spc(d, s); return; }
    int h = n / 2, nl = 0;
    while (nl < nf && fz[nl] < (unsigned)h) nl++;
    unsigned* rf = (unsigned*)malloc(sizeof(unsigned) * (nf - nl + 1));
    for (int i = nl; i < nf; ++i) rf[i - nl] = fz[i] - h;
    for (int p = 0; p < d->P; ++p) {
        int8_t *in = AT(d->llr, p, s), *out = AT(d->llr, p, s - 1);
        for (int i = 0; i < h; ++i) out[i] = f8(in[i], in[i + h]);
    }
    node(d, s - 1, fz, nl);
    for (int p = 0; p < d->P; ++p) { /* prepareRightDecoding: swap Bit <-> LeftBit at the child stage */
        int8_t *b = AT(d->bit, p, s - 1), *lb = AT(d->lbit, p, s - 1);
        for (int i = 0; i < h; ++i) { int8_t t = b[i]; b[i] = lb[i]; lb[i] = t; }
        int8_t *in = AT(d->llr, p, s), *out = AT(d->llr, p, s - 1);
        for (int i = 0; i < h; ++i) out[i] = g8(in[i], in[i + h], lb[i]);
    }
    node(d, s - 1, rf, nf - nl);
    for (int p = 0; p < d->P; ++p) { /* CombineBits, fip_char.h:173-185 */
        int8_t *lb = AT(d->lbit, p, s - 1), *b = AT(d->bit, p, s - 1), *o = AT(d->bit, p, s);
        for (int i = 0; i < h; ++i) { o[i] = lb[i] ^ b[i]; o[i + h] = b[i]; }
    }
    free(rf);
}

/* SclFipChar::decode :801-855. Returns the detector verdict (decoderSuccess). */
int pco_scl_i8_decode(int N, const unsigned* frozen, int nf, int L, int crc_kind, int systematic, const int8_t* llr,
                      uint8_t* info_out)
{
    scl D, *d = &D;
    int S = 0;
    while ((1 << S) < N) S++;
    d->N = N; d->S = S; d->L = L; d->P = 1; d->frozen = frozen; d->nf = nf;
    size_t sz = (size_t)L * 2 * N;
    d->llr = (int8_t*)calloc(sz, 1); d->bit = (int8_t*)calloc(sz, 1); d->lbit = (int8_t*)calloc(sz, 1);
    d->nllr = (int8_t*)calloc(sz, 1); d->nbit = (int8_t*)calloc(sz, 1); d->nlbit = (int8_t*)calloc(sz, 1);
    d->metric = (long long*)calloc(L, sizeof(long long)); d->nmetric = (long long*)calloc(L, sizeof(long long));
    memcpy(AT(d->llr, 0, S), llr, N);
    node(d, S, frozen, nf);
    int K = N - nf, kb = (K + 7) / 8, ok = 0;
    uint8_t* soft = (uint8_t*)malloc(N);
    uint8_t* cw = (uint8_t*)malloc(N / 8 + 1);
    for (int p = 0; p <= d->P && !ok; ++p) {
        int path = p < d->P ? p : 0; /* fall back to the ML path if no candidate passes */
        const int8_t* b = AT(d->bit, path, S);
        memset(info_out, 0, kb);
        if (systematic) {
            int k = 0, fc = 0;
            for (int i = 0; i < N; ++i) {
                if (fc < nf && frozen[fc] == (unsigned)i) { fc++; continue; }
                if (b[i] & 0x80) info_out[k / 8] |= (uint8_t)(0x80 >> (k % 8));
                k++;
            }
        } else {
            for (int i = 0; i < N; ++i) soft[i] = (b[i] & 0x80) ? 1 : 0;
            for (int dd = 1; dd < N; dd <<= 1)
                for (int g = 0; g < N; g += 2 * dd)
                    for (int i = g; i < g + dd; ++i) soft[i] ^= soft[i + dd];
            int k = 0, fc = 0;
            for (int i = 0; i < N; ++i) {
                if (fc < nf && frozen[fc] == (unsigned)i) { fc++; continue; }
                if (soft[i]) info_out[k / 8] |= (uint8_t)(0x80 >> (k % 8));
                k++;
            }
        }
        if (p < d->P) ok = pco_crc_check(crc_kind, info_out, kb);
        else break;
    }
    free(soft); free(cw);
    free(d->llr); free(d->bit); free(d->lbit); free(d->nllr); free(d->nbit); free(d->nlbit);
    free(d->metric); free(d->nmetric);
    return ok;
}
