// Host-only code construction: runs once per code, results must be VALUE-identical to the reference
// because the frozen set decides the whole decoder tree.
//   Bhattacharyya  reference src/polarcode/construction/bhattacharrya.cpp:39-78 (float design parameter,
//                  double recursion, stable descending insertion sort src/polarcode/arrayfuncs.cpp:108-124)
//   beta expansion reference src/polarcode/construction/betaexpansion.cpp:42-82 (beta = 2^(1/4), ascending argsort)
// Compile with -ffp-contract=off: the recursion's T*T and 2T - T*T must round separately.
#include "pc_construct.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <numeric>
#include <string>

namespace pc {

static bool pow2(int N) { return N >= 2 && (N & (N - 1)) == 0; }

static void bhattacharyya(int N, int K, float dsnr, std::vector<unsigned>& frozen)
{
    const float lin = (float)std::pow(10.0, dsnr / 10.0);
    const float init = (float)std::exp(-2.0 * lin * (size_t)K / (size_t)N);
    std::vector<double> z(N, 0.0);
    std::vector<unsigned> perm(N);
    int n = 0;
    while ((1 << n) < N) n++;
    z[0] = init;
    for (int stage = n - 1; stage >= 0; --stage) {
        const int step = 1 << stage;
        for (int j = 0; j < N; j += 2 * step) {
            const double t = z[j];
            const double t2 = t * t;
            z[j + step] = t2;
            z[j] = 2 * t - t2;
        }
    }
    std::iota(perm.begin(), perm.end(), 0u);
    // worst channels first; equal parameters keep index order (stable)
    std::stable_sort(perm.begin(), perm.end(), [&](unsigned a, unsigned b) { return z[a] > z[b]; });
    frozen.assign(perm.begin(), perm.begin() + (N - K));
    std::sort(frozen.begin(), frozen.end());
}

static void beta_expansion(int N, int K, std::vector<unsigned>& frozen)
{
    int n = 0;
    while ((1 << n) < N) n++;
    const double beta = std::pow(2.0, 1.0 / 4.0);
    std::vector<double> w(n), weight(N);
    for (int j = 0; j < n; ++j) w[j] = std::pow(beta, (double)j);
    for (int i = 0; i < N; ++i) {
        double s = 0.0;
        for (int j = 0; j < n; ++j) s += w[j] * ((i >> j) & 1);
        weight[i] = s;
    }
    std::vector<unsigned> perm(N);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&](unsigned a, unsigned b) { return weight[a] < weight[b]; });
    frozen.assign(perm.begin(), perm.begin() + (N - K));
    std::sort(frozen.begin(), frozen.end());
}

#include "pc_5g_table.inc"

// FiveGList (reference src/polarcode/construction/fiveGList.cpp:30-43): the N - K least reliable positions of the 3GPP
// polar sequence. The reference takes the first N - K entries of the Nmax = 1024 table UNFILTERED, which is the
// standard's rule only for N = 1024; for N < 1024 it returns indices >= N (82 of the 384 at N = 512, K = 128). Here the
// sequence is first restricted to the indices below N, as TS 38.212 section 5.3.1.2 prescribes (Q_0^{N-1} is the
// subsequence of Q_0^{Nmax-1} with values less than N), so N = 1024 is value-identical to the reference and shorter
// codes are the NR construction the reference's python/polar_5g_parameters.py computes.
static void five_g(int N, int K, std::vector<unsigned>& frozen)
{
    frozen.clear();
    for (int i = 0; i < 1024 && (int)frozen.size() < N - K; ++i)
        if ((int)k5gSequence[i] < N) frozen.push_back(k5gSequence[i]);
    std::sort(frozen.begin(), frozen.end());
}

int frozen_bits(int N, int K, float design_snr_db, const char* type, std::vector<unsigned>& frozen, const char** msg)
{
    static const char* e_n = "new blockLength is not a power of 2!";
    static const char* e_k = "Invalid polar code(N, K)";
    static const char* e_5g = "5G standard does not allow for block size N > 1024!";
    std::string t(type ? type : "BB");
    std::transform(t.begin(), t.end(), t.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    if (!pow2(N)) {
        if (msg) *msg = e_n;
        return 1;
    }
    if (K < 0 || K > N) {
        if (msg) *msg = e_k;
        return 1;
    }
    if (t.find("be") != std::string::npos) {
        beta_expansion(N, K, frozen);
    } else if (t.find("5g") != std::string::npos) {
        if (N > 1024) { // fiveGList.cpp:22-25 std::invalid_argument
            if (msg) *msg = e_5g;
            return 1;
        }
        five_g(N, K, frozen);
    } else {
        bhattacharyya(N, K, design_snr_db, frozen);
    }
    return 0;
}

} // namespace pc
