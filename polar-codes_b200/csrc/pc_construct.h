// Host-only code construction (see pc_construct.cpp).
#ifndef PC_CONSTRUCT_H
#define PC_CONSTRUCT_H
#include <vector>

namespace pc {
// returns 0 on success, 1 invalid argument, 3 unsupported; frozen = N-K sorted indices
int frozen_bits(int N, int K, float design_snr_db, const char* type, std::vector<unsigned>& frozen, const char** msg);
} // namespace pc
#endif
