// C-ABI entry points other than decode: encoder, detectors, quantiser, channel simulation.
#include "pc_internal.h"

using namespace pc;

extern "C" {

int pc_encode_batch(pc_code*, const uint8_t*, uint8_t*, uint8_t*, size_t, void*)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_encode_batch: not implemented in this build");
}
int pc_encode_batch_host(pc_code*, const uint8_t*, uint8_t*, uint8_t*, size_t)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_encode_batch_host: not implemented in this build");
}
int pc_crc_generate_batch(int, uint8_t*, int, size_t, void*)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_crc_generate_batch: not implemented in this build");
}
int pc_crc_check_batch(int, const uint8_t*, int, uint8_t*, size_t, void*)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_crc_check_batch: not implemented in this build");
}
uint64_t pc_crc_calculate(int, const uint8_t*, size_t) { return 0; }
int pc_quantize_batch(int, const float*, int8_t*, size_t, void*)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_quantize_batch: not implemented in this build");
}
int pc_awgn_generate(pc_code*, uint64_t, uint64_t, float, float, int, void*, uint8_t*, size_t, void*)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_awgn_generate: not implemented in this build");
}
int pc_count_errors(const uint8_t*, const uint8_t*, const uint8_t*, int, size_t, unsigned long long*, void*)
{
    return fail(PC_ERR_UNSUPPORTED, "pc_count_errors: not implemented in this build");
}

} // extern "C"
