// TEST HARNESS: exercises the C++ class mirror (include/polarcode_b200.hpp) the way a caller of the reference's
// classes would, behind extern "C" so that the python tests can drive it with ctypes. Status codes: 0 ok,
// 1 std::invalid_argument, 2 std::logic_error, 3 std::runtime_error, 4 anything else.
#include <cstring>
#include <memory>

#include "../include/polarcode_b200.hpp"

using namespace PolarCode;

template <class F>
static int guarded(F&& f)
{
    try {
        f();
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    } catch (const std::logic_error&) {
        return 2;
    } catch (const std::runtime_error&) {
        return 3;
    } catch (...) {
        return 4;
    }
}

extern "C" {

int cxx_frozen_bits(int N, int K, float dsnr, const char* type, unsigned* out, int* n)
{
    return guarded([&] {
        auto v = Construction::frozen_bits(N, K, dsnr, type);
        *n = (int)v.size();
        std::memcpy(out, v.data(), 4 * v.size());
    });
}

// Detector: `bytes` payload bytes followed by the checksum; generate(data, bytes + crc) as in the reference
// (errordetector.h:52-57: the size includes the checksum, which replaces the last bytes), then check; mode string out
int cxx_detector(unsigned size, const char* type, unsigned char* data, int bytes, int* check_ok, int* check_corrupt,
                 unsigned long long* calc, char* mode, int* bits)
{
    return guarded([&] {
        std::unique_ptr<ErrorDetection::Detector> d(ErrorDetection::create(size, type));
        const int cb = (int)((d->getCheckBitCount() + 7) / 8);
        d->generate(data, bytes + cb);
        *check_ok = d->check(data, bytes + cb) ? 1 : 0;
        data[0] ^= 0x10;
        *check_corrupt = d->check(data, bytes + cb) ? 1 : 0;
        data[0] ^= 0x10;
        void* arrs[2] = { data + 1, data };
        if (cb && bytes > 4 && d->multiCheck(arrs, 2, bytes + cb) != 1 && size != 0) *check_ok = 0;
        *calc = d->calculate(data, 8 * (size_t)bytes);
        std::strcpy(mode, d->getType().c_str());
        *bits = (int)d->getCheckBitCount();
    });
}

// ButterflyFipPacked: encode_vector per row (info is modified in place like the reference when a detector is set)
int cxx_encode(int N, const unsigned* frozen, int nf, unsigned crc_size, const char* crc_type, int systematic,
               unsigned char* info, unsigned char* code, long B, int batch, char* mode)
{
    return guarded([&] {
        std::vector<unsigned> fz(frozen, frozen + nf);
        Encoding::ButterflyFipPacked enc((size_t)N, fz);
        std::unique_ptr<ErrorDetection::Detector> det(ErrorDetection::create(crc_size, crc_type));
        enc.setErrorDetection(det.get());
        enc.setSystematic(systematic != 0);
        const size_t kb = (enc.infoLength() + 7) / 8;
        if (batch) {
            std::vector<unsigned char> with_crc(kb * B);
            enc.encode_batch(info, code, (size_t)B, with_crc.data());
            std::memcpy(info, with_crc.data(), kb * B);
        } else {
            for (long i = 0; i < B; ++i) enc.encode_vector(info + i * kb, code + i * (N / 8));
        }
        std::strcpy(mode, enc.getErrorDetectionMode().c_str());
    });
}

// Decoding::create + decode_vector per row (or one decode_batch call)
int cxx_decode(int N, int L, const unsigned* frozen, int nf, const char* kind, int crc_size, const char* crc_type,
               int systematic, int is_float, const void* llr, long B, unsigned char* out, unsigned char* ok, int batch,
               char* mode, int* list_size)
{
    return guarded([&] {
        std::vector<unsigned> fz(frozen, frozen + nf);
        std::unique_ptr<Decoding::Decoder> dec(Decoding::create((size_t)N, (size_t)L, fz, kind));
        std::unique_ptr<ErrorDetection::Detector> det;
        if (crc_size >= 0) {
            det.reset(ErrorDetection::create((unsigned)crc_size, crc_type));
            dec->setErrorDetection(det.get());
        }
        dec->setSystematic(systematic != 0);
        const size_t kb = (dec->infoLength() + 7) / 8;
        if (batch) {
            if (is_float)
                dec->decode_batch(static_cast<const float*>(llr), out, ok, (size_t)B);
            else
                dec->decode_batch(static_cast<const char*>(llr), out, ok, (size_t)B);
        } else {
            for (long i = 0; i < B; ++i) {
                bool r = is_float ? dec->decode_vector(static_cast<const float*>(llr) + i * N, out + i * kb)
                                  : dec->decode_vector(static_cast<const char*>(llr) + i * N, out + i * kb);
                ok[i] = r ? 1 : 0;
            }
        }
        std::strcpy(mode, dec->getErrorDetectionMode().c_str());
        *list_size = (int)dec->getListSize();
    });
}
// Puncturer (puncturer.h:37-101): positions, then puncture<float>, puncturePacked, depuncture<float> of one vector
// each; status 5 = std::out_of_range (checked before logic_error, its base class)
// Decoder::getSoftCodeword / getSoftInformation after decode_vector, per row (the container type follows the decoder:
// int8 for 'char', float for 'float'); elem = bytes per container element
int cxx_soft(int N, int L, const unsigned* frozen, int nf, const char* kind, int is_float, const void* llr, long B, void* soft_code,
             void* soft_info, int elem)
{
    return guarded([&] {
        std::vector<unsigned> fz(frozen, frozen + nf);
        std::unique_ptr<Decoding::Decoder> dec(Decoding::create((size_t)N, (size_t)L, fz, kind));
        const size_t K = dec->infoLength();
        std::vector<unsigned char> out((K + 7) / 8 + 8);
        for (long i = 0; i < B; ++i) {
            if (is_float)
                dec->decode_vector(static_cast<const float*>(llr) + i * N, out.data());
            else
                dec->decode_vector(static_cast<const char*>(llr) + i * N, out.data());
            dec->getSoftCodeword(static_cast<char*>(soft_code) + (size_t)i * N * elem);
            dec->getSoftInformation(static_cast<char*>(soft_info) + (size_t)i * K * elem);
        }
    });
}

// Encoder::setCodeword / setCharCodeword / setFloatCodeword (+ clearFrozenBits) followed by encode() and getEncodedData:
// form 0 packed bytes, 1 chars (MSB = bit), 2 floats (sign = bit); `clear` calls clearFrozenBits() before encode()
int cxx_encode_codeword(int N, const unsigned* frozen, int nf, int systematic, int form, int clear, void* in, unsigned char* out)
{
    return guarded([&] {
        std::vector<unsigned> fz(frozen, frozen + nf);
        Encoding::ButterflyFipPacked enc((size_t)N, fz);
        enc.setSystematic(systematic != 0);
        if (form == 0)
            enc.setCodeword(in);
        else if (form == 1)
            enc.setCharCodeword(in);
        else
            enc.setFloatCodeword(in);
        if (clear) enc.clearFrozenBits();
        enc.encode();
        enc.getEncodedData(out);
    });
}

int cxx_puncturer(long blockLength, const unsigned* frozen, int nf, unsigned* positions, long* parent, const float* in,
                  float* punctured, const unsigned char* packed_in, unsigned char* packed_out, float* depunctured)
{
    try {
        Puncturer p(blockLength, std::vector<unsigned>(frozen, frozen + nf));
        *parent = (long)p.parentBlockLength();
        auto pos = p.blockOutputPositions();
        std::memcpy(positions, pos.data(), 4 * pos.size());
        p.puncture<float>(punctured, in);
        p.puncturePacked(packed_out, packed_in);
        p.depuncture<float>(depunctured, punctured);
        return (long)p.blockLength() == blockLength ? 0 : 4;
    } catch (const std::out_of_range&) {
        return 5;
    } catch (const std::runtime_error&) {
        return 3;
    } catch (...) {
        return 4;
    }
}
}
