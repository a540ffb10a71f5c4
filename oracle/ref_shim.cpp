// TEST ORACLE / CPU BASELINE ONLY -- never linked or imported by the product path.
//
// Thin extern "C" batch wrapper over the UNMODIFIED reference classes (compiled against the headers where
// they lie under /root/reference/include; output goes to oracle/_ref/libref_shim.so). It lets tests/ and
// bench.py's cpu_baseline leg drive the reference's own per-codeword API over a batch:
//   PolarCode::Decoding::create            (reference src/polarcode/decoding/decoder.cpp:26-52)
//   Decoder::decode_vector(char*/float*)   (decoder.cpp:154-181)
//   Encoding::ButterflyFipPacked           (src/polarcode/encoding/butterfly_fip_packed.cpp:45-59)
//   ErrorDetection::create                 (src/polarcode/errordetection/errordetector.cpp:27-88)
//   Construction::frozen_bits              (src/polarcode/construction/constructor.cpp:41-63)
// Threading mirrors pcsim's worker model (src/simulation/simulator.cpp:87-116): one private codec object
// per std::thread, frames split into contiguous slices.
#include <polarcode/construction/constructor.h>
#include <polarcode/decoding/adaptive_char.h>
#include <polarcode/decoding/adaptive_mixed.h>
#include <polarcode/decoding/decoder.h>
#include <polarcode/decoding/depth_first.h>
#include <polarcode/encoding/butterfly_fip_packed.h>
#include <polarcode/errordetection/errordetector.h>
#include <polarcode/puncturer.h>

#include <chrono>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

using namespace PolarCode;

// "adaptive_char" / "adaptive_mixed": the two wrappers pcsim constructs by class (simulator.cpp:721-733); everything
// else goes through the reference factory. (They carry "char" / "mixed" in their names, so test them first.)
static Decoding::Decoder* make_decoder(int N, int L, const std::vector<unsigned>& fz, const std::string& kind)
{
    if (kind == "adaptive_char") return new Decoding::AdaptiveChar(N, L, fz);
    if (kind == "adaptive_mixed") return new Decoding::AdaptiveMixed(N, L, fz);
    if (kind == "depthfirst") return new Decoding::DepthFirst(N, L, fz); // pcsim's SC-Flip decoder (simulator.cpp:715-716), L = trial limit
    return Decoding::create(N, L, fz, kind);
}

static std::vector<unsigned> mkfrozen(const unsigned* f, int nf) { return std::vector<unsigned>(f, f + nf); }

extern "C" {

// returns number of frozen positions written (N-K) or -1
int ref_frozen_bits(int N, int K, float dsnr, const char* type, unsigned* out)
{
    try {
        auto v = Construction::frozen_bits(N, K, dsnr, std::string(type));
        for (size_t i = 0; i < v.size(); ++i) out[i] = v[i];
        return (int)v.size();
    } catch (...) {
        return -1;
    }
}

// kind: "char" / "float" / "mixed" / "adaptive_char" / "adaptive_mixed" (the last two take float LLRs only); crc_bits<0 => keep the factory default (CRC-8, decoder.cpp:85);
// is_float: llr points to float[B][N] else int8[B][N]. Returns seconds spent inside the decode loops
// (max over threads), or a negative value on error.
// fresh != 0: a NEW decoder object for every frame. SclAvxFloat never resets the metric of path 0 between decode()
// calls (PathList::setFirstPath, scl_avx_float.cpp:103-109, assigns no metric; only the constructor does, :31,35), so
// with one long-lived object every frame starts from the previous frame's best metric: harmless in exact arithmetic,
// but float rounding then depends on the decoder's history. A fresh object per frame is the history-free behaviour.
double ref_decode_batch2(int N, int L, const unsigned* frozen, int nf, const char* kind, int crc_bits,
                         const char* crc_type, int systematic, int is_float, const void* llr, long B,
                         unsigned char* out, unsigned char* ok, int threads, int repeat, int fresh);
double ref_decode_batch(int N, int L, const unsigned* frozen, int nf, const char* kind, int crc_bits,
                        const char* crc_type, int systematic, int is_float, const void* llr, long B,
                        unsigned char* out, unsigned char* ok, int threads, int repeat)
{
    return ref_decode_batch2(N, L, frozen, nf, kind, crc_bits, crc_type, systematic, is_float, llr, B, out, ok, threads,
                             repeat, 0);
}
double ref_decode_batch2(int N, int L, const unsigned* frozen, int nf, const char* kind, int crc_bits,
                         const char* crc_type, int systematic, int is_float, const void* llr, long B,
                         unsigned char* out, unsigned char* ok, int threads, int repeat, int fresh)
{
    try {
        const int K = N - nf;
        const int kb = (K + 7) / 8;
        if (threads < 1) threads = 1;
        if (repeat < 1) repeat = 1;
        std::vector<double> secs(threads, 0.0);
        std::vector<std::thread> pool;
        auto fz = mkfrozen(frozen, nf);
        for (int t = 0; t < threads; ++t) {
            pool.emplace_back([&, t]() {
                Decoding::Decoder* dec = make_decoder(N, L, fz, std::string(kind));
                ErrorDetection::Detector* det = nullptr;
                if (crc_bits >= 0) {
                    det = ErrorDetection::create(crc_bits, std::string(crc_type));
                    dec->setErrorDetection(det);
                }
                dec->setSystematic(systematic != 0);
                long lo = B * t / threads, hi = B * (t + 1) / threads;
                std::vector<unsigned char> tmp(kb + 64);
                auto t0 = std::chrono::high_resolution_clock::now();
                for (int r = 0; r < repeat; ++r) {
                    for (long i = lo; i < hi; ++i) {
                        bool res;
                        if (fresh && i > lo) {
                            delete dec;
                            dec = make_decoder(N, L, fz, std::string(kind));
                            if (det) dec->setErrorDetection(det);
                            dec->setSystematic(systematic != 0);
                        }
                        if (is_float)
                            res = dec->decode_vector(reinterpret_cast<const float*>(llr) + i * N, tmp.data());
                        else
                            res = dec->decode_vector(reinterpret_cast<const char*>(llr) + i * N, tmp.data());
                        memcpy(out + i * kb, tmp.data(), kb);
                        if (ok) ok[i] = res ? 1 : 0;
                    }
                }
                auto t1 = std::chrono::high_resolution_clock::now();
                secs[t] = std::chrono::duration<double>(t1 - t0).count();
                delete dec;
                delete det;
            });
        }
        for (auto& th : pool) th.join();
        double mx = 0;
        for (double s : secs) mx = s > mx ? s : mx;
        return mx;
    } catch (...) {
        return -1.0;
    }
}

// Decoder::getSoftCodeword / getSoftInformation (decoder.cpp:147-152) of the 'char' Fast-SSC decoder after decode():
// soft_code int8[B][N], soft_info int8[B][K] (either may be null). Returns 0 or -1.
int ref_decode_soft(int N, const unsigned* frozen, int nf, int systematic, const signed char* llr, long B,
                    signed char* soft_code, signed char* soft_info)
{
    try {
        auto fz = mkfrozen(frozen, nf);
        const int K = N - nf;
        Decoding::Decoder* dec = Decoding::create(N, 1, fz, std::string("char"));
        dec->setSystematic(systematic != 0);
        std::vector<unsigned char> tmp((K + 7) / 8 + 64);
        for (long i = 0; i < B; ++i) {
            dec->decode_vector(reinterpret_cast<const char*>(llr) + i * N, tmp.data());
            if (soft_code) dec->getSoftCodeword(soft_code + i * N);
            if (soft_info) dec->getSoftInformation(soft_info + i * K);
        }
        delete dec;
        return 0;
    } catch (...) {
        return -1;
    }
}

// The same getters of the 'float' Fast-SSC decoder (FloatContainer::getSoftBits / getSoftInformation,
// bitcontainer.cpp:294-297, 331-339): soft_code float[B][N], soft_info float[B][K] (either may be null). Returns 0 or -1.
int ref_decode_soft_f32(int N, const unsigned* frozen, int nf, int systematic, const float* llr, long B, float* soft_code,
                        float* soft_info)
{
    try {
        auto fz = mkfrozen(frozen, nf);
        const int K = N - nf;
        Decoding::Decoder* dec = Decoding::create(N, 1, fz, std::string("float"));
        dec->setSystematic(systematic != 0);
        std::vector<unsigned char> tmp((K + 7) / 8 + 64);
        std::vector<float> sc(N < 8 ? 8 : N);
        for (long i = 0; i < B; ++i) {
            dec->decode_vector(llr + i * N, tmp.data());
            if (soft_code) {
                dec->getSoftCodeword(sc.data());
                memcpy(soft_code + i * N, sc.data(), sizeof(float) * N);
            }
            if (soft_info) dec->getSoftInformation(soft_info + i * K);
        }
        delete dec;
        return 0;
    } catch (...) {
        return -1;
    }
}

// Encoder::setCodeword / setCharCodeword / setFloatCodeword (+ clearFrozenBits), encode(), getEncodedData (encoder.cpp:56-77).
// form 0 packed bytes, 1 chars, 2 floats. Returns 0 or -1.
int ref_encode_codeword(int N, const unsigned* frozen, int nf, int systematic, int form, int clear, void* in, unsigned char* out)
{
    try {
        auto fz = mkfrozen(frozen, nf);
        Encoding::ButterflyFipPacked enc(N, fz);
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
        return 0;
    } catch (...) {
        return -1;
    }
}

// info: uint8[B][ceil(K/8)] (a private copy is made per frame because encode() writes the CRC in place,
// butterfly_fip_packed.cpp:47-49; the copy incl. CRC is returned through info_out when non-null).
double ref_encode_batch(int N, const unsigned* frozen, int nf, int crc_bits, const char* crc_type,
                        int systematic, const unsigned char* info, long B, unsigned char* code,
                        unsigned char* info_out, int threads, int repeat)
{
    try {
        const int K = N - nf;
        const int kb = (K + 7) / 8;
        if (threads < 1) threads = 1;
        if (repeat < 1) repeat = 1;
        auto fz = mkfrozen(frozen, nf);
        std::vector<double> secs(threads, 0.0);
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) {
            pool.emplace_back([&, t]() {
                Encoding::ButterflyFipPacked enc(N, fz);
                ErrorDetection::Detector* det = nullptr;
                if (crc_bits >= 0) {
                    det = ErrorDetection::create(crc_bits, std::string(crc_type));
                    enc.setErrorDetection(det);
                }
                enc.setSystematic(systematic != 0);
                long lo = B * t / threads, hi = B * (t + 1) / threads;
                std::vector<unsigned char> tmp(kb + 64);
                auto t0 = std::chrono::high_resolution_clock::now();
                for (int r = 0; r < repeat; ++r) {
                    for (long i = lo; i < hi; ++i) {
                        memcpy(tmp.data(), info + i * kb, kb);
                        enc.encode_vector(tmp.data(), code + i * (N / 8));
                        if (info_out) memcpy(info_out + i * kb, tmp.data(), kb);
                    }
                }
                auto t1 = std::chrono::high_resolution_clock::now();
                secs[t] = std::chrono::duration<double>(t1 - t0).count();
                delete det;
            });
        }
        for (auto& th : pool) th.join();
        double mx = 0;
        for (double s : secs) mx = s > mx ? s : mx;
        return mx;
    } catch (...) {
        return -1.0;
    }
}

// op: 0 generate (writes CRC into the tail of each row), 1 check (ok[i]), 2 calculate over `bits` bits (val[i])
int ref_crc_batch(int crc_bits, const char* crc_type, int op, unsigned char* data, int bytes, long bits, long B,
                  unsigned char* ok, unsigned long long* val)
{
    try {
        ErrorDetection::Detector* det = ErrorDetection::create(crc_bits, std::string(crc_type));
        for (long i = 0; i < B; ++i) {
            unsigned char* row = data + i * bytes;
            if (op == 0)
                det->generate(row, bytes);
            else if (op == 1)
                ok[i] = det->check(row, bytes) ? 1 : 0;
            else
                val[i] = det->calculate(row, bits);
        }
        delete det;
        return 0;
    } catch (...) {
        return -1;
    }
}

// float -> int8 quantiser of the char decoders' input container (src/polarcode/bitcontainer.cpp:468-516)
int ref_quantize(int N, const float* in, long B, signed char* out)
{
    try {
        CharContainer c(N);
        for (long i = 0; i < B; ++i) {
            c.insertLlr(in + i * N);
            memcpy(out + i * N, c.data(), N);
        }
        return 0;
    } catch (...) {
        return -1;
    }
}

// PolarCode::Puncturer (src/polarcode/puncturer.cpp, include/polarcode/puncturer.h) over a batch.
// mode 0: blockOutputPositions -> out (unsigned[M]); 1: puncturePacked; 2/3/4: puncture<uint8_t/float/double>;
// 5/6/7: depuncture<uint8_t/float/double>. Returns parentBlockLength, -2 for std::out_of_range, -1 otherwise.
long ref_puncturer(long blockLength, const unsigned* frozen, int nf, int mode, const void* in, long B, void* out)
{
    try {
        Puncturer p(blockLength, mkfrozen(frozen, nf));
        const long M = p.blockLength(), Np = p.parentBlockLength();
        if (mode == 0) {
            auto pos = p.blockOutputPositions();
            memcpy(out, pos.data(), sizeof(unsigned) * pos.size());
            return Np;
        }
        for (long b = 0; b < B; ++b) {
            switch (mode) {
            case 1: p.puncturePacked((unsigned char*)out + b * (M / 8), (const unsigned char*)in + b * (Np / 8)); break;
            case 2: p.puncture<unsigned char>((unsigned char*)out + b * M, (const unsigned char*)in + b * Np); break;
            case 3: p.puncture<float>((float*)out + b * M, (const float*)in + b * Np); break;
            case 4: p.puncture<double>((double*)out + b * M, (const double*)in + b * Np); break;
            case 5: p.depuncture<unsigned char>((unsigned char*)out + b * Np, (const unsigned char*)in + b * M); break;
            case 6: p.depuncture<float>((float*)out + b * Np, (const float*)in + b * M); break;
            case 7: p.depuncture<double>((double*)out + b * Np, (const double*)in + b * M); break;
            default: return -1;
            }
        }
        return Np;
    } catch (const std::out_of_range&) {
        return -2;
    } catch (...) {
        return -1;
    }
}

} // extern "C"
