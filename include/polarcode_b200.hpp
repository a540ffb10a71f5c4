// polarcode_b200.hpp -- the reference's C++ class interfaces for the codec path, implemented over the C-ABI of
// polar_b200.h (header-only; link libpolar_b200.so). A caller of the reference switches by including this file
// instead of <polarcode/...> and keeps its code: same namespaces, class names, method names, argument meaning,
// defaults and error behaviour. What each class replaces (paths relative to the reference tree):
//
//   PolarCode::ErrorDetection::Detector / create      include/polarcode/errordetection/errordetector.h:24-89
//   PolarCode::Encoding::Encoder                      include/polarcode/encoding/encoder.h:24-157
//   PolarCode::Encoding::ButterflyFipPacked           include/polarcode/encoding/butterfly_fip_packed.h:25-49
//   PolarCode::Decoding::Decoder / create / makeDecoder   include/polarcode/decoding/decoder.h:40-211
//   PolarCode::Construction::frozen_bits              include/polarcode/construction/constructor.h:58-62
//
// Added on top (the reason this library exists): decode_batch / encode_batch, which take B codewords per call, with
// host or device buffers. The single-codeword methods are batches of one through the pinned-staging host entry points
// (correctness path, not the performance path). There is no CPU fallback: constructing an encoder or decoder without
// a usable GPU throws std::runtime_error.
//
// Soft outputs: getSoftCodeword / getSoftInformation are served for the 'char' Fast-SSC decoder by a second kernel that
// keeps the bit container as bytes (pc_fastssc_i8_soft.cu), the fp32 Fast-SSC one as float patterns (k_fastssc_f32_soft);
// the list decoders' containers are not kept.
// Encoder::setCodeword / setCharCodeword / setFloatCodeword / clearFrozenBits (the hooks the reference's own decoders use
// for non-systematic output) are mirrored on top of a frozen-set-free code whose encoder is the bare polar transform.
// Not carried over: the BitContainer accessors (inputContainer / outputContainer), the 'scan' kind and CMAC detectors.
#ifndef POLARCODE_B200_HPP
#define POLARCODE_B200_HPP
#include <algorithm>
#include <cctype>
#include <chrono>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "polar_b200.h"

namespace PolarCode {

namespace detail {
inline void check(int status)
{
    if (status == PC_OK) return;
    const std::string msg = pc_last_error();
    if (status == PC_ERR_ARG) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
}
} // namespace detail

namespace Construction {
// constructor.h:58-62; std::invalid_argument for a block length that is not a power of two (constructor.cpp:25-31)
inline std::vector<unsigned> frozen_bits(const int blockLength, const int infoLength, const float designSNR,
                                         const std::string& type = "BB")
{
    std::vector<unsigned> out(blockLength > 0 ? (size_t)blockLength : 1);
    const int n = pc_frozen_bits(blockLength, infoLength, designSNR, type.c_str(), out.data());
    if (n < 0) throw std::invalid_argument(pc_last_error());
    out.resize((size_t)n);
    return out;
}
} // namespace Construction

// puncturer.h:37-101. Per-vector calls copy through the device (host pointers, like the reference); the *_batch
// members take device pointers and run one launch.
class Puncturer
{
protected:
    pc_puncturer* mHandle = nullptr;
    size_t mBlockLength = 0, mParentBlockLength = 0;
    std::vector<unsigned> mOutputPositions;

    template <typename T>
    void run(T* pOutput, const T* pInput, size_t nIn, size_t nOut, bool scatter)
    {
        static_assert(sizeof(T) == 1 || sizeof(T) == 4 || sizeof(T) == 8, "element size must be 1, 4 or 8 bytes");
        void* dIn = pc_device_alloc(nIn * sizeof(T));
        void* dOut = pc_device_alloc(nOut * sizeof(T));
        if (!dIn || !dOut) {
            pc_device_free(dIn);
            pc_device_free(dOut);
            throw std::runtime_error(pc_last_error());
        }
        int st = pc_memcpy_h2d(dIn, pInput, nIn * sizeof(T));
        if (st == PC_OK)
            st = scatter ? pc_depuncture_batch(mHandle, dIn, dOut, (int)sizeof(T), 1, nullptr)
                         : pc_puncture_batch(mHandle, dIn, dOut, (int)sizeof(T), 1, nullptr);
        if (st == PC_OK) st = pc_memcpy_d2h(pOutput, dOut, nOut * sizeof(T));
        pc_device_free(dIn);
        pc_device_free(dOut);
        detail::check(st);
    }

public:
    Puncturer(const size_t blockLength, const std::vector<unsigned> frozenBitPositions)
    {
        mHandle = pc_puncturer_create(blockLength, frozenBitPositions.data(), (int)frozenBitPositions.size());
        if (!mHandle) {
            const std::string msg = pc_last_error();
            if (msg.find("exceeds frozen") != std::string::npos) throw std::out_of_range(msg); // puncturer.cpp:57-60
            throw std::runtime_error(msg);
        }
        mBlockLength = pc_puncturer_block_length(mHandle);
        mParentBlockLength = pc_puncturer_parent_block_length(mHandle);
        mOutputPositions.resize(mBlockLength);
        pc_puncturer_positions(mHandle, mOutputPositions.data());
    }
    Puncturer(const Puncturer&) = delete;
    Puncturer& operator=(const Puncturer&) = delete;
    virtual ~Puncturer() { pc_puncturer_destroy(mHandle); }

    size_t blockLength() { return mBlockLength; }
    size_t parentBlockLength() { return mParentBlockLength; }
    std::vector<unsigned> blockOutputPositions() { return mOutputPositions; }

    template <typename T>
    void puncture(T* pOutput, const T* pInput)
    {
        run(pOutput, pInput, mParentBlockLength, mBlockLength, false);
    }
    template <typename T>
    void depuncture(T* pOutput, const T* pInput)
    {
        run(pOutput, pInput, mBlockLength, mParentBlockLength, true);
    }
    void puncturePacked(unsigned char* pOutput, const unsigned char* pInput)
    {
        void* dIn = pc_device_alloc(mParentBlockLength / 8);
        void* dOut = pc_device_alloc(mBlockLength / 8);
        int st = (dIn && dOut) ? pc_memcpy_h2d(dIn, pInput, mParentBlockLength / 8) : PC_ERR_CUDA;
        if (st == PC_OK) st = pc_puncture_packed_batch(mHandle, (const uint8_t*)dIn, (uint8_t*)dOut, 1, nullptr);
        if (st == PC_OK) st = pc_memcpy_d2h(pOutput, dOut, mBlockLength / 8);
        pc_device_free(dIn);
        pc_device_free(dOut);
        detail::check(st);
    }
    // batch extensions: device pointers, asynchronous on `stream`
    void puncture_batch(void* dOut, const void* dIn, int elemBytes, size_t B, void* stream = nullptr)
    {
        detail::check(pc_puncture_batch(mHandle, dIn, dOut, elemBytes, B, stream));
    }
    void depuncture_batch(void* dOut, const void* dIn, int elemBytes, size_t B, void* stream = nullptr)
    {
        detail::check(pc_depuncture_batch(mHandle, dIn, dOut, elemBytes, B, stream));
    }
    void puncturePacked_batch(uint8_t* dOut, const uint8_t* dIn, size_t B, void* stream = nullptr)
    {
        detail::check(pc_puncture_packed_batch(mHandle, dIn, dOut, B, stream));
    }
};

namespace ErrorDetection {

// errordetector.h:24-78. One concrete class covers the CRC family and the dummy. The single-block methods are
// batches of one through the same kernels as the batch methods.
class Detector
{
    int mKind;
    unsigned mBits;

    size_t crcBytes() const { return (size_t)pc_crc_bytes(mKind); }
    // one row through the batch kernels (device scratch, synchronous): the per-call API of the reference is kept, but
    // like every compute entry point of this library it runs on the GPU and throws when there is none
    struct DeviceRow {
        uint8_t* d = nullptr;
        explicit DeviceRow(size_t bytes) : d(static_cast<uint8_t*>(pc_device_alloc(bytes + 8)))
        {
            if (!d) throw std::runtime_error(pc_last_error());
        }
        ~DeviceRow() { pc_device_free(d); }
    };

public:
    explicit Detector(int kind = PC_CRC_NONE, unsigned bits = 0) : mKind(kind), mBits(bits) {}
    virtual ~Detector() {}
    int kind() const { return mKind; } ///< PC_CRC_* id for the C-ABI

    virtual unsigned getCheckBitCount() { return mBits; }
    virtual std::string getType()
    {
        switch (mKind) {
        case PC_CRC_NONE: return "DUMMY";
        case PC_CRC_8:
        case PC_CRC_16:
        case PC_CRC_32C: return "CRC";
        default: return "CRCNR";
        }
    }
    virtual uint64_t calculate(void* data, size_t bits)
    {
        return pc_crc_calculate(mKind, static_cast<const uint8_t*>(data), bits);
    }
    /// errordetector.h:52-57: `bytes` is the block size INCLUDING the checksum; the last crcBytes() bytes of the block
    /// are replaced (crc8.cpp:59-64). Nothing outside [data, data + bytes) is touched.
    virtual void generate(void* data, int bytes)
    {
        if (mKind == PC_CRC_NONE) return;
        if (bytes < (int)crcBytes()) throw std::invalid_argument("Detector::generate: block shorter than the checksum");
        DeviceRow row((size_t)bytes);
        detail::check(pc_memcpy_h2d(row.d, data, (size_t)bytes));
        detail::check(pc_crc_generate_batch(mKind, row.d, bytes, 1, nullptr));
        detail::check(pc_memcpy_d2h(data, row.d, (size_t)bytes));
    }
    /// CRC over the first bytes-c bytes against the trailing c bytes
    virtual bool check(void* data, int bytes)
    {
        if (mKind == PC_CRC_NONE) return true;
        if (bytes < (int)crcBytes()) return false;
        DeviceRow row((size_t)bytes + 8);
        uint8_t ok = 0;
        detail::check(pc_memcpy_h2d(row.d, data, (size_t)bytes));
        detail::check(pc_crc_check_batch(mKind, row.d, bytes, row.d + bytes, 1, nullptr));
        detail::check(pc_memcpy_d2h(&ok, row.d + bytes, 1));
        return ok != 0;
    }
    /// first array that passes, or -1
    virtual int multiCheck(void** data, int nArrays, int nBytes)
    {
        for (int i = 0; i < nArrays; ++i)
            if (check(data[i], nBytes)) return i;
        return -1;
    }
    /// B rows of `bytes` bytes each (checksum included, like generate()), DEVICE memory
    void generate_batch_device(uint8_t* d_rows, int bytes, size_t B, void* stream = nullptr)
    {
        detail::check(pc_crc_generate_batch(mKind, d_rows, bytes, B, stream));
    }
    void check_batch_device(const uint8_t* d_rows, int bytes, uint8_t* d_ok, size_t B, void* stream = nullptr)
    {
        detail::check(pc_crc_check_batch(mKind, d_rows, bytes, d_ok, B, stream));
    }
};

// errordetector.cpp:27-88: std::logic_error("CRC INVALID SIZE!") for sizes the factory rejects
inline Detector* create(unsigned size, std::string type)
{
    const int k = pc_crc_kind(size, type.c_str());
    if (k < 0) {
        std::string t = type;
        std::transform(t.begin(), t.end(), t.begin(), [](unsigned char c) { return (char)std::tolower(c); });
        if (t.find("cmac") != std::string::npos)
            throw std::runtime_error("CMAC detectors are host-side crypto and out of scope of the GPU path");
        throw std::logic_error(pc_last_error());
    }
    return new Detector(k, size);
}

inline Detector& globalDummy()
{
    static Detector d(PC_CRC_NONE, 0);
    return d;
}

} // namespace ErrorDetection

namespace detail {
// shared handle logic: the C-ABI code object carries N, the frozen set, the decoder kind, the detector and the
// systematic flag; the classes below keep the reference's member semantics on top of it
class CodeHandle
{
protected:
    pc_code* mCode = nullptr;
    int mKind = PC_DEC_FASTSSC_I8;
    size_t mListSize = 1;
    void create(size_t N, const std::vector<unsigned>& frozen, int crcKind, bool systematic)
    {
        release();
        mCode = pc_code_create((int)N, frozen.data(), (int)frozen.size(), mKind, (int)mListSize);
        if (!mCode) {
            const std::string msg = pc_last_error();
            if (msg.find("CUDA") != std::string::npos) throw std::runtime_error(msg);
            throw std::invalid_argument(msg);
        }
        check(pc_code_set_crc(mCode, crcKind));
        check(pc_code_set_systematic(mCode, systematic ? 1 : 0));
    }
    void release()
    {
        if (mCode) pc_code_destroy(mCode);
        mCode = nullptr;
    }
    ~CodeHandle() { release(); }

public:
    pc_code* handle() const { return mCode; } ///< for direct C-ABI calls (streams, device buffers)
};

// a code of block length N without frozen bits, non-systematic, no detector: its encoder is the bare polar transform
class XformHandle
{
    pc_code* mCode = nullptr;
    size_t mN = 0;

public:
    XformHandle() {}
    XformHandle(const XformHandle&) {}
    XformHandle& operator=(const XformHandle&) { return *this; }
    ~XformHandle()
    {
        if (mCode) pc_code_destroy(mCode);
    }
    pc_code* get(size_t N)
    {
        if (mCode && mN == N) return mCode;
        if (mCode) pc_code_destroy(mCode);
        mCode = pc_code_create((int)N, nullptr, 0, PC_DEC_FASTSSC_I8, 1);
        if (!mCode) throw std::runtime_error(pc_last_error());
        mN = N;
        check(pc_code_set_crc(mCode, PC_CRC_NONE));
        check(pc_code_set_systematic(mCode, 0));
        return mCode;
    }
};
} // namespace detail

namespace Encoding {

// encoder.h:24-157
class Encoder : public detail::CodeHandle
{
    size_t mEncoderDuration = 0;

protected:
    ErrorDetection::Detector* mErrorDetector = &ErrorDetection::globalDummy(); // encoder.cpp:18: a bare encoder keeps Dummy
    size_t mBlockLength = 0;
    bool mSystematic = true;
    bool mCodewordReady = false;
    unsigned char* xmInputData = nullptr;
    std::vector<unsigned char> mCodeword; // packed, N/8 bytes
    std::vector<unsigned char> mInfoOut;  // info block as encoded (check bytes filled in)
    std::vector<unsigned> mFrozenBits;
    size_t informationByteSize() const { return (infoLength() + 7) / 8; }
    void rebuild() { create(mBlockLength, mFrozenBits, mErrorDetector->kind(), mSystematic); }

public:
    Encoder() {}
    virtual ~Encoder() {}
    virtual void encode() = 0;
    virtual void initialize(size_t blockLength, const std::vector<unsigned>& frozenBits) = 0;

    /// encoder.cpp:79-90: setInformation -> encode -> getEncodedData
    void encode_vector(void* pInfo, void* pCode)
    {
        const auto t0 = std::chrono::high_resolution_clock::now();
        setInformation(pInfo);
        encode();
        getEncodedData(pCode);
        mEncoderDuration = (size_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
                               std::chrono::high_resolution_clock::now() - t0).count();
    }
    /// B info blocks of ceil(K/8) bytes -> B codewords of N/8 bytes, HOST buffers. The caller's info blocks are not
    /// modified; h_info_crc (optional) receives them with the detector's check bytes filled in.
    void encode_batch(const void* h_info, void* h_code, size_t B, void* h_info_crc = nullptr)
    {
        detail::check(pc_encode_batch_host(mCode, static_cast<const uint8_t*>(h_info), static_cast<uint8_t*>(h_code),
                                           static_cast<uint8_t*>(h_info_crc), B));
    }
    /// same with DEVICE buffers, asynchronous on `stream`
    void encode_batch_device(const uint8_t* d_info, uint8_t* d_code, size_t B, void* stream = nullptr,
                             uint8_t* d_info_crc = nullptr)
    {
        detail::check(pc_encode_batch(mCode, d_info, d_code, d_info_crc, B, stream));
    }
    size_t duration_ns() { return mEncoderDuration; }
    size_t infoLength() const { return mBlockLength - mFrozenBits.size(); }
    size_t blockLength() { return mBlockLength; }
    std::vector<unsigned> frozenBits() { return mFrozenBits; }
    /// the detector is NOT owned (raw pointer, as in the reference)
    void setErrorDetection(ErrorDetection::Detector* pDetector)
    {
        mErrorDetector = pDetector;
        if (mCode) detail::check(pc_code_set_crc(mCode, pDetector->kind()));
    }
    std::string getErrorDetectionMode()
    {
        return std::string(mErrorDetector->getType() + "-" + std::to_string(mErrorDetector->getCheckBitCount()));
    }
    void setSystematic(bool sys)
    {
        mSystematic = sys;
        if (mCode) detail::check(pc_code_set_systematic(mCode, sys ? 1 : 0));
    }
    bool isSystematic() { return mSystematic; }
    void setInformation(void* pData)
    {
        xmInputData = static_cast<unsigned char*>(pData);
        mCodewordReady = false;
    }
    void getInformation(void* pData) { std::memcpy(pData, mInfoOut.data(), informationByteSize()); }
    void getEncodedData(void* pData) { std::memcpy(pData, mCodeword.data(), mBlockLength / 8); }

    /// encoder.cpp:56-77: the "codeword ready" hooks the reference's decoders use for non-systematic output. The next
    /// encode() skips detector and information scatter and transforms the N bits given here (x F^(x)n; systematic:
    /// transform, clear the frozen positions, transform again -- butterfly_fip_packed.cpp:45-59).
    /// setCodeword: N / 8 packed bytes (MSB first); setCharCodeword: N bytes whose MSB is the bit
    /// (PackedContainer::insertCharBits bitcontainer.cpp:959-973); setFloatCodeword: N floats whose sign is the bit
    /// (insertLlr :975-992). (For N < 256 the reference's insertCharBits writes to the wrong end of its 256-bit vector;
    /// that defect is not reproduced: all three forms mean "bit i of the codeword" here.)
    void setCodeword(void* pData)
    {
        xmInputData = static_cast<unsigned char*>(pData);
        std::memcpy(mCodeword.data(), pData, mBlockLength / 8);
        mCodewordReady = true;
    }
    void setCharCodeword(void* cData)
    {
        const unsigned char* in = static_cast<const unsigned char*>(cData);
        for (size_t byte = 0; byte < mBlockLength / 8; ++byte) {
            unsigned char b = 0;
            for (unsigned bit = 0; bit < 8; ++bit) b |= (unsigned char)((in[byte * 8 + bit] & 0x80) >> bit);
            mCodeword[byte] = b;
        }
        mCodewordReady = true;
    }
    void setFloatCodeword(void* fData)
    {
        const unsigned char* in = static_cast<const unsigned char*>(fData); // sign bit = MSB of byte 3 (little endian)
        for (size_t byte = 0; byte < mBlockLength / 8; ++byte) {
            unsigned char b = 0;
            for (unsigned bit = 0; bit < 8; ++bit) b |= (unsigned char)((in[4 * (byte * 8 + bit) + 3] & 0x80) >> bit);
            mCodeword[byte] = b;
        }
        mCodewordReady = true;
    }
    /// PackedContainer::resetFrozenBits (:1004-1010) on the held codeword
    void clearFrozenBits()
    {
        for (unsigned f : mFrozenBits) mCodeword[f >> 3] &= (unsigned char)~(0x80u >> (f & 7));
    }

protected:
    // x -> x F^(x)n of the held codeword on the device: a non-systematic encode of a code without frozen bits
    detail::XformHandle mXform;
    void transformHeld()
    {
        std::vector<unsigned char> out(mBlockLength / 8 + 1);
        detail::check(pc_encode_batch_host(mXform.get(mBlockLength), mCodeword.data(), out.data(), nullptr, 1));
        std::memcpy(mCodeword.data(), out.data(), mBlockLength / 8);
    }
    void encodeHeld()
    {
        transformHeld();
        if (mSystematic) {
            clearFrozenBits();
            transformHeld();
        }
        mCodewordReady = false;
    }
};

// butterfly_fip_packed.h:25-49
class ButterflyFipPacked : public Encoder
{
public:
    ButterflyFipPacked() {}
    ButterflyFipPacked(size_t blockLength, const std::vector<unsigned>& frozenBits) { initialize(blockLength, frozenBits); }
    ~ButterflyFipPacked() {}
    void initialize(size_t blockLength, const std::vector<unsigned>& frozenBits)
    {
        if (blockLength == mBlockLength && frozenBits == mFrozenBits && mCode) return;
        mBlockLength = blockLength;
        mFrozenBits = frozenBits;
        rebuild();
        mCodeword.assign(blockLength / 8 + 1, 0);
        mInfoOut.assign(informationByteSize() + 8, 0);
    }
    /// butterfly_fip_packed.cpp:45-59. The reference writes the check bytes into the caller's buffer
    /// (:47-49); this does the same so that a caller that reads them back keeps working.
    void encode()
    {
        if (mCodewordReady) return encodeHeld();
        if (!xmInputData) throw std::logic_error("ButterflyFipPacked::encode(): no information set");
        detail::check(pc_encode_batch_host(mCode, xmInputData, mCodeword.data(), mInfoOut.data(), 1));
        std::memcpy(xmInputData, mInfoOut.data(), informationByteSize());
    }
};

} // namespace Encoding

namespace Decoding {

// decoder.h:40-180. One concrete class: the kind (PC_DEC_*) picks the kernel the way makeDecoder picks the subclass.
class Decoder : public detail::CodeHandle
{
    size_t mDecoderDuration = 0;

protected:
    ErrorDetection::Detector* mErrorDetector = &ErrorDetection::globalDummy();
    size_t mBlockLength = 0;
    bool mSystematic = true;
    std::vector<unsigned char> mOutputContainer; // ceil(K/8) packed info bytes of the last decode()
    std::vector<unsigned> mFrozenBits;
    std::vector<float> mFloatLlr;
    std::vector<char> mCharLlr;
    bool mSignalIsFloat = false;

    // pData: N (code) / K (info) int8 values for the 'char' Fast-SSC decoder, float values for the 'float' one
    void soft(void* code, void* info)
    {
        if (mKind != PC_DEC_FASTSSC_I8 && mKind != PC_DEC_FASTSSC_F32)
            throw std::logic_error("soft outputs are available for the 'char' and 'float' Fast-SSC decoders only");
        const void* llr = mSignalIsFloat ? static_cast<const void*>(mFloatLlr.data()) : static_cast<const void*>(mCharLlr.data());
        if ((mSignalIsFloat ? mFloatLlr.size() : mCharLlr.size()) != mBlockLength) throw std::logic_error("no signal set");
        if (mKind == PC_DEC_FASTSSC_I8)
            detail::check(pc_decode_soft_batch_host(mCode, llr, mSignalIsFloat ? 1 : 0, static_cast<int8_t*>(code),
                                                    static_cast<int8_t*>(info), 1));
        else
            detail::check(pc_decode_soft_batch_f32_host(mCode, llr, mSignalIsFloat ? 1 : 0, static_cast<float*>(code),
                                                        static_cast<float*>(info), 1));
    }

public:
    Decoder() {}
    Decoder(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int kind)
    {
        mKind = kind;
        mListSize = listSize;
        initialize(blockLength, frozenBits);
    }
    virtual ~Decoder() {}

    virtual void initialize(size_t blockLength, const std::vector<unsigned>& frozenBits)
    {
        if (blockLength == mBlockLength && frozenBits == mFrozenBits && mCode) return;
        mBlockLength = blockLength;
        mFrozenBits = frozenBits;
        create(mBlockLength, mFrozenBits, mErrorDetector->kind(), mSystematic);
        mOutputContainer.assign((infoLength() + 7) / 8 + 8, 0);
    }
    /// decodes the signal set by setSignal; returns the detector's verdict. A failed check is not an error: the
    /// best-effort bits are still available (decoder.cpp:154-181).
    virtual bool decode()
    {
        uint8_t ok = 0;
        const void* llr = mSignalIsFloat ? static_cast<const void*>(mFloatLlr.data()) : static_cast<const void*>(mCharLlr.data());
        detail::check(pc_decode_batch_host(mCode, llr, mSignalIsFloat ? 1 : 0, mOutputContainer.data(), &ok, 1));
        return ok != 0;
    }
    bool decode_vector(const float* pLlr, void* pData)
    {
        const auto t0 = std::chrono::high_resolution_clock::now();
        setSignal(pLlr);
        const bool r = decode();
        getDecodedInformationBits(pData);
        stamp(t0);
        return r;
    }
    bool decode_vector(const char* pLlr, void* pData)
    {
        const auto t0 = std::chrono::high_resolution_clock::now();
        setSignal(pLlr);
        const bool r = decode();
        getDecodedInformationBits(pData);
        stamp(t0);
        return r;
    }
    /// B codewords per call, HOST buffers: llr[B][N] (float or int8) -> info[B][ceil(K/8)], ok[B] (may be null)
    void decode_batch(const float* h_llr, void* h_info, unsigned char* h_ok, size_t B)
    {
        detail::check(pc_decode_batch_host(mCode, h_llr, 1, static_cast<uint8_t*>(h_info), h_ok, B));
    }
    void decode_batch(const char* h_llr, void* h_info, unsigned char* h_ok, size_t B)
    {
        detail::check(pc_decode_batch_host(mCode, h_llr, 0, static_cast<uint8_t*>(h_info), h_ok, B));
    }
    /// same with DEVICE buffers, asynchronous on `stream`
    void decode_batch_device(const void* d_llr, bool llrIsFloat, uint8_t* d_info, uint8_t* d_ok, size_t B, void* stream = nullptr)
    {
        detail::check(pc_decode_batch(mCode, d_llr, llrIsFloat ? 1 : 0, d_info, d_ok, B, stream));
    }
    /// decoder.cpp:147-152: the bit container after the last decode(): N int8 values for the 'char' Fast-SSC decoder
    /// (CharContainer::getSoftBits), N floats for the 'float' one (FloatContainer::getSoftBits -- inner words are XORs of
    /// float patterns, only their sign is meaningful). The list decoders throw std::logic_error: their containers belong
    /// to paths the kernels do not keep. Recomputed on the device from the signal of the last setSignal()/decode_vector().
    void getSoftCodeword(void* pData) { soft(pData, nullptr); }
    /// the container values at the information positions (K values)
    void getSoftInformation(void* pData) { soft(nullptr, pData); }
    /// batch forms, DEVICE buffers: d_soft_code[B][N], d_soft_info[B][K] (either may be null); int8 for 'char', float for 'float'
    void soft_batch_device(const void* d_llr, bool llrIsFloat, int8_t* d_soft_code, int8_t* d_soft_info, size_t B, void* stream = nullptr)
    {
        detail::check(pc_decode_soft_batch(mCode, d_llr, llrIsFloat ? 1 : 0, d_soft_code, d_soft_info, B, stream));
    }
    void soft_batch_device(const void* d_llr, bool llrIsFloat, float* d_soft_code, float* d_soft_info, size_t B, void* stream = nullptr)
    {
        detail::check(pc_decode_soft_batch_f32(mCode, d_llr, llrIsFloat ? 1 : 0, d_soft_code, d_soft_info, B, stream));
    }
    size_t duration_ns() { return mDecoderDuration; }
    std::vector<unsigned> frozenBits() { return mFrozenBits; }
    size_t blockLength() { return mBlockLength; }
    size_t infoLength() { return mBlockLength - mFrozenBits.size(); }
    unsigned char* packedOutput() { return mOutputContainer.data(); }
    virtual void setSystematic(bool sys)
    {
        mSystematic = sys;
        if (mCode) detail::check(pc_code_set_systematic(mCode, sys ? 1 : 0));
    }
    bool isSystematic() { return mSystematic; }
    /// the detector is NOT owned (raw pointer, decoder.cpp:133-136)
    virtual void setErrorDetection(ErrorDetection::Detector* pDetector)
    {
        mErrorDetector = pDetector;
        if (mCode) detail::check(pc_code_set_crc(mCode, pDetector->kind()));
    }
    std::string getErrorDetectionMode()
    {
        return std::string(mErrorDetector->getType() + "-" + std::to_string(mErrorDetector->getCheckBitCount()));
    }
    virtual size_t getListSize() { return mListSize; }
    virtual void setSignal(const float* pLlr)
    {
        mFloatLlr.assign(pLlr, pLlr + mBlockLength);
        mSignalIsFloat = true;
    }
    void setSignal(const char* pLlr)
    {
        mCharLlr.assign(pLlr, pLlr + mBlockLength);
        mSignalIsFloat = false;
    }
    void getDecodedInformationBits(void* pData) { std::memcpy(pData, mOutputContainer.data(), (infoLength() + 7) / 8); }

private:
    void stamp(std::chrono::high_resolution_clock::time_point t0)
    {
        mDecoderDuration = (size_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
                               std::chrono::high_resolution_clock::now() - t0).count();
    }
};

// The reference's concrete decoder classes as thin subclasses that only pick the kind, so that code which constructs
// them directly (pcsim's SimulationWorker::setCoders, src/simulation/simulator.cpp:703-765) compiles against this
// header unchanged. Unlike makeDecoder, the reference constructors install no detector (the Dummy stays).
#define PC_B200_DECODER_CLASS_1(NAME, KIND)                                                                        \
    class NAME : public Decoder                                                                                    \
    {                                                                                                              \
    public:                                                                                                        \
        NAME(size_t blockLength, const std::vector<unsigned>& frozenBits) : Decoder(blockLength, 1, frozenBits, KIND) {} \
    };
#define PC_B200_DECODER_CLASS_L(NAME, KIND)                                                                        \
    class NAME : public Decoder                                                                                    \
    {                                                                                                              \
    public:                                                                                                        \
        NAME(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits)                         \
            : Decoder(blockLength, listSize, frozenBits, KIND)                                                     \
        {                                                                                                          \
        }                                                                                                          \
    };
PC_B200_DECODER_CLASS_1(FastSscFipChar, PC_DEC_FASTSSC_I8)   // fastssc_fip_char.h
PC_B200_DECODER_CLASS_1(FastSscAvxFloat, PC_DEC_FASTSSC_F32) // fastssc_avx_float.h
PC_B200_DECODER_CLASS_L(SclFipChar, PC_DEC_SCL_I8)           // scl_fip_char.h
PC_B200_DECODER_CLASS_L(SclAvxFloat, PC_DEC_SCL_F32)         // scl_avx_float.h
PC_B200_DECODER_CLASS_L(AdaptiveFloat, PC_DEC_ADAPTIVE_F32)  // adaptive_float.h
PC_B200_DECODER_CLASS_L(AdaptiveChar, PC_DEC_ADAPTIVE_I8)    // adaptive_char.h
PC_B200_DECODER_CLASS_L(AdaptiveMixed, PC_DEC_ADAPTIVE_MIXED) // adaptive_mixed.h
PC_B200_DECODER_CLASS_L(DepthFirst, PC_DEC_SCFLIP_F32)        // depth_first.h: DepthFirst(N, trialLimit, frozenBits), SC-Flip
#undef PC_B200_DECODER_CLASS_1
#undef PC_B200_DECODER_CLASS_L

// decoder.cpp:54-87. decoder_impl: 0 char, 1 float, 2 mixed (3 = scan is rejected: there is no CPU fallback).
// Installs a CRC-8 detector like the reference (:85; the object is leaked there too).
inline Decoder* makeDecoder(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, int decoder_impl = 0)
{
    int kind;
    if (decoder_impl == 3) throw std::logic_error("Unknown PolarDecoder type!");
    if (listSize == 1)
        kind = decoder_impl == 1 ? PC_DEC_FASTSSC_F32 : PC_DEC_FASTSSC_I8;
    else
        kind = decoder_impl == 1 ? PC_DEC_SCL_F32 : (decoder_impl == 2 ? PC_DEC_ADAPTIVE_F32 : PC_DEC_SCL_I8);
    Decoder* dec = new Decoder(blockLength, listSize, frozenBits, kind);
    dec->setErrorDetection(ErrorDetection::create(8, "crc"));
    return dec;
}

// decoder.cpp:26-52: case-insensitive substring match on 'char' / 'float' / 'mixed' / 'scan'
inline Decoder* create(size_t blockLength, size_t listSize, const std::vector<unsigned>& frozenBits, std::string decoderType)
{
    std::transform(decoderType.begin(), decoderType.end(), decoderType.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    int flag;
    if (decoderType.find("char") != std::string::npos)
        flag = 0;
    else if (decoderType.find("float") != std::string::npos)
        flag = 1;
    else if (decoderType.find("mixed") != std::string::npos)
        flag = 2;
    else
        throw std::logic_error("Unknown PolarDecoder type!"); // includes 'scan': out of scope, no CPU fallback
    if (listSize < 2 && flag != 0) flag = 1;
    return makeDecoder(blockLength, listSize, frozenBits, flag);
}

} // namespace Decoding
} // namespace PolarCode
#endif
