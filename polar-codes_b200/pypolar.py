"""Host-side mirror of the reference's `pypolar` module on top of the C-ABI (libpolar_b200.so).

Same names, argument meaning and error behaviour as reference python/bindings/*.cc:
    PolarEncoder   encoder_python.cc:28-61     (Encoding::ButterflyFipPacked)
    PolarDecoder   decoder_python.cc:21-74     (Decoding::create, kinds 'char' / 'float' / 'mixed')
    Detector       detector_python.cc:23-76    (ErrorDetection::create)
    frozen_bits    frozen_bits_python.cc:24-29 (Construction::frozen_bits)
plus the batch entry points this framework adds: `decode_batch`, `encode_batch` (numpy arrays travel
through pinned staging buffers; torch CUDA tensors are used in place).

All arithmetic happens in the CUDA kernels behind the C-ABI; nothing here computes on the CPU.
"""
import ctypes

import numpy as np

from . import capi
from .capi import check, lib, ptr


def frozen_bits(blockLength, infoLength, designSNR, constructorType="BB"):
    """Construction::frozen_bits (reference include/polarcode/construction/constructor.h:58-62)"""
    N = int(blockLength)
    if N < 1 or (N & (N - 1)):
        raise ValueError("block length must be a power of two")  # constructor.cpp:25-31 std::invalid_argument
    out = np.zeros(N, dtype=np.uint32)
    n = lib().pc_frozen_bits(N, int(infoLength), float(designSNR), constructorType.encode(), ptr(out))
    if n < 0:
        raise ValueError(capi.last_error())
    return [int(v) for v in out[:n]]


def _kind_from_string(decoderType, listSize):
    # extension: an int selects a PC_DEC_* kind directly (AdaptiveChar / AdaptiveMixed / DepthFirst have no type string
    # in the reference factory; pcsim constructs them by class, simulator.cpp:715-733). For PC_DEC_SCFLIP_F32 the list
    # size argument is DepthFirst's trial limit.
    if isinstance(decoderType, int):
        if not 0 <= decoderType <= capi.PC_DEC_SCFLIP_F32:
            raise RuntimeError("Unknown PolarDecoder type!")
        return decoderType
    # Decoding::create (reference src/polarcode/decoding/decoder.cpp:26-52): lower-case substring match
    t = decoderType.lower()
    if "char" in t:
        impl = 0
    elif "float" in t:
        impl = 1
    elif "mixed" in t:
        impl = 2
    elif "scan" in t:
        # out of scope: no CPU fallback exists, so the kind is rejected like any unknown one
        raise RuntimeError("Unknown PolarDecoder type!")
    else:
        raise RuntimeError("Unknown PolarDecoder type!")
    if listSize < 2 and impl != 0:
        impl = 1  # decoder.cpp:48-50
    # makeDecoder (decoder.cpp:54-87)
    if listSize == 1:
        return capi.PC_DEC_FASTSSC_F32 if impl == 1 else capi.PC_DEC_FASTSSC_I8
    return {1: capi.PC_DEC_SCL_F32, 2: capi.PC_DEC_ADAPTIVE_F32}.get(impl, capi.PC_DEC_SCL_I8)


# Detector::getType() + "-" + getCheckBitCount() (reference include/polarcode/errordetection/*.h, dummy.cpp:20-22)
_CRC_NAMES = {0: "DUMMY-0", 6: "CRCNR-6", 8: "CRC-8", 11: "CRCNR-11", 16: "CRC-16", 17: "CRCNR-16", 24: "CRCNR-24", 32: "CRC-32"}


class _Code:
    """shared handle logic of PolarEncoder / PolarDecoder"""

    def __init__(self, N, frozen, kind, L, default_crc):
        fz = np.ascontiguousarray(np.asarray(frozen, dtype=np.uint32))
        self._fz = [int(v) for v in fz]
        self._h = lib().pc_code_create(int(N), ptr(fz), len(fz), kind, int(L))
        if not self._h:
            msg = capi.last_error()
            raise (capi.PolarError if "CUDA" in msg else ValueError)(msg)
        self._N = lib().pc_code_N(self._h)
        self._K = lib().pc_code_K(self._h)
        self._systematic = True
        self._crc = default_crc
        check(lib().pc_code_set_crc(self._h, default_crc))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                lib().pc_code_destroy(h)
            except Exception:
                pass

    def blockLength(self):
        return self._N

    def infoLength(self):
        return self._K

    def frozenBits(self):
        return list(self._fz)

    def isSystematic(self):
        return self._systematic

    def setSystematic(self, flag):
        self._systematic = bool(flag)
        check(lib().pc_code_set_systematic(self._h, int(self._systematic)))

    def getErrorDetectionMode(self):
        return _CRC_NAMES[self._crc]

    def setErrorDetection(self, size=0, type="crc"):
        k = lib().pc_crc_kind(int(size), type.encode())
        if k < 0:
            if "cmac" in type.lower():
                raise RuntimeError("CMAC detectors are host-side crypto and out of scope of the GPU path")
            raise ValueError(capi.last_error())  # std::logic_error("CRC INVALID SIZE!")
        self._crc = k
        check(lib().pc_code_set_crc(self._h, k))


class PolarDecoder(_Code):
    def __init__(self, blockLength, listSize, frozenBitPositions, decoderType):
        kind = _kind_from_string(decoderType, int(listSize))
        self._L = int(listSize)
        self._kind = kind
        # makeDecoder installs CRC-8 (decoder.cpp:85)
        super().__init__(blockLength, frozenBitPositions, kind, listSize, 8)

    def listSize(self):
        return self._L

    def decode_batch(self, llrs, return_ok=False):
        """llrs: [B, N] int8 or float32. numpy in -> numpy out (host path, copies inside the call);
        torch CUDA tensor in -> torch CUDA tensor out (device path, asynchronous on the current stream)."""
        kb = (self._K + 7) // 8
        if hasattr(llrs, "is_cuda"):
            import torch
            if not llrs.is_cuda:
                llrs = llrs.numpy()
            else:
                if llrs.dim() != 2 or llrs.shape[1] != self._N:
                    raise RuntimeError("Input must be [B, blockLength]")
                isf = llrs.dtype == torch.float32
                if not isf and llrs.dtype != torch.int8:
                    raise RuntimeError("LLR dtype must be int8 or float32")
                llrs = llrs.contiguous()
                B = llrs.shape[0]
                out = torch.empty((B, kb), dtype=torch.uint8, device=llrs.device)
                ok = torch.empty((B,), dtype=torch.uint8, device=llrs.device)
                st = torch.cuda.current_stream(llrs.device).cuda_stream
                check(lib().pc_decode_batch(self._h, ptr(llrs), int(isf), ptr(out), ptr(ok), B, ctypes.c_void_p(st)))
                return (out, ok) if return_ok else out
        a = np.asarray(llrs)
        if a.ndim != 2 or a.shape[1] != self._N:
            raise RuntimeError("Input must be [B, blockLength]")
        if a.dtype != np.int8:
            a = a.astype(np.float32, copy=False)
        a = np.ascontiguousarray(a)
        B = a.shape[0]
        out = np.empty((B, kb), dtype=np.uint8)
        ok = np.empty(B, dtype=np.uint8)
        check(lib().pc_decode_batch_host(self._h, ptr(a), int(a.dtype == np.float32), ptr(out), ptr(ok), B))
        return (out, ok) if return_ok else out

    def decode_soft_batch(self, llrs):
        """[B, N] LLRs -> (soft codeword [B, N], soft information [B, K]): Decoder::getSoftCodeword / getSoftInformation
        (decoder.cpp:147-152) after decode(), for every row. 'char' Fast-SSC (L = 1): int8 containers (float32 input is
        quantised like decode_vector does); 'float' Fast-SSC (L = 1): float32 containers, bit for bit the reference's
        patterns (inner words are XORs of float patterns -- compare them through .view(uint32)). Other kinds raise.
        numpy in -> numpy out; torch CUDA tensor in -> CUDA tensors out."""
        isf = self._kind == capi.PC_DEC_FASTSSC_F32
        fn_dev = lib().pc_decode_soft_batch_f32 if isf else lib().pc_decode_soft_batch
        fn_host = lib().pc_decode_soft_batch_f32_host if isf else lib().pc_decode_soft_batch_host
        if hasattr(llrs, "is_cuda") and llrs.is_cuda:
            import torch
            llrs = llrs.contiguous()
            B = llrs.shape[0]
            dt = torch.float32 if isf else torch.int8
            sc = torch.empty((B, self._N), dtype=dt, device=llrs.device)
            si = torch.empty((B, self._K), dtype=dt, device=llrs.device)
            st = torch.cuda.current_stream(llrs.device).cuda_stream
            check(fn_dev(self._h, ptr(llrs), int(llrs.dtype == torch.float32), ptr(sc), ptr(si), B, ctypes.c_void_p(st)))
            return sc, si
        a = np.asarray(llrs)
        if a.ndim != 2 or a.shape[1] != self._N:
            raise RuntimeError("Input must be [B, blockLength]")
        if a.dtype != np.int8:
            a = a.astype(np.float32, copy=False)
        a = np.ascontiguousarray(a)
        B = a.shape[0]
        dt = np.float32 if isf else np.int8
        sc, si = np.empty((B, self._N), dt), np.empty((B, self._K), dt)
        check(fn_host(self._h, ptr(a), int(a.dtype == np.float32), ptr(sc), ptr(si), B))
        return sc, si

    def decode_vector(self, array):
        """decoder_python.cc:41-74: float32[N] or int8[N] -> uint8[K/8] (one codeword = a batch of one)"""
        a = np.asarray(array)
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        if a.size != self._N:
            raise RuntimeError("Input vector size != blockSize // 8!")
        return self.decode_batch(a.reshape(1, -1))[0]


class PolarEncoder(_Code):
    def __init__(self, blockLength, frozenBitPositions):
        # a bare ButterflyFipPacked keeps the base-class Dummy detector (encoder.cpp:18)
        super().__init__(blockLength, frozenBitPositions, capi.PC_DEC_FASTSSC_I8, 1, 0)

    def encode_batch(self, info, return_info=False):
        """info: [B, ceil(K/8)] uint8 -> code words [B, N/8] uint8. The detector's check bytes are written into a
        copy of the info block (returned when return_info), not into the caller's array."""
        kb = (self._K + 7) // 8
        if hasattr(info, "is_cuda") and info.is_cuda:
            import torch
            info = info.contiguous()
            B = info.shape[0]
            code = torch.empty((B, self._N // 8), dtype=torch.uint8, device=info.device)
            icrc = torch.empty_like(info)
            st = torch.cuda.current_stream(info.device).cuda_stream
            check(lib().pc_encode_batch(self._h, ptr(info), ptr(code), ptr(icrc), B, ctypes.c_void_p(st)))
            return (code, icrc) if return_info else code
        a = np.ascontiguousarray(np.asarray(info, dtype=np.uint8))
        if a.ndim != 2 or a.shape[1] != kb:
            raise RuntimeError("Input must be [B, ceil(infoLength / 8)]")
        B = a.shape[0]
        code = np.empty((B, max(self._N // 8, 1)), dtype=np.uint8)
        icrc = np.empty_like(a)
        check(lib().pc_encode_batch_host(self._h, ptr(a), ptr(code), ptr(icrc), B))
        return (code, icrc) if return_info else code

    def encode_vector(self, array):
        """encoder_python.cc:45-61: uint8[K/8] -> uint8[N/8]. Like the reference, a detector's check bytes are written
        INTO the caller's array (ButterflyFipPacked::encode runs Detector::generate on the input buffer,
        butterfly_fip_packed.cpp:47-49; the binding hands it the numpy buffer itself): qa_pypolar_decoder.py compares
        the decoder's output with the array it passed in."""
        a = np.asarray(array, dtype=np.uint8)
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        if a.size != self._K // 8:
            raise RuntimeError("Input vector size != infoLength // 8!")
        code, icrc = self.encode_batch(a.reshape(1, -1), return_info=True)
        if self._crc != 0 and isinstance(array, np.ndarray) and array.dtype == np.uint8 and array.flags.writeable:
            array[...] = icrc[0]
        return code[0]


class Detector:
    """ErrorDetection::create(size, type) (reference src/polarcode/errordetection/errordetector.cpp:27-88)"""

    def __init__(self, size, type="crc"):
        k = lib().pc_crc_kind(int(size), type.encode())
        if k < 0:
            if "cmac" in type.lower():
                raise RuntimeError("CMAC detectors are out of scope of the GPU path")
            raise ValueError(capi.last_error())
        self._kind = k
        self._bits = int(size)
        self._type = _CRC_NAMES[k].split("-")[0]  # Detector::getType(): "CRC", "CRCNR" or "DUMMY"

    def getCheckBitCount(self):
        return self._bits

    def getType(self):
        return self._type

    def _crc_bytes(self):
        return lib().pc_crc_bytes(self._kind)

    def calculate(self, data, bitsize):
        a = np.ascontiguousarray(np.asarray(data, dtype=np.uint8))
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        if a.size * 8 < bitsize:
            raise RuntimeError("Provided number of bits is smaller than stated significant bits!")
        return int(lib().pc_crc_calculate(self._kind, ptr(a), int(bitsize)))

    def generate_batch(self, data):
        """data [B, bytes] -> [B, bytes + crc_bytes] with the check bytes appended (device kernel)"""
        a = np.asarray(data, dtype=np.uint8)
        B, nb = a.shape
        cb = self._crc_bytes()
        rows = np.zeros((B, nb + cb), dtype=np.uint8)
        rows[:, :nb] = a
        if B:
            _run_rows(lambda d, st: lib().pc_crc_generate_batch(self._kind, d, nb + cb, B, st), rows, writeback=True)
        return rows

    def check_batch(self, data):
        a = np.ascontiguousarray(np.asarray(data, dtype=np.uint8))
        B, nb = a.shape
        ok = np.zeros(B, dtype=np.uint8)
        if B:
            d_ok = lib().pc_device_alloc(B)
            try:
                _run_rows(lambda d, st: lib().pc_crc_check_batch(self._kind, d, nb, ctypes.c_void_p(d_ok), B, st), a)
                check(lib().pc_memcpy_d2h(ptr(ok), ctypes.c_void_p(d_ok), B))
            finally:
                lib().pc_device_free(ctypes.c_void_p(d_ok))
        return ok.astype(bool)

    def generate(self, array):
        a = np.asarray(array, dtype=np.uint8)
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        return self.generate_batch(a.reshape(1, -1))[0]

    def check(self, array):
        a = np.asarray(array, dtype=np.uint8)
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        return bool(self.check_batch(a.reshape(1, -1))[0])


def _run_rows(fn, rows, writeback=False):
    nbytes = rows.size
    d = lib().pc_device_alloc(nbytes)
    if not d:
        raise capi.PolarError(capi.last_error())
    try:
        check(lib().pc_memcpy_h2d(ctypes.c_void_p(d), ptr(rows), nbytes))
        check(fn(ctypes.c_void_p(d), None))
        check(lib().pc_device_synchronize())
        if writeback:
            check(lib().pc_memcpy_d2h(ptr(rows), ctypes.c_void_p(d), nbytes))
    finally:
        lib().pc_device_free(ctypes.c_void_p(d))


class Puncturer:
    """pypolar.Puncturer (reference python/bindings/puncturer_python.cc:19-150 over src/polarcode/puncturer.cpp):
    same methods, argument checks and dtypes (uint8 / float32 / float64) for one vector; `*_batch` take [B, n] numpy
    arrays or torch CUDA tensors and run one gather / scatter launch."""
    _DTYPES = {np.dtype(np.uint8): 1, np.dtype(np.int8): 1, np.dtype(np.float32): 4, np.dtype(np.float64): 8}

    def __init__(self, blockLength, frozenBitPositions):
        f = np.ascontiguousarray(np.asarray(frozenBitPositions, dtype=np.uint32))
        self._h = lib().pc_puncturer_create(int(blockLength), ptr(f), int(f.size))
        if not self._h:
            msg = capi.last_error()
            if "exceeds frozen" in msg:
                raise IndexError(msg)                      # std::out_of_range -> IndexError through pybind11
            raise RuntimeError(msg)
        self._h = ctypes.c_void_p(self._h)
        self._M = int(lib().pc_puncturer_block_length(self._h))
        self._Np = int(lib().pc_puncturer_parent_block_length(self._h))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                lib().pc_puncturer_destroy(h)
            except Exception:
                pass

    def blockLength(self):
        return self._M

    def parentBlockLength(self):
        return self._Np

    def blockOutputPositions(self):
        out = np.zeros(self._M, dtype=np.uint32)
        lib().pc_puncturer_positions(self._h, ptr(out))
        return [int(v) for v in out]

    # ---- batch entry points ------------------------------------------------------------------------------------
    def _run(self, fn, x, n_in, n_out, elem=None):
        if hasattr(x, "is_cuda") and x.is_cuda:            # torch CUDA tensor: used in place, asynchronous
            import torch
            if x.dim() != 2 or x.shape[1] != n_in:
                raise RuntimeError("Input must be [B, %d]" % n_in)
            x = x.contiguous()
            out = torch.empty((x.shape[0], n_out), dtype=x.dtype, device=x.device)
            st = ctypes.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            e = x.element_size() if elem is None else elem
            check(fn(ptr(x), ptr(out), e, x.shape[0], st))
            return out
        # host arrays: staged through the C-ABI's own device helpers (no torch on this path)
        a = np.ascontiguousarray(x.numpy() if hasattr(x, "numpy") else x)
        if a.ndim != 2 or a.shape[1] != n_in:
            raise RuntimeError("Input must be [B, %d]" % n_in)
        if a.dtype not in self._DTYPES:
            raise TypeError("unsupported dtype %s" % a.dtype)
        B = a.shape[0]
        out = np.empty((B, n_out), dtype=a.dtype)
        if B == 0:
            return out
        e = a.dtype.itemsize if elem is None else elem
        d_in, d_out = lib().pc_device_alloc(a.nbytes), lib().pc_device_alloc(out.nbytes)
        try:
            if not d_in or not d_out:
                raise capi.PolarError(capi.last_error())
            check(lib().pc_memcpy_h2d(ctypes.c_void_p(d_in), ptr(a), a.nbytes))
            check(fn(ctypes.c_void_p(d_in), ctypes.c_void_p(d_out), e, B, None))
            check(lib().pc_device_synchronize())
            check(lib().pc_memcpy_d2h(ptr(out), ctypes.c_void_p(d_out), out.nbytes))
        finally:
            lib().pc_device_free(ctypes.c_void_p(d_in))
            lib().pc_device_free(ctypes.c_void_p(d_out))
        return out

    def puncture_batch(self, x):
        return self._run(lambda i, o, e, B, st: lib().pc_puncture_batch(self._h, i, o, e, B, st), x, self._Np, self._M)

    def depuncture_batch(self, x):
        return self._run(lambda i, o, e, B, st: lib().pc_depuncture_batch(self._h, i, o, e, B, st), x, self._M, self._Np)

    def puncturePacked_batch(self, x):
        return self._run(lambda i, o, e, B, st: lib().pc_puncture_packed_batch(self._h, i, o, B, st), x, self._Np // 8,
                         self._M // 8, elem=1)

    # ---- the reference's per-vector calls ---------------------------------------------------------------------------
    @staticmethod
    def _vector(array, size, what):
        a = np.asarray(array)
        if a.dtype not in (np.uint8, np.float32, np.float64):
            a = a.astype(np.float64)                       # pybind11 overload order: double first, forcecast
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        if a.size != size:
            raise RuntimeError(what)
        return a

    def puncture(self, array):
        a = self._vector(array, self._Np, "Input vector size != parentBlockSize!")
        return self.puncture_batch(a.reshape(1, -1))[0]

    def depuncture(self, array):
        a = self._vector(array, self._M, "Input vector size != blockSize!")
        return self.depuncture_batch(a.reshape(1, -1))[0]

    def puncturePacked(self, array):
        a = np.asarray(array, dtype=np.uint8)
        if a.ndim != 1:
            raise RuntimeError("Only ONE-dimensional vectors allowed!")
        if a.size != self._Np // 8:
            raise RuntimeError("Input vector size != parentBlockSize!")
        return self.puncturePacked_batch(a.reshape(1, -1))[0]
