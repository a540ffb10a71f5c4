// C-ABI of libpolar_b200.so (include/polar_b200.h): handle management, host-side tree compilation,
// launch planning. No CPU fallback: every compute entry point needs a usable CUDA device.
#include <algorithm>
#include <atomic>
#include <climits>
#include <cstdlib>
#include <cstring>

#include "pc_construct.h"
#include "pc_internal.h"

namespace pc {

static thread_local std::string g_err;
static std::atomic<unsigned long long> g_launches{ 0 };

void set_error(const std::string& s) { g_err = s; }
int fail(int status, const std::string& s)
{
    g_err = s;
    return status;
}
int cuda_fail(cudaError_t e, const char* what)
{
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return PC_ERR_CUDA;
}
void count_launch(unsigned long long n) { g_launches += n; }

// ---- constant-pool allocator (first fit over kConstWords, shared by all live codes) ---------------
static std::mutex g_pool_mu;
static std::vector<std::pair<int, int>> g_pool_used; // (offset, words), sorted by offset

static int pool_alloc(int words)
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    int pos = 0;
    size_t i = 0;
    for (; i < g_pool_used.size(); ++i) {
        if (g_pool_used[i].first - pos >= words) break;
        pos = g_pool_used[i].first + g_pool_used[i].second;
    }
    if (pos + words > kConstWords) return -1;
    g_pool_used.insert(g_pool_used.begin() + i, std::make_pair(pos, words));
    return pos;
}
static void pool_free(int offset)
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (size_t i = 0; i < g_pool_used.size(); ++i)
        if (g_pool_used[i].first == offset) {
            g_pool_used.erase(g_pool_used.begin() + i);
            return;
        }
}

static int device_props(int* sms, size_t* smem_optin)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    int v = 0;
    e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    *sms = v;
    e = cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    *smem_optin = (size_t)v;
    return PC_OK;
}

// Chooses threads per CTA, what lives in shared memory and the grid for the Fast-SSC int8 kernel.
static int plan_fastssc(pc_code* c)
{
    int sms = 0;
    size_t smem_max = 0;
    int st = device_props(&sms, &smem_max);
    if (st != PC_OK) return st;
    FastSscLaunch& P = c->plan;
    const FastSscProgram& g = c->prog;
    P.S = g.S;
    P.K = g.K;
    P.nruns = (int)g.runs.size();
    P.staged_root = g.staged_root ? 1 : 0;
    P.systematic = c->systematic;
    crc_spec(c->crc_kind, P.crc);
    P.const_code = c->const_code;
    P.const_runs = c->const_runs;
    P.g_code = c->d_code;
    P.g_runs = c->d_runs;
    P.g_crc_table = c->d_crc_table;
    P.nbw = (c->N + 31) / 32;
    const int top = g.staged_root ? g.S : g.S - 1;              // highest stage kept in the arena
    const int all_chunks = top >= 4 ? (1 << (top - 3)) - 1 : 0; // chunks 1 .. 2^(top-3)-1
    const size_t budget = smem_max;                              // per CTA, includes the 1 KiB CRC table
    const size_t sm_smem = 233472 - 1024;                        // per SM (228 KiB minus the per-CTA reservation)
    // Placement: stages >= sg live in the L2-resident HBM arena, stages < sg in shared memory. Each candidate
    // split is scored by resident threads per SM; PC_FASTSSC_SPLIT / PC_FASTSSC_THREADS override (tuning).
    int force_sg = -1, force_T = -1;
    if (const char* e = getenv("PC_FASTSSC_SPLIT")) force_sg = atoi(e);
    if (const char* e = getenv("PC_FASTSSC_THREADS")) force_T = atoi(e);
    if (force_T != -1 && (force_T < 32 || force_T > 256 || (force_T & 31))) force_T = -1;
    struct Cand { int sg, T, per_sm, bits_smem; size_t per_thread; };
    Cand best{ -1, 0, 0, 1, 0 };
    for (int sg = top + 1; sg >= 4; --sg) {
        if (force_sg != -1 && sg != (force_sg > top + 1 ? top + 1 : force_sg) && !(force_sg < 4 && sg == 4)) continue;
        const int sm_chunks = sg > top ? all_chunks : (1 << (sg - 4)) - 1;
        for (int bits_smem = 1; bits_smem >= 0; --bits_smem) {
            const size_t per_thread = 16 + 16 * (size_t)sm_chunks + (bits_smem ? 4 * (size_t)P.nbw : 0);
            for (int T = 256; T >= 32; T -= 32) {
                if (force_T != -1 && T != force_T) continue;
                const size_t cta = 1024 + per_thread * T;
                if (cta > budget) continue;
                int per_sm = (int)(sm_smem / (cta + 1024));
                if (per_sm < 1) per_sm = 1;
                if (per_sm * T > 1024) per_sm = 1024 / T; // more threads than this never helped (register file, L2 arena size)
                if (per_sm < 1) continue;
                const int threads = per_sm * T;
                // prefer: more resident threads up to 512; then everything in smem; then larger CTAs
                const int score = (threads > 512 ? 512 : threads) * 8 + (sg > top ? 4 : 0) + bits_smem * 2 + (T >= 128);
                const int bscore = best.sg < 0 ? -1
                                               : ((best.T * best.per_sm > 512 ? 512 : best.T * best.per_sm) * 8 +
                                                  (best.sg > top ? 4 : 0) + best.bits_smem * 2 + (best.T >= 128));
                if (score > bscore) best = Cand{ sg, T, per_sm, bits_smem, per_thread };
            }
        }
        if (force_sg == -1 && best.sg > top && best.T * best.per_sm >= 128) break; // whole tree fits on chip
    }
    if (best.sg < 0) return fail(PC_ERR_UNSUPPORTED, "fastssc_i8: no launch configuration fits shared memory");
    P.bits_in_smem = best.bits_smem;
    if (best.sg > top) {
        P.gsplit = INT_MAX;
        P.g_chunks = 0;
        P.smem_chunks = all_chunks;
    } else {
        P.gsplit = 1 << (best.sg - 4);
        P.smem_chunks = P.gsplit - 1;
        P.g_chunks = all_chunks + 1 - P.gsplit;
    }
    const int T = best.T;
    P.threads = T;
    P.smem_bytes = 1024 + best.per_thread * T;
    P.grid = sms * best.per_sm;
    const size_t need = ((size_t)P.g_chunks * 16 + (P.bits_in_smem ? 0 : 4 * (size_t)P.nbw)) * (size_t)P.grid * T;
    if (need > c->arena_bytes) {
        if (c->arena) cudaFree(c->arena);
        c->arena = nullptr;
        c->arena_bytes = 0;
        cudaError_t e = cudaMalloc(&c->arena, need);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(arena)");
        c->arena_bytes = need;
    }
    P.arena = c->arena;
    c->plan_valid = true;
    return PC_OK;
}

template <class T>
static int upload(T** d, const std::vector<T>& h)
{
    if (*d) cudaFree(*d);
    *d = nullptr;
    if (h.empty()) return PC_OK;
    cudaError_t e = cudaMalloc(d, h.size() * sizeof(T));
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    e = cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy");
    return PC_OK;
}

static int upload_crc(pc_code* c)
{
    CrcSpec s;
    if (!crc_spec(c->crc_kind, s)) return fail(PC_ERR_ARG, "CRC INVALID SIZE!");
    std::vector<uint32_t> table(256);
    crc_build_table(s, table.data());
    return upload(&c->d_crc_table, table);
}

} // namespace pc

using namespace pc;

extern "C" {

const char* pc_last_error(void) { return g_err.c_str(); }
const char* pc_version(void) { return "polar_b200 0.1 (sm_100a)"; }
unsigned long long pc_launch_count(void) { return g_launches.load(); }

int pc_frozen_bits(int N, int K, float design_snr_db, const char* type, unsigned* out)
{
    std::vector<unsigned> fz;
    const char* msg = "";
    int st = frozen_bits(N, K, design_snr_db, type, fz, &msg);
    if (st != 0) {
        set_error(msg);
        return -st;
    }
    for (size_t i = 0; i < fz.size(); ++i) out[i] = fz[i];
    return (int)fz.size();
}

int pc_crc_kind(unsigned size, const char* type)
{
    // ErrorDetection::create (reference src/polarcode/errordetection/errordetector.cpp:27-88)
    std::string t(type ? type : "crc");
    for (auto& ch : t) ch = (char)tolower((unsigned char)ch);
    if (t.find("cmac") != std::string::npos) return -PC_ERR_UNSUPPORTED;
    switch (size) {
    case 0: return PC_CRC_NONE;
    case 6: return PC_CRC_6NR;
    case 8: return PC_CRC_8;
    case 11: return PC_CRC_11NR;
    case 16: return t.find("nr") != std::string::npos ? PC_CRC_16NR : PC_CRC_16;
    case 24: return PC_CRC_24NRC;
    case 32: return PC_CRC_32C;
    }
    set_error("CRC INVALID SIZE!");
    return -PC_ERR_ARG;
}

int pc_crc_bytes(int crc_kind)
{
    CrcSpec s;
    if (!crc_spec(crc_kind, s)) return -PC_ERR_ARG;
    return s.bytes;
}

pc_code* pc_code_create(int N, const unsigned* frozen, int n_frozen, int kind, int L)
{
    if (kind < PC_DEC_FASTSSC_I8 || kind > PC_DEC_ADAPTIVE_F32) {
        set_error("Unknown PolarDecoder type!");
        return nullptr;
    }
    if (L < 1) L = 1;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error(std::string("no usable CUDA device (this library has no CPU path): ") +
                  (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0"));
        return nullptr;
    }
    pc_code* c = new pc_code();
    const char* msg = "";
    if (!compile_fastssc_i8(N, frozen, n_frozen, c->prog, &msg)) {
        set_error(msg);
        delete c;
        return nullptr;
    }
    c->N = N;
    c->K = c->prog.K;
    c->S = c->prog.S;
    c->kind = kind;
    c->L = L;
    c->frozen.assign(frozen, frozen + n_frozen);
    std::sort(c->frozen.begin(), c->frozen.end());
    cudaGetDevice(&c->device);
    int st = upload(&c->d_code, c->prog.code);
    if (st == PC_OK) st = upload(&c->d_runs, c->prog.runs);
    if (st == PC_OK) st = upload(&c->d_lut, c->prog.lut);
    if (st == PC_OK) st = upload(&c->d_info_mask, c->prog.info_mask);
    if (st == PC_OK) st = upload_crc(c);
    if (st == PC_OK) {
        const int words = (int)(c->prog.code.size() + c->prog.runs.size());
        int off = pool_alloc(words);
        if (off >= 0) {
            c->const_code = off;
            c->const_runs = off + (int)c->prog.code.size();
            c->const_words = words;
            e = const_pool_upload(c->const_code, c->prog.code.data(), (int)c->prog.code.size(), nullptr);
            if (e == cudaSuccess && !c->prog.runs.empty())
                e = const_pool_upload(c->const_runs, c->prog.runs.data(), (int)c->prog.runs.size(), nullptr);
            if (e == cudaSuccess) e = cudaDeviceSynchronize();
            if (e != cudaSuccess) st = cuda_fail(e, "constant pool upload");
        }
    }
    if (st != PC_OK) {
        pc_code_destroy(c);
        return nullptr;
    }
    return c;
}

void pc_code_destroy(pc_code* c)
{
    if (!c) return;
    if (c->const_code >= 0) pool_free(c->const_code);
    cudaFree(c->d_code);
    cudaFree(c->d_runs);
    cudaFree(c->d_crc_table);
    cudaFree(c->d_lut);
    cudaFree(c->d_info_mask);
    cudaFree(c->arena);
    cudaFree(c->d_stage_in);
    cudaFree(c->d_stage_out);
    cudaFree(c->d_stage_ok);
    for (auto& s : c->streams)
        if (s) cudaStreamDestroy(s);
    delete c;
}

int pc_code_set_crc(pc_code* c, int crc_kind)
{
    if (!c) return fail(PC_ERR_ARG, "null code");
    CrcSpec s;
    if (!crc_spec(crc_kind, s)) return fail(PC_ERR_ARG, "CRC INVALID SIZE!");
    c->crc_kind = crc_kind;
    c->plan_valid = false;
    return upload_crc(c);
}

int pc_code_set_systematic(pc_code* c, int systematic)
{
    if (!c) return fail(PC_ERR_ARG, "null code");
    c->systematic = systematic ? 1 : 0;
    c->plan_valid = false;
    return PC_OK;
}

int pc_code_N(const pc_code* c) { return c ? c->N : -1; }
int pc_code_K(const pc_code* c) { return c ? c->K : -1; }
int pc_code_info_bytes(const pc_code* c) { return c ? (c->K + 7) / 8 : -1; }
int pc_code_program_length(const pc_code* c) { return c ? (int)c->prog.code.size() : -1; }
int pc_code_program(const pc_code* c, uint32_t* out, int capacity)
{
    if (!c) return -1;
    int n = (int)c->prog.code.size();
    for (int i = 0; i < n && i < capacity; ++i) out[i] = c->prog.code[i];
    return n;
}
int pc_code_launch_info(const pc_code* cc, int* threads_per_cta, int* smem_bytes, int* frames_per_cta)
{
    pc_code* c = const_cast<pc_code*>(cc);
    if (!c) return fail(PC_ERR_ARG, "null code");
    if (!c->plan_valid) {
        int st = plan_fastssc(c);
        if (st != PC_OK) return st;
    }
    if (threads_per_cta) *threads_per_cta = c->plan.threads;
    if (smem_bytes) *smem_bytes = (int)c->plan.smem_bytes;
    if (frames_per_cta) *frames_per_cta = c->plan.threads;
    return PC_OK;
}

} // extern "C"

// ---- decode -------------------------------------------------------------------------------------
namespace pc {
int decode_device(pc_code* c, const void* d_llr, int llr_is_float, uint8_t* d_info, uint8_t* d_ok, size_t B,
                  cudaStream_t st);
}

#include "pc_api_decode.inc"
