// C-ABI of libpolar_b200.so (include/polar_b200.h): handle management, host-side tree compilation,
// launch planning. No CPU fallback: every compute entry point needs a usable CUDA device.
#include <algorithm>
#include <cstdio>
#include <atomic>
#include <climits>
#include <cstdlib>
#include <cstring>

#include "pc_construct.h"
#include "pc_internal.h"
#include "pc_fastssc_i8_core.h"

namespace pc {

void encode2_forget(const pc_code* c); // pc_encode2.cu: drop the code from the encoder's constant-table cache

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
int check_code_device(const pc_code* c)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev != c->device)
        return fail(PC_ERR_ARG, "pc_code was created on CUDA device " + std::to_string(c->device) + " but device " +
                                    std::to_string(dev) + " is current (one code object per device)");
    return PC_OK;
}
int check_aligned(const void* p, size_t align, const char* what)
{
    if (reinterpret_cast<uintptr_t>(p) & (align - 1))
        return fail(PC_ERR_ARG, std::string(what) + " must be " + std::to_string(align) + "-byte aligned");
    return PC_OK;
}

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

// Launch planner of the Fast-SSC int8 kernel: threads per CTA, CTAs per SM and the memory every LLR stage and the
// decision words live in (shared memory, tensor memory, L2-resident arena). One thread needs about N + N/8 bytes
// of private storage; resident warps per SM (latency hiding for the integer pipes) are bounded by that.
// Greedy placement per candidate geometry: stages ascending into shared memory while they fit, the rest into
// tensor memory while it fits, the rest into the arena; decision words likewise.
// Environment overrides (tuning only): PC_FASTSSC_THREADS, PC_FASTSSC_CTAS, PC_FASTSSC_NOTMEM=1.
// cap_T: largest CTA size considered. 1024 for batches that fill the GPU; smaller for small batches, so that every SM gets
// a CTA (N=16384, 65 536 frames: 64 CTAs of 1024 threads 76 Gbit/s, 128 CTAs of 512 threads 102)
static int plan_fastssc(pc_code* c, int cap_T = 1024)
{
    int sms = 0;
    size_t smem_max = 0;
    int st = device_props(&sms, &smem_max);
    if (st != PC_OK) return st;
    FastSscLaunch& P = cap_T >= 1024 ? c->plan : c->plan_small;
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
    const int top = g.staged_root ? g.S : g.S - 1; // highest stage stored by the kernel
    int force_T = -1, force_ctas = -1, no_tmem = 0;
    if (const char* e = getenv("PC_FASTSSC_THREADS")) force_T = atoi(e);
    if (const char* e = getenv("PC_FASTSSC_CTAS")) force_ctas = atoi(e);
    if (const char* e = getenv("PC_FASTSSC_NOTMEM")) no_tmem = atoi(e);
    const bool no_flip = getenv("PC_FASTSSC_NOFLIP") != nullptr; // experiment knob, see DESIGN.md 3.1 (open item)
    const size_t sm_smem = 233472; // 228 KiB per SM, 1 KiB of it reserved per resident CTA
    struct Plan {
        int T = 0, per_sm = 0, smem_chunks = 0, tm_alloc = 0, tm_cols = 0, arena_chunks = 0, bits_kind = 0, bits_base = 0;
        uint32_t place[24] = { 0 };
        size_t smem_bytes = 0;
        long score = -1;
    } best;
    for (int per_sm = 1; per_sm <= 2; ++per_sm) {
        if (force_ctas != -1 && per_sm != force_ctas) continue;
        for (int T = cap_T; T >= 32; T -= 32) {
            if (force_T != -1 ? T != force_T : (T != 32 && T != 64 && (T % 128) != 0)) continue;
            if (T * per_sm > 1024) continue; // 64 registers per thread: more would not be resident (and the grid would run in two rounds)
            const size_t cta_budget = std::min(smem_max, sm_smem / per_sm - 1024);
            if (cta_budget < 1024 + 64 * (size_t)T) continue;
            const size_t per_thread_budget = (cta_budget - 1024) / T;
            const int warps = T / 32, stack = (warps + 3) / 4;
            int tm_alloc = no_tmem ? 0 : 512 / per_sm; // power of two
            int tm_budget = tm_alloc / stack;           // columns per thread
            Plan q;
            q.T = T;
            q.per_sm = per_sm;
            size_t used = 16; // the four small words
            int tm_used = 0, ar_used = 0, sm_chunks = 0;
            bool smem_open = true;
            for (int s = 4; s <= top && s < 24; ++s) {
                const int nch = 1 << (s - 4);
                if (smem_open && (s <= 5 || used + 16 * (size_t)nch <= per_thread_budget)) {
                    q.place[s] = mk_place(PK_SMEM, (uint32_t)sm_chunks);
                    sm_chunks += nch;
                    used += 16 * (size_t)nch;
                    continue;
                }
                smem_open = false; // keep the stage order monotone: small stages on chip, large ones further out
                if (tm_used + 4 * nch <= tm_budget) {
                    q.place[s] = mk_place(PK_TMEM, (uint32_t)tm_used);
                    tm_used += 4 * nch;
                } else {
                    q.place[s] = mk_place(PK_ARENA, (uint32_t)ar_used);
                    ar_used += nch;
                }
            }
            if (used > per_thread_budget) continue;
            if (used + 4 * (size_t)P.nbw <= per_thread_budget) {
                q.bits_kind = PK_SMEM;
                used += 4 * (size_t)P.nbw;
            } else if (tm_used + P.nbw <= tm_budget) {
                q.bits_kind = PK_TMEM;
                q.bits_base = tm_used;
                tm_used += P.nbw;
            } else {
                q.bits_kind = PK_ARENA;
            }
            q.smem_chunks = sm_chunks;
            q.arena_chunks = ar_used;
            q.tm_cols = tm_budget;
            q.tm_alloc = tm_used ? tm_alloc : 0;
            q.smem_bytes = 1024 + 16 * (size_t)T + 16 * (size_t)sm_chunks * T + (q.bits_kind == PK_SMEM ? 4 * (size_t)P.nbw * T : 0);
            const int resident = T * per_sm;
            // Measured on B200 (N=1024, tools/sweep.sh, whole rounds of ~1M frames, 10 steps, gpurun_out/sweep5.log):
            // 1 x 1024 threads 200.9 Gbit/s, 1 x 640 192.5, 1 x 960 176.2, 2 x 512 175.8, 1 x 512 167.5, 1 x 768 155.6,
            // 1 x 896 136.4 (the 1 x 1024 and 2 x 512 plans place every stage identically). So: maximise resident
            // threads up to the register-file limit (64 regs x 1024), prefer ONE CTA per SM, and charge only a little
            // for arena use. Full grids (2 x 148 x 1024 frames per launch), forced shapes: N=4096 1 x 1024 169 Gbit/s,
            // 2 x 512 159; N=8192 155 / 152; N=16384 143 / 140, but 1 x 1024 showed single runs at 118 (N=8192) and
            // 113 (N=16384) on the same box, so from N=8192 on (arena chunks >= 400) two CTAs of 512 stay preferred.
            q.score = (long)std::min(resident, 1024) * 100 - (ar_used ? 1000 + 4L * std::min(ar_used, 512) : 0) -
                      (q.bits_kind == PK_ARENA ? 4000 : 0) - (per_sm - 1) * ((ar_used >= 400 && !no_flip) ? -2000 : 2000);
            if (q.score > best.score) best = q;
        }
    }
    if (best.score < 0) {
        if (getenv("PC_DEBUG_OCC")) fprintf(stderr, "plan_fastssc: cap_T %d top %d smem_max %zu nbw %d\n", cap_T, top, smem_max, P.nbw);
        return fail(PC_ERR_UNSUPPORTED, "fastssc_i8: no launch configuration fits the on-chip memories");
    }
    if (getenv("PC_DEBUG_OCC"))
        fprintf(stderr, "plan_fastssc: N=%d cap %d -> %d x %d threads, smem %zu B, arena chunks %d, tmem cols %d, score %ld\n", c->N, cap_T,
                best.per_sm, best.T, best.smem_bytes, best.arena_chunks, best.tm_alloc, best.score);
    P.threads = best.T;
    P.ctas_per_sm = best.per_sm;
    P.grid = sms * best.per_sm;
    P.smem_bytes = best.smem_bytes;
    for (int i = 0; i < 24; ++i) P.place[i] = best.place[i];
    P.smem_chunks = best.smem_chunks;
    P.tmem_alloc_cols = best.tm_alloc;
    P.tmem_cols_thread = best.tm_cols;
    P.arena_chunks = best.arena_chunks;
    P.bits_kind = best.bits_kind;
    P.bits_base = best.bits_base;
    const int T = P.threads;
    const size_t need = ((size_t)P.arena_chunks * 16 + (P.bits_kind == PK_ARENA ? 4 * (size_t)P.nbw : 0)) * (size_t)P.grid * T;
    if (need > c->arena_bytes) {
        if (c->arena) cudaFree(c->arena);
        c->arena = nullptr;
        c->arena_bytes = 0;
        cudaError_t e = cudaMalloc(&c->arena, need);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(arena)");
        c->arena_bytes = need;
    }
    P.arena = c->arena; // refreshed at every launch (the other plan may have grown the buffer)
    if (cap_T >= 1024)
        c->plan_valid = true;
    else
        c->plan_small_cap = cap_T;
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
    if (kind < PC_DEC_FASTSSC_I8 || kind > PC_DEC_SCFLIP_F32) {
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
    if (st == PC_OK) st = build_encoder_tables(c);
    if (st == PC_OK && (kind == PC_DEC_FASTSSC_F32 || kind == PC_DEC_ADAPTIVE_F32)) {
        if (!compile_fastssc_f32(N, frozen, n_frozen, c->f32_prog, &msg)) {
            set_error(msg);
            st = PC_ERR_ARG;
        } else {
            st = upload(&c->d_f32_code, c->f32_prog.code);
        }
    }
    if (st == PC_OK && (kind == PC_DEC_SCL_I8 || (kind == PC_DEC_ADAPTIVE_I8 && L > 1))) {
        if (!compile_scl(N, frozen, n_frozen, c->scl_prog, &msg)) {
            set_error(msg);
            st = PC_ERR_ARG;
        } else {
            st = upload(&c->d_scl_code, c->scl_prog.code);
        }
    }
    if (st == PC_OK && (kind == PC_DEC_SCL_F32 || ((kind == PC_DEC_ADAPTIVE_F32 || kind == PC_DEC_ADAPTIVE_MIXED) && L > 1))) {
        if (!compile_scl(N, frozen, n_frozen, c->sclf_prog, &msg, true)) {
            set_error(msg);
            st = PC_ERR_ARG;
        } else {
            st = upload(&c->d_sclf_code, c->sclf_prog.code);
        }
    }
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
    encode2_forget(c);
    if (c->const_code >= 0) pool_free(c->const_code);
    cudaFree(c->d_code);
    cudaFree(c->d_runs);
    cudaFree(c->d_crc_table);
    cudaFree(c->d_lut);
    cudaFree(c->d_info_mask);
    cudaFree(c->d_enc_runs);
    cudaFree(c->d_enc_first);
    cudaFree(c->d_enc_tab);
    cudaFree(c->d_enc_info);
    cudaFree(c->d_chan_code);
    cudaFree(c->d_quant);
    cudaFree(c->d_scl_code);
    cudaFree(c->scl_arena);
    cudaFree(c->d_f32_code);
    cudaFree(c->d_sclf_code);
    cudaFree(c->sclf_arena);
    cudaFree(c->d_fail_idx);
    cudaFree(c->d_fail_count);
    cudaFree(c->d_ok_tmp);
    cudaFree(c->f32_arena);
    cudaFree(c->d_f32_in);
    cudaFree(c->arena);
    cudaFree(c->soft_arena);
    cudaFree(c->scf_arena);
    cudaFree(c->d_stage_in);
    cudaFree(c->d_stage_out);
    cudaFree(c->d_stage_ok);
    for (auto& s : c->streams)
        if (s) cudaStreamDestroy(s);
    if (c->ev_decode) cudaEventDestroy(c->ev_decode);
    delete c;
}

int pc_code_set_crc(pc_code* c, int crc_kind)
{
    if (!c) return fail(PC_ERR_ARG, "null code");
    CrcSpec s;
    if (!crc_spec(crc_kind, s)) return fail(PC_ERR_ARG, "CRC INVALID SIZE!");
    if (int ds = check_code_device(c)) return ds;
    c->crc_kind = crc_kind;
    c->plan_valid = false;
    c->plan_small_cap = 0;
    c->scl_plan_valid = false;
    c->f32_plan_valid = false;
    c->sclf_plan_valid = false;
    return upload_crc(c);
}

int pc_code_set_systematic(pc_code* c, int systematic)
{
    if (!c) return fail(PC_ERR_ARG, "null code");
    c->systematic = systematic ? 1 : 0;
    c->plan_valid = false;
    c->plan_small_cap = 0;
    c->scl_plan_valid = false;
    c->f32_plan_valid = false;
    c->sclf_plan_valid = false;
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
