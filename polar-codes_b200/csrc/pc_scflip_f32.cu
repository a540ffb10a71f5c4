// SC-Flip ("DepthFirst") float decoder for sm_100a: PolarCode::Decoding::DepthFirst of the reference
// (src/polarcode/decoding/depth_first.cpp:556-640, node classes :185-545, Manager :20-178), B codewords per launch.
//
// What the reference does per codeword: an UNPRUNED float SC decoder (createDecoder :497-552 only builds leaves of
// length 1: RateZero for frozen bits, RateOne for information bits; the REP / SPC leaves are commented out) whose
// information leaves report a reliability |LLR| and can be told to flip their decision for one run (setOption). A
// Manager keeps a FIFO of "configurations" (which leaves to flip): after the first run it ranks the leaves by
// reliability and queues "flip the r-th weakest" for r = 0, 1, ... (:29-82); every further run pops the configuration
// just decoded, queues two children that additionally fix the weakest not-yet-configured leaf to 0 / 1 (:85-148),
// configures the next one in the queue and decodes the whole tree again. DepthFirst::decode (:599-640) repeats that
// until the detector accepts or `trialLimit` runs are used; the output is the last run's decision.
//
// Here: one THREAD per codeword (the state machine is per frame and data dependent), the LLR stages, decision bits and
// leaf reliabilities in a thread-interleaved HBM arena (coalesced), the element-wise steps shared with the fp32 Fast-SSC
// core (same IEEE operations as _mm256_polarf_ps / _mm256_polarg_ps, so the leaf LLRs -- and with them the ranking -- are
// the reference's values bit for bit). Two launches per batch like the adaptive decoders: every frame once with one run,
// then only the frames the detector rejected (k_compact_failed) with the full trial limit.
// Tie rule: the reference ranks with std::sort; leaves with EXACTLY equal |LLR| fall where libstdc++'s introsort puts
// them. Here equal reliabilities rank by leaf order. (Continuous channel LLRs make exact ties a measure-zero event; the
// parity tests use such inputs.)
#include "pc_fastssc_f32_core.h"
#include "pc_internal.h"

namespace pc {

constexpr int kScfMaxTrials = 32;
constexpr int kScfMaxDepth = 6; // nodes fixed by one configuration beyond the first; runs are breadth first, so depth stays 1-2

struct ScfParams {
    const float* llr;
    uint8_t* info;
    uint8_t* ok;
    unsigned long long B;
    const unsigned* frame_idx;       // optional indirection (second launch: the rejected frames)
    const unsigned long long* count; // optional device-side frame count
    int S, K, nruns, systematic, trials;
    CrcSpec crc;
    const uint32_t* runs;
    const uint32_t* crc_table;
    const uint32_t* info_mask;       // bit i%32 of word i/32 set for information positions
    int llr_off[24];                 // element offset of LLR stage s (s < S)
    int elems;                       // floats per thread: the LLR stages
    int nbw;                         // decision words per frame
    size_t stride;                   // threads the arena rows are laid out for
    float* arena;                    // [elems | nbw | K reliabilities | K/32 flip words][stride]
};

struct ScfConf {
    unsigned short node[kScfMaxDepth + 1]; // information-leaf indices (0..K-1) fixed by this configuration, in order
    unsigned char depth, opts;             // opts bit d = option of node[d] (1 = flip)
};

// one full SC run over the tree with the flips in `flip` (bit j = information leaf j flips); fills the decision bits
// (codeword domain, like the reference's bit container after the root's combine) and the leaf reliabilities
__device__ __forceinline__ void scf_run(const FCtx& c, const ScfParams& p, const float* row, float* rel, size_t rs,
                                        const uint32_t* flip, size_t fs)
{
    const int S = p.S, N = 1 << S;
    int j = 0; // information-leaf counter
    for (int i = 0; i < N; ++i) {
        // the steps above leaf i: the g step of the node whose right half starts at i (none for i = 0), then f steps down
        int top = S;
        if (i) {
            top = __ffs(i); // trailing zeros + 1 = log2 of that node's size
            const FView in = (top == S) ? FView{ row, 1 } : c.stage(top);
            fop_g(c, in, top, i - (1 << (top - 1)));
            top -= 1;
        }
        for (int s = top; s >= 1; --s) {
            const FView in = (s == S) ? FView{ row, 1 } : c.stage(s);
            fop_f(c, in, s);
        }
        // leaf (RateZeroDecoder :324-328 / RateOneDecoder :371-398 with mBlockLength == 1)
        uint32_t bit = 0;
        if ((__ldg(p.info_mask + (i >> 5)) >> (i & 31)) & 1u) {
            const float l = S ? c.stage(0)[0] : row[0];
            rel[(size_t)j * rs] = fabs_(l);
            bit = fsign(l) ^ ((flip[(size_t)(j >> 5) * fs] >> (j & 31)) & 1u);
            ++j;
        }
        const uint32_t w = c.bld(i >> 5);
        c.bst(i >> 5, (w & ~(1u << (i & 31))) | (bit << (i & 31)));
        // combine every node whose right half ends at leaf i (RateRNode::decode :255-262)
        for (int s = 1; s <= S && ((i >> (s - 1)) & 1); ++s) fop_comb(c, s, (i + 1) - (1 << s), false);
    }
}

__global__ void __launch_bounds__(128) k_scflip_f32(const __grid_constant__ ScfParams p)
{
    __shared__ uint32_t crc_table[256];
    const int T = blockDim.x, t = threadIdx.x;
    if (p.crc.width)
        for (int i = t; i < 256; i += T) crc_table[i] = p.crc_table[i];
    __syncthreads();
    const size_t GT = (size_t)gridDim.x * T, gt = (size_t)blockIdx.x * T + t, AS = p.stride;
    FCtx c;
    c.S = p.S;
    c.llr_off = p.llr_off;
    c.llr = p.arena + gt;
    c.ls = AS;
    c.bits = reinterpret_cast<uint32_t*>(p.arena + (size_t)p.elems * AS) + gt;
    c.bs = AS;
    float* rel = p.arena + (size_t)(p.elems + p.nbw) * AS + gt;
    uint32_t* flip = reinterpret_cast<uint32_t*>(p.arena + (size_t)(p.elems + p.nbw + p.K) * AS) + gt;
    const int fw = (p.K + 31) >> 5;
    Ctx ic; // the int8 core's decision-word helpers (transform, gather) on the same words
    ic.S = p.S;
    ic.bits_kind = PK_ARENA;
    ic.bits = c.bits;
    ic.bws = AS;
    const int N = 1 << p.S, kb = (p.K + 7) >> 3, K = p.K;
    const unsigned long long nfr = p.count ? *p.count : p.B;
    for (unsigned long long fi = gt; fi < nfr; fi += GT) {
        const unsigned long long f = p.frame_idx ? p.frame_idx[fi] : fi;
        const float* row = p.llr + f * N;
        uint8_t* out_row = p.info + f * kb;
        for (int w = 0; w < fw; ++w) flip[(size_t)w * AS] = 0u;
        scf_run(c, p, row, rel, AS, flip, AS);
        if (!p.systematic) transform_bits_inplace(ic);
        bool success = extract_info(ic, p.runs, p.nruns, K, out_row, p.crc, crc_table, true);
        if (!success && p.trials > 1) {
            // ---- Manager::decode (:29-82): rank the leaves, queue "flip the r-th weakest"
            ScfConf q[3 * kScfMaxTrials + 2];
            int qh = 0, qt = 0;
            unsigned short ranked[kScfMaxTrials + 1];
            int nranked = 0;
            const int want = p.trials + 1 < K ? p.trials + 1 : K;
            float last = -1.0f;
            int lastj = -1;
            for (int r = 0; r < want; ++r) { // selection of the next smallest (reliability, leaf index) after (last, lastj)
                float bv = INFINITY;
                int bj = -1;
                for (int jj = 0; jj < K; ++jj) {
                    const float v = rel[(size_t)jj * AS];
                    const bool after = v > last || (v == last && jj > lastj);
                    if (after && (bj < 0 || v < bv)) {
                        bv = v;
                        bj = jj;
                    }
                }
                if (bj < 0) break;
                ranked[nranked++] = (unsigned short)bj;
                last = bv;
                lastj = bj;
            }
            int nodeRank = 0;
            do {
                ScfConf cf;
                cf.depth = 0;
                cf.node[0] = ranked[nodeRank];
                for (int o = (nodeRank == 0 ? 0 : 1); o < 2; ++o) {
                    cf.opts = (unsigned char)o;
                    q[qt++] = cf;
                }
                nodeRank++;
            } while (nodeRank < nranked &&
                     (nodeRank < p.trials * 2 / 3 || (double)rel[(size_t)ranked[nodeRank] * AS] < 2.1972245773362196 /* log(9) */) &&
                     (qt - qh) < p.trials);
            bool firstRun = true;
            // ---- DepthFirst::decode (:599-640) / Manager::decodeNext (:85-148)
            for (int run = 1; !success && run < p.trials; ++run) {
                const ScfConf cur = q[qh++];
                if ((qt - qh) < p.trials && cur.depth < kScfMaxDepth) {
                    int weakest;
                    if (firstRun) {
                        weakest = nranked > 1 ? ranked[1] : -1; // hintList[depth + 1] of the ranked list
                    } else {
                        // the weakest leaf, by the reliabilities of the run just made, among those this configuration
                        // has not fixed (std::sort of hintList[depth + 1 ..], :105-109)
                        float bv = INFINITY;
                        weakest = -1;
                        for (int jj = 0; jj < K; ++jj) {
                            bool fixed = false;
                            for (int d = 0; d <= cur.depth; ++d) fixed |= cur.node[d] == jj;
                            const float v = rel[(size_t)jj * AS];
                            if (!fixed && (weakest < 0 || v < bv)) {
                                bv = v;
                                weakest = jj;
                            }
                        }
                    }
                    if (weakest >= 0) {
                        ScfConf ch = cur;
                        ch.depth = (unsigned char)(cur.depth + 1);
                        ch.node[ch.depth] = (unsigned short)weakest;
                        for (int o = 0; o < 2; ++o) {
                            ch.opts = (unsigned char)(cur.opts | (o << ch.depth));
                            q[qt++] = ch;
                        }
                    }
                }
                firstRun = false;
                if (qh == qt) break; // cannot happen for trial limits >= 2 (two entries are queued before the first pop)
                const ScfConf& nx = q[qh]; // configure the next one in the queue (it stays queued until the next call)
                for (int w = 0; w < fw; ++w) flip[(size_t)w * AS] = 0u;
                for (int d = 0; d <= nx.depth; ++d)
                    if ((nx.opts >> d) & 1) flip[(size_t)(nx.node[d] >> 5) * AS] |= 1u << (nx.node[d] & 31);
                scf_run(c, p, row, rel, AS, flip, AS);
                if (!p.systematic) transform_bits_inplace(ic);
                success = extract_info(ic, p.runs, p.nruns, K, out_row, p.crc, crc_table, true);
            }
        }
        if (p.ok) p.ok[f] = success ? 1 : 0;
    }
}

int scflip_decode_device(pc_code* c, const float* d_llr, uint8_t* d_info, uint8_t* d_ok, size_t B, const unsigned* d_frame_idx,
                         const unsigned long long* d_count, int trials, cudaStream_t st)
{
    ScfParams p;
    p.llr = d_llr;
    p.info = d_info;
    p.ok = d_ok;
    p.B = B;
    p.frame_idx = d_frame_idx;
    p.count = d_count;
    p.S = c->S;
    p.K = c->K;
    p.nruns = (int)c->prog.runs.size();
    p.systematic = c->systematic;
    p.trials = trials < 1 ? 1 : (trials > kScfMaxTrials ? kScfMaxTrials : trials);
    crc_spec(c->crc_kind, p.crc);
    p.runs = c->d_runs;
    p.crc_table = c->d_crc_table;
    p.info_mask = c->d_info_mask;
    int w = 0;
    for (int s = 0; s < 24; ++s) p.llr_off[s] = 0;
    for (int s = 0; s < c->S; ++s) {
        p.llr_off[s] = w;
        w += 1 << s;
    }
    p.elems = w > 0 ? w : 1;
    p.nbw = (c->N + 31) / 32;
    const int threads = 128;
    const size_t per_thread = (size_t)4 * (p.elems + p.nbw + c->K + (c->K + 31) / 32);
    size_t ctas = (B + threads - 1) / threads;
    size_t cap = ((size_t)768 << 20) / (per_thread * threads); // arena <= 768 MiB
    if (cap > 148 * 8) cap = 148 * 8;
    if (cap < 1) cap = 1;
    if (ctas > cap) ctas = cap;
    p.stride = cap * threads; // fixed layout: both launches of a batch use the same rows
    const size_t need = per_thread * p.stride;
    int s = ensure_bytes(reinterpret_cast<void**>(&c->scf_arena), &c->scf_arena_bytes, need);
    if (s != PC_OK) return s;
    p.arena = c->scf_arena;
    k_scflip_f32<<<(int)ctas, threads, 0, st>>>(p);
    count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? PC_OK : cuda_fail(e, "scflip launch");
}

} // namespace pc
