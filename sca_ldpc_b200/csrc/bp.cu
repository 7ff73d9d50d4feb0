// libscaldpc -- binary product-sum BP decoder: graph upload, workspace, decode pipeline and C ABI
// (include/scaldpc.h).  Replaces ldpc.bp_decoder at the call sites simulate/hqc.py:694-708 and
// simulate/decode.py:155-171.  No CPU compute path exists here.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "bp_graph.h"
#include "bp_kernels.cuh"
#include "bp_build.cuh"
#include "common.h"

using namespace scaldpc;

// One set of per-slot state.  Two sets exist (ping-pong): compaction gathers the running trials of the current set
// into the other one; set 1 only ever receives <= half of the slots.
struct BpSet {
    int cap_tiles = 0;
    float *msg = nullptr, *prior_t = nullptr;
    uint32_t *synd_w = nullptr, *valid = nullptr, *done = nullptr, *unsat = nullptr;
    int *orig = nullptr;
};

struct scaldpc_bp {
    int m = 0, n = 0, E = 0, dc_max = 0, dv_max = 0;
    int max_iter = 0, device = 0, num_sms = 148;
    int class_end[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // segmented CSR: G graphs of the same shape (scaldpc_bp_create_multi); E is then the LARGEST edge count (table stride)
    int G = 1;
    int class_max[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // largest size of every degree class over the graphs
    int *d_class_end = nullptr;   // [G][kBpVnClasses+1]
    int *d_tile_graph = nullptr;  // [cap_tile_graph] graph of every tile of the running decode
    int *d_slot_orig = nullptr;   // staging of the slot -> trial map
    size_t cap_tile_graph = 0;
    const int *cur_tile_graph = nullptr;
    int *d_row_ptr = nullptr, *d_col = nullptr, *d_col_ptr = nullptr, *d_vn_order = nullptr, *d_vn_desc = nullptr;
    uint32_t *d_csc_edge = nullptr;
    int *d_edge_row = nullptr;     // E: check of every edge (first pass with shared priors)
    float *d_first_tab = nullptr;  // E: first-iteration check-to-bit messages for syndrome bit 0 (shared priors)
    float *d_csc_tab = nullptr;    // the same values in CSC order (read by the FIRST bit pass next to csc_off)
    int *d_csc_row = nullptr, *d_edge_csc = nullptr;  // check of every CSC entry; CSC position of every edge
    // workspace
    int cap_tiles = 0;
    bool cap_post = false, cap_prior_t = false;
    BpSet set[2];
    float *d_prior = nullptr, *d_post = nullptr;
    uint32_t *d_dec_w = nullptr;
    int *d_iters = nullptr, *d_conv = nullptr, *d_counts = nullptr, *d_src = nullptr, *d_nact = nullptr;
    int cap_counts = 0;
    int *h_pin = nullptr;  // pinned: [0] running trials
    // staging for host-pointer calls
    size_t cap_stage_T = 0;
    bool stage_probs_t = false, stage_llr = false;
    uint8_t *d_synd = nullptr, *d_dec = nullptr;
    int32_t *d_nrows = nullptr;        // per-trial active-check counts of a nested-graph decode (device copy, T entries)
    size_t cap_nrows = 0;
    const int *cur_nrows = nullptr;    // what the kernels of the running decode see (null outside nested decodes)
    int32_t *d_truth = nullptr, *d_stats = nullptr;  // scaldpc_bp_decode_scored staging (host-pointer calls)
    size_t cap_truth = 0, cap_stats_T = 0;
    double *d_probs = nullptr;
    float *d_llr = nullptr;
    int cn_warps = 8, cn_occ[2] = {1, 1};
    size_t smem_cn = 0;
    int vn_occ[8] = {0};  // bit pass: resident CTAs per SM of each degree-class kernel
    int vn_pipe_occ[8] = {0};  // ... of the cp.async-pipelined kernels
    bool vn_pipe = false;      // SCALDPC_BP_VN_PIPE=1: cp.async-pipelined bit pass.  Measured SLOWER than the register-prefetch
                               // kernels (hqc128 413 k vs 454 k trials/s, profiles/r2o_*): off, kept for experiments
    int vn_single_tiles = 64;  // at most this many tiles: one any-degree launch per bit pass (SCALDPC_BP_VN_SINGLE=n)
    long long vn_single_items = 65536;  // ... and at most this many (tile, variable) items (SCALDPC_BP_VN_SINGLE_ITEMS=n)
    // (round 1 also had k_bp_vn_tma: column gather by per-lane bulk-async copies into a 4-stage ring.  Measured on hqc128 /
    // hqc256: 1.41 / 1.98 ms against 1.35 / 1.91 ms for the register-prefetch kernel -- one UBLKCP per lane is serialised --
    // so it was removed in round 2.)
    int vn_gen_wpc = 1;   // warps per CTA of the shared-memory kernel for heavier columns
    size_t smem_vn = 0;
    std::mutex mu;
    long long launches = 0;
    int last_iters_run = 0, last_compactions = 0;
    // host path: device-to-host copy of the finished rows while the stragglers still iterate (bp_decode_impl)
    cudaStream_t s_copy = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int early_copy_max = 256;  // start it at the first compaction that leaves at most this many running (SCALDPC_BP_EARLY_COPY)
    int last_early_copy_rows = -1;
};

namespace {

template <typename T>
int upload_vec(T **dst, const std::vector<T> &src)
{
    SCALDPC_CUDA(cudaMalloc((void **)dst, std::max<size_t>(1, src.size()) * sizeof(T)));
    if (!src.empty()) SCALDPC_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return 0;
}

void free_set(BpSet &s)
{
    cudaFree(s.msg);
    cudaFree(s.prior_t);
    cudaFree(s.synd_w);
    cudaFree(s.valid);
    cudaFree(s.done);
    cudaFree(s.unsat);
    cudaFree(s.orig);
    s = BpSet();
}

void free_ws(scaldpc_bp *h)
{
    free_set(h->set[0]);
    free_set(h->set[1]);
    cudaFree(h->d_post);
    cudaFree(h->d_dec_w);
    cudaFree(h->d_iters);
    cudaFree(h->d_conv);
    cudaFree(h->d_src);
    h->d_post = nullptr;
    h->d_dec_w = nullptr;
    h->d_iters = h->d_conv = h->d_src = nullptr;
    h->cap_tiles = 0;
    h->cap_post = h->cap_prior_t = false;
}

void free_stage(scaldpc_bp *h)
{
    cudaFree(h->d_synd);
    cudaFree(h->d_dec);
    cudaFree(h->d_probs);
    cudaFree(h->d_llr);
    h->d_synd = h->d_dec = nullptr;
    h->d_probs = nullptr;
    h->d_llr = nullptr;
    h->cap_stage_T = 0;
    h->stage_probs_t = h->stage_llr = false;
}

int alloc_set(scaldpc_bp *h, BpSet &s, int t, bool prior_t)
{
    const size_t lanes = (size_t)t * 32;
    SCALDPC_CUDA(cudaMalloc((void **)&s.msg, std::max<size_t>(1, lanes * h->E) * sizeof(float)));
    if (prior_t) SCALDPC_CUDA(cudaMalloc((void **)&s.prior_t, lanes * h->n * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&s.synd_w, (size_t)t * h->m * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&s.valid, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&s.done, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&s.unsat, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&s.orig, lanes * sizeof(int)));
    s.cap_tiles = t;
    return 0;
}

int ensure_ws(scaldpc_bp *h, int tiles, bool post, bool prior_t)
{
    if (tiles <= h->cap_tiles && (!post || h->cap_post) && (!prior_t || h->cap_prior_t)) return 0;
    const int t = std::max(tiles, h->cap_tiles);
    post = post || h->cap_post;
    prior_t = prior_t || h->cap_prior_t;
    free_ws(h);
    const size_t lanes = (size_t)t * 32;
    int rc;
    if ((rc = alloc_set(h, h->set[0], t, prior_t))) return rc;
    if ((rc = alloc_set(h, h->set[1], (t + 1) / 2, prior_t))) return rc;  // compaction target: <= half the slots
    if (post) SCALDPC_CUDA(cudaMalloc((void **)&h->d_post, lanes * h->n * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_dec_w, (size_t)t * h->n * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_iters, lanes * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_conv, lanes * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_src, lanes * sizeof(int)));
    h->cap_tiles = t;
    h->cap_post = post;
    h->cap_prior_t = prior_t;
    return 0;
}

int ensure_stage(scaldpc_bp *h, size_t T, bool probs_t, bool llr)
{
    if (T <= h->cap_stage_T && (!probs_t || h->stage_probs_t) && (!llr || h->stage_llr)) return 0;
    const size_t t = std::max(T, h->cap_stage_T);
    probs_t = probs_t || h->stage_probs_t;
    llr = llr || h->stage_llr;
    free_stage(h);
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_synd, t * h->m));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_dec, t * h->n));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_probs, (probs_t ? t : 1) * h->n * sizeof(double)));
    if (llr) SCALDPC_CUDA(cudaMalloc((void **)&h->d_llr, t * h->n * sizeof(float)));
    h->cap_stage_T = t;
    h->stage_probs_t = probs_t;
    h->stage_llr = llr;
    return 0;
}

BpDev view(const scaldpc_bp *h, int cur, bool per_trial_prior, bool want_post)
{
    const BpSet &s = h->set[cur];
    BpDev p;
    p.m = h->m;
    p.n = h->n;
    p.E = h->E;
    p.dc_max = h->dc_max;
    p.dv_max = h->dv_max;
    p.per_trial_prior = per_trial_prior;
    p.row_ptr = h->d_row_ptr;
    p.col = h->d_col;
    p.col_ptr = h->d_col_ptr;
    p.csc_off = h->d_csc_edge;
    p.vn_order = h->d_vn_order;
    p.vn_desc = (const int4 *)h->d_vn_desc;
    p.msg = s.msg;
    p.prior = per_trial_prior ? s.prior_t : h->d_prior;
    p.synd_w = s.synd_w;
    p.dec_w = h->d_dec_w;
    p.post = want_post ? h->d_post : nullptr;
    p.done = s.done;
    p.valid = s.valid;
    p.unsat = s.unsat;
    p.orig = s.orig;
    p.iters = h->d_iters;
    p.conv = h->d_conv;
    p.counts = h->d_counts;
    p.nrows = h->cur_nrows;
    p.first_tab = h->d_first_tab;
    p.edge_row = h->d_edge_row;
    p.csc_tab = h->d_csc_tab;
    p.csc_row = h->d_csc_row;
    p.edge_csc = h->d_edge_csc;
    p.g_rp = h->m + 1;
    p.g_e = h->E;
    p.g_cp = h->n + 1;
    p.g_n = h->n;
    if (h->G > 1) {
        p.tile_graph = h->cur_tile_graph;
        p.class_end = h->d_class_end;
    }
    return p;
}

// Check pass: persistent grid, `cn_warps` warps per CTA, as many CTAs per SM as the staging buffers allow.
int launch_cn(scaldpc_bp *h, const BpDev &p, int tiles, bool first, cudaStream_t st)
{
    const long long items = (long long)tiles * h->m;
    const int ctas = (int)std::min<long long>((long long)h->num_sms * h->cn_occ[first], (items + h->cn_warps - 1) / h->cn_warps);
    if (first)
        k_bp_cn<true><<<std::max(1, ctas), h->cn_warps * 32, h->smem_cn, st>>>(p, tiles);
    else
        k_bp_cn<false><<<std::max(1, ctas), h->cn_warps * 32, h->smem_cn, st>>>(p, tiles);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

typedef void (*bp_vn_fn)(const BpDev, const int, const int, const int, const int);
bp_vn_fn vn_class_fn(int c, bool first = false, bool shared = false)
{
#define CLS(A_, B_)                                                                                  \
    return shared ? (first ? k_bp_vn<A_, B_, true, true> : k_bp_vn<A_, B_, false, true>)            \
                  : (first ? k_bp_vn<A_, B_, true, false> : k_bp_vn<A_, B_, false, false>)
    switch (c) {
    case 0: CLS(0, 1);
    case 1: CLS(1, 4);
    case 2: CLS(4, 8);
    case 3: CLS(8, 12);
    case 4: CLS(12, 16);
    case 5: CLS(16, 24);
    case 6: CLS(24, 32);
    default: return nullptr;
    }
#undef CLS
}

// The cp.async-pipelined bit pass of iterations >= 2 (k_bp_vn_pipe) and its shared memory per CTA.
bp_vn_fn vn_pipe_fn(int c)
{
    switch (c) {
    case 1: return k_bp_vn_pipe<1, 4>;
    case 2: return k_bp_vn_pipe<4, 8>;
    case 3: return k_bp_vn_pipe<8, 12>;
    case 4: return k_bp_vn_pipe<12, 16>;
    case 5: return k_bp_vn_pipe<16, 24>;
    case 6: return k_bp_vn_pipe<24, 32>;
    default: return nullptr;  // degree <= 1: nothing to pipeline
    }
}
size_t vn_pipe_smem(int c)
{
    switch (c) {
    case 1: return BpVnPipe<4>::smem_bytes;
    case 2: return BpVnPipe<8>::smem_bytes;
    case 3: return BpVnPipe<12>::smem_bytes;
    case 4: return BpVnPipe<16>::smem_bytes;
    case 5: return BpVnPipe<24>::smem_bytes;
    case 6: return BpVnPipe<32>::smem_bytes;
    default: return 0;
    }
}

// Bit pass: one persistent launch per non-empty degree class (variables are sorted by degree); columns heavier than
// the largest register class go through the shared-memory kernel.
int launch_vn(scaldpc_bp *h, const BpDev &p, int tiles, cudaStream_t st, bool first = false)
{
    auto generic = first ? k_bp_vn_generic<true> : k_bp_vn_generic<false>;
    if (tiles <= h->vn_single_tiles && (long long)tiles * h->n <= h->vn_single_items) {
        // few (tile, variable) items (the non-converging tail after compaction): the iteration is bound by kernel launches,
        // so ONE launch of the any-degree kernel over all variables replaces one launch per degree class (same arithmetic).
        // Only while the launch latency dominates: per item the any-degree kernel is ~2x slower than the class kernels
        // (hqc256, 4 tiles x 65 637 variables: 230 us against ~120 us -- profiles/r2l), so larger tails keep the classes.
        generic<<<dim3((h->n + h->vn_gen_wpc - 1) / h->vn_gen_wpc, tiles), h->vn_gen_wpc * 32, h->smem_vn, st>>>(p, 0, h->n, -1);
        ++h->launches;
        SCALDPC_CUDA(cudaGetLastError());
        return 0;
    }
    const bool multi = h->G > 1;  // every tile takes its own graph's slice of the class: [0, largest slice) is launched
    int lo = 0;
    for (int c = 0; c <= kBpVnClasses; ++c) {
        const int hi = multi ? h->class_max[c] : h->class_end[c];
        if (multi) lo = 0;
        if (hi > lo) {
            if (c < kBpVnClasses) {
                const long long items = (long long)tiles * (hi - lo);
                const bool pipe = !first && h->vn_pipe && vn_pipe_fn(c);
                const int occ = pipe ? h->vn_pipe_occ[c] : h->vn_occ[c];
                const int ctas = (int)std::min<long long>((long long)h->num_sms * occ, (items + kBpVnWarps - 1) / kBpVnWarps);
                if (pipe)
                    vn_pipe_fn(c)<<<std::max(1, ctas), kBpVnWarps * 32, vn_pipe_smem(c), st>>>(p, lo, hi, tiles, multi ? c : -1);
                else
                    vn_class_fn(c, first, !p.per_trial_prior && !p.post)<<<std::max(1, ctas), kBpVnWarps * 32, 0, st>>>(p, lo, hi, tiles, multi ? c : -1);
            } else {
                generic<<<dim3((hi - lo + h->vn_gen_wpc - 1) / h->vn_gen_wpc, tiles), h->vn_gen_wpc * 32, h->smem_vn, st>>>(
                    p, lo, hi, multi ? c : -1);
            }
            ++h->launches;
        }
        lo = hi;
    }
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

// Outputs of the slots selected by `mask` (per tile) of set `cur` -> the caller's arrays (original trial order).
int write_outputs(scaldpc_bp *h, int cur, int tiles, const uint32_t *mask, uint8_t *ddec, float *dllr, cudaStream_t st)
{
    k_bp_unpack_dec<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(h->d_dec_w, ddec, h->set[cur].orig, mask, h->n);
    ++h->launches;
    if (dllr) {
        k_bp_untile_f32<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(h->d_post, dllr, h->set[cur].orig, mask, h->n);
        ++h->launches;
    }
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

// Packs the `n_run` running trials of set `cur` (cur_tiles tiles) into set cur^1; finished trials are written out first.
int compact(scaldpc_bp *h, int cur, int cur_tiles, int n_run, bool probs_t, uint8_t *ddec, float *dllr, cudaStream_t st)
{
    BpSet &a = h->set[cur], &b = h->set[cur ^ 1];
    const int new_tiles = (n_run + 31) / 32;
    if (new_tiles > b.cap_tiles) return fail(SCALDPC_ERR_ARG, "internal: compaction target too small");
    int rc;
    if ((rc = write_outputs(h, cur, cur_tiles, a.done, ddec, dllr, st))) return rc;
    k_bp_compact_plan<<<1, kBpPlanThreads, 0, st>>>(a.valid, a.done, a.orig, cur_tiles, h->d_src, b.orig, b.valid, b.done,
                                                    b.unsat, h->d_nact);
    k_bp_compact_move<<<dim3((h->E + 63) / 64, new_tiles), 256, 0, st>>>(a.msg, b.msg, h->d_src, h->E);
    if (probs_t) {
        k_bp_compact_move<<<dim3((h->n + 63) / 64, new_tiles), 256, 0, st>>>(a.prior_t, b.prior_t, h->d_src, h->n);
        ++h->launches;
    }
    k_bp_compact_bits<<<dim3((h->m + 7) / 8, new_tiles), 256, 0, st>>>(a.synd_w, b.synd_w, h->d_src, h->m);
    h->launches += 3;
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

static int bp_finish_create(scaldpc_bp *h);

// Degree-class bounds of all graphs -> the handle's launch parameters; shared-memory budgets of the kernels whose scratch
// depends on the largest row / column degree (h->dc_max, h->dv_max).
static int bp_size_kernels(scaldpc_bp *h, const std::vector<int> &class_end_all)
{
    for (int c = 0; c <= kBpVnClasses; ++c) {
        h->class_end[c] = class_end_all[c];
        h->class_max[c] = 0;
    }
    for (int g = 0; g < h->G; ++g)
        for (int c = 0; c <= kBpVnClasses; ++c) {
            const int *ce = &class_end_all[(size_t)g * (kBpVnClasses + 1)];
            h->class_max[c] = std::max(h->class_max[c], ce[c] - (c ? ce[c - 1] : 0));
        }
    // check pass: per warp a double-buffered row stage + the prefix array, 3 x dc_max x 128 B
    const size_t per_warp = (size_t)3 * h->dc_max * 32 * sizeof(float);
    h->cn_warps = (int)std::min<size_t>(kBpCnWarpsMax, (216 * 1024) / per_warp);
    if (h->cn_warps < 1)
        return fail(SCALDPC_ERR_UNSUPPORTED, "row degree %d too large for the shared-memory staging", h->dc_max);
    h->smem_cn = per_warp * h->cn_warps;
    // shared-memory bit-pass kernel: heavy columns, and the whole pass of small (launch-bound) batches
    const size_t pw = (size_t)2 * h->dv_max * 32 * sizeof(float);
    h->vn_gen_wpc = (int)std::min<size_t>(kBpVnWarps, (216 * 1024) / pw);
    if (h->vn_gen_wpc < 1)
        return fail(SCALDPC_ERR_UNSUPPORTED, "column degree %d too large for the shared-memory scratch", h->dv_max);
    h->smem_vn = pw * h->vn_gen_wpc;
    return 0;
}

// G >= 1 graphs of shape m x n: row_ptr = G x (m+1) (every segment starts at 0), col = the G column arrays back to back.
static int bp_create_impl(int G, int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int device,
                          scaldpc_bp **out)
{
    if (!out) return fail(SCALDPC_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (max_iter < 0) return fail(SCALDPC_ERR_ARG, "max_iter must be >= 0");
    if (G < 1 || !row_ptr || !col) return fail(SCALDPC_ERR_ARG, "bad parity-check matrix");
    std::vector<BpHostGraph> gs(G);
    int Emax = 0, dc_max = 0, dv_max = 0;
    {
        size_t coff = 0;
        for (int g = 0; g < G; ++g) {
            std::string err;
            const int rc = build_bp_graph(m, n, row_ptr + (size_t)g * (m + 1), col + coff, gs[g], err);
            if (rc) return G > 1 ? fail(rc, "graph %d: %s", g, err.c_str()) : fail(rc, "%s", err.c_str());
            coff += (size_t)gs[g].E;
            Emax = std::max(Emax, gs[g].E);
            dc_max = std::max(dc_max, gs[g].dc_max);
            dv_max = std::max(dv_max, gs[g].dv_max);
        }
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(SCALDPC_ERR_CUDA, "CUDA device %d not available (%d visible) -- libscaldpc has no CPU path", device,
                    ndev);
    std::unique_ptr<scaldpc_bp, void (*)(scaldpc_bp *)> h(new scaldpc_bp, scaldpc_bp_destroy);
    h->m = m;
    h->n = n;
    h->E = Emax;
    h->G = G;
    h->dc_max = std::max(1, dc_max);
    h->dv_max = std::max(1, dv_max);
    h->max_iter = max_iter == 0 ? n : max_iter;  // upstream: max_iter = 0 means n
    h->device = device;
    std::vector<int> class_end_all((size_t)G * (kBpVnClasses + 1));
    for (int g = 0; g < G; ++g)
        for (int c = 0; c <= kBpVnClasses; ++c) class_end_all[(size_t)g * (kBpVnClasses + 1) + c] = gs[g].class_end[c];
    int rc;
    if ((rc = bp_size_kernels(h.get(), class_end_all))) return rc;
    SCALDPC_CUDA(cudaSetDevice(device));
    // the per-graph tables, concatenated with the fixed strides m+1 / Emax / n+1 / n (BpDev g_rp / g_e / g_cp / g_n);
    // a graph with fewer edges than Emax is padded: its padding edges belong to "row m" (never active) and are never
    // reached through row_ptr / col_ptr
    std::vector<int> a_row_ptr((size_t)G * (m + 1)), a_col((size_t)G * std::max(1, Emax), 0), a_col_ptr((size_t)G * (n + 1));
    std::vector<int> a_vn_order((size_t)G * n), a_vn_desc((size_t)G * 4 * std::max(1, n)), a_edge_row((size_t)G * std::max(1, Emax), m);
    std::vector<uint32_t> a_csc_off((size_t)G * std::max(1, Emax), 0u);
    std::vector<int> a_edge_csc((size_t)G * std::max(1, Emax), 0), a_csc_row((size_t)G * std::max(1, Emax), m);
    for (int g = 0; g < G; ++g) {
        const BpHostGraph &q = gs[g];
        std::copy(q.row_ptr.begin(), q.row_ptr.end(), a_row_ptr.begin() + (size_t)g * (m + 1));
        std::copy(q.col.begin(), q.col.end(), a_col.begin() + (size_t)g * Emax);
        std::copy(q.col_ptr.begin(), q.col_ptr.end(), a_col_ptr.begin() + (size_t)g * (n + 1));
        std::copy(q.vn_order.begin(), q.vn_order.end(), a_vn_order.begin() + (size_t)g * n);
        std::copy(q.vn_desc.begin(), q.vn_desc.begin() + (size_t)4 * n, a_vn_desc.begin() + (size_t)g * 4 * n);
        std::copy(q.csc_off.begin(), q.csc_off.begin() + q.E, a_csc_off.begin() + (size_t)g * Emax);
        for (int r = 0; r < m; ++r)
            for (int e = q.row_ptr[r]; e < q.row_ptr[r + 1]; ++e) a_edge_row[(size_t)g * Emax + e] = r;
        for (int i = 0; i < q.E; ++i) {
            const int e = q.csc_edge[i];
            a_edge_csc[(size_t)g * Emax + e] = i;
            a_csc_row[(size_t)g * Emax + i] = a_edge_row[(size_t)g * Emax + e];
        }
    }
    if ((rc = upload_vec(&h->d_row_ptr, a_row_ptr)) || (rc = upload_vec(&h->d_col, a_col)) ||
        (rc = upload_vec(&h->d_col_ptr, a_col_ptr)) || (rc = upload_vec(&h->d_csc_edge, a_csc_off)) ||
        (rc = upload_vec(&h->d_vn_order, a_vn_order)) || (rc = upload_vec(&h->d_vn_desc, a_vn_desc)) ||
        (rc = upload_vec(&h->d_edge_row, a_edge_row)) || (rc = upload_vec(&h->d_class_end, class_end_all)) ||
        (rc = upload_vec(&h->d_edge_csc, a_edge_csc)) || (rc = upload_vec(&h->d_csc_row, a_csc_row)))
        return rc;
    if ((rc = bp_finish_create(h.get()))) return rc;
    *out = h.release();
    return 0;
}

// Everything of a handle that does not depend on where its graph tables came from: work buffers, kernel attributes,
// occupancies.  The tables (d_row_ptr ... d_csc_row, d_class_end) are on the device, bp_size_kernels() has run.
static int bp_finish_create(scaldpc_bp *h)
{
    const int G = h->G, Emax = h->E, n = h->n, device = h->device;
    SCALDPC_CUDA(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_first_tab, (size_t)G * std::max(1, Emax) * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_csc_tab, (size_t)G * std::max(1, Emax) * sizeof(float)));
    SCALDPC_CUDA(cudaMemset(h->d_csc_tab, 0, (size_t)G * std::max(1, Emax) * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_prior, (size_t)n * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_nact, sizeof(int)));
    h->cap_counts = h->max_iter + 2;
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_counts, (size_t)h->cap_counts * sizeof(int)));
    SCALDPC_CUDA(cudaMallocHost((void **)&h->h_pin, 2 * sizeof(int)));
    SCALDPC_CUDA(cudaFuncSetAttribute(k_bp_cn<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_cn));
    SCALDPC_CUDA(cudaFuncSetAttribute(k_bp_cn<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_cn));
    if (h->smem_vn)
    {
        SCALDPC_CUDA(cudaFuncSetAttribute(k_bp_vn_generic<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_vn));
        SCALDPC_CUDA(cudaFuncSetAttribute(k_bp_vn_generic<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_vn));
    }
    if (const char *e = getenv("SCALDPC_BP_VN_SINGLE")) h->vn_single_tiles = atoi(e);
    if (const char *e = getenv("SCALDPC_BP_VN_SINGLE_ITEMS")) h->vn_single_items = atoll(e);
    if (const char *e = getenv("SCALDPC_BP_VN_PIPE")) h->vn_pipe = atoi(e) != 0;
    if (const char *e = getenv("SCALDPC_BP_EARLY_COPY")) h->early_copy_max = atoi(e);
    for (int c = 0; c < kBpVnClasses; ++c) {
        int occ_shared = 0;
        SCALDPC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->vn_occ[c], vn_class_fn(c), kBpVnWarps * 32, 0));
        SCALDPC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_shared, vn_class_fn(c, false, true), kBpVnWarps * 32, 0));
        h->vn_occ[c] = std::max(1, std::min(h->vn_occ[c], occ_shared));
        if (vn_pipe_fn(c)) {
            SCALDPC_CUDA(cudaFuncSetAttribute(vn_pipe_fn(c), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vn_pipe_smem(c)));
            SCALDPC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->vn_pipe_occ[c], vn_pipe_fn(c), kBpVnWarps * 32,
                                                                       vn_pipe_smem(c)));
            h->vn_pipe_occ[c] = std::max(1, h->vn_pipe_occ[c]);
        }
    }
    SCALDPC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->cn_occ[0], k_bp_cn<false>, h->cn_warps * 32, h->smem_cn));
    SCALDPC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->cn_occ[1], k_bp_cn<true>, h->cn_warps * 32, h->smem_cn));
    h->cn_occ[0] = std::max(1, h->cn_occ[0]);
    h->cn_occ[1] = std::max(1, h->cn_occ[1]);
    return 0;
}

int scaldpc_bp_create(int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int device,
                      scaldpc_bp **out)
{
    return bp_create_impl(1, m, n, row_ptr, col, max_iter, device, out);
}

int scaldpc_bp_create_multi(int n_graphs, int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int device,
                            scaldpc_bp **out)
{
    if (n_graphs < 1) return fail(SCALDPC_ERR_ARG, "n_graphs must be >= 1");
    return bp_create_impl(n_graphs, m, n, row_ptr, col, max_iter, device, out);
}

void scaldpc_bp_destroy(scaldpc_bp *h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    free_ws(h);
    free_stage(h);
    cudaFree(h->d_truth);
    cudaFree(h->d_stats);
    cudaFree(h->d_nrows);
    cudaFree(h->d_row_ptr);
    cudaFree(h->d_col);
    cudaFree(h->d_col_ptr);
    cudaFree(h->d_csc_edge);
    cudaFree(h->d_vn_order);
    cudaFree(h->d_vn_desc);
    cudaFree(h->d_edge_row);
    cudaFree(h->d_class_end);
    cudaFree(h->d_tile_graph);
    cudaFree(h->d_slot_orig);
    cudaFree(h->d_first_tab);
    cudaFree(h->d_csc_tab);
    cudaFree(h->d_csc_row);
    cudaFree(h->d_edge_csc);
    cudaFree(h->d_prior);
    cudaFree(h->d_nact);
    cudaFree(h->d_counts);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    if (h->s_copy) cudaStreamDestroy(h->s_copy);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    delete h;
}

int scaldpc_bp_info(const scaldpc_bp *h, int *m, int *n, int *edges, int *max_iter)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (m) *m = h->m;
    if (n) *n = h->n;
    if (edges) *edges = h->E;
    if (max_iter) *max_iter = h->max_iter;
    return 0;
}

long long scaldpc_bp_launch_count(const scaldpc_bp *h) { return h ? h->launches : 0; }

// Scoring request of scaldpc_bp_decode_scored (hqc.py:709-758).
struct BpScore {
    const int32_t *truth;  // T x weight: support of the true message per trial
    int weight, n_msg;
    int32_t *stats;        // T x 6
};

static int bp_decode_impl(scaldpc_bp *h, int T, const uint8_t *synd, const double *channel_probs, uint8_t *decoding,
                          float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream,
                          const BpScore *score, const int32_t *n_checks, const int32_t *graph_of = nullptr)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (T < 0) return fail(SCALDPC_ERR_ARG, "T < 0");
    if (T == 0) return 0;
    const bool dev_ptrs = flags & SCALDPC_DEVICE_PTRS;
    if (!synd || !channel_probs || (!decoding && (dev_ptrs || !score)))
        return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    if (score) {
        if (!score->truth || !score->stats || score->weight < 0) return fail(SCALDPC_ERR_ARG, "NULL truth/stats pointer");
        if (score->n_msg < 0 || score->n_msg + h->m != h->n)
            return fail(SCALDPC_ERR_ARG, "scoring needs H = [Hin | I]: n_msg + m must equal n (hqc.py:680)");
    }
    const bool probs_t = flags & SCALDPC_BP_PROBS_PER_TRIAL;
    const bool early = !(flags & SCALDPC_BP_NO_EARLY_STOP);
    const bool multi = h->G > 1;
    // segmented CSR: a tile's lanes must share a graph, so trials of different graphs are never packed together
    const bool compacting = early && !(flags & SCALDPC_BP_NO_COMPACT) && !multi;
    const bool want_post = llr != nullptr;
    cudaStream_t st = (cudaStream_t)stream;

    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    int tiles = (T + 31) / 32;
    int rc;
    std::vector<int> slot_orig, tile_graph;  // segmented CSR: every graph's trials get tiles of their own
    if (multi) {
        if (!graph_of) return fail(SCALDPC_ERR_ARG, "this handle holds %d graphs: decode with scaldpc_bp_decode_multi", h->G);
        std::vector<int> cnt(h->G, 0), first(h->G, 0);
        for (int t = 0; t < T; ++t) {
            if (graph_of[t] < 0 || graph_of[t] >= h->G)
                return fail(SCALDPC_ERR_ARG, "graph_of[%d] = %d is outside 0 .. %d", t, graph_of[t], h->G - 1);
            ++cnt[graph_of[t]];
        }
        tiles = 0;
        for (int g = 0; g < h->G; ++g) {
            first[g] = tiles;
            tiles += (cnt[g] + 31) / 32;
        }
        tile_graph.resize(std::max(1, tiles));
        for (int g = 0; g < h->G; ++g)
            for (int k = 0; k < (cnt[g] + 31) / 32; ++k) tile_graph[first[g] + k] = g;
        slot_orig.assign((size_t)std::max(1, tiles) * 32, -1);
        std::fill(cnt.begin(), cnt.end(), 0);
        for (int t = 0; t < T; ++t) {
            const int g = graph_of[t];
            slot_orig[(size_t)first[g] * 32 + cnt[g]++] = t;  // trial order preserved inside a graph
        }
        if ((size_t)tiles > h->cap_tile_graph) {
            cudaFree(h->d_tile_graph);
            cudaFree(h->d_slot_orig);
            h->d_tile_graph = h->d_slot_orig = nullptr;
            h->cap_tile_graph = 0;
            SCALDPC_CUDA(cudaMalloc((void **)&h->d_tile_graph, (size_t)tiles * sizeof(int)));
            SCALDPC_CUDA(cudaMalloc((void **)&h->d_slot_orig, (size_t)tiles * 32 * sizeof(int)));
            h->cap_tile_graph = (size_t)tiles;
        }
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_tile_graph, tile_graph.data(), (size_t)tiles * sizeof(int), cudaMemcpyHostToDevice, st));
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_slot_orig, slot_orig.data(), (size_t)tiles * 32 * sizeof(int), cudaMemcpyHostToDevice, st));
        h->cur_tile_graph = h->d_tile_graph;
    }
    if ((rc = ensure_ws(h, tiles, want_post, probs_t))) return rc;
    struct NrowsGuard {  // the kernels see the per-trial row limits only while this decode runs
        scaldpc_bp *h;
        ~NrowsGuard() { h->cur_nrows = nullptr; }
    } nrows_guard{h};
    h->cur_nrows = nullptr;
    if (n_checks) {
        if (!dev_ptrs)
            for (int t = 0; t < T; ++t)
                if (n_checks[t] < 0 || n_checks[t] > h->m)
                    return fail(SCALDPC_ERR_ARG, "n_checks[%d] = %d is outside 0 .. m = %d", t, n_checks[t], h->m);
        if ((size_t)T > h->cap_nrows) {
            cudaFree(h->d_nrows);
            h->d_nrows = nullptr;
            h->cap_nrows = 0;
            SCALDPC_CUDA(cudaMalloc((void **)&h->d_nrows, (size_t)T * sizeof(int32_t)));
            h->cap_nrows = (size_t)T;
        }
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_nrows, n_checks, (size_t)T * sizeof(int32_t),
                                     dev_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
        h->cur_nrows = h->d_nrows;
    }

    const uint8_t *dsynd = synd;
    const double *dprobs = channel_probs;
    uint8_t *ddec = decoding;
    float *dllr = llr;
    if (!dev_ptrs) {
        if ((rc = ensure_stage(h, (size_t)T, probs_t, want_post))) return rc;
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_synd, synd, (size_t)T * h->m, cudaMemcpyHostToDevice, st));
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_probs, channel_probs, (size_t)(probs_t ? T : 1) * h->n * sizeof(double),
                                     cudaMemcpyHostToDevice, st));
        dsynd = h->d_synd;
        dprobs = h->d_probs;
        ddec = h->d_dec;
        dllr = want_post ? h->d_llr : nullptr;
    }

    // ---- inputs -> tile layout (always into set 0)
    BpSet &s0 = h->set[0];
    const int tb = (tiles + 255) / 256;
    const int *slot_map = nullptr;  // identity unless the trials were grouped by graph
    if (multi) {
        SCALDPC_CUDA(cudaMemcpyAsync(s0.orig, h->d_slot_orig, (size_t)tiles * 32 * sizeof(int), cudaMemcpyDeviceToDevice, st));
        k_bp_reset_slots<<<tb, 256, 0, st>>>(s0.valid, s0.done, s0.unsat, s0.orig, h->d_iters, h->d_conv, tiles, h->max_iter);
        slot_map = s0.orig;
    } else {
        k_bp_reset<<<tb, 256, 0, st>>>(s0.valid, s0.done, s0.unsat, s0.orig, h->d_iters, h->d_conv, tiles, T, h->max_iter);
    }
    if (probs_t)
        k_bp_prior_tiled<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(dprobs, s0.prior_t, T, h->n, slot_map);
    else
        k_bp_prior_shared<<<(h->n + 255) / 256, 256, 0, st>>>(dprobs, h->d_prior, h->n);
    k_bp_pack_synd<<<dim3((h->m + 31) / 32, tiles), 256, 0, st>>>(dsynd, s0.synd_w, T, h->m, slot_map);
    SCALDPC_CUDA(cudaMemsetAsync(h->d_dec_w, 0, (size_t)tiles * h->n * sizeof(uint32_t), st));
    SCALDPC_CUDA(cudaMemsetAsync(h->d_counts, 0, (size_t)h->cap_counts * sizeof(int), st));
    h->launches += 3;
    SCALDPC_CUDA(cudaGetLastError());

    // ---- iterations.  The host reads the number of running trials (one 4-byte D2H + sync) after the first
    // iterations, then every 4th (every iteration while the batch is large enough to hide the sync); it stops
    // launching at zero and compacts the running trials when they fill at most half of the current slots.
    int cur = 0, cur_tiles = tiles;
    h->last_compactions = 0;
    h->last_early_copy_rows = -1;
    bool early_copy = false, early_copy_ok = false;
    std::vector<int> tail_orig;
    if (!dev_ptrs && decoding && compacting && !score && h->early_copy_max > 0 && (long long)T * h->n >= (8 << 20)) {
        // only into page-locked memory: a copy into pageable memory blocks the host thread that must keep launching
        cudaPointerAttributes pa;
        bool pinned = cudaPointerGetAttributes(&pa, decoding) == cudaSuccess && pa.type == cudaMemoryTypeHost;
        if (pinned && llr) pinned = cudaPointerGetAttributes(&pa, llr) == cudaSuccess && pa.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (pinned) {
            if (!h->s_copy) {
                SCALDPC_CUDA(cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking));
                SCALDPC_CUDA(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
                SCALDPC_CUDA(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
            }
            early_copy_ok = true;
        }
    }
    int it = 1;
    for (; it <= h->max_iter; ++it) {
        const bool last = it == h->max_iter;
        const BpDev p = view(h, cur, probs_t, want_post);
        bool vn_first = false;
        if (it == 1 && !probs_t && !getenv("SCALDPC_BP_NO_FIRST_TABLE")) {
            // shared priors: the first check-to-bit messages differ between trials only in sign (bp_kernels.cuh): one
            // table per graph; the first bit pass forms +-tab[e] itself (k_bp_vn<.., FIRST>), so iteration 1 writes the
            // message array once and never reads it.  SCALDPC_BP_FIRST_FILL=1: round 1's way (k_bp_first_fill streams
            // the messages out, the ordinary bit pass reads them back) -- kept for the bit-identity test.
            k_bp_first_table<<<dim3((h->m + 127) / 128, h->G), 128, 0, st>>>(p, h->d_first_tab);
            ++h->launches;
            if (getenv("SCALDPC_BP_FIRST_FILL")) {
                k_bp_first_fill<<<dim3((h->E + 63) / 64, cur_tiles), 256, 0, st>>>(p, h->d_first_tab, h->d_edge_row, cur_tiles);
                ++h->launches;
            } else {
                vn_first = true;
            }
            SCALDPC_CUDA(cudaGetLastError());
        } else if ((rc = launch_cn(h, p, cur_tiles, it == 1, st))) {
            return rc;
        }
        if ((rc = launch_vn(h, p, cur_tiles, st, vn_first))) return rc;
        k_bp_syndrome<<<dim3((h->m + 255) / 256, cur_tiles), 256, 0, st>>>(p);
        k_bp_finish_iter<<<(cur_tiles + 255) / 256, 256, 0, st>>>(p, cur_tiles, it, early, last);
        h->launches += 2;
        SCALDPC_CUDA(cudaGetLastError());
        // polling = one 4-byte D2H + a stream sync (the GPU idles for the round trip, ~30 us): every iteration while most
        // trials converge (the first four) or while the batch is large enough to hide it, then every 4th / 8th --
        // kernels launched for tiles that have retired meanwhile exit at once
        // a batch of a few tiles (the reference-shaped single decode, hqc.py:694-708) is bound by exactly these round trips: it
        // polls after iterations 3 and 5 instead of 1, 2, 3, 4 -- the launches of a converged tile exit at once, a sync costs more
        const bool poll = cur_tiles <= 4 ? (it == 3 || it == 5 || (it >= 8 && (it <= 16 ? it % 4 == 0 : it % 8 == 0)))
                                         : (it <= 4 || cur_tiles >= 512 || (it <= 16 ? it % 4 == 0 : it % 8 == 0));
        if (early && !last && poll) {
            // the count comes home through a store into page-locked host memory, not through the copy engine: a 4-byte
            // copy would queue behind the early copy of the finished rows (below) and stall the stragglers for its duration
            k_bp_publish<<<1, 1, 0, st>>>(h->d_counts + it, h->h_pin);
            ++h->launches;
            SCALDPC_CUDA(cudaStreamSynchronize(st));
            const int n_run = *(volatile int *)h->h_pin;
            if (n_run == 0) break;
            if (compacting && cur_tiles >= 2 && (long long)n_run * 2 <= (long long)cur_tiles * 32) {
                if ((rc = compact(h, cur, cur_tiles, n_run, probs_t, ddec, dllr, st))) return rc;
                cur ^= 1;
                cur_tiles = (n_run + 31) / 32;
                ++h->last_compactions;
                if (early_copy_ok && !early_copy && n_run <= h->early_copy_max) {
                    // Every row of d_dec / d_llr but those of the n_run stragglers is final (compact() wrote the finished
                    // trials out): send the whole arrays home on a side stream while the stragglers iterate -- the tail
                    // lasts tens of iterations on a few tiles, the copy of T x n bytes a few ms -- and send the stragglers'
                    // rows again at the end.
                    tail_orig.resize((size_t)cur_tiles * 32);
                    SCALDPC_CUDA(cudaMemcpyAsync(tail_orig.data(), h->set[cur].orig, tail_orig.size() * sizeof(int),
                                                 cudaMemcpyDeviceToHost, st));
                    SCALDPC_CUDA(cudaEventRecord(h->ev_fork, st));
                    SCALDPC_CUDA(cudaStreamSynchronize(st));
                    SCALDPC_CUDA(cudaStreamWaitEvent(h->s_copy, h->ev_fork, 0));
                    SCALDPC_CUDA(cudaMemcpyAsync(decoding, h->d_dec, (size_t)T * h->n, cudaMemcpyDeviceToHost, h->s_copy));
                    if (llr) SCALDPC_CUDA(cudaMemcpyAsync(llr, h->d_llr, (size_t)T * h->n * sizeof(float), cudaMemcpyDeviceToHost, h->s_copy));
                    SCALDPC_CUDA(cudaEventRecord(h->ev_join, h->s_copy));
                    early_copy = true;
                    h->last_early_copy_rows = n_run;
                }
            }
        }
    }
    h->last_iters_run = std::min(it, h->max_iter);

    // ---- outputs of every trial still resident (finished since the last compaction, or never converged)
    if ((rc = write_outputs(h, cur, cur_tiles, h->set[cur].valid, ddec, dllr, st))) return rc;
    const cudaMemcpyKind kind = dev_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    if (score) {
        const int32_t *dtruth = score->truth;
        int32_t *dstats = score->stats;
        const size_t nt = std::max<size_t>(1, (size_t)T * score->weight);
        if (!dev_ptrs) {
            if (nt > h->cap_truth || (size_t)T > h->cap_stats_T) {
                cudaFree(h->d_truth);
                cudaFree(h->d_stats);
                h->d_truth = h->d_stats = nullptr;
                h->cap_truth = h->cap_stats_T = 0;
                SCALDPC_CUDA(cudaMalloc((void **)&h->d_truth, nt * sizeof(int32_t)));
                SCALDPC_CUDA(cudaMalloc((void **)&h->d_stats, (size_t)T * 6 * sizeof(int32_t)));
                h->cap_truth = nt;
                h->cap_stats_T = (size_t)T;
            }
            SCALDPC_CUDA(cudaMemcpyAsync(h->d_truth, score->truth, (size_t)T * score->weight * sizeof(int32_t),
                                         cudaMemcpyHostToDevice, st));
            dtruth = h->d_truth;
            dstats = h->d_stats;
        }
        k_bp_score<<<(T + 7) / 8, 256, 0, st>>>(ddec, dsynd, dtruth, h->cur_nrows, dstats, T, h->n, score->n_msg, score->weight);
        ++h->launches;
        SCALDPC_CUDA(cudaGetLastError());
        if (!dev_ptrs)
            SCALDPC_CUDA(cudaMemcpyAsync(score->stats, h->d_stats, (size_t)T * 6 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    }
    if (!dev_ptrs && early_copy) {
        SCALDPC_CUDA(cudaStreamWaitEvent(st, h->ev_join, 0));  // the stragglers' rows are rewritten AFTER the early copy
        for (int o : tail_orig) {
            if (o < 0) continue;
            SCALDPC_CUDA(cudaMemcpyAsync(decoding + (size_t)o * h->n, h->d_dec + (size_t)o * h->n, (size_t)h->n, kind, st));
            if (llr)
                SCALDPC_CUDA(cudaMemcpyAsync(llr + (size_t)o * h->n, h->d_llr + (size_t)o * h->n, (size_t)h->n * sizeof(float), kind, st));
        }
    } else if (!dev_ptrs) {
        if (decoding) SCALDPC_CUDA(cudaMemcpyAsync(decoding, h->d_dec, (size_t)T * h->n, kind, st));
        if (llr) SCALDPC_CUDA(cudaMemcpyAsync(llr, h->d_llr, (size_t)T * h->n * sizeof(float), kind, st));
    }
    if (iters) SCALDPC_CUDA(cudaMemcpyAsync(iters, h->d_iters, (size_t)T * sizeof(int), kind, st));
    if (converged) SCALDPC_CUDA(cudaMemcpyAsync(converged, h->d_conv, (size_t)T * sizeof(int), kind, st));
    if (!dev_ptrs) SCALDPC_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int scaldpc_bp_decode(scaldpc_bp *h, int T, const uint8_t *synd, const double *channel_probs, uint8_t *decoding,
                      float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream)
{
    return bp_decode_impl(h, T, synd, channel_probs, decoding, llr, iters, converged, flags, stream, nullptr, nullptr);
}

int scaldpc_bp_decode_nested(scaldpc_bp *h, int T, const int32_t *n_checks, const uint8_t *synd, const double *channel_probs,
                             uint8_t *decoding, float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream)
{
    if (!n_checks) return fail(SCALDPC_ERR_ARG, "n_checks is NULL");
    return bp_decode_impl(h, T, synd, channel_probs, decoding, llr, iters, converged, flags, stream, nullptr, n_checks);
}

int scaldpc_bp_decode_scored(scaldpc_bp *h, int T, int n_msg, const int32_t *n_checks, const uint8_t *synd,
                             const double *channel_probs, const int32_t *truth, int weight, int32_t *stats, uint8_t *decoding,
                             int32_t *iters, int32_t *converged, unsigned flags, void *stream)
{
    const BpScore score{truth, weight, n_msg, stats};
    return bp_decode_impl(h, T, synd, channel_probs, decoding, nullptr, iters, converged, flags, stream, &score, n_checks);
}

int scaldpc_bp_create_circulant(int n_graphs, int n_msg, int weight, int m, const int32_t *first_col, const int32_t *rows,
                                int max_iter, int device, scaldpc_bp **out)
{
    if (!out) return fail(SCALDPC_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (max_iter < 0) return fail(SCALDPC_ERR_ARG, "max_iter must be >= 0");
    if (n_graphs < 1 || n_msg < 1 || m < 1 || weight < 1 || weight > 128 || weight > n_msg || !first_col || !rows)
        return fail(SCALDPC_ERR_ARG, "bad circulant description (1 <= weight <= 128)");
    const long long E = (long long)m * (weight + 1);
    if (E * 128 > 0xffffffffLL) return fail(SCALDPC_ERR_UNSUPPORTED, "too many edges");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(SCALDPC_ERR_CUDA, "CUDA device %d not available (%d visible) -- libscaldpc has no CPU path", device,
                    ndev);
    std::unique_ptr<scaldpc_bp, void (*)(scaldpc_bp *)> h(new scaldpc_bp, scaldpc_bp_destroy);
    const int G = n_graphs, n = n_msg + m;
    h->m = m;
    h->n = n;
    h->E = (int)E;
    h->G = G;
    h->dc_max = weight + 1;
    h->max_iter = max_iter == 0 ? n : max_iter;
    h->device = device;
    SCALDPC_CUDA(cudaSetDevice(device));
    // every table of bp_create_impl, allocated empty and filled by the kernels of bp_build.cuh
    const size_t ge = (size_t)G * E;
    int *d_first = nullptr, *d_rows = nullptr, *d_cursor = nullptr, *d_err = nullptr, *d_csc_edge = nullptr, *d_bound = nullptr;
    struct Scratch {
        int **p[6];
        ~Scratch() { for (auto q : p) if (q && *q) cudaFree(*q); }
    } scratch{{&d_first, &d_rows, &d_cursor, &d_err, &d_csc_edge, &d_bound}};
    SCALDPC_CUDA(cudaMalloc((void **)&d_first, (size_t)G * weight * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&d_rows, (size_t)G * m * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&d_cursor, (size_t)G * n * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&d_err, 2 * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&d_csc_edge, ge * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&d_bound, kBpVnClasses * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_row_ptr, (size_t)G * (m + 1) * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_col, ge * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_col_ptr, (size_t)G * (n + 1) * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_csc_edge, ge * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_vn_order, (size_t)G * n * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_vn_desc, (size_t)G * 4 * n * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_edge_row, ge * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_edge_csc, ge * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_csc_row, ge * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_class_end, (size_t)G * (kBpVnClasses + 1) * sizeof(int)));
    SCALDPC_CUDA(cudaMemcpy(d_first, first_col, (size_t)G * weight * sizeof(int), cudaMemcpyHostToDevice));
    SCALDPC_CUDA(cudaMemcpy(d_rows, rows, (size_t)G * m * sizeof(int), cudaMemcpyHostToDevice));
    SCALDPC_CUDA(cudaMemcpy(d_bound, kBpVnClass, kBpVnClasses * sizeof(int), cudaMemcpyHostToDevice));
    SCALDPC_CUDA(cudaMemset(d_cursor, 0, (size_t)G * n * sizeof(int)));
    SCALDPC_CUDA(cudaMemset(d_err, 0, 2 * sizeof(int)));
    CircDev d;
    d.G = G; d.N = n_msg; d.W = weight; d.m = m; d.n = n; d.E = (int)E;
    d.first_col = d_first; d.rows = d_rows;
    d.row_ptr = h->d_row_ptr; d.col = h->d_col; d.col_ptr = h->d_col_ptr; d.vn_order = h->d_vn_order; d.vn_desc = h->d_vn_desc;
    d.edge_row = h->d_edge_row; d.edge_csc = h->d_edge_csc; d.csc_row = h->d_csc_row; d.class_end = h->d_class_end;
    d.csc_off = h->d_csc_edge; d.cursor = d_cursor; d.err = d_err;
    k_circ_sort_support<<<(G + 63) / 64, 64>>>(d);
    k_circ_rows<<<(unsigned)((ge + 255) / 256), 256>>>(d);
    k_circ_scan<<<G, 1024>>>(d);
    k_circ_fill<<<(unsigned)((ge + 255) / 256), 256>>>(d, d_csc_edge);
    k_circ_order<<<(unsigned)(((size_t)G * n + 255) / 256), 256>>>(d, d_csc_edge);
    int herr[2] = {0, 0};
    SCALDPC_CUDA(cudaMemcpy(herr, d_err, sizeof herr, cudaMemcpyDeviceToHost));  // also the first synchronisation point
    if (herr[0]) return fail(SCALDPC_ERR_ARG, "first_col needs %d distinct positions in 0 .. n_msg-1 per graph, rows positions in 0 .. n_msg-1", weight);
    h->dv_max = std::max(1, herr[1]);
    k_circ_degsort<<<G, 1024>>>(d, h->dv_max, d_bound, kBpVnClasses);
    h->launches += 6;
    SCALDPC_CUDA(cudaGetLastError());
    std::vector<int> class_end_all((size_t)G * (kBpVnClasses + 1));
    SCALDPC_CUDA(cudaMemcpy(class_end_all.data(), h->d_class_end, class_end_all.size() * sizeof(int), cudaMemcpyDeviceToHost));
    int rc;
    if ((rc = bp_size_kernels(h.get(), class_end_all)) || (rc = bp_finish_create(h.get()))) return rc;
    *out = h.release();
    return 0;
}

int scaldpc_bp_decode_multi(scaldpc_bp *h, int T, const int32_t *graph_of, const int32_t *n_checks, const uint8_t *synd,
                            const double *channel_probs, const int32_t *truth, int weight, int32_t *stats, uint8_t *decoding,
                            float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream)
{
    if (!graph_of) return fail(SCALDPC_ERR_ARG, "graph_of is NULL");
    if (h && h->G == 1)
        for (int t = 0; t < T; ++t)
            if (graph_of[t] != 0) return fail(SCALDPC_ERR_ARG, "graph_of[%d] = %d, but the handle holds one graph", t, graph_of[t]);
    if (stats) {
        const BpScore score{truth, weight, h ? h->n - h->m : 0, stats};
        return bp_decode_impl(h, T, synd, channel_probs, decoding, llr, iters, converged, flags, stream, &score, n_checks, graph_of);
    }
    return bp_decode_impl(h, T, synd, channel_probs, decoding, llr, iters, converged, flags, stream, nullptr, n_checks, graph_of);
}

int scaldpc_bp_last_run(const scaldpc_bp *h, int *iterations_launched, int *compactions)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (iterations_launched) *iterations_launched = h->last_iters_run;
    if (compactions) *compactions = h->last_compactions;
    return 0;
}

int scaldpc_bp_last_early_copy(const scaldpc_bp *h, int *rows)
{
    if (!h || !rows) return fail(SCALDPC_ERR_ARG, "NULL argument");
    *rows = h->last_early_copy_rows;
    return 0;
}

int scaldpc_bp_time_kernel(scaldpc_bp *h, int which, int T, unsigned flags, int reps, void *stream, float *avg_ms)
{
    if (!h || !avg_ms || reps <= 0 || T <= 0) return fail(SCALDPC_ERR_ARG, "bad argument");
    if (h->G > 1) return fail(SCALDPC_ERR_UNSUPPORTED, "the timing probe covers single-graph handles");
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    const int tiles = (T + 31) / 32;
    if (tiles > h->cap_tiles) return fail(SCALDPC_ERR_ARG, "time_kernel needs a previous decode of >= T trials");
    cudaStream_t st = (cudaStream_t)stream;
    const bool probs_t = flags & SCALDPC_BP_PROBS_PER_TRIAL;
    if (probs_t && !h->cap_prior_t) return fail(SCALDPC_ERR_ARG, "no per-trial priors resident");
    // every slot of set 0 runs again: reset the masks and rebuild an iteration-1 message state from the priors
    // (the syndrome words are whatever the previous decode left there; the kernels' work does not depend on them)
    BpSet &s0 = h->set[0];
    k_bp_reset<<<(tiles + 255) / 256, 256, 0, st>>>(s0.valid, s0.done, s0.unsat, s0.orig, h->d_iters, h->d_conv, tiles, T,
                                                    h->max_iter);
    const BpDev p = view(h, 0, probs_t, false);
    int rc;
    if ((rc = launch_cn(h, p, tiles, true, st)) || (rc = launch_vn(h, p, tiles, st))) return rc;
    cudaEvent_t e0, e1;
    SCALDPC_CUDA(cudaEventCreate(&e0));
    SCALDPC_CUDA(cudaEventCreate(&e1));
    for (int i = -1; i < reps && !rc; ++i) {
        if (i == 0) SCALDPC_CUDA(cudaEventRecord(e0, st));
        rc = which == 0 ? launch_cn(h, p, tiles, false, st) : launch_vn(h, p, tiles, st);
    }
    SCALDPC_CUDA(cudaEventRecord(e1, st));
    SCALDPC_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    SCALDPC_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *avg_ms = ms / reps;
    return rc;
}

}  // extern "C"

// -----------------------------------------------------------------------------------------------------------
// scaldpc_bp_decode_static: the decode of scaldpc_bp_decode for callers that must not allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every per-decode buffer (messages, syndrome / decision bit-planes, slot masks, iteration counters) is carved
// out of ONE file-scope device array of kBpStaticBytes; the graph tables and the [n] prior / first-pass tables belong to the
// handle and were allocated when it was created.  Every GPU that loads the library holds the array until exit.
// In host code `g_bp_static` is a host-side stand-in, not a device address: bp_static_base() resolves the address once per
// device with cudaGetSymbolAddress (that device current) and only that pointer is handed to kernels.
// All calls (all handles, all streams) share the array and it keeps the last call's contents: whatever must start at zero is
// cleared here, and two calls that overlap on the GPU race.
// -----------------------------------------------------------------------------------------------------------
constexpr size_t kBpStaticBytes = (size_t)SCALDPC_BP_STATIC_BYTES;
static __device__ __align__(256) unsigned char g_bp_static[kBpStaticBytes];

namespace {
constexpr int kBpStaticMaxDevices = 64;
std::mutex g_bp_static_mu;
unsigned char *g_bp_static_ptr[kBpStaticMaxDevices] = {nullptr};

// The device address of g_bp_static on `device` (which must be current).
int bp_static_base(int device, unsigned char **base)
{
    if (device < 0 || device >= kBpStaticMaxDevices) return fail(SCALDPC_ERR_ARG, "device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_bp_static_mu);
    if (!g_bp_static_ptr[device]) {
        void *ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_bp_static);
        if (e != cudaSuccess || !ptr)
            return fail(SCALDPC_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device, cudaGetErrorString(e));
        g_bp_static_ptr[device] = (unsigned char *)ptr;
    }
    *base = g_bp_static_ptr[device];
    return 0;
}

// Carves `bytes` (rounded up to 256) from the scratch; `used` runs ahead even when base is null (size query).
template <class T>
T *bp_static_take(unsigned char *base, size_t &used, size_t count)
{
    T *p = base ? (T *)(base + used) : nullptr;
    used += (count * sizeof(T) + 255) & ~(size_t)255;
    return p;
}
}  // namespace

extern "C" {

static size_t bp_static_bytes_needed(const scaldpc_bp *h, int T)  // not exported
{
    if (!h || T <= 0) return 0;
    const size_t tiles = ((size_t)T + 31) / 32;
    size_t used = 0;
    bp_static_take<float>(nullptr, used, tiles * 32 * std::max(1, h->E));  // messages
    bp_static_take<uint32_t>(nullptr, used, tiles * h->m);                  // syndrome bit-planes
    bp_static_take<uint32_t>(nullptr, used, tiles * h->n);                  // decision bit-planes
    for (int i = 0; i < 3; ++i) bp_static_take<uint32_t>(nullptr, used, tiles);  // valid / done / unsat
    for (int i = 0; i < 3; ++i) bp_static_take<int>(nullptr, used, tiles * 32);  // slot -> trial, iterations, converged
    return used;
}

int scaldpc_bp_decode_static(scaldpc_bp *h, int T, const uint8_t *synd, const double *channel_probs, uint8_t *decoding,
                             int32_t *iters, int32_t *converged, void *stream)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (T < 0) return fail(SCALDPC_ERR_ARG, "T < 0");
    if (T == 0) return 0;
    if (!synd || !channel_probs || !decoding) return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    if (h->G > 1) return fail(SCALDPC_ERR_UNSUPPORTED, "scaldpc_bp_decode_static covers single-graph handles");
    const size_t need = bp_static_bytes_needed(h, T);
    if (need > kBpStaticBytes)
        return fail(SCALDPC_ERR_ARG,
                    "scaldpc_bp_decode_static: %d trials on this graph need %zu bytes of scratch, the static array holds %zu "
                    "(SCALDPC_BP_STATIC_BYTES) -- use scaldpc_bp_decode",
                    T, need, kBpStaticBytes);
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    unsigned char *base = nullptr;
    int rc;
    if ((rc = bp_static_base(h->device, &base))) return rc;

    const int tiles = (T + 31) / 32;
    size_t used = 0;
    float *msg = bp_static_take<float>(base, used, (size_t)tiles * 32 * std::max(1, h->E));
    uint32_t *synd_w = bp_static_take<uint32_t>(base, used, (size_t)tiles * h->m);
    uint32_t *dec_w = bp_static_take<uint32_t>(base, used, (size_t)tiles * h->n);
    uint32_t *valid = bp_static_take<uint32_t>(base, used, tiles);
    uint32_t *done = bp_static_take<uint32_t>(base, used, tiles);
    uint32_t *unsat = bp_static_take<uint32_t>(base, used, tiles);
    int *orig = bp_static_take<int>(base, used, (size_t)tiles * 32);
    int *d_iters = bp_static_take<int>(base, used, (size_t)tiles * 32);
    int *d_conv = bp_static_take<int>(base, used, (size_t)tiles * 32);

    // the handle's view with every per-decode buffer replaced by its slice of the static array
    struct NrowsGuard {
        scaldpc_bp *h;
        const int32_t *saved;
        ~NrowsGuard() { h->cur_nrows = saved; }
    } guard{h, h->cur_nrows};
    h->cur_nrows = nullptr;
    BpDev p = view(h, 0, false, false);
    p.msg = msg;
    p.prior = h->d_prior;
    p.synd_w = synd_w;
    p.dec_w = dec_w;
    p.post = nullptr;
    p.done = done;
    p.valid = valid;
    p.unsat = unsat;
    p.orig = orig;
    p.iters = d_iters;
    p.conv = d_conv;

    // ---- inputs -> tile layout; the array keeps the last call's contents: masks, counters and decision words start clean
    const int tb = (tiles + 255) / 256;
    k_bp_reset<<<tb, 256, 0, st>>>(valid, done, unsat, orig, d_iters, d_conv, tiles, T, h->max_iter);
    k_bp_prior_shared<<<(h->n + 255) / 256, 256, 0, st>>>(channel_probs, h->d_prior, h->n);
    k_bp_pack_synd<<<dim3((h->m + 31) / 32, tiles), 256, 0, st>>>(synd, synd_w, T, h->m, nullptr);
    SCALDPC_CUDA(cudaMemsetAsync(dec_w, 0, (size_t)tiles * h->n * sizeof(uint32_t), st));
    SCALDPC_CUDA(cudaMemsetAsync(h->d_counts, 0, (size_t)h->cap_counts * sizeof(int), st));
    h->launches += 3;
    SCALDPC_CUDA(cudaGetLastError());

    // ---- iterations: converged trials freeze on the device (k_bp_finish_iter), finished tiles exit at the top of every
    // kernel; the host never reads anything back, so the call is a fixed sequence of launches
    for (int it = 1; it <= h->max_iter; ++it) {
        bool vn_first = false;
        if (it == 1) {
            k_bp_first_table<<<dim3((h->m + 127) / 128, h->G), 128, 0, st>>>(p, h->d_first_tab);
            ++h->launches;
            vn_first = true;
            SCALDPC_CUDA(cudaGetLastError());
        } else if ((rc = launch_cn(h, p, tiles, false, st))) {
            return rc;
        }
        if ((rc = launch_vn(h, p, tiles, st, vn_first))) return rc;
        k_bp_syndrome<<<dim3((h->m + 255) / 256, tiles), 256, 0, st>>>(p);
        k_bp_finish_iter<<<(tiles + 255) / 256, 256, 0, st>>>(p, tiles, it, 1, it == h->max_iter);
        h->launches += 2;
        SCALDPC_CUDA(cudaGetLastError());
    }
    h->last_iters_run = h->max_iter;
    h->last_compactions = 0;

    k_bp_unpack_dec<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(dec_w, decoding, orig, valid, h->n);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    if (iters) SCALDPC_CUDA(cudaMemcpyAsync(iters, d_iters, (size_t)T * sizeof(int), cudaMemcpyDeviceToDevice, st));
    if (converged) SCALDPC_CUDA(cudaMemcpyAsync(converged, d_conv, (size_t)T * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return 0;
}

}  // extern "C"
