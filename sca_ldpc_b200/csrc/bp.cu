// libscaldpc -- binary product-sum BP decoder: graph upload, workspace, decode pipeline and C ABI
// (include/scaldpc.h).  Replaces ldpc.bp_decoder at the call sites simulate/hqc.py:694-708 and
// simulate/decode.py:155-171.  No CPU compute path exists here.
#include <algorithm>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "bp_graph.h"
#include "bp_kernels.cuh"
#include "common.h"

using namespace scaldpc;

struct scaldpc_bp {
    int m = 0, n = 0, E = 0, dc_max = 0, dv_max = 0;
    int max_iter = 0, device = 0;
    int *d_row_ptr = nullptr, *d_col = nullptr, *d_col_ptr = nullptr, *d_csc_edge = nullptr;
    // workspace
    int cap_tiles = 0;
    bool cap_post = false, cap_prior_t = false;
    float *d_msg = nullptr, *d_prior = nullptr, *d_prior_t = nullptr, *d_post = nullptr;
    uint32_t *d_synd_w = nullptr, *d_dec_w = nullptr, *d_done = nullptr, *d_valid = nullptr, *d_unsat = nullptr;
    int *d_iters = nullptr, *d_conv = nullptr, *d_n_active = nullptr;
    int *h_n_active = nullptr;  // pinned
    // staging for host-pointer calls
    size_t cap_stage_T = 0;
    bool stage_probs_t = false, stage_llr = false;
    uint8_t *d_synd = nullptr, *d_dec = nullptr;
    double *d_probs = nullptr;
    float *d_llr = nullptr;
    size_t smem_cn = 0, smem_vn = 0;
    std::mutex mu;
    long long launches = 0;
    int last_iters_run = 0;
};

namespace {

template <typename T>
int upload_vec(T **dst, const std::vector<T> &src)
{
    SCALDPC_CUDA(cudaMalloc((void **)dst, std::max<size_t>(1, src.size()) * sizeof(T)));
    if (!src.empty()) SCALDPC_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return 0;
}

void free_ws(scaldpc_bp *h)
{
    cudaFree(h->d_msg);
    cudaFree(h->d_prior_t);
    cudaFree(h->d_post);
    cudaFree(h->d_synd_w);
    cudaFree(h->d_dec_w);
    cudaFree(h->d_done);
    cudaFree(h->d_valid);
    cudaFree(h->d_unsat);
    cudaFree(h->d_iters);
    cudaFree(h->d_conv);
    h->d_msg = h->d_prior_t = h->d_post = nullptr;
    h->d_synd_w = h->d_dec_w = h->d_done = h->d_valid = h->d_unsat = nullptr;
    h->d_iters = h->d_conv = nullptr;
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

int ensure_ws(scaldpc_bp *h, int tiles, bool post, bool prior_t)
{
    if (tiles <= h->cap_tiles && (!post || h->cap_post) && (!prior_t || h->cap_prior_t)) return 0;
    const int t = std::max(tiles, h->cap_tiles);
    post = post || h->cap_post;
    prior_t = prior_t || h->cap_prior_t;
    free_ws(h);
    const size_t lanes = (size_t)t * 32;
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_msg, std::max<size_t>(1, lanes * h->E) * sizeof(float)));
    if (prior_t) SCALDPC_CUDA(cudaMalloc((void **)&h->d_prior_t, lanes * h->n * sizeof(float)));
    if (post) SCALDPC_CUDA(cudaMalloc((void **)&h->d_post, lanes * h->n * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_synd_w, (size_t)t * h->m * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_dec_w, (size_t)t * h->n * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_done, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_valid, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_unsat, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_iters, lanes * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_conv, lanes * sizeof(int)));
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

BpDev view(const scaldpc_bp *h, bool per_trial_prior, bool want_post)
{
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
    p.csc_edge = h->d_csc_edge;
    p.msg = h->d_msg;
    p.prior = per_trial_prior ? h->d_prior_t : h->d_prior;
    p.synd_w = h->d_synd_w;
    p.dec_w = h->d_dec_w;
    p.post = want_post ? h->d_post : nullptr;
    p.done = h->d_done;
    p.valid = h->d_valid;
    p.unsat = h->d_unsat;
    p.iters = h->d_iters;
    p.conv = h->d_conv;
    p.n_active = h->d_n_active;
    return p;
}

int launch_cn(scaldpc_bp *h, const BpDev &p, int tiles, bool first, cudaStream_t st)
{
    dim3 grid((h->m + kBpWarps - 1) / kBpWarps, tiles);
    k_bp_cn<<<grid, kBpWarps * 32, h->smem_cn, st>>>(p, first);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

int launch_vn(scaldpc_bp *h, const BpDev &p, int tiles, cudaStream_t st)
{
    dim3 grid((h->n + kBpWarps - 1) / kBpWarps, tiles);
    k_bp_vn<<<grid, kBpWarps * 32, h->smem_vn, st>>>(p);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int scaldpc_bp_create(int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int device,
                      scaldpc_bp **out)
{
    if (!out) return fail(SCALDPC_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (max_iter < 0) return fail(SCALDPC_ERR_ARG, "max_iter must be >= 0");
    BpHostGraph g;
    {
        std::string err;
        const int rc = build_bp_graph(m, n, row_ptr, col, g, err);
        if (rc) return fail(rc, "%s", err.c_str());
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(SCALDPC_ERR_CUDA, "CUDA device %d not available (%d visible) -- libscaldpc has no CPU path", device,
                    ndev);
    std::unique_ptr<scaldpc_bp, void (*)(scaldpc_bp *)> h(new scaldpc_bp, scaldpc_bp_destroy);
    h->m = m;
    h->n = n;
    h->E = g.E;
    h->dc_max = std::max(1, g.dc_max);
    h->dv_max = std::max(1, g.dv_max);
    h->max_iter = max_iter == 0 ? n : max_iter;  // upstream: max_iter = 0 means n
    h->device = device;
    h->smem_cn = (size_t)kBpWarps * 2 * h->dc_max * 32 * sizeof(float);
    h->smem_vn = (size_t)kBpWarps * 2 * h->dv_max * 32 * sizeof(float);
    if (h->smem_cn > 227 * 1024 || h->smem_vn > 227 * 1024)
        return fail(SCALDPC_ERR_UNSUPPORTED, "node degree too large for the shared-memory scratch (dc=%d, dv=%d)",
                    h->dc_max, h->dv_max);
    SCALDPC_CUDA(cudaSetDevice(device));
    int rc;
    if ((rc = upload_vec(&h->d_row_ptr, g.row_ptr)) || (rc = upload_vec(&h->d_col, g.col)) ||
        (rc = upload_vec(&h->d_col_ptr, g.col_ptr)) || (rc = upload_vec(&h->d_csc_edge, g.csc_edge)))
        return rc;
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_prior, (size_t)n * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_n_active, sizeof(int)));
    SCALDPC_CUDA(cudaMallocHost((void **)&h->h_n_active, sizeof(int)));
    SCALDPC_CUDA(cudaFuncSetAttribute(k_bp_cn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_cn));
    SCALDPC_CUDA(cudaFuncSetAttribute(k_bp_vn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_vn));
    *out = h.release();
    return 0;
}

void scaldpc_bp_destroy(scaldpc_bp *h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    free_ws(h);
    free_stage(h);
    cudaFree(h->d_row_ptr);
    cudaFree(h->d_col);
    cudaFree(h->d_col_ptr);
    cudaFree(h->d_csc_edge);
    cudaFree(h->d_prior);
    cudaFree(h->d_n_active);
    if (h->h_n_active) cudaFreeHost(h->h_n_active);
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

int scaldpc_bp_decode(scaldpc_bp *h, int T, const uint8_t *synd, const double *channel_probs, uint8_t *decoding,
                      float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (T < 0) return fail(SCALDPC_ERR_ARG, "T < 0");
    if (T == 0) return 0;
    if (!synd || !channel_probs || !decoding) return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    const bool dev_ptrs = flags & SCALDPC_DEVICE_PTRS;
    const bool probs_t = flags & SCALDPC_BP_PROBS_PER_TRIAL;
    const bool early = !(flags & SCALDPC_BP_NO_EARLY_STOP);
    const bool want_post = llr != nullptr;
    cudaStream_t st = (cudaStream_t)stream;

    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    const int tiles = (T + 31) / 32;
    int rc;
    if ((rc = ensure_ws(h, tiles, want_post, probs_t))) return rc;

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

    // ---- inputs -> tile layout
    const int tb = (tiles + 255) / 256;
    k_bp_reset<<<tb, 256, 0, st>>>(h->d_valid, h->d_done, h->d_unsat, h->d_iters, h->d_conv, tiles, T, h->max_iter);
    if (probs_t)
        k_bp_prior_tiled<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(dprobs, h->d_prior_t, T, h->n);
    else
        k_bp_prior_shared<<<(h->n + 255) / 256, 256, 0, st>>>(dprobs, h->d_prior, h->n);
    k_bp_pack_synd<<<dim3((h->m + 31) / 32, tiles), 256, 0, st>>>(dsynd, h->d_synd_w, T, h->m);
    SCALDPC_CUDA(cudaMemsetAsync(h->d_dec_w, 0, (size_t)tiles * h->n * sizeof(uint32_t), st));
    h->launches += 3;
    SCALDPC_CUDA(cudaGetLastError());

    // ---- iterations
    const BpDev p = view(h, probs_t, want_post);
    // The host looks at the number of running tiles every `poll` iterations (one small D2H + sync) and stops
    // launching when it reaches zero; in between, retired tiles exit at the top of every kernel.
    const int poll = 4;
    int it = 1;
    for (; it <= h->max_iter; ++it) {
        const bool last = it == h->max_iter;
        if ((rc = launch_cn(h, p, tiles, it == 1, st))) return rc;
        if ((rc = launch_vn(h, p, tiles, st))) return rc;
        k_bp_syndrome<<<dim3((h->m + 255) / 256, tiles), 256, 0, st>>>(p);
        SCALDPC_CUDA(cudaMemsetAsync(h->d_n_active, 0, sizeof(int), st));
        k_bp_finish_iter<<<tb, 256, 0, st>>>(p, tiles, it, early, last);
        h->launches += 2;
        SCALDPC_CUDA(cudaGetLastError());
        if (early && !last && (it % poll == 0 || it <= 3)) {
            SCALDPC_CUDA(cudaMemcpyAsync(h->h_n_active, h->d_n_active, sizeof(int), cudaMemcpyDeviceToHost, st));
            SCALDPC_CUDA(cudaStreamSynchronize(st));
            if (*h->h_n_active == 0) break;
        }
    }
    h->last_iters_run = std::min(it, h->max_iter);

    // ---- outputs
    k_bp_unpack_dec<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(h->d_dec_w, ddec, T, h->n);
    ++h->launches;
    if (dllr) {
        k_bp_untile_f32<<<dim3((h->n + 31) / 32, tiles), 256, 0, st>>>(h->d_post, dllr, T, h->n);
        ++h->launches;
    }
    SCALDPC_CUDA(cudaGetLastError());
    const cudaMemcpyKind kind = dev_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    if (!dev_ptrs) {
        SCALDPC_CUDA(cudaMemcpyAsync(decoding, h->d_dec, (size_t)T * h->n, kind, st));
        if (llr) SCALDPC_CUDA(cudaMemcpyAsync(llr, h->d_llr, (size_t)T * h->n * sizeof(float), kind, st));
    }
    if (iters) SCALDPC_CUDA(cudaMemcpyAsync(iters, h->d_iters, (size_t)T * sizeof(int), kind, st));
    if (converged) SCALDPC_CUDA(cudaMemcpyAsync(converged, h->d_conv, (size_t)T * sizeof(int), kind, st));
    if (!dev_ptrs) SCALDPC_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int scaldpc_bp_time_kernel(scaldpc_bp *h, int which, int T, unsigned flags, int reps, void *stream, float *avg_ms)
{
    if (!h || !avg_ms || reps <= 0 || T <= 0) return fail(SCALDPC_ERR_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    const int tiles = (T + 31) / 32;
    if (tiles > h->cap_tiles) return fail(SCALDPC_ERR_ARG, "time_kernel needs a previous decode of >= T trials");
    cudaStream_t st = (cudaStream_t)stream;
    const bool probs_t = flags & SCALDPC_BP_PROBS_PER_TRIAL;
    if (probs_t && !h->cap_prior_t) return fail(SCALDPC_ERR_ARG, "no per-trial priors resident");
    const BpDev p = view(h, probs_t, false);
    // every tile must run: clear the done masks (the message state of a finished decode is a valid BP state)
    SCALDPC_CUDA(cudaMemsetAsync(h->d_done, 0, (size_t)tiles * sizeof(uint32_t), st));
    cudaEvent_t e0, e1;
    SCALDPC_CUDA(cudaEventCreate(&e0));
    SCALDPC_CUDA(cudaEventCreate(&e1));
    int rc = 0;
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
