// libscaldpc -- q-ary min-sum decoder: graph build, workspace, decode pipeline and the C ABI
// (include/scaldpc.h).  Replaces the reference's Rust decoder core behind the same boundary
// (simulate_rs/src/pydecoder.rs, decoder.rs, decoder_special.rs).  No CPU compute path exists here.
#include <algorithm>
#include <condition_variable>
#include <deque>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <tuple>
#include <vector>

#include "common.h"
#include "qary_graph.h"
#include "qary_kernels.cuh"
#include "qary_sim.cuh"

namespace scaldpc {

std::string &last_error()
{
    static thread_local std::string s;
    return s;
}

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
    return code;
}

}  // namespace scaldpc

using namespace scaldpc;

struct scaldpc_qary {
    int R = 0, N = 0, bvars = 0, B = 0, BSUM = 0, Q = 0, QS = 0, E = 0;
    int max_iter = 1, device = 0;
    int S = 0, SL = 0, SLS = 0;
    int lmax = 0;      // max row degree
    int wmax = 0;      // max DP table width over the checks
    int num_sms = 148;
    int fast_sw = 0;   // >0: every row is SW B-variables + 1 sum-variable with BSUM == SW*B and a template exists
    bool uniform_ok = false;   // generic decoder with a k_cn_uniform<Q> instantiation whose tables fit shared memory
    int max_resident = 65536;  // trials decoded per pass of the device-pointer path (workspace bound: ~0.1-0.2 MB each)
    int vn_reg_d = 0;  // >0: every B-variable has exactly this degree and a register-resident k_vn_reg exists
    // device graph
    int *d_row_ptr = nullptr, *d_col = nullptr, *d_edge_off = nullptr, *d_col_ptr = nullptr;
    int8_t *d_sign = nullptr;
    int2 *d_csc = nullptr;
    // workspace (grown on demand)
    int cap_tiles = 0;
    bool cap_post = false;
    float *d_msg = nullptr, *d_llr = nullptr, *d_llr_sum = nullptr, *d_post = nullptr, *d_post_sum = nullptr;
    int8_t *d_hard_t = nullptr;
    int *d_status = nullptr;   // 2 x lanes: per-trial status, then scratch for the re-decode pass
    // at-risk detection + exact-order re-decode of the fast path
    uint32_t *d_risk = nullptr;   // [tiles] flag words
    int *d_risk_list = nullptr;   // [tiles*32] flagged trial indices, ascending
    int *d_scan_out = nullptr;    // {count, worst status, list offset} of the last scan; [4..5] = call-wide {count, worst}
    int *h_scan_out = nullptr;    // pinned mirror (6 ints)
    int *d_glist = nullptr;       // flagged trials of the whole call (batch-wide indices), re-decoded once at the end
    size_t cap_glist = 0;
    // host pipeline: priors of the flagged trials, put aside before their chunk's staging buffer is reused
    size_t cap_retry = 0;         // rows
    float *d_retry_in = nullptr, *d_retry_in_sum = nullptr;
    uint16_t *d_retry_obs = nullptr, *d_retry_obs_sum = nullptr;
    float risk_eps = 1e-2f;
    long long last_trials = 0, last_redecoded = 0;  // of the last decode / simulate call
    long long last_work = 0;      // trial-iterations launched by the last call (restarts and re-decodes included)
    // early termination (SCALDPC_QARY_EARLY_STOP)
    uint32_t *d_done = nullptr;   // [tiles] frozen-trial words
    int *d_iters = nullptr;       // [tiles*32] last iteration of every slot of the current round
    int *d_iters_all = nullptr;   // [cap_iters_T] per-trial iteration counts of the last call
    size_t cap_iters_T = 0;
    int es_restart_min = 256;     // batches (chunks) at least this large restart their stragglers in a compacted second round
    // staging for host-pointer calls: the batch is cut into chunks and pipelined (H2D of chunk i+1 and D2H of chunk
    // i-1 overlap the decode of chunk i), so every staging buffer exists twice
    size_t cap_stage_T = 0;  // trials per output staging buffer
    size_t cap_in_T = 0;     // trials per prior-vector input staging buffer
    size_t cap_status_T = 0;
    float *d_in[2] = {nullptr, nullptr}, *d_in_sum[2] = {nullptr, nullptr};
    float *d_opost[2] = {nullptr, nullptr}, *d_opost_sum[2] = {nullptr, nullptr};
    int8_t *d_ohard[2] = {nullptr, nullptr};
    int *d_status_all = nullptr;
    // observation-indexed priors (scaldpc_qary_decode_observed): index staging + LLR tables
    size_t cap_obs_T = 0, cap_tab = 0, cap_tab_sum = 0;
    uint16_t *d_obs[2] = {nullptr, nullptr}, *d_obs_sum[2] = {nullptr, nullptr};
    float *d_tab_in = nullptr, *d_tab_sum_in = nullptr, *d_tab_llr = nullptr, *d_tab_sum_llr = nullptr;
    int *d_tab_st = nullptr, *d_tab_sum_st = nullptr;
    // on-device trial synthesis (scaldpc_qary_simulate): per-pass trial data + code tables + counters
    size_t cap_sim_T = 0;
    int8_t *d_sim_secret = nullptr, *d_sim_hard = nullptr;
    int16_t *d_sim_ssum = nullptr;
    uint16_t *d_sim_code = nullptr, *d_sim_code_sum = nullptr;
    unsigned long long *d_sim_cnt = nullptr;
    // reference-shaped single calls (host buffers, T <= 32 = one tile): the whole decode -- copies in, ~25 launches,
    // copies out -- is ONE CUDA graph per (T, iterations, flags, posteriors wanted), replayed on an internal stream
    std::map<std::tuple<int, int, unsigned, int>, std::pair<cudaGraphExec_t, int>> small_graphs;  // -> (graph, launches)
    cudaStream_t s_small = nullptr;
    float *h_pin_in = nullptr, *h_pin_in_sum = nullptr, *h_pin_post = nullptr, *h_pin_post_sum = nullptr;
    int8_t *h_pin_hard = nullptr;
    int *h_pin_status = nullptr;
    bool small_graph_off = false;  // SCALDPC_NO_GRAPH=1
    int exact_split_items = 0;     // k_cn_exact: items up to which the split variant runs (SCALDPC_EXACT_SPLIT_ITEMS)
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
    std::mutex mu;
    // scaldpc_qary_decode_one: concurrent single decodes are combined into one batch (leader / followers)
    struct OneReq {
        const float *in, *in_sum;
        int8_t *hard;
        int status = 0;   // per-trial status, or the batch's error code
        bool done = false;
    };
    std::mutex one_mu;
    std::condition_variable one_cv;
    std::deque<OneReq *> one_q;
    bool one_leading = false;
    std::vector<float> one_in, one_in_sum;
    std::vector<int8_t> one_hard;
    long long one_batches = 0, one_requests = 0;
    long long launches = 0;
};

namespace {

template <typename T>
int upload(T **dst, const std::vector<T> &src)
{
    SCALDPC_CUDA(cudaMalloc((void **)dst, std::max<size_t>(1, src.size()) * sizeof(T)));
    if (!src.empty()) SCALDPC_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return 0;
}

void free_workspace(scaldpc_qary *h)
{
    cudaFree(h->d_msg);
    cudaFree(h->d_llr);
    cudaFree(h->d_llr_sum);
    cudaFree(h->d_post);
    cudaFree(h->d_post_sum);
    cudaFree(h->d_hard_t);
    cudaFree(h->d_status);
    cudaFree(h->d_risk);
    cudaFree(h->d_risk_list);
    cudaFree(h->d_done);
    cudaFree(h->d_iters);
    h->d_done = nullptr;
    h->d_iters = nullptr;
    h->d_msg = h->d_llr = h->d_llr_sum = h->d_post = h->d_post_sum = nullptr;
    h->d_hard_t = nullptr;
    h->d_status = nullptr;
    h->d_risk = nullptr;
    h->d_risk_list = nullptr;
    h->cap_tiles = 0;
    h->cap_post = false;
}

void free_stage(scaldpc_qary *h)
{
    for (int i = 0; i < 2; ++i) {
        cudaFree(h->d_in[i]);
        cudaFree(h->d_in_sum[i]);
        cudaFree(h->d_opost[i]);
        cudaFree(h->d_opost_sum[i]);
        cudaFree(h->d_ohard[i]);
        h->d_in[i] = h->d_in_sum[i] = h->d_opost[i] = h->d_opost_sum[i] = nullptr;
        h->d_ohard[i] = nullptr;
    }
    h->cap_stage_T = 0;
    h->cap_in_T = 0;
}

void free_observed(scaldpc_qary *h)
{
    for (int i = 0; i < 2; ++i) {
        cudaFree(h->d_obs[i]);
        cudaFree(h->d_obs_sum[i]);
        h->d_obs[i] = h->d_obs_sum[i] = nullptr;
    }
    cudaFree(h->d_tab_in);
    cudaFree(h->d_tab_sum_in);
    cudaFree(h->d_tab_llr);
    cudaFree(h->d_tab_sum_llr);
    cudaFree(h->d_tab_st);
    cudaFree(h->d_tab_sum_st);
    h->d_tab_in = h->d_tab_sum_in = h->d_tab_llr = h->d_tab_sum_llr = nullptr;
    h->d_tab_st = h->d_tab_sum_st = nullptr;
    h->cap_obs_T = h->cap_tab = h->cap_tab_sum = 0;
}

void free_sim(scaldpc_qary *h)
{
    cudaFree(h->d_sim_secret);
    cudaFree(h->d_sim_hard);
    cudaFree(h->d_sim_ssum);
    cudaFree(h->d_sim_code);
    cudaFree(h->d_sim_code_sum);
    cudaFree(h->d_sim_cnt);
    h->d_sim_secret = h->d_sim_hard = nullptr;
    h->d_sim_ssum = nullptr;
    h->d_sim_code = h->d_sim_code_sum = nullptr;
    h->d_sim_cnt = nullptr;
    h->cap_sim_T = 0;
}

int ensure_sim(scaldpc_qary *h, size_t chunk)
{
    if (!h->d_sim_cnt) {
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_sim_cnt, 8 * sizeof(unsigned long long)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_sim_code, (size_t)h->Q * sizeof(uint16_t)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_sim_code_sum, (size_t)h->QS * sizeof(uint16_t)));
    }
    if (chunk <= h->cap_sim_T) return 0;
    cudaFree(h->d_sim_secret);
    cudaFree(h->d_sim_hard);
    cudaFree(h->d_sim_ssum);
    h->d_sim_secret = h->d_sim_hard = nullptr;
    h->d_sim_ssum = nullptr;
    h->cap_sim_T = 0;
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_sim_secret, chunk * h->bvars));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_sim_hard, chunk * h->N));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_sim_ssum, chunk * h->R * sizeof(int16_t)));
    h->cap_sim_T = chunk;
    return 0;
}

int ensure_workspace(scaldpc_qary *h, int tiles, bool want_post)
{
    if (tiles <= h->cap_tiles && (!want_post || h->cap_post)) return 0;
    const int t = std::max(tiles, h->cap_tiles);
    const bool post = want_post || h->cap_post;
    free_workspace(h);
    const size_t lanes = (size_t)t * 32;
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_msg, lanes * h->S * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_llr, lanes * h->SL * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_llr_sum, std::max<size_t>(1, lanes * h->SLS) * sizeof(float)));
    if (post) {
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_post, lanes * h->SL * sizeof(float)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_post_sum, std::max<size_t>(1, lanes * h->SLS) * sizeof(float)));
    }
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_hard_t, lanes * h->N));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_status, 2 * lanes * sizeof(int)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_risk, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_risk_list, 2 * lanes * sizeof(int)));  // two lists (rounds of early termination)
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_done, (size_t)t * sizeof(uint32_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_iters, lanes * sizeof(int)));
    if (!h->d_scan_out) {
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_scan_out, 8 * sizeof(int)));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_scan_out, 8 * sizeof(int), cudaHostAllocDefault));
    }
    h->cap_tiles = t;
    h->cap_post = post;
    return 0;
}

// Streams / events of the chunked host pipeline, the whole-batch status array and the per-chunk OUTPUT staging.
int ensure_pipeline(scaldpc_qary *h, size_t chunk, size_t T, bool want_post)
{
    if (!h->s_in) {
        SCALDPC_CUDA(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
        SCALDPC_CUDA(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            SCALDPC_CUDA(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
            SCALDPC_CUDA(cudaEventCreateWithFlags(&h->ev_comp[i], cudaEventDisableTiming));
            SCALDPC_CUDA(cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming));
        }
    }
    if (T > h->cap_status_T) {
        cudaFree(h->d_status_all);
        h->d_status_all = nullptr;
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_status_all, T * sizeof(int)));
        h->cap_status_T = T;
    }
    if (chunk <= h->cap_stage_T && (!want_post || h->d_opost[0])) return 0;
    const size_t t = std::max(chunk, h->cap_stage_T);
    const bool post = want_post || h->d_opost[0];
    for (int i = 0; i < 2; ++i) {
        cudaFree(h->d_opost[i]);
        cudaFree(h->d_opost_sum[i]);
        cudaFree(h->d_ohard[i]);
        h->d_opost[i] = h->d_opost_sum[i] = nullptr;
        h->d_ohard[i] = nullptr;
    }
    h->cap_stage_T = 0;
    for (int i = 0; i < 2; ++i) {
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_ohard[i], t * h->N));
        if (post) {
            SCALDPC_CUDA(cudaMalloc((void **)&h->d_opost[i], t * h->SL * sizeof(float)));
            SCALDPC_CUDA(cudaMalloc((void **)&h->d_opost_sum[i], std::max<size_t>(1, t * h->SLS) * sizeof(float)));
        }
    }
    h->cap_stage_T = t;
    return 0;
}

// Per-chunk INPUT staging of prior vectors (scaldpc_qary_decode with host pointers).
int ensure_stage_priors(scaldpc_qary *h, size_t chunk)
{
    if (chunk <= h->cap_in_T) return 0;
    for (int i = 0; i < 2; ++i) {
        cudaFree(h->d_in[i]);
        cudaFree(h->d_in_sum[i]);
        h->d_in[i] = h->d_in_sum[i] = nullptr;
    }
    h->cap_in_T = 0;
    for (int i = 0; i < 2; ++i) {
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_in[i], chunk * h->SL * sizeof(float)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_in_sum[i], std::max<size_t>(1, chunk * h->SLS) * sizeof(float)));
    }
    h->cap_in_T = chunk;
    return 0;
}

// Per-chunk INPUT staging of observation indices (scaldpc_qary_decode_observed with host pointers).
int ensure_stage_observed(scaldpc_qary *h, size_t chunk)
{
    if (chunk <= h->cap_obs_T) return 0;
    for (int i = 0; i < 2; ++i) {
        cudaFree(h->d_obs[i]);
        cudaFree(h->d_obs_sum[i]);
        h->d_obs[i] = h->d_obs_sum[i] = nullptr;
    }
    h->cap_obs_T = 0;
    for (int i = 0; i < 2; ++i) {
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_obs[i], chunk * h->bvars * sizeof(uint16_t)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_obs_sum[i], std::max<size_t>(1, chunk * h->R) * sizeof(uint16_t)));
    }
    h->cap_obs_T = chunk;
    return 0;
}

// Host pipeline: room for the priors of `rows` flagged trials (see finish_verify).
int ensure_retry(scaldpc_qary *h, size_t rows, bool observed)
{
    const bool have = observed ? h->d_retry_obs != nullptr : h->d_retry_in != nullptr;
    if (rows <= h->cap_retry && have) return 0;
    const size_t r = std::max(rows, h->cap_retry);
    cudaFree(h->d_retry_in);
    cudaFree(h->d_retry_in_sum);
    cudaFree(h->d_retry_obs);
    cudaFree(h->d_retry_obs_sum);
    h->d_retry_in = h->d_retry_in_sum = nullptr;
    h->d_retry_obs = h->d_retry_obs_sum = nullptr;
    h->cap_retry = 0;
    // both kinds are (re)allocated together so that cap_retry describes every buffer that exists
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_retry_in, r * h->SL * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_retry_in_sum, std::max<size_t>(1, r * h->SLS) * sizeof(float)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_retry_obs, r * h->bvars * sizeof(uint16_t)));
    SCALDPC_CUDA(cudaMalloc((void **)&h->d_retry_obs_sum, std::max<size_t>(1, r * h->R) * sizeof(uint16_t)));
    h->cap_retry = r;
    return 0;
}

// Device copies of the prior tables and their LLR form.
int ensure_tables(scaldpc_qary *h, size_t n_rows, size_t n_rows_sum)
{
    if (n_rows > h->cap_tab) {
        cudaFree(h->d_tab_in);
        cudaFree(h->d_tab_llr);
        cudaFree(h->d_tab_st);
        h->d_tab_in = h->d_tab_llr = nullptr;
        h->d_tab_st = nullptr;
        h->cap_tab = 0;
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_tab_in, n_rows * h->Q * sizeof(float)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_tab_llr, n_rows * h->Q * sizeof(float)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_tab_st, n_rows * sizeof(int)));
        h->cap_tab = n_rows;
    }
    if (n_rows_sum > h->cap_tab_sum) {
        cudaFree(h->d_tab_sum_in);
        cudaFree(h->d_tab_sum_llr);
        cudaFree(h->d_tab_sum_st);
        h->d_tab_sum_in = h->d_tab_sum_llr = nullptr;
        h->d_tab_sum_st = nullptr;
        h->cap_tab_sum = 0;
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_tab_sum_in, n_rows_sum * h->QS * sizeof(float)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_tab_sum_llr, n_rows_sum * h->QS * sizeof(float)));
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_tab_sum_st, n_rows_sum * sizeof(int)));
        h->cap_tab_sum = n_rows_sum;
    }
    return 0;
}

QaryDev dev_view(const scaldpc_qary *h, bool want_post)
{
    QaryDev p;
    p.R = h->R;
    p.N = h->N;
    p.bvars = h->bvars;
    p.B = h->B;
    p.BSUM = h->BSUM;
    p.Q = h->Q;
    p.QS = h->QS;
    p.S = h->S;
    p.SL = h->SL;
    p.SLS = h->SLS;
    p.row_ptr = h->d_row_ptr;
    p.col = h->d_col;
    p.sign = h->d_sign;
    p.edge_off = h->d_edge_off;
    p.col_ptr = h->d_col_ptr;
    p.csc = h->d_csc;
    p.msg = h->d_msg;
    p.llr = h->d_llr;
    p.llr_sum = h->d_llr_sum;
    p.post = want_post ? h->d_post : nullptr;
    p.post_sum = want_post ? h->d_post_sum : nullptr;
    p.hard_t = h->d_hard_t;
    p.status = h->d_status;
    return p;
}

// ---- kernel dispatch tables ------------------------------------------------------------------------------
typedef void (*cn_fast_fn)(const QaryDev, const int);
typedef void (*vn_fn)(const QaryDev, const int);

// The instantiations of the persistent check-node kernel for one shape, indexed by cn_mode_index().
struct CnFast {
    cn_fast_fn fn[8];
    size_t smem;
};
constexpr int kCnFastVariants = 8;

// Iteration `it` of `iters` -> kernel variant (qary_cn_fast.cuh MODE bits): the sum-variable's c2v is written only by
// the second-to-last iteration and read only by the last one.
int cn_mode_index(int it, int iters, bool every_iteration_decides)
{
    // early termination: every iteration emits decisions (LAST) and the sum-variable's message makes the full round trip
    // (WSUM now, RSUM next time) -- plain forward/backward flooding, oracle algo 2
    if (every_iteration_decides) return it == 1 ? 6 : 7;
    const bool first = it == 1, last = it == iters, penult = it == iters - 1;
    if (first) return last ? 1 : (penult ? 2 : 0);
    return last ? 5 : (penult ? 4 : 3);
}

template <int SW, int B>
const CnFast *cn_fast_table()
{
    static const CnFast t = {{k_cn_fast<SW, B, kCnFirst>, k_cn_fast<SW, B, kCnFirst | kCnLast>,
                              k_cn_fast<SW, B, kCnFirst | kCnWsum>, k_cn_fast<SW, B, 0>, k_cn_fast<SW, B, kCnWsum>,
                              k_cn_fast<SW, B, kCnLast | kCnRsum>, k_cn_fast<SW, B, kCnFirst | kCnLast | kCnWsum>,
                              k_cn_fast<SW, B, kCnLast | kCnRsum | kCnWsum>},
                             CnFastShape<SW, B>::smem_bytes};
    return &t;
}

const CnFast *pick_cn_fast(int sw, int b)
{
#define CASE(SW_, B_) \
    if (sw == SW_ && b == B_) return cn_fast_table<SW_, B_>();
    CASE(6, 2)  // Kyber-768, sum_weight 6  (lib.rs:54-75 DecoderN1024R256SW6 / DecoderN1280R512SW6)
    CASE(6, 3)  // Kyber-512, eta1 = 3      (BASELINE.json configs[1])
    CASE(4, 2)  // Kyber-1024, sum_weight 4
    CASE(8, 2)  // Kyber-1024, sum_weight 8
    CASE(3, 2)  // kyber.py:385-392 DecoderN1024R256SW3
#undef CASE
    return nullptr;
}

// The exact-order check-node kernel for the same shapes (qary_cn_exact.cuh).
typedef void (*cn_exact_fn)(const QaryDev, const int, const int, const int);
struct CnExact {
    cn_exact_fn fn;
    size_t smem;
    cn_exact_fn fn_split;  // one CTA of `nr` warps per item (few items: the re-decode of one or two tiles)
    int nr;
};

const CnExact *pick_cn_exact(int sw, int b)
{
#define CASE(SW_, B_)                                                                            \
    if (sw == SW_ && b == B_) {                                                                  \
        static const CnExact t = {k_cn_exact<SW_, B_, false>, CnExactShape<SW_, B_>::smem_bytes, \
                                  k_cn_exact<SW_, B_, true>, CnExactSplit<SW_>::NR};             \
        return &t;                                                                               \
    }
    CASE(6, 2) CASE(6, 3) CASE(4, 2) CASE(8, 2) CASE(3, 2)
#undef CASE
    return nullptr;
}

vn_fn pick_vn(int q)
{
    switch (q) {
#define CASE(Q_) \
    case Q_:     \
        return k_vn<Q_>;
        CASE(3) CASE(5) CASE(7) CASE(9) CASE(11) CASE(13) CASE(15) CASE(17) CASE(19) CASE(21) CASE(23) CASE(25)
        CASE(27) CASE(29) CASE(31)
#undef CASE
    default:
        return nullptr;
    }
}

vn_fn pick_vn_reg(int q, int d)
{
#define CASE(Q_, D_) \
    if (q == Q_ && d == D_) return k_vn_reg<Q_, D_>;
    CASE(5, 1) CASE(5, 2) CASE(5, 3) CASE(5, 4) CASE(7, 1) CASE(7, 2) CASE(7, 3) CASE(7, 4) CASE(7, 6) CASE(3, 3)
#undef CASE
    return nullptr;
}

typedef void (*cn_uniform_fn)(const QaryDev, const int, const int);
cn_uniform_fn pick_cn_uniform(int q)
{
    switch (q) {
#define CASE(Q_) \
    case Q_:     \
        return k_cn_uniform<Q_>;
        CASE(3) CASE(5) CASE(7) CASE(9) CASE(11) CASE(13) CASE(15)
#undef CASE
    default:
        return nullptr;
    }
}
size_t uniform_smem_bytes(const scaldpc_qary *h)
{
    return (size_t)cn_uniform_rows(h->lmax, h->Q, (h->wmax - 1) / 2) * 32 * sizeof(float);
}

size_t generic_smem_bytes(const scaldpc_qary *h, CnGenericArgs *g)
{
    g->QM = std::max(h->Q, h->QS);
    g->W = h->wmax;
    g->LMAX = h->lmax;
    const size_t floats = (size_t)g->LMAX * g->QM + (size_t)(g->LMAX + 1) * g->W + 2 * (size_t)g->W;
    return floats * 32 * sizeof(float);  // == generic_scratch_floats(graph) * 128
}

// One check-node launch / one variable-node launch over `tiles` tiles.
int launch_cn(scaldpc_qary *h, const QaryDev &p, int tiles, int it, int iters, unsigned flags, cudaStream_t st)
{
    const bool decide = flags & SCALDPC_QARY_EARLY_STOP;  // every iteration emits hard decisions
    const bool first = it == 1, last = it == iters || decide;
    const bool exact = flags & SCALDPC_CN_EXACT;
    const bool force_generic = flags & SCALDPC_CN_GENERIC;
    const CnFast *fast = (!exact && !force_generic && h->fast_sw) ? pick_cn_fast(h->fast_sw, h->B) : nullptr;
    if (fast) {
        // persistent: one CTA of kCnWarps warps per SM, warps stride over the (tile, check) items
        const int n_items = tiles * h->R;
        const int grid = std::max(1, std::min(h->num_sms, n_items));  // few items: a warp each on as many SMs as possible
        fast->fn[cn_mode_index(it, iters, decide)]<<<grid, kCnWarps * 32, fast->smem, st>>>(p, n_items);
        ++h->launches;
    } else if (exact && !force_generic && h->fast_sw) {
        // exact-order dynamic programme in registers, one warp per (tile, check)
        const CnExact *ex = pick_cn_exact(h->fast_sw, h->B);
        const int n_items = tiles * h->R;
        if (n_items <= h->exact_split_items)  // too few items to fill the GPU with a warp each: a CTA of nr warps per item
            ex->fn_split<<<n_items, ex->nr * 32, ex->smem, st>>>(p, n_items, first, last);
        else
            ex->fn<<<(n_items + kCnExactWarps - 1) / kCnExactWarps, kCnExactWarps * 32, ex->smem, st>>>(p, n_items, first, last);
        ++h->launches;
    } else if (h->BSUM == 0 && !exact && !force_generic && h->uniform_ok) {
        // generic decoder, one alphabet: the unrolled / padded forward-backward kernel
        if (first) {
            dim3 grid((h->E + 7) / 8, tiles);
            k_init_v2c<<<grid, 256, 0, st>>>(p, h->E);
            ++h->launches;
        }
        pick_cn_uniform(h->Q)<<<dim3(h->R, tiles), 32, uniform_smem_bytes(h), st>>>(p, h->lmax, (h->wmax - 1) / 2);
        ++h->launches;
    } else {
        if (first) {
            dim3 grid((h->E + 7) / 8, tiles);
            k_init_v2c<<<grid, 256, 0, st>>>(p, h->E);
            ++h->launches;
        }
        CnGenericArgs g;
        g.exact = exact;
        g.first = first;
        g.last = last;
        const size_t smem = generic_smem_bytes(h, &g);
        dim3 grid(h->R, tiles);
        k_cn_generic<<<grid, 32, smem, st>>>(p, g);
        ++h->launches;
    }
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

int launch_vn(scaldpc_qary *h, const QaryDev &p, int tiles, bool last, unsigned flags, cudaStream_t st)
{
    vn_fn vn = (h->vn_reg_d && !(flags & SCALDPC_CN_GENERIC)) ? pick_vn_reg(h->Q, h->vn_reg_d) : pick_vn(h->Q);
    dim3 grid((h->bvars + kVnWarps - 1) / kVnWarps, tiles);
    vn<<<grid, kVnWarps * 32, 0, st>>>(p, last);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

int load_priors(scaldpc_qary *h, const float *in, float *out, int T, int tiles, int rows, int q, bool is_llr,
                cudaStream_t st, int *status, const int *map)
{
    int rpb = std::max(1, 256 / q);
    // Few tiles (the reference-shaped single decode, the re-decode of the flagged trials): the kernel's time is the serial chain
    // of division + logarithm a thread walks through, rpb / 8 rows of q entries -- cut the rows into more, smaller CTAs until
    // the machine has a couple of CTAs per SM.  Large batches keep 256 / q rows per CTA (fewer, fuller CTAs).
    const long long want = 2LL * h->num_sms;
    if ((long long)tiles * ((rows + rpb - 1) / rpb) < want) {
        const int per_tile = (int)((want + tiles - 1) / tiles);
        rpb = std::max(1, std::min(rpb, (rows + per_tile - 1) / per_tile));
    }
    const int pitch = (rpb * q) | 1;
    dim3 grid((rows + rpb - 1) / rpb, tiles), block(32, kTrRows);
    k_load_priors<<<grid, block, (size_t)32 * pitch * sizeof(float), st>>>(in, out, T, rows, q, rpb, is_llr, status,
                                                                           map);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

// ============================================================================================================
// C ABI
// ============================================================================================================
extern "C" {

const char *scaldpc_last_error(void) { return scaldpc::last_error().c_str(); }
const char *scaldpc_version(void) { return "scaldpc 0.1.0 sm_100a"; }

int scaldpc_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return fail(SCALDPC_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return n;
}

static int qary_create_from_graph(const HostGraph &g, int R, int N, int b, int bsum, int max_iter, int device,
                                  scaldpc_qary **out);

int scaldpc_qary_create(const int8_t *H, int R, int N, int b, int bsum, int dv, int dc, int max_iter, int device,
                        scaldpc_qary **out)
{
    if (!out) return fail(SCALDPC_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (max_iter < 0) return fail(SCALDPC_ERR_ARG, "iterations must be >= 0");
    HostGraph g;
    std::string err;
    const int grc = build_graph(H, R, N, b, bsum, dv, dc, g, err);
    if (grc) return fail(grc, "%s", err.c_str());
    return qary_create_from_graph(g, R, N, b, bsum, max_iter, device, out);
}

int scaldpc_qary_create_csr(int R, int N, const int32_t *row_ptr, const int32_t *col, const int8_t *val, int b, int bsum, int dv,
                            int dc, int max_iter, int device, scaldpc_qary **out)
{
    if (!out) return fail(SCALDPC_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (max_iter < 0) return fail(SCALDPC_ERR_ARG, "iterations must be >= 0");
    HostGraph g;
    std::string err;
    const int grc = build_graph_csr(R, N, row_ptr, col, val, b, bsum, dv, dc, g, err);
    if (grc) return fail(grc, "%s", err.c_str());
    return qary_create_from_graph(g, R, N, b, bsum, max_iter, device, out);
}

}  // extern "C"

static int qary_create_from_graph(const HostGraph &g, int R, int N, int b, int bsum, int max_iter, int device,
                                  scaldpc_qary **out)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(SCALDPC_ERR_CUDA, "CUDA device %d not available (%d visible) -- libscaldpc has no CPU path", device,
                    ndev);

    std::unique_ptr<scaldpc_qary, void (*)(scaldpc_qary *)> h(new scaldpc_qary, scaldpc_qary_destroy);
    h->R = R;
    h->N = N;
    h->B = b;
    h->BSUM = bsum;
    h->Q = g.Q;
    h->QS = g.QS;
    h->bvars = g.bvars;
    h->max_iter = max_iter;
    h->device = device;
    if (const char *e = getenv("SCALDPC_MAX_RESIDENT_TRIALS")) h->max_resident = std::max(32, atoi(e));
    if (const char *e = getenv("SCALDPC_RISK_MARGIN")) h->risk_eps = std::max(0.0f, (float)atof(e));
    if (const char *e = getenv("SCALDPC_ES_RESTART_MIN")) h->es_restart_min = std::max(1, atoi(e));
    if (const char *e = getenv("SCALDPC_NO_GRAPH")) h->small_graph_off = atoi(e) != 0;

    h->E = g.E;
    h->S = g.S;
    h->SL = g.SL;
    h->SLS = g.SLS;
    h->lmax = g.lmax;
    h->wmax = g.wmax;
    if (g.regular_sw && pick_cn_fast(g.regular_sw, b)) h->fast_sw = g.regular_sw;
    {  // regular columns (all B-variables of one degree, edges of variable v at csc[v*D ..]): register-resident kernel
        int d0 = g.bvars > 0 ? g.col_ptr[1] - g.col_ptr[0] : 0;
        for (int v = 0; v < g.bvars && d0 > 0; ++v)
            if (g.col_ptr[v + 1] - g.col_ptr[v] != d0) d0 = 0;
        if (d0 > 0 && pick_vn_reg(g.Q, d0)) h->vn_reg_d = d0;
    }
    if (!pick_vn(h->Q)) return fail(SCALDPC_ERR_UNSUPPORTED, "no variable-node kernel for Q=%d", h->Q);
    std::vector<int2> csc(g.E);
    for (int e = 0; e < g.E; ++e) csc[e] = make_int2(g.csc[e].off, g.csc[e].sign);
    const std::vector<int> &row_ptr = g.row_ptr, &col = g.col, &edge_off = g.edge_off, &col_ptr = g.col_ptr;
    const std::vector<int8_t> &sign = g.sign;

    SCALDPC_CUDA(cudaSetDevice(device));
    SCALDPC_CUDA(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device));
    h->exact_split_items = 2 * h->num_sms;  // one wave of 2 CTAs per SM
    if (const char *e = getenv("SCALDPC_EXACT_SPLIT_ITEMS")) h->exact_split_items = atoi(e);
    if (h->fast_sw) {
        const CnFast *f = pick_cn_fast(h->fast_sw, b);
        for (int i = 0; i < kCnFastVariants; ++i)
            SCALDPC_CUDA(cudaFuncSetAttribute(f->fn[i], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->smem));
        const CnExact *ex = pick_cn_exact(h->fast_sw, b);
        SCALDPC_CUDA(cudaFuncSetAttribute(ex->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ex->smem));
        SCALDPC_CUDA(cudaFuncSetAttribute(ex->fn_split, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ex->smem));
    }
    int rc;
    if ((rc = upload(&h->d_row_ptr, row_ptr)) || (rc = upload(&h->d_col, col)) || (rc = upload(&h->d_sign, sign)) ||
        (rc = upload(&h->d_edge_off, edge_off)) || (rc = upload(&h->d_col_ptr, col_ptr)) || (rc = upload(&h->d_csc, csc)))
        return rc;
    if (h->BSUM == 0 && pick_cn_uniform(h->Q) && uniform_smem_bytes(h.get()) <= 227 * 1024) {
        h->uniform_ok = true;
        SCALDPC_CUDA(cudaFuncSetAttribute(pick_cn_uniform(h->Q), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)uniform_smem_bytes(h.get())));
    }
    {
        CnGenericArgs g;
        const size_t smem = generic_smem_bytes(h.get(), &g);
        if (smem > 227 * 1024)
            return fail(SCALDPC_ERR_UNSUPPORTED, "check-node scratch of %zu bytes exceeds shared memory", smem);
        SCALDPC_CUDA(cudaFuncSetAttribute(k_cn_generic<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    *out = h.release();
    return 0;
}

extern "C" {

void scaldpc_qary_destroy(scaldpc_qary *h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    free_workspace(h);
    free_stage(h);
    free_observed(h);
    free_sim(h);
    cudaFree(h->d_status_all);
    for (auto &kv : h->small_graphs) cudaGraphExecDestroy(kv.second.first);
    h->small_graphs.clear();
    if (h->s_small) cudaStreamDestroy(h->s_small);
    cudaFreeHost(h->h_pin_in);
    cudaFreeHost(h->h_pin_in_sum);
    cudaFreeHost(h->h_pin_post);
    cudaFreeHost(h->h_pin_post_sum);
    cudaFreeHost(h->h_pin_hard);
    cudaFreeHost(h->h_pin_status);
    cudaFree(h->d_scan_out);
    cudaFree(h->d_glist);
    cudaFree(h->d_retry_in);
    cudaFree(h->d_retry_in_sum);
    cudaFree(h->d_retry_obs);
    cudaFree(h->d_retry_obs_sum);
    cudaFree(h->d_iters_all);
    if (h->h_scan_out) cudaFreeHost(h->h_scan_out);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    for (int i = 0; i < 2; ++i) {
        if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]);
        if (h->ev_comp[i]) cudaEventDestroy(h->ev_comp[i]);
        if (h->ev_out[i]) cudaEventDestroy(h->ev_out[i]);
    }
    cudaFree(h->d_row_ptr);
    cudaFree(h->d_col);
    cudaFree(h->d_sign);
    cudaFree(h->d_edge_off);
    cudaFree(h->d_col_ptr);
    cudaFree(h->d_csc);
    delete h;
}

int scaldpc_qary_info(const scaldpc_qary *h, int *R, int *N, int *bvars, int *q, int *qs, int *edges, int *max_iter)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (R) *R = h->R;
    if (N) *N = h->N;
    if (bvars) *bvars = h->bvars;
    if (q) *q = h->Q;
    if (qs) *qs = h->QS;
    if (edges) *edges = h->E;
    if (max_iter) *max_iter = h->max_iter;
    return 0;
}

int scaldpc_qary_set_max_iter(scaldpc_qary *h, int max_iter)
{
    if (!h || max_iter < 0) return fail(SCALDPC_ERR_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    h->max_iter = max_iter;
    return 0;
}

long long scaldpc_qary_launch_count(const scaldpc_qary *h) { return h ? h->launches : 0; }

int scaldpc_qary_last_run(const scaldpc_qary *h, long long *trials, long long *redecoded, long long *trial_iterations)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (trials) *trials = h->last_trials;
    if (redecoded) *redecoded = h->last_redecoded;
    if (trial_iterations) *trial_iterations = h->last_work;
    return 0;
}

int scaldpc_qary_last_iterations(scaldpc_qary *h, int32_t *iters, int T, unsigned flags, void *stream)
{
    if (!h || !iters || T < 0) return fail(SCALDPC_ERR_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->d_iters_all || (size_t)T > h->cap_iters_T)
        return fail(SCALDPC_ERR_ARG, "the last decode on this handle did not use SCALDPC_QARY_EARLY_STOP on >= T trials");
    SCALDPC_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const bool dev = flags & SCALDPC_DEVICE_PTRS;
    SCALDPC_CUDA(cudaMemcpyAsync(iters, h->d_iters_all, (size_t)T * sizeof(int),
                                 dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
    if (!dev) SCALDPC_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int scaldpc_qary_set_risk_margin(scaldpc_qary *h, float eps)
{
    if (!h || !(eps >= 0.0f)) return fail(SCALDPC_ERR_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    h->risk_eps = eps;
    return 0;
}

// The most severe per-trial status of a batch -> return value + error text.
static int status_error(int worst)
{
    if (worst == SCALDPC_ERR_PROB_SUM) return fail(worst, "a prior row does not sum to 1 +- 0.001 (decoder.rs:683-684)");
    if (worst == SCALDPC_ERR_NO_MAX) return fail(worst, "No maximum probability found (decoder.rs:680)");
    if (worst == SCALDPC_ERR_NO_CONFIG) return fail(worst, "a check has no valid configuration (decoder.rs:618)");
    if (worst == SCALDPC_ERR_ARG) return fail(worst, "an observation index is outside its prior table");
    if (worst < 0) return fail(worst, "a trial failed with status %d", worst);
    return 0;
}

// Where the priors of a batch come from.  Vectors: in / in_sum as scaldpc_qary_decode documents them.  Observed:
// row indices into prior tables (scaldpc_qary_decode_observed); d_tab* are the handle's device LLR tables.
struct PriorSrc {
    bool observed = false;
    const float *in = nullptr, *in_sum = nullptr;
    const uint16_t *obs = nullptr, *obs_sum = nullptr;
    int n_rows = 0, n_rows_sum = 0;
};

// Priors of `T` trials (device, row-major; slot t = row map[t] when `map` is given) -> channel LLRs in the tile layout.
static int stage_priors(scaldpc_qary *h, int T, const PriorSrc &src, unsigned flags, int *status, const int *map,
                        cudaStream_t st)
{
    const bool special = h->BSUM > 0;
    const bool is_llr = flags & SCALDPC_IN_LLR;
    const int tiles = (T + 31) / 32;
    int rc;
    if (src.observed) {
        k_load_observed<<<dim3((h->bvars + kObsRows - 1) / kObsRows, tiles), 256, 0, st>>>(
            src.obs, h->d_tab_llr, h->d_tab_st, h->d_llr, T, h->bvars, h->Q, src.n_rows, status, map);
        ++h->launches;
        if (special) {
            k_load_observed<<<dim3((h->R + kObsRows - 1) / kObsRows, tiles), 256, 0, st>>>(
                src.obs_sum, h->d_tab_sum_llr, h->d_tab_sum_st, h->d_llr_sum, T, h->R, h->QS, src.n_rows_sum, status, map);
            ++h->launches;
        }
        SCALDPC_CUDA(cudaGetLastError());
    } else {
        if ((rc = load_priors(h, src.in, h->d_llr, T, tiles, h->bvars, h->Q, is_llr, st, status, map))) return rc;
        if (special && (rc = load_priors(h, src.in_sum, h->d_llr_sum, T, tiles, h->R, h->QS, is_llr, st, status, map)))
            return rc;
    }
    return 0;
}

// max(1, max_iter) flooding iterations over `tiles` tiles (decoder.rs:575-663: `it` starts at 1, exit when it >= max_iter).
static int run_iterations(scaldpc_qary *h, const QaryDev &p, int tiles, unsigned flags, cudaStream_t st)
{
    const int iters = std::max(1, h->max_iter);
    int rc;
    for (int it = 1; it <= iters; ++it) {
        if ((rc = launch_cn(h, p, tiles, it, iters, flags, st))) return rc;
        if ((rc = launch_vn(h, p, tiles, it == iters, flags, st))) return rc;
    }
    return 0;
}

// Tile-layout results of `T` trials -> the caller's row-major arrays (slot t -> row map[t] when `map` is given).
static int emit_outputs(scaldpc_qary *h, int T, const int *map, int8_t *dhard, float *dpost, float *dpost_sum,
                        cudaStream_t st)
{
    const int tiles = (T + 31) / 32;
    dim3 block(32, kTrRows);
    k_untile<int8_t><<<dim3((h->N + 31) / 32, tiles), block, 0, st>>>(h->d_hard_t, dhard, T, h->N, map);
    ++h->launches;
    if (dpost) {
        k_untile<float><<<dim3((h->SL + 31) / 32, tiles), block, 0, st>>>(h->d_post, dpost, T, h->SL, map);
        ++h->launches;
    }
    if (dpost_sum && h->BSUM > 0) {
        k_untile<float><<<dim3((h->SLS + 31) / 32, tiles), block, 0, st>>>(h->d_post_sum, dpost_sum, T, h->SLS, map);
        ++h->launches;
    }
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

// Flagged / still-running trials of the `n` slots -> ascending list + count (and, with `status`, the most severe per-trial
// status of slots 0 .. n-1).  Synchronises `st`.
static int scan_flags(scaldpc_qary *h, int n, const int *status, int *list, cudaStream_t st, int *count, int *worst)
{
    k_risk_scan<<<1, 1024, 0, st>>>(h->d_risk, (n + 31) / 32, n, status, list, h->d_scan_out, nullptr, nullptr, 0);
    ++h->launches;
    SCALDPC_CUDA(cudaGetLastError());
    SCALDPC_CUDA(cudaMemcpyAsync(h->h_scan_out, h->d_scan_out, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    SCALDPC_CUDA(cudaStreamSynchronize(st));
    *count = h->h_scan_out[0];
    if (worst) *worst = h->h_scan_out[1];
    return 0;
}

// Exact-order decode (the reference's arithmetic, fixed iteration count) of the `nf` trials listed in `list`, staged
// compactly into the leading tiles of the workspace; their outputs overwrite rows list[i] of the caller's arrays.
static int redecode_exact(scaldpc_qary *h, int nf, const int *list, const PriorSrc &src, int8_t *dhard, float *dpost,
                          float *dpost_sum, unsigned flags, cudaStream_t st, const int *out_list)
{
    const bool want_post = dpost != nullptr || dpost_sum != nullptr;
    int rc;
    h->last_redecoded += nf;
    h->last_work += (long long)nf * std::max(1, h->max_iter);
    if ((rc = stage_priors(h, nf, src, flags, nullptr, list, st))) return rc;
    QaryDev p2 = dev_view(h, want_post);
    p2.status = h->d_status + (size_t)h->cap_tiles * 32;  // scratch: the fast pass has already recorded every status
    const unsigned f2 = (flags | SCALDPC_CN_EXACT) & ~(SCALDPC_CN_GENERIC | SCALDPC_QARY_EARLY_STOP);
    if ((rc = run_iterations(h, p2, (nf + 31) / 32, f2, st))) return rc;
    return emit_outputs(h, nf, out_list, dhard, dpost, dpost_sum, st);
}

// One round of early termination over the `n` trials staged in the workspace: after every iteration each trial whose
// hard decisions satisfy every row (and whose posteriors are clear of ties) is frozen; the round ends when nobody is
// running, at the iteration cap, or -- `may_restart` -- as soon as at most 1/8 of the slots are still running, so that the
// caller can re-stage those stragglers compactly instead of dragging 7/8 idle lanes along.  Synchronises once per iteration
// (4 bytes).  Leaves risk[] = the slots still running.
static int early_stop_round(scaldpc_qary *h, int n, bool want_post, unsigned flags, bool may_restart, cudaStream_t st,
                            int *iterations, int *running)
{
    const int tiles = (n + 31) / 32;
    const int iters = std::max(1, h->max_iter);
    int rc;
    SCALDPC_CUDA(cudaMemsetAsync(h->d_done, 0, (size_t)tiles * sizeof(uint32_t), st));
    SCALDPC_CUDA(cudaMemsetAsync(h->d_iters, 0, (size_t)tiles * 32 * sizeof(int), st));
    QaryDev p = dev_view(h, want_post);
    p.risk = h->d_risk;
    p.risk_eps = h->risk_eps;
    p.done = h->d_done;
    int it = 0, run = n;
    while (it < iters && run > 0) {
        ++it;
        SCALDPC_CUDA(cudaMemsetAsync(h->d_risk, 0, (size_t)tiles * sizeof(uint32_t), st));
        SCALDPC_CUDA(cudaMemsetAsync(h->d_scan_out, 0, sizeof(int), st));
        if ((rc = launch_cn(h, p, tiles, it, iters, flags, st))) return rc;
        if ((rc = launch_vn(h, p, tiles, true, flags, st))) return rc;
        k_qary_syndrome<<<dim3((h->R + 7) / 8, tiles), 256, 0, st>>>(p);
        k_qary_finish<<<(tiles + 255) / 256, 256, 0, st>>>(h->d_risk, h->d_done, h->d_iters, tiles, n, it, h->d_scan_out);
        h->launches += 2;
        SCALDPC_CUDA(cudaGetLastError());
        SCALDPC_CUDA(cudaMemcpyAsync(h->h_scan_out, h->d_scan_out, sizeof(int), cudaMemcpyDeviceToHost, st));
        SCALDPC_CUDA(cudaStreamSynchronize(st));
        run = h->h_scan_out[0];
        if (may_restart && it < iters && (long long)run * 8 <= n) break;
    }
    h->last_work += (long long)n * it;
    *iterations = it;
    *running = run;
    return 0;
}

// The device pipeline of one batch: priors (device, row-major) -> hard decisions (device, row-major), all on `st`.
//
// Default mode = the fast forward/backward check node, VERIFIED: the last iteration flags every trial
// whose decisions could depend on the rounding order (a posterior with two near-equal smallest entries, or decisions that
// violate a parity check -- i.e. every trial that did not converge), the flagged trials are decoded AGAIN from their priors
// with the exact-order check node (the reference's arithmetic bit for bit, decoder_special.rs:539-553) and their outputs
// replace the fast ones.  That needs the flagged count on the host: ONE stream synchronisation per batch, which also
// returns the most severe per-trial status (*worst, may be null).  SCALDPC_NO_VERIFY / SCALDPC_CN_EXACT skip all of it.
//
// SCALDPC_QARY_EARLY_STOP (an extension, off by default): round 1 decodes everybody and freezes trials as
// they converge; if it was cut short (<= 1/8 still running) the stragglers are re-staged compactly and decoded from
// scratch in round 2, freezing again; what is still running at the iteration cap keeps its last decisions or -- verified
// mode -- is decoded once more with the exact-order check node and the reference's fixed iteration count.
// `iters_out` (device, may be null): per-trial iteration counts.
static int decode_on_device(scaldpc_qary *h, int T, const PriorSrc &src, int8_t *dhard, float *dpost, float *dpost_sum,
                            unsigned flags, cudaStream_t st, int *worst, int *iters_out, int base, bool stash)
{
    const bool want_post = dpost != nullptr || dpost_sum != nullptr;
    const bool exact = flags & SCALDPC_CN_EXACT;
    const bool early = (flags & SCALDPC_QARY_EARLY_STOP) && !exact;
    const bool verify = !(flags & (SCALDPC_CN_EXACT | SCALDPC_NO_VERIFY));
    const int tiles = (T + 31) / 32;
    const int iters = std::max(1, h->max_iter);
    int rc;
    if (worst) *worst = 0;
    h->last_trials += T;
    SCALDPC_CUDA(cudaMemsetAsync(h->d_status, 0, (size_t)tiles * 32 * sizeof(int), st));
    if ((rc = stage_priors(h, T, src, flags, h->d_status, nullptr, st))) return rc;
    int *list_a = h->d_risk_list, *list_b = h->d_risk_list + (size_t)h->cap_tiles * 32;

    if (early) {
        int its = 0, run = 0;
        if ((rc = early_stop_round(h, T, want_post, flags, T >= h->es_restart_min, st, &its, &run))) return rc;
        if ((rc = emit_outputs(h, T, nullptr, dhard, dpost, dpost_sum, st))) return rc;
        if (iters_out) {
            k_scatter_int<<<(T + 255) / 256, 256, 0, st>>>(h->d_iters, nullptr, iters_out, T);
            ++h->launches;
        }
        int nf = 0;
        if ((rc = scan_flags(h, T, h->d_status, list_a, st, &nf, worst))) return rc;  // also fetches the statuses
        if (nf > 0 && its < iters) {  // round 2: the stragglers, compacted, from scratch
            const int n2 = nf;
            if ((rc = stage_priors(h, n2, src, flags, nullptr, list_a, st))) return rc;
            if ((rc = early_stop_round(h, n2, want_post, flags, false, st, &its, &run))) return rc;
            if ((rc = emit_outputs(h, n2, list_a, dhard, dpost, dpost_sum, st))) return rc;
            if (iters_out) {
                k_scatter_int<<<(n2 + 255) / 256, 256, 0, st>>>(h->d_iters, list_a, iters_out, n2);
                ++h->launches;
            }
            nf = 0;
            if (run > 0 && verify) {  // slots of round 2 still running -> trial indices of the batch
                if ((rc = scan_flags(h, n2, nullptr, list_b, st, &nf, nullptr))) return rc;
                k_compose_map<<<(nf + 255) / 256, 256, 0, st>>>(list_b, list_a, h->d_scan_out);
                ++h->launches;
                SCALDPC_CUDA(cudaGetLastError());
                if (nf > 0) return redecode_exact(h, nf, list_b, src, dhard, dpost, dpost_sum, flags, st, list_b);
            }
            return 0;
        }
        if (nf > 0 && verify) return redecode_exact(h, nf, list_a, src, dhard, dpost, dpost_sum, flags, st, list_a);
        return 0;
    }

    QaryDev p = dev_view(h, want_post);
    if (verify) {
        SCALDPC_CUDA(cudaMemsetAsync(h->d_risk, 0, (size_t)tiles * sizeof(uint32_t), st));
        p.risk = h->d_risk;
        p.risk_eps = h->risk_eps;
    }
    if ((rc = run_iterations(h, p, tiles, flags, st))) return rc;
    h->last_work += (long long)T * iters;
    if (verify) {
        k_qary_syndrome<<<dim3((h->R + 7) / 8, tiles), 256, 0, st>>>(p);
        ++h->launches;
    }
    if ((rc = emit_outputs(h, T, nullptr, dhard, dpost, dpost_sum, st))) return rc;
    if (!verify) return 0;
    // The flagged trials join the call-wide list (no synchronisation here); finish_verify() decodes them again, once, after
    // the last chunk / pass -- a re-decode per chunk costs its fixed latency (~0.3 ms of small launches) every time.
    k_risk_scan<<<1, 1024, 0, st>>>(h->d_risk, tiles, T, h->d_status, list_a, h->d_scan_out, h->d_glist, h->d_scan_out + 4,
                                    base);
    ++h->launches;
    if (stash) {  // host pipeline: the staging buffer holding these priors is about to be overwritten
        const dim3 grid(8, std::min(T, 64));  // x: slices of a row, y: flagged rows (grid-stride)
        if (src.observed) {
            k_gather_rows<uint16_t><<<grid, 256, 0, st>>>(src.obs, h->d_retry_obs, list_a, h->d_scan_out, h->bvars, (int)h->cap_retry);
            if (h->BSUM > 0)
                k_gather_rows<uint16_t><<<grid, 256, 0, st>>>(src.obs_sum, h->d_retry_obs_sum, list_a, h->d_scan_out, h->R, (int)h->cap_retry);
        } else {
            k_gather_rows<float><<<grid, 256, 0, st>>>(src.in, h->d_retry_in, list_a, h->d_scan_out, h->SL, (int)h->cap_retry);
            if (h->BSUM > 0)
                k_gather_rows<float><<<grid, 256, 0, st>>>(src.in_sum, h->d_retry_in_sum, list_a, h->d_scan_out, h->SLS, (int)h->cap_retry);
        }
        h->launches += h->BSUM > 0 ? 2 : 1;
    }
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

// Start of a call: the call-wide list of flagged trials is empty.
static int begin_verify(scaldpc_qary *h, int T, cudaStream_t st)
{
    if ((size_t)T > h->cap_glist) {
        cudaFree(h->d_glist);
        h->d_glist = nullptr;
        h->cap_glist = 0;
        SCALDPC_CUDA(cudaMalloc((void **)&h->d_glist, (size_t)T * sizeof(int)));
        h->cap_glist = (size_t)T;
    }
    SCALDPC_CUDA(cudaMemsetAsync(h->d_scan_out + 4, 0, 2 * sizeof(int), st));
    return 0;
}

// End of a call (default, verified mode): ONE synchronisation to learn how many trials the chunks / passes flagged, then
// their exact-order re-decode.  Device mode: priors are read from the caller's arrays `src` through the
// list, results are written to rows list[i] of the caller's device arrays.  Host mode (`host` = the caller's HOST arrays,
// src = null): the first cap_retry flagged trials' priors were put aside on the device by decode_on_device(stash), the
// others are uploaded again; results come back through the output staging buffers and are scattered on the host.
static int finish_verify(scaldpc_qary *h, const PriorSrc *src, const PriorSrc *host, int8_t *hard, float *post,
                         float *post_sum, unsigned flags, cudaStream_t st, int *worst)
{
    int rc;
    SCALDPC_CUDA(cudaMemcpyAsync(h->h_scan_out + 4, h->d_scan_out + 4, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    SCALDPC_CUDA(cudaStreamSynchronize(st));
    const int count = h->h_scan_out[4];
    if (worst) *worst = std::min(*worst, h->h_scan_out[5]);
    if (count <= 0) return 0;
    const int room = h->cap_tiles * 32;  // trials the workspace holds
    if (src) {
        for (int o = 0; o < count; o += room) {
            const int nb = std::min(room, count - o);
            if ((rc = redecode_exact(h, nb, h->d_glist + o, *src, hard, post, post_sum, flags, st, h->d_glist + o))) return rc;
        }
        return 0;
    }
    // ---- host mode
    const bool special = h->BSUM > 0;
    std::vector<int> hl(count);
    SCALDPC_CUDA(cudaMemcpy(hl.data(), h->d_glist, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost));
    const int cap = (int)std::min<size_t>(h->cap_retry, std::min<size_t>((size_t)room, h->cap_stage_T));
    std::vector<int8_t> th;
    std::vector<float> tp, tps;
    for (int o = 0; o < count;) {
        const int resident = (int)std::min<size_t>((size_t)count, h->cap_retry);  // rows 0 .. resident-1 are on the device
        int nb, first_row;
        if (o < resident) {  // a batch of resident rows, read in place
            nb = std::min(cap, resident - o);
            first_row = o;
        } else {  // not resident (more flagged trials than the retry buffers hold): upload them again, row by row
            nb = std::min(cap, count - o);
            first_row = 0;
            for (int i = 0; i < nb; ++i) {
                const size_t t = (size_t)hl[o + i];
                if (host->observed) {
                    SCALDPC_CUDA(cudaMemcpyAsync(h->d_retry_obs + (size_t)i * h->bvars, host->obs + t * h->bvars,
                                                 (size_t)h->bvars * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
                    if (special)
                        SCALDPC_CUDA(cudaMemcpyAsync(h->d_retry_obs_sum + (size_t)i * h->R, host->obs_sum + t * h->R,
                                                     (size_t)h->R * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
                } else {
                    SCALDPC_CUDA(cudaMemcpyAsync(h->d_retry_in + (size_t)i * h->SL, host->in + t * h->SL,
                                                 (size_t)h->SL * sizeof(float), cudaMemcpyHostToDevice, st));
                    if (special)
                        SCALDPC_CUDA(cudaMemcpyAsync(h->d_retry_in_sum + (size_t)i * h->SLS, host->in_sum + t * h->SLS,
                                                     (size_t)h->SLS * sizeof(float), cudaMemcpyHostToDevice, st));
                }
            }
        }
        PriorSrc rs = *host;
        rs.in = h->d_retry_in ? h->d_retry_in + (size_t)first_row * h->SL : nullptr;
        rs.in_sum = h->d_retry_in_sum ? h->d_retry_in_sum + (size_t)first_row * h->SLS : nullptr;
        rs.obs = h->d_retry_obs ? h->d_retry_obs + (size_t)first_row * h->bvars : nullptr;
        rs.obs_sum = h->d_retry_obs_sum ? h->d_retry_obs_sum + (size_t)first_row * h->R : nullptr;
        if ((rc = redecode_exact(h, nb, nullptr, rs, h->d_ohard[0], post ? h->d_opost[0] : nullptr,
                                 (post_sum && special) ? h->d_opost_sum[0] : nullptr, flags, st, nullptr)))
            return rc;
        th.resize((size_t)nb * h->N);
        SCALDPC_CUDA(cudaMemcpyAsync(th.data(), h->d_ohard[0], th.size(), cudaMemcpyDeviceToHost, st));
        if (post) {
            tp.resize((size_t)nb * h->SL);
            SCALDPC_CUDA(cudaMemcpyAsync(tp.data(), h->d_opost[0], tp.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
        }
        if (post_sum && special) {
            tps.resize((size_t)nb * h->SLS);
            SCALDPC_CUDA(cudaMemcpyAsync(tps.data(), h->d_opost_sum[0], tps.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
        }
        SCALDPC_CUDA(cudaStreamSynchronize(st));
        for (int i = 0; i < nb; ++i) {
            const size_t t = (size_t)hl[o + i];
            memcpy(hard + t * h->N, th.data() + (size_t)i * h->N, (size_t)h->N);
            if (post) memcpy(post + t * h->SL, tp.data() + (size_t)i * h->SL, (size_t)h->SL * sizeof(float));
            if (post_sum && special) memcpy(post_sum + t * h->SLS, tps.data() + (size_t)i * h->SLS, (size_t)h->SLS * sizeof(float));
        }
        o += nb;
    }
    return 0;
}

// Host buffers, one tile (T <= 32), prior vectors, default stream: the reference-shaped call min_sum(channel_output[,
// channel_output_sum]) (pydecoder.rs:53-65, 125-145; decode.py:262 calls it once per decode).  ~25 dependent launches of
// ~2-5 us of work each are launch-latency-bound (0.25 ms per call in round 1), so the whole sequence -- H2D of the priors,
// into_llr, the iterations, the row test, the flag scan, D2H of decisions / posteriors / statuses / flagged count -- is
// captured once into a CUDA graph and replayed.  Flagged trials (the graph's last copy brings their count) are decoded
// again outside the graph, as in every other path.
static int decode_small_graph(scaldpc_qary *h, int T, const PriorSrc &src, int8_t *hard, float *post, float *post_sum,
                              int32_t *status, unsigned flags)
{
    const bool special = h->BSUM > 0;
    const bool want_post = post != nullptr || post_sum != nullptr;
    const bool deferred = !(flags & (SCALDPC_CN_EXACT | SCALDPC_NO_VERIFY));
    int rc;
    if (!h->s_small) {
        SCALDPC_CUDA(cudaStreamCreateWithFlags(&h->s_small, cudaStreamNonBlocking));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_pin_in, (size_t)32 * h->SL * sizeof(float), cudaHostAllocDefault));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_pin_in_sum, std::max<size_t>(1, (size_t)32 * h->SLS) * sizeof(float), cudaHostAllocDefault));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_pin_post, (size_t)32 * h->SL * sizeof(float), cudaHostAllocDefault));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_pin_post_sum, std::max<size_t>(1, (size_t)32 * h->SLS) * sizeof(float), cudaHostAllocDefault));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_pin_hard, (size_t)32 * h->N, cudaHostAllocDefault));
        SCALDPC_CUDA(cudaHostAlloc((void **)&h->h_pin_status, 32 * sizeof(int), cudaHostAllocDefault));
    }
    cudaStream_t st = h->s_small;
    if ((rc = ensure_workspace(h, 1, want_post))) return rc;
    if ((rc = ensure_pipeline(h, 32, 32, want_post))) return rc;
    if ((rc = ensure_stage_priors(h, 32))) return rc;
    if (deferred && (rc = begin_verify(h, 32, st))) return rc;  // allocation only matters here; the memset is in the graph
    memcpy(h->h_pin_in, src.in, (size_t)T * h->SL * sizeof(float));
    if (special) memcpy(h->h_pin_in_sum, src.in_sum, (size_t)T * h->SLS * sizeof(float));

    PriorSrc dsrc = src;
    dsrc.in = h->d_in[0];
    dsrc.in_sum = h->d_in_sum[0];
    const auto key = std::make_tuple(T, std::max(1, h->max_iter), flags, want_post ? 1 : 0);
    auto it = h->small_graphs.find(key);
    if (it == h->small_graphs.end()) {
        const long long l0 = h->launches, t0 = h->last_trials, w0 = h->last_work;
        SCALDPC_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        rc = 0;
        cudaError_t ce = cudaSuccess;
        do {
            if ((ce = cudaMemcpyAsync(h->d_in[0], h->h_pin_in, (size_t)T * h->SL * sizeof(float), cudaMemcpyHostToDevice, st))) break;
            if (special && (ce = cudaMemcpyAsync(h->d_in_sum[0], h->h_pin_in_sum, (size_t)T * h->SLS * sizeof(float), cudaMemcpyHostToDevice, st))) break;
            if (deferred && (ce = cudaMemsetAsync(h->d_scan_out + 4, 0, 2 * sizeof(int), st))) break;
            if ((rc = decode_on_device(h, T, dsrc, h->d_ohard[0], want_post ? h->d_opost[0] : nullptr,
                                       (want_post && special) ? h->d_opost_sum[0] : nullptr, flags, st, nullptr, nullptr, 0, false)))
                break;
            if ((ce = cudaMemcpyAsync(h->h_pin_hard, h->d_ohard[0], (size_t)T * h->N, cudaMemcpyDeviceToHost, st))) break;
            if (want_post && (ce = cudaMemcpyAsync(h->h_pin_post, h->d_opost[0], (size_t)T * h->SL * sizeof(float), cudaMemcpyDeviceToHost, st))) break;
            if (want_post && special && (ce = cudaMemcpyAsync(h->h_pin_post_sum, h->d_opost_sum[0], (size_t)T * h->SLS * sizeof(float), cudaMemcpyDeviceToHost, st))) break;
            if ((ce = cudaMemcpyAsync(h->h_pin_status, h->d_status, (size_t)T * sizeof(int), cudaMemcpyDeviceToHost, st))) break;
            if (deferred && (ce = cudaMemcpyAsync(h->h_scan_out + 4, h->d_scan_out + 4, 2 * sizeof(int), cudaMemcpyDeviceToHost, st))) break;
        } while (0);
        cudaGraph_t g = nullptr;
        const cudaError_t ee = cudaStreamEndCapture(st, &g);
        const int n_launch = (int)(h->launches - l0);
        h->launches = l0;  // counted per replay below
        h->last_trials = t0;
        h->last_work = w0;
        if (rc || ce != cudaSuccess || ee != cudaSuccess || !g) {
            if (g) cudaGraphDestroy(g);
            cudaGetLastError();
            h->small_graph_off = true;  // fall back to plain launches from now on
            return rc ? rc : fail(SCALDPC_ERR_CUDA, "CUDA graph capture of the single-decode path failed: %s",
                                  cudaGetErrorString(ce != cudaSuccess ? ce : ee));
        }
        cudaGraphExec_t ge = nullptr;
        const cudaError_t ie = cudaGraphInstantiate(&ge, g, 0);
        cudaGraphDestroy(g);
        if (ie != cudaSuccess) {
            h->small_graph_off = true;
            return fail(SCALDPC_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ie));
        }
        it = h->small_graphs.emplace(key, std::make_pair(ge, n_launch)).first;
    }
    SCALDPC_CUDA(cudaGraphLaunch(it->second.first, st));
    h->launches += it->second.second;
    h->last_trials += T;
    h->last_work += (long long)T * std::max(1, h->max_iter);
    SCALDPC_CUDA(cudaStreamSynchronize(st));
    int worst = 0;
    if (deferred && h->h_scan_out[4] > 0) {  // flagged trials: exact-order re-decode, results into the staging rows
        const int count = h->h_scan_out[4];
        if ((rc = redecode_exact(h, count, h->d_glist, dsrc, h->d_ohard[0], want_post ? h->d_opost[0] : nullptr,
                                 (want_post && special) ? h->d_opost_sum[0] : nullptr, flags, st, h->d_glist)))
            return rc;
        SCALDPC_CUDA(cudaMemcpyAsync(h->h_pin_hard, h->d_ohard[0], (size_t)T * h->N, cudaMemcpyDeviceToHost, st));
        if (want_post) {
            SCALDPC_CUDA(cudaMemcpyAsync(h->h_pin_post, h->d_opost[0], (size_t)T * h->SL * sizeof(float), cudaMemcpyDeviceToHost, st));
            if (special)
                SCALDPC_CUDA(cudaMemcpyAsync(h->h_pin_post_sum, h->d_opost_sum[0], (size_t)T * h->SLS * sizeof(float), cudaMemcpyDeviceToHost, st));
        }
        SCALDPC_CUDA(cudaStreamSynchronize(st));
    }
    memcpy(hard, h->h_pin_hard, (size_t)T * h->N);
    if (post) memcpy(post, h->h_pin_post, (size_t)T * h->SL * sizeof(float));
    if (post_sum && special) memcpy(post_sum, h->h_pin_post_sum, (size_t)T * h->SLS * sizeof(float));
    for (int t = 0; t < T; ++t) worst = std::min(worst, h->h_pin_status[t]);
    if (status) memcpy(status, h->h_pin_status, (size_t)T * sizeof(int));
    return status_error(worst);
}

// Shared body of scaldpc_qary_decode / scaldpc_qary_decode_observed.  `src` holds the caller's pointers (host or
// device); tables (observed mode) are already resident in the handle as LLRs.
static int decode_batch(scaldpc_qary *h, int T, const PriorSrc &src, int8_t *hard, float *post, float *post_sum,
                        int32_t *status, unsigned flags, cudaStream_t st)
{
    const bool special = h->BSUM > 0;
    const bool want_post = post != nullptr || post_sum != nullptr;
    const bool dev_ptrs = flags & SCALDPC_DEVICE_PTRS;
    int rc, worst = 0;
    h->last_trials = h->last_redecoded = h->last_work = 0;
    if (flags & SCALDPC_QARY_EARLY_STOP) {  // per-trial iteration counts of this call (scaldpc_qary_last_iterations)
        if ((size_t)T > h->cap_iters_T) {
            cudaFree(h->d_iters_all);
            h->d_iters_all = nullptr;
            h->cap_iters_T = 0;
            SCALDPC_CUDA(cudaMalloc((void **)&h->d_iters_all, (size_t)T * sizeof(int)));
            h->cap_iters_T = (size_t)T;
        }
    } else if (h->d_iters_all) {
        cudaFree(h->d_iters_all);
        h->d_iters_all = nullptr;
        h->cap_iters_T = 0;
    }

    // default (verified) mode without early termination: the chunks / passes only LIST their at-risk trials, the exact
    // re-decode runs once at the end of the call (finish_verify)
    const bool deferred = !(flags & (SCALDPC_CN_EXACT | SCALDPC_NO_VERIFY | SCALDPC_QARY_EARLY_STOP));

    if (!dev_ptrs && T <= 32 && !src.observed && st == nullptr && !(flags & SCALDPC_QARY_EARLY_STOP) && !h->small_graph_off) {
        rc = decode_small_graph(h, T, src, hard, post, post_sum, status, flags);
        if (!h->small_graph_off || rc != SCALDPC_ERR_CUDA) return rc;
        // capture is not possible here (e.g. another capture in progress): plain launches below
    }

    if (dev_ptrs) {  // asynchronous on `st` (one synchronisation per call unless SCALDPC_NO_VERIFY); at most `max_resident` trials' messages live in the workspace at a time
        const int cap = std::max(32, h->max_resident / 32 * 32);
        if ((rc = ensure_workspace(h, (std::min(T, cap) + 31) / 32, want_post))) return rc;
        if (deferred && (rc = begin_verify(h, T, st))) return rc;
        for (int t0 = 0; t0 < T; t0 += cap) {
            const int tc = std::min(cap, T - t0);
            PriorSrc part = src;
            if (src.observed) {
                part.obs = src.obs + (size_t)t0 * h->bvars;
                part.obs_sum = src.obs_sum ? src.obs_sum + (size_t)t0 * h->R : nullptr;
            } else {
                part.in = src.in + (size_t)t0 * h->SL;
                part.in_sum = src.in_sum ? src.in_sum + (size_t)t0 * h->SLS : nullptr;
            }
            int w = 0;
            if ((rc = decode_on_device(h, tc, part, hard + (size_t)t0 * h->N, post ? post + (size_t)t0 * h->SL : nullptr,
                                       post_sum ? post_sum + (size_t)t0 * h->SLS : nullptr, flags, st, &w,
                                       h->d_iters_all ? h->d_iters_all + t0 : nullptr, t0, false)))
                return rc;
            worst = std::min(worst, w);
            if (status)
                SCALDPC_CUDA(cudaMemcpyAsync(status + t0, h->d_status, (size_t)tc * sizeof(int), cudaMemcpyDeviceToDevice, st));
        }
        if (deferred && (rc = finish_verify(h, &src, nullptr, hard, post, post_sum, flags, st, &worst))) return rc;
        return status_error(worst);
    }

    // ---- host buffers: chunked pipeline.  Chunk c is copied in on s_in, decoded on `st`, copied out on s_out; the
    // two staging buffers alternate, so the H2D of chunk c+1 and the D2H of chunk c-1 overlap the decode of chunk c
    // (truly asynchronous only for pinned host memory; pageable memory degrades to the serial order, still correct).
    // prior vectors: 8 chunks keep the PCIe link busy behind the decode; observations are ~35x smaller, so two large
    // chunks (better-filled persistent kernels) are enough to overlap the copies
    int chunk = T;
    static const int n_prior_chunks = [] {  // experiment knob: SCALDPC_QARY_CHUNKS=<n> (default 8)
        const char *e = getenv("SCALDPC_QARY_CHUNKS");
        const int v = e ? atoi(e) : 0;
        return v >= 1 && v <= 64 ? v : 8;
    }();
    if (T >= 4096) chunk = src.observed ? (T + 1) / 2 : std::max(512, (T + n_prior_chunks - 1) / n_prior_chunks);
    chunk = (chunk + 31) / 32 * 32;
    const int nchunks = (T + chunk - 1) / chunk;
    if ((rc = ensure_workspace(h, chunk / 32, want_post))) return rc;
    if ((rc = ensure_pipeline(h, (size_t)chunk, (size_t)T, want_post))) return rc;
    if ((rc = src.observed ? ensure_stage_observed(h, (size_t)chunk) : ensure_stage_priors(h, (size_t)chunk))) return rc;
    if (deferred && ((rc = ensure_retry(h, (size_t)chunk, src.observed)) || (rc = begin_verify(h, T, st)))) return rc;
    // Host-side enqueue order per chunk c: H2D(c+1), decode(c), D2H(c-1).  A D2H into pageable memory blocks the
    // host until decode(c-1) is done; by then decode(c) and H2D(c+1) are already queued, so the device never idles.
    auto h2d = [&](int c) -> int {
        const int b = c & 1, t0 = c * chunk, tc = std::min(chunk, T - t0);
        if (c >= 2) SCALDPC_CUDA(cudaStreamWaitEvent(h->s_in, h->ev_comp[b], 0));  // decode(c-2) has consumed the buffer
        if (src.observed) {
            SCALDPC_CUDA(cudaMemcpyAsync(h->d_obs[b], src.obs + (size_t)t0 * h->bvars, (size_t)tc * h->bvars * sizeof(uint16_t),
                                         cudaMemcpyHostToDevice, h->s_in));
            if (special)
                SCALDPC_CUDA(cudaMemcpyAsync(h->d_obs_sum[b], src.obs_sum + (size_t)t0 * h->R,
                                             (size_t)tc * h->R * sizeof(uint16_t), cudaMemcpyHostToDevice, h->s_in));
        } else {
            SCALDPC_CUDA(cudaMemcpyAsync(h->d_in[b], src.in + (size_t)t0 * h->SL, (size_t)tc * h->SL * sizeof(float),
                                         cudaMemcpyHostToDevice, h->s_in));
            if (special)
                SCALDPC_CUDA(cudaMemcpyAsync(h->d_in_sum[b], src.in_sum + (size_t)t0 * h->SLS,
                                             (size_t)tc * h->SLS * sizeof(float), cudaMemcpyHostToDevice, h->s_in));
        }
        SCALDPC_CUDA(cudaEventRecord(h->ev_in[b], h->s_in));
        return 0;
    };
    auto d2h = [&](int c) -> int {
        const int b = c & 1, t0 = c * chunk, tc = std::min(chunk, T - t0);
        SCALDPC_CUDA(cudaStreamWaitEvent(h->s_out, h->ev_comp[b], 0));
        SCALDPC_CUDA(cudaMemcpyAsync(hard + (size_t)t0 * h->N, h->d_ohard[b], (size_t)tc * h->N, cudaMemcpyDeviceToHost, h->s_out));
        if (post)
            SCALDPC_CUDA(cudaMemcpyAsync(post + (size_t)t0 * h->SL, h->d_opost[b], (size_t)tc * h->SL * sizeof(float),
                                         cudaMemcpyDeviceToHost, h->s_out));
        if (post_sum && special)
            SCALDPC_CUDA(cudaMemcpyAsync(post_sum + (size_t)t0 * h->SLS, h->d_opost_sum[b],
                                         (size_t)tc * h->SLS * sizeof(float), cudaMemcpyDeviceToHost, h->s_out));
        SCALDPC_CUDA(cudaEventRecord(h->ev_out[b], h->s_out));
        return 0;
    };
    if ((rc = h2d(0))) return rc;
    for (int c = 0; c < nchunks; ++c) {
        const int b = c & 1, t0 = c * chunk, tc = std::min(chunk, T - t0);
        if (c + 1 < nchunks && (rc = h2d(c + 1))) return rc;
        SCALDPC_CUDA(cudaStreamWaitEvent(st, h->ev_in[b], 0));
        if (c >= 2) SCALDPC_CUDA(cudaStreamWaitEvent(st, h->ev_out[b], 0));  // d_ohard[b] of chunk c-2 has been copied out
        PriorSrc dsrc = src;
        dsrc.in = h->d_in[b];
        dsrc.in_sum = h->d_in_sum[b];
        dsrc.obs = h->d_obs[b];
        dsrc.obs_sum = h->d_obs_sum[b];
        if ((rc = decode_on_device(h, tc, dsrc, h->d_ohard[b], post ? h->d_opost[b] : nullptr,
                                   post_sum ? h->d_opost_sum[b] : nullptr, flags, st, nullptr,
                                   h->d_iters_all ? h->d_iters_all + t0 : nullptr, t0, deferred)))
            return rc;
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_status_all + t0, h->d_status, (size_t)tc * sizeof(int), cudaMemcpyDeviceToDevice, st));
        SCALDPC_CUDA(cudaEventRecord(h->ev_comp[b], st));
        if (c >= 1 && (rc = d2h(c - 1))) return rc;
    }
    if ((rc = d2h(nchunks - 1))) return rc;
    std::vector<int> hs(T);
    SCALDPC_CUDA(cudaMemcpyAsync(hs.data(), h->d_status_all, (size_t)T * sizeof(int), cudaMemcpyDeviceToHost, st));
    SCALDPC_CUDA(cudaStreamSynchronize(st));
    SCALDPC_CUDA(cudaStreamSynchronize(h->s_out));
    for (int t = 0; t < T; ++t) worst = std::min(worst, hs[t]);
    if (status) memcpy(status, hs.data(), (size_t)T * sizeof(int));
    // every chunk's fast results are in the caller's arrays; the flagged trials' exact-order results overwrite theirs
    if (deferred && (rc = finish_verify(h, nullptr, &src, hard, post, post_sum, flags, st, nullptr))) return rc;
    return status_error(worst);
}

int scaldpc_qary_decode(scaldpc_qary *h, int T, const float *in, const float *in_sum, int8_t *hard, float *post,
                        float *post_sum, int32_t *status, unsigned flags, void *stream)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (T < 0) return fail(SCALDPC_ERR_ARG, "T < 0");
    if (T == 0) return 0;
    const bool special = h->BSUM > 0;
    if (!in || !hard || (special && !in_sum)) return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    PriorSrc src;
    src.in = in;
    src.in_sum = in_sum;
    return decode_batch(h, T, src, hard, post, post_sum, status, flags, (cudaStream_t)stream);
}

// One decode, the reference's call shape (pydecoder.rs:53-65, 125-145: min_sum(channel_output[, channel_output_sum]) with the
// GIL released, called from a thread pool at simulate/decode.py:247-262).  Concurrent callers on one handle are COMBINED: the
// thread that finds no batch in progress becomes the leader, decodes every request queued so far (at most 32 = one tile: they
// cost what one costs, and the T <= 32 path is a CUDA graph replay) and wakes the others; a request left over elects the next
// leader.  Returns this trial's status: 0 or SCALDPC_ERR_PROB_SUM / NO_MAX / NO_CONFIG (the reference panics in THAT call only).
int scaldpc_qary_decode_one(scaldpc_qary *h, const float *in, const float *in_sum, int8_t *hard)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    const bool special = h->BSUM > 0;
    if (!in || !hard || (special && !in_sum)) return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    scaldpc_qary::OneReq me;
    me.in = in;
    me.in_sum = in_sum;
    me.hard = hard;
    std::unique_lock<std::mutex> lk(h->one_mu);
    h->one_q.push_back(&me);
    while (!me.done) {
        if (h->one_leading) {
            h->one_cv.wait(lk);
            continue;
        }
        h->one_leading = true;
        int rounds = 0;
    next_batch:
        scaldpc_qary::OneReq *batch[32];
        int T = 0;
        while (T < 32 && !h->one_q.empty()) {
            batch[T++] = h->one_q.front();
            h->one_q.pop_front();
        }
        lk.unlock();
        int rc = 0;
        int32_t status[32] = {0};
        {
            std::lock_guard<std::mutex> dk(h->mu);
            h->one_in.resize((size_t)32 * h->SL);
            h->one_in_sum.resize(std::max<size_t>(1, (size_t)32 * h->SLS));
            h->one_hard.resize((size_t)32 * h->N);
            for (int i = 0; i < T; ++i) {
                memcpy(h->one_in.data() + (size_t)i * h->SL, batch[i]->in, (size_t)h->SL * sizeof(float));
                if (special) memcpy(h->one_in_sum.data() + (size_t)i * h->SLS, batch[i]->in_sum, (size_t)h->SLS * sizeof(float));
            }
            if (cudaSetDevice(h->device) != cudaSuccess) {
                rc = fail(SCALDPC_ERR_CUDA, "cudaSetDevice(%d) failed", h->device);
            } else {
                PriorSrc src;
                src.in = h->one_in.data();
                src.in_sum = h->one_in_sum.data();
                rc = decode_batch(h, T, src, h->one_hard.data(), nullptr, nullptr, status, 0u, nullptr);
            }
            const bool per_trial = rc == 0 || rc == SCALDPC_ERR_PROB_SUM || rc == SCALDPC_ERR_NO_MAX || rc == SCALDPC_ERR_NO_CONFIG;
            for (int i = 0; i < T; ++i) {
                batch[i]->status = per_trial ? status[i] : rc;
                if (per_trial && status[i] == 0) memcpy(batch[i]->hard, h->one_hard.data() + (size_t)i * h->N, (size_t)h->N);
            }
            ++h->one_batches;
            h->one_requests += T;
        }
        lk.lock();
        for (int i = 0; i < T; ++i) batch[i]->done = true;
        h->one_cv.notify_all();
        // requests that arrived meanwhile: the leader goes on (handing over costs a wake-up during which the GPU idles),
        // a bounded number of times so that its own caller is not starved
        if (!h->one_q.empty() && ++rounds < 8) goto next_batch;
        h->one_leading = false;
    }
    lk.unlock();
    if (me.status == 0) return 0;
    const int st = status_error(me.status);
    return st ? st : fail(me.status, "the batch this decode was combined into failed with status %d", me.status);
}

// (requests combined so far, batches they were decoded in) of scaldpc_qary_decode_one on this handle
int scaldpc_qary_one_stats(scaldpc_qary *h, long long *requests, long long *batches)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    std::lock_guard<std::mutex> lk(h->one_mu);
    if (requests) *requests = h->one_requests;
    if (batches) *batches = h->one_batches;
    return 0;
}

// ---- fp64 verification mode -------------------------------------------------------------------------------
typedef void (*vn64_fn)(const QaryDevT<double>, const int);
static vn64_fn pick_vn64(int q)
{
    switch (q) {
#define CASE(Q_) \
    case Q_:     \
        return k_vn<Q_, double>;
        CASE(3) CASE(5) CASE(7) CASE(9) CASE(11) CASE(13) CASE(15) CASE(17) CASE(19) CASE(21) CASE(23) CASE(25)
        CASE(27) CASE(29) CASE(31)
#undef CASE
    default:
        return nullptr;
    }
}

// The decode of scaldpc_qary_decode with every message, LLR and posterior in DOUBLE (north_star: "fp32 log-domain; fp64 option
// for verification").  Reference semantics with f64 in place of f32: into_llr (decoder.rs:668-692), the check node in the
// reference's order of additions (decoder.rs:600-629 / decoder_special.rs:539-553; exact-order dynamic programme, k_cn_generic),
// the variable node (decoder.rs:634-658 / decoder_special.rs:566-609).  A checker, not a product path: run-time-shape kernels,
// buffers allocated and freed per call, at most 512 trials resident, synchronous, host pointers.
int scaldpc_qary_decode_f64(scaldpc_qary *h, int T, const float *in, const float *in_sum, int8_t *hard, double *post,
                            double *post_sum, int32_t *status, unsigned flags)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (T < 0) return fail(SCALDPC_ERR_ARG, "T < 0");
    if (T == 0) return 0;
    const bool special = h->BSUM > 0;
    if (!in || !hard || (special && !in_sum)) return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    if (flags & ~(unsigned)SCALDPC_IN_LLR) return fail(SCALDPC_ERR_ARG, "scaldpc_qary_decode_f64 takes SCALDPC_IN_LLR only");
    const vn64_fn vn = pick_vn64(h->Q);
    if (!vn) return fail(SCALDPC_ERR_UNSUPPORTED, "no fp64 variable-node kernel for q = %d", h->Q);
    CnGenericArgs g;
    const size_t smem = 2 * generic_smem_bytes(h, &g);  // the float figure, in doubles
    if (smem > 227 * 1024) return fail(SCALDPC_ERR_UNSUPPORTED, "fp64 check-node scratch of %zu bytes exceeds shared memory", smem);
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    SCALDPC_CUDA(cudaFuncSetAttribute(k_cn_generic<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int cap = 512, cap_tiles = cap / 32;
    const int iters = std::max(1, h->max_iter);
    const bool is_llr = flags & SCALDPC_IN_LLR;
    struct Bufs {
        void *p[9] = {nullptr};
        ~Bufs() { for (void *q : p) cudaFree(q); }
    } b;
    const size_t sizes[9] = {(size_t)cap * h->SL * sizeof(float),      std::max<size_t>(1, (size_t)cap * h->SLS) * sizeof(float),
                             (size_t)cap * h->SL * sizeof(double),     std::max<size_t>(1, (size_t)cap * h->SLS) * sizeof(double),
                             (size_t)cap * h->S * sizeof(double),      (size_t)cap * h->SL * sizeof(double),
                             std::max<size_t>(1, (size_t)cap * h->SLS) * sizeof(double), (size_t)cap * h->N,
                             (size_t)cap * (sizeof(int) + h->N)};
    for (int i = 0; i < 9; ++i) SCALDPC_CUDA(cudaMalloc(&b.p[i], sizes[i]));
    float *d_in = (float *)b.p[0], *d_in_sum = (float *)b.p[1];
    double *d_llr = (double *)b.p[2], *d_llr_sum = (double *)b.p[3], *d_msg = (double *)b.p[4], *d_post = (double *)b.p[5],
           *d_post_sum = (double *)b.p[6];
    int8_t *d_hard_t = (int8_t *)b.p[7];
    int *d_status = (int *)b.p[8];
    int8_t *d_hard = (int8_t *)(d_status + cap);
    // row-major staging of the posteriors reuses the LLR buffers once the iterations are over
    cudaStream_t st = nullptr;
    int worst = 0;
    std::vector<int> hstatus(cap);
    for (int t0 = 0; t0 < T; t0 += cap) {
        const int tc = std::min(cap, T - t0), tiles = (tc + 31) / 32;
        SCALDPC_CUDA(cudaMemcpyAsync(d_in, in + (size_t)t0 * h->SL, (size_t)tc * h->SL * sizeof(float), cudaMemcpyHostToDevice, st));
        if (special)
            SCALDPC_CUDA(cudaMemcpyAsync(d_in_sum, in_sum + (size_t)t0 * h->SLS, (size_t)tc * h->SLS * sizeof(float), cudaMemcpyHostToDevice, st));
        SCALDPC_CUDA(cudaMemsetAsync(d_status, 0, (size_t)cap * sizeof(int), st));
        k_load_priors_f64<<<dim3((h->bvars + 7) / 8, tiles), 256, 0, st>>>(d_in, d_llr, tc, h->bvars, h->Q, is_llr, d_status);
        if (special) k_load_priors_f64<<<dim3((h->R + 7) / 8, tiles), 256, 0, st>>>(d_in_sum, d_llr_sum, tc, h->R, h->QS, is_llr, d_status);
        QaryDevT<double> p;
        p.R = h->R; p.N = h->N; p.bvars = h->bvars; p.B = h->B; p.BSUM = h->BSUM; p.Q = h->Q; p.QS = h->QS;
        p.S = h->S; p.SL = h->SL; p.SLS = h->SLS;
        p.row_ptr = h->d_row_ptr; p.col = h->d_col; p.sign = h->d_sign; p.edge_off = h->d_edge_off; p.col_ptr = h->d_col_ptr;
        p.csc = h->d_csc;
        p.msg = d_msg; p.llr = d_llr; p.llr_sum = d_llr_sum; p.post = d_post; p.post_sum = special ? d_post_sum : nullptr;
        p.hard_t = d_hard_t; p.status = d_status;
        for (int it = 1; it <= iters; ++it) {
            g.exact = 1;
            g.first = it == 1;
            g.last = it == iters;
            if (it == 1) k_init_v2c<double><<<dim3((h->E + 7) / 8, tiles), 256, 0, st>>>(p, h->E);
            k_cn_generic<double><<<dim3(h->R, tiles), 32, smem, st>>>(p, g);
            vn<<<dim3((h->bvars + kVnWarps - 1) / kVnWarps, tiles), kVnWarps * 32, 0, st>>>(p, it == iters);
            h->launches += 2 + (it == 1);
        }
        SCALDPC_CUDA(cudaGetLastError());
        const dim3 block(32, kTrRows);
        k_untile<int8_t><<<dim3((h->N + 31) / 32, tiles), block, 0, st>>>(d_hard_t, d_hard, tc, h->N, nullptr);
        SCALDPC_CUDA(cudaMemcpyAsync(hard + (size_t)t0 * h->N, d_hard, (size_t)tc * h->N, cudaMemcpyDeviceToHost, st));
        if (post) {
            k_untile<double><<<dim3((h->SL + 31) / 32, tiles), block, 0, st>>>(d_post, d_llr, tc, h->SL, nullptr);
            SCALDPC_CUDA(cudaMemcpyAsync(post + (size_t)t0 * h->SL, d_llr, (size_t)tc * h->SL * sizeof(double), cudaMemcpyDeviceToHost, st));
        }
        if (post_sum && special) {
            k_untile<double><<<dim3((h->SLS + 31) / 32, tiles), block, 0, st>>>(d_post_sum, d_llr_sum, tc, h->SLS, nullptr);
            SCALDPC_CUDA(cudaMemcpyAsync(post_sum + (size_t)t0 * h->SLS, d_llr_sum, (size_t)tc * h->SLS * sizeof(double), cudaMemcpyDeviceToHost, st));
        }
        SCALDPC_CUDA(cudaMemcpyAsync(hstatus.data(), d_status, (size_t)tc * sizeof(int), cudaMemcpyDeviceToHost, st));
        SCALDPC_CUDA(cudaStreamSynchronize(st));
        SCALDPC_CUDA(cudaGetLastError());
        for (int t = 0; t < tc; ++t) worst = std::min(worst, hstatus[t]);
        if (status) memcpy(status + t0, hstatus.data(), (size_t)tc * sizeof(int));
    }
    return status_error(worst);
}

// tables -> device LLR tables (into_llr once per row, decoder.rs:668-692), on the compute stream
static int upload_tables(scaldpc_qary *h, const float *table, int n_rows, const float *table_sum, int n_rows_sum,
                         unsigned flags, cudaStream_t st)
{
    const bool special = h->BSUM > 0;
    int rc;
    if ((rc = ensure_tables(h, (size_t)n_rows, special ? (size_t)n_rows_sum : 0))) return rc;
    const bool dev_ptrs = flags & SCALDPC_DEVICE_PTRS;
    const bool is_llr = flags & SCALDPC_IN_LLR;
    const cudaMemcpyKind kind = dev_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    SCALDPC_CUDA(cudaMemcpyAsync(h->d_tab_in, table, (size_t)n_rows * h->Q * sizeof(float), kind, st));
    k_table_llr<<<(n_rows + 127) / 128, 128, 0, st>>>(h->d_tab_in, h->d_tab_llr, h->d_tab_st, n_rows, h->Q, is_llr);
    ++h->launches;
    if (special) {
        SCALDPC_CUDA(cudaMemcpyAsync(h->d_tab_sum_in, table_sum, (size_t)n_rows_sum * h->QS * sizeof(float), kind, st));
        k_table_llr<<<(n_rows_sum + 127) / 128, 128, 0, st>>>(h->d_tab_sum_in, h->d_tab_sum_llr, h->d_tab_sum_st, n_rows_sum,
                                                              h->QS, is_llr);
        ++h->launches;
    }
    SCALDPC_CUDA(cudaGetLastError());
    return 0;
}

int scaldpc_qary_decode_observed(scaldpc_qary *h, int T, const float *table, int n_rows, const float *table_sum,
                                 int n_rows_sum, const uint16_t *obs, const uint16_t *obs_sum, int8_t *hard, float *post,
                                 float *post_sum, int32_t *status, unsigned flags, void *stream)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (T < 0) return fail(SCALDPC_ERR_ARG, "T < 0");
    const bool special = h->BSUM > 0;
    if (!table || n_rows <= 0 || n_rows > 65535 || (special && (!table_sum || n_rows_sum <= 0 || n_rows_sum > 65535)))
        return fail(SCALDPC_ERR_ARG, "prior tables must have 1 .. 65535 rows");
    if (T == 0) return 0;
    if (!obs || !hard || (special && !obs_sum)) return fail(SCALDPC_ERR_ARG, "NULL input/output pointer");
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    int rc;
    if ((rc = upload_tables(h, table, n_rows, table_sum, n_rows_sum, flags, st))) return rc;
    PriorSrc src;
    src.observed = true;
    src.obs = obs;
    src.obs_sum = obs_sum;
    src.n_rows = n_rows;
    src.n_rows_sum = n_rows_sum;
    return decode_batch(h, T, src, hard, post, post_sum, status, flags, st);
}

int scaldpc_qary_simulate(scaldpc_qary *h, int T, uint64_t seed, uint64_t first_trial, const float *table, int n_rows,
                          const float *table_sum, int n_rows_sum, const uint16_t *code, int nbits, const uint16_t *code_sum,
                          int nbits_sum, double accuracy, int eta, int64_t *counters, int8_t *secret, int16_t *ssum,
                          uint16_t *obs, uint16_t *obs_sum, int8_t *hard, unsigned flags, void *stream)
{
    if (!h) return fail(SCALDPC_ERR_ARG, "handle is NULL");
    if (h->BSUM <= 0) return fail(SCALDPC_ERR_UNSUPPORTED, "trial synthesis needs a decoder with sum variables (BSUM > 0)");
    if (flags & SCALDPC_DEVICE_PTRS) return fail(SCALDPC_ERR_ARG, "scaldpc_qary_simulate takes host pointers");
    if (T < 0 || !counters || !table || !table_sum || !code || !code_sum)
        return fail(SCALDPC_ERR_ARG, "NULL pointer or T < 0");
    if (n_rows <= 0 || n_rows > 65535 || n_rows_sum <= 0 || n_rows_sum > 65535)
        return fail(SCALDPC_ERR_ARG, "prior tables must have 1 .. 65535 rows");
    if (nbits < 1 || nbits > 16 || nbits_sum < 1 || nbits_sum > 16 || (1 << nbits) > n_rows || (1 << nbits_sum) > n_rows_sum)
        return fail(SCALDPC_ERR_ARG, "code words of nbits / nbits_sum bits need tables of 2^nbits / 2^nbits_sum rows");
    if (eta < 1 || eta > 16 || eta > h->B) return fail(SCALDPC_ERR_ARG, "eta must be in 1 .. B");
    if (!(accuracy >= 0.0 && accuracy <= 1.0)) return fail(SCALDPC_ERR_ARG, "accuracy must be in [0, 1]");
    for (int i = 0; i < h->Q; ++i)
        if (code[i] >> nbits) return fail(SCALDPC_ERR_ARG, "code word %d has more than nbits bits", i);
    for (int i = 0; i < h->QS; ++i)
        if (code_sum[i] >> nbits_sum) return fail(SCALDPC_ERR_ARG, "sum code word %d has more than nbits_sum bits", i);
    for (int i = 0; i < 5; ++i) counters[i] = 0;
    if (T == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(h->mu);
    h->last_trials = h->last_redecoded = h->last_work = 0;
    SCALDPC_CUDA(cudaSetDevice(h->device));
    int rc;
    if ((rc = upload_tables(h, table, n_rows, table_sum, n_rows_sum, flags, st))) return rc;
    const int cap = std::max(32, h->max_resident / 32 * 32);
    const int chunk = (std::min(T, cap) + 31) / 32 * 32;
    if ((rc = ensure_sim(h, (size_t)chunk))) return rc;
    if ((rc = ensure_stage_observed(h, (size_t)chunk))) return rc;
    if ((rc = ensure_workspace(h, chunk / 32, false))) return rc;
    SCALDPC_CUDA(cudaMemcpyAsync(h->d_sim_code, code, (size_t)h->Q * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
    SCALDPC_CUDA(cudaMemcpyAsync(h->d_sim_code_sum, code_sum, (size_t)h->QS * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
    SCALDPC_CUDA(cudaMemsetAsync(h->d_sim_cnt, 0, 8 * sizeof(unsigned long long), st));
    SimParams sp;
    sp.seed = seed;
    sp.eta = eta;
    sp.nbits = nbits;
    sp.nbits_sum = nbits_sum;
    sp.flip_thr = accuracy >= 1.0 ? (1ull << 32) : (uint64_t)(accuracy * 4294967296.0);
    const QaryDev p = dev_view(h, false);
    PriorSrc src;
    src.observed = true;
    src.obs = h->d_obs[0];
    src.obs_sum = h->d_obs_sum[0];
    src.n_rows = n_rows;
    src.n_rows_sum = n_rows_sum;
    int worst = 0;
    std::vector<int> hs;
    for (int t0 = 0; t0 < T; t0 += cap) {
        const int tc = std::min(cap, T - t0);
        sp.first_trial = first_trial + (uint64_t)t0;
        const int grid = h->num_sms * 8;
        k_sim_secret<<<grid, 256, 0, st>>>(sp, h->d_sim_code, h->d_sim_secret, h->d_obs[0], tc, h->bvars, h->B);
        // decode_on_device clears d_status first, so the range check of k_sim_sum reports through the counters' slot 5
        k_sim_sum<<<grid, 256, 0, st>>>(sp, p, h->d_sim_code_sum, h->d_sim_secret, h->d_sim_ssum, h->d_obs_sum[0],
                                        (int *)(h->d_sim_cnt + 5), tc);
        h->launches += 2;
        SCALDPC_CUDA(cudaGetLastError());
        const bool deferred = !(flags & (SCALDPC_CN_EXACT | SCALDPC_NO_VERIFY | SCALDPC_QARY_EARLY_STOP));
        if (deferred && (rc = begin_verify(h, tc, st))) return rc;
        if ((rc = decode_on_device(h, tc, src, h->d_sim_hard, nullptr, nullptr, flags, st, nullptr, nullptr, 0, false))) return rc;
        if (deferred && (rc = finish_verify(h, &src, nullptr, h->d_sim_hard, nullptr, nullptr, flags, st, nullptr))) return rc;
        k_sim_count<<<grid, 256, 0, st>>>(h->d_sim_hard, h->d_sim_secret, h->d_sim_ssum, h->d_sim_cnt, tc, h->N, h->bvars);
        ++h->launches;
        SCALDPC_CUDA(cudaGetLastError());
        // optional per-trial outputs (parity tests, debugging); production calls pass NULL and copy nothing
        if (secret)
            SCALDPC_CUDA(cudaMemcpyAsync(secret + (size_t)t0 * h->bvars, h->d_sim_secret, (size_t)tc * h->bvars, cudaMemcpyDeviceToHost, st));
        if (ssum)
            SCALDPC_CUDA(cudaMemcpyAsync(ssum + (size_t)t0 * h->R, h->d_sim_ssum, (size_t)tc * h->R * sizeof(int16_t),
                                         cudaMemcpyDeviceToHost, st));
        if (obs)
            SCALDPC_CUDA(cudaMemcpyAsync(obs + (size_t)t0 * h->bvars, h->d_obs[0], (size_t)tc * h->bvars * sizeof(uint16_t),
                                         cudaMemcpyDeviceToHost, st));
        if (obs_sum)
            SCALDPC_CUDA(cudaMemcpyAsync(obs_sum + (size_t)t0 * h->R, h->d_obs_sum[0], (size_t)tc * h->R * sizeof(uint16_t),
                                         cudaMemcpyDeviceToHost, st));
        if (hard)
            SCALDPC_CUDA(cudaMemcpyAsync(hard + (size_t)t0 * h->N, h->d_sim_hard, (size_t)tc * h->N, cudaMemcpyDeviceToHost, st));
        hs.resize(tc);
        SCALDPC_CUDA(cudaMemcpyAsync(hs.data(), h->d_status, (size_t)tc * sizeof(int), cudaMemcpyDeviceToHost, st));
        SCALDPC_CUDA(cudaStreamSynchronize(st));  // the pass's buffers are reused by the next one
        for (int t = 0; t < tc; ++t) worst = std::min(worst, hs[t]);
    }
    unsigned long long cnt[8];
    SCALDPC_CUDA(cudaMemcpyAsync(cnt, h->d_sim_cnt, sizeof(cnt), cudaMemcpyDeviceToHost, st));
    SCALDPC_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < 5; ++i) counters[i] = (int64_t)cnt[i];
    if (cnt[5]) return fail(SCALDPC_ERR_ARG, "a check sum left [-BSUM, BSUM]: BSUM is smaller than row weight * eta");
    if (worst == SCALDPC_ERR_PROB_SUM) return fail(worst, "a prior row does not sum to 1 +- 0.001 (decoder.rs:683-684)");
    if (worst == SCALDPC_ERR_NO_MAX) return fail(worst, "No maximum probability found (decoder.rs:680)");
    if (worst == SCALDPC_ERR_NO_CONFIG) return fail(worst, "a check has no valid configuration (decoder.rs:618)");
    if (worst < 0) return fail(worst, "trial synthesis failed with status %d", worst);
    return 0;
}

int scaldpc_qary_time_kernel(scaldpc_qary *h, int which, int T, unsigned flags, int reps, void *stream, float *avg_ms)
{
    if (!h || !avg_ms || reps <= 0 || T <= 0) return fail(SCALDPC_ERR_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    SCALDPC_CUDA(cudaSetDevice(h->device));
    const int tiles = (T + 31) / 32;
    if (tiles > h->cap_tiles) return fail(SCALDPC_ERR_ARG, "time_kernel needs a previous decode of >= T trials");
    cudaStream_t st = (cudaStream_t)stream;
    const QaryDev p = dev_view(h, false);
    cudaEvent_t e0, e1;
    SCALDPC_CUDA(cudaEventCreate(&e0));
    SCALDPC_CUDA(cudaEventCreate(&e1));
    int rc = 0;
    // one untimed launch, then `reps` timed ones (state stays valid: middle-iteration kernels map a valid
    // message state to a valid message state)
    for (int i = -1; i < reps && !rc; ++i) {
        if (i == 0) SCALDPC_CUDA(cudaEventRecord(e0, st));
        rc = which == 0 ? launch_cn(h, p, tiles, 2, 4, flags, st) : launch_vn(h, p, tiles, false, flags, st);  // middle iteration
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
