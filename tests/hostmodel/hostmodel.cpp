// TEST INFRASTRUCTURE ONLY.  Compiles the CUDA kernels of sca_ldpc_b200/csrc/qary_kernels.cuh with g++ and runs
// their threads one after another on the host, so that the kernel LOGIC (index arithmetic, DP tables, NaN/inf
// handling, in-place message updates) can be checked against the oracle in the `-m "not gpu"` test tier, on a
// machine without a GPU.  This file is never linked into libscaldpc.so and the product has no CPU path.
//
// Only kernels whose threads are independent are modelled (k_cn_fast, k_cn_generic, k_vn, k_init_v2c, k_bp_cn,
// k_bp_vn, k_bp_syndrome, k_bp_finish_iter -- their warp ballots are replaced by per-lane read-modify-writes under
// SCALDPC_HOST_MODEL); the layout-conversion kernels cooperate through shared memory and are replaced by plain
// loops here.
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#define SCALDPC_HOST_MODEL 1
struct hm_dim3 {
    unsigned x = 1, y = 0, z = 0;
};
static hm_dim3 threadIdx, blockIdx, blockDim, gridDim;
static float *scaldpc_host_model_smem = nullptr;
struct int2 {
    int x, y;
};
static inline int2 make_int2(int x, int y) { return int2{x, y}; }
struct int4 {
    int x, y, z, w;
};
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __restrict__
#define __shared__ static
template <typename T>
static inline T __ldg(const T *p) { return *p; }
static inline int atomicMin(int *p, int v)
{
    int o = *p;
    *p = std::min(o, v);
    return o;
}
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline void __syncthreads() {}
using std::max;
using std::min;

#include "../../sca_ldpc_b200/csrc/qary_graph.h"
#include "../../sca_ldpc_b200/csrc/qary_kernels.cuh"
#include "../../sca_ldpc_b200/csrc/bp_graph.h"
#include "../../sca_ldpc_b200/csrc/bp_kernels.cuh"

using namespace scaldpc;

namespace {

template <typename K, typename... A>
void run_grid(unsigned gx, unsigned gy, unsigned threads, K kernel, A... args)
{
    gridDim.x = gx;
    gridDim.y = gy;
    blockDim.x = threads;
    for (unsigned by = 0; by < gy; ++by)
        for (unsigned bx = 0; bx < gx; ++bx)
            for (unsigned t = 0; t < threads; ++t) {
                blockIdx.x = bx;
                blockIdx.y = by;
                threadIdx.x = t;
                kernel(args...);
            }
}

typedef void (*cn_fast_fn)(const QaryDev, const int);
typedef void (*vn_fn)(const QaryDev, const int);

// iteration `it` of `iters` -> kernel variant, as cn_mode_index() / cn_fast_table() in qary.cu
template <int SW, int B>
cn_fast_fn cn_fast_variant(int it, int iters, bool decide)
{
    if (decide)  // early termination: every iteration emits decisions, full sum-variable round trip
        return it == 1 ? k_cn_fast<SW, B, kCnFirst | kCnLast | kCnWsum> : k_cn_fast<SW, B, kCnLast | kCnRsum | kCnWsum>;
    const bool first = it == 1, last = it == iters, penult = it == iters - 1;
    if (first)
        return last ? k_cn_fast<SW, B, kCnFirst | kCnLast>
                    : (penult ? k_cn_fast<SW, B, kCnFirst | kCnWsum> : k_cn_fast<SW, B, kCnFirst>);
    return last ? k_cn_fast<SW, B, kCnLast | kCnRsum> : (penult ? k_cn_fast<SW, B, kCnWsum> : k_cn_fast<SW, B, 0>);
}

cn_fast_fn pick_cn_fast(int sw, int b, int it = 2, int iters = 4, bool decide = false)
{
#define CASE(SW_, B_) \
    if (sw == SW_ && b == B_) return cn_fast_variant<SW_, B_>(it, iters, decide);
    CASE(6, 2) CASE(6, 3) CASE(4, 2) CASE(8, 2) CASE(3, 2)
#undef CASE
    return nullptr;
}

typedef void (*cn_exact_fn)(const QaryDev, const int, const int, const int);
cn_exact_fn pick_cn_exact(int sw, int b)
{
#define CASE(SW_, B_) \
    if (sw == SW_ && b == B_) return k_cn_exact<SW_, B_>;
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

}  // namespace

static int g_last_redecoded = 0;
static long long g_last_work = 0;
static std::vector<int> g_last_iters;
constexpr unsigned kHmNoRestart = 0x40000000u;  // internal: second round of early termination

static int hm_qary_decode(const int8_t *H, int R, int N, int b, int bsum, int dv, int dc, int max_iter, int T,
                          const float *llr_in, const float *llr_sum_in, int8_t *hard, float *post, float *post_sum,
                          int32_t *status, unsigned flags, int *used_fast, char *errbuf, int errlen, int *iters_out);

extern "C" {

// Trials the verification of the last hostmodel_qary_decode call re-decoded with the exact-order check node.
int hostmodel_last_redecoded(void) { return g_last_redecoded; }
// Trial-iterations the last call launched (restarts and re-decodes counted again) and its per-trial iteration counts.
long long hostmodel_last_work(void) { return g_last_work; }
void hostmodel_last_iterations(int32_t *out, int T)
{
    for (int t = 0; t < T && t < (int)g_last_iters.size(); ++t) out[t] = g_last_iters[t];
}

// sc_logf (sca_ldpc_b200/csrc/sc_logf.cuh), the device restatement of glibc's logf, evaluated on the host.
void hostmodel_logf(const float *x, float *y, long n)
{
    for (long i = 0; i < n; ++i) y[i] = sc_logf(x[i], sc_logf_tab);
}

// Full decode of T trials from LLRs through the modelled kernels.  Same argument meaning as
// scaldpc_qary_decode with SCALDPC_IN_LLR; flags: SCALDPC_CN_EXACT / SCALDPC_CN_GENERIC.
// Returns 0 or a (negative) status code; *used_fast tells whether k_cn_fast ran.
int hostmodel_qary_decode(const int8_t *H, int R, int N, int b, int bsum, int dv, int dc, int max_iter, int T,
                          const float *llr_in, const float *llr_sum_in, int8_t *hard, float *post, float *post_sum,
                          int32_t *status, unsigned flags, int *used_fast, char *errbuf, int errlen)
{
    g_last_redecoded = 0;
    g_last_work = 0;
    g_last_iters.assign(T, 0);
    return hm_qary_decode(H, R, N, b, bsum, dv, dc, max_iter, T, llr_in, llr_sum_in, hard, post, post_sum, status, flags,
                          used_fast, errbuf, errlen, g_last_iters.data());
}

}  // extern "C"

static int hm_qary_decode(const int8_t *H, int R, int N, int b, int bsum, int dv, int dc, int max_iter, int T,
                          const float *llr_in, const float *llr_sum_in, int8_t *hard, float *post, float *post_sum,
                          int32_t *status, unsigned flags, int *used_fast, char *errbuf, int errlen, int *iters_out)
{
    HostGraph g;
    std::string err;
    int rc = build_graph(H, R, N, b, bsum, dv, dc, g, err);
    if (rc) {
        if (errbuf) snprintf(errbuf, errlen, "%s", err.c_str());
        return rc;
    }
    const int tiles = (T + 31) / 32;
    const size_t lanes = (size_t)tiles * 32;
    std::vector<float> msg(lanes * g.S), llr(lanes * g.SL), llrs(std::max<size_t>(1, lanes * g.SLS));
    std::vector<float> tpost(lanes * g.SL), tposts(std::max<size_t>(1, lanes * g.SLS));
    std::vector<int8_t> hard_t(lanes * g.N);
    std::vector<int> st(lanes, 0);
    std::vector<int2> csc(g.E);
    for (int e = 0; e < g.E; ++e) csc[e] = make_int2(g.csc[e].off, g.csc[e].sign);
    // tile the inputs (what k_load_priors does with is_llr = 1)
    for (size_t t = 0; t < lanes; ++t) {
        const size_t tile = t / 32, lane = t % 32;
        for (int i = 0; i < g.SL; ++i)
            llr[(tile * g.SL + i) * 32 + lane] = t < (size_t)T ? llr_in[t * g.SL + i] : 0.0f;
        for (int i = 0; i < g.SLS; ++i)
            llrs[(tile * g.SLS + i) * 32 + lane] = t < (size_t)T ? llr_sum_in[t * g.SLS + i] : 0.0f;
    }
    QaryDev p;
    p.R = g.R; p.N = g.N; p.bvars = g.bvars; p.B = g.B; p.BSUM = g.BSUM; p.Q = g.Q; p.QS = g.QS;
    p.S = g.S; p.SL = g.SL; p.SLS = g.SLS;
    p.row_ptr = g.row_ptr.data(); p.col = g.col.data(); p.sign = g.sign.data(); p.edge_off = g.edge_off.data();
    p.col_ptr = g.col_ptr.data(); p.csc = csc.data();
    p.msg = msg.data(); p.llr = llr.data(); p.llr_sum = llrs.data();
    p.post = tpost.data(); p.post_sum = tposts.data(); p.hard_t = hard_t.data(); p.status = st.data();

    const bool exact = flags & SCALDPC_CN_EXACT, force_generic = flags & SCALDPC_CN_GENERIC;
    const bool verify = !(flags & (SCALDPC_CN_EXACT | SCALDPC_NO_VERIFY));  // as decode_on_device (qary.cu)
    const bool early = (flags & SCALDPC_QARY_EARLY_STOP) && !exact;
    int restart_min = 256;
    if (const char *e = getenv("SCALDPC_ES_RESTART_MIN")) restart_min = std::max(1, atoi(e));
    const bool may_restart = early && !(flags & kHmNoRestart) && T >= restart_min;
    std::vector<uint32_t> risk(tiles, 0u), done(tiles, 0u);
    std::vector<int> slot_iters(lanes, 0);
    if (verify || early) {
        p.risk = risk.data();
        p.risk_eps = 1e-2f;
    }
    if (early) p.done = done.data();
    cn_fast_fn fast = (!exact && !force_generic && g.regular_sw) ? pick_cn_fast(g.regular_sw, b) : nullptr;
    vn_fn vn = pick_vn(g.Q);
    if (!vn) return SCALDPC_ERR_UNSUPPORTED;
    {  // regular columns: the register-resident kernel, as scaldpc_qary_create chooses it (qary.cu)
        int d0 = g.bvars > 0 ? g.col_ptr[1] - g.col_ptr[0] : 0;
        for (int v = 0; v < g.bvars && d0 > 0; ++v)
            if (g.col_ptr[v + 1] - g.col_ptr[v] != d0) d0 = 0;
        if (d0 > 0 && !force_generic && pick_vn_reg(g.Q, d0)) vn = pick_vn_reg(g.Q, d0);
    }
    if (used_fast) *used_fast = fast != nullptr;
    std::vector<float> scratch(generic_scratch_floats(g) * 32);
    scaldpc_host_model_smem = scratch.data();

    const int iters = std::max(1, max_iter);
    int it_run = 0, running = T;
    for (int it = 1; it <= iters && running > 0; ++it) {
        it_run = it;
        const int first = it == 1, last = it == iters || early;
        if (early) std::fill(risk.begin(), risk.end(), 0u);
        if (fast) {
            // persistent kernel: a deliberately small grid, so every warp strides over several items
            run_grid(3, 1, kCnWarps * 32, pick_cn_fast(g.regular_sw, b, it, iters, early), p, tiles * g.R);
        } else if (exact && !force_generic && g.regular_sw && pick_cn_exact(g.regular_sw, b)) {
            // exact-order DP in registers (qary_cn_exact.cuh), as launch_cn chooses it (qary.cu)
            const int n_items = tiles * g.R;
            run_grid((n_items + kCnExactWarps - 1) / kCnExactWarps, 1, kCnExactWarps * 32, pick_cn_exact(g.regular_sw, b), p,
                     n_items, first, last);
        } else if (g.BSUM == 0 && !exact && !force_generic && (g.Q == 3 || g.Q == 15)) {
            // generic decoder, one alphabet: k_cn_uniform<Q>, as launch_cn chooses it (qary.cu)
            if (first) run_grid((g.E + 7) / 8, tiles, 256, k_init_v2c<float>, p, g.E);
            std::vector<float> us(((size_t)g.lmax * g.Q + (size_t)(g.lmax + 2 + 1) * (g.wmax + 2 * g.B)) * 32);
            scaldpc_host_model_smem = us.data();
            if (g.Q == 3) run_grid(g.R, tiles, 32, k_cn_uniform<3>, p, g.lmax, (g.wmax - 1) / 2);
            else run_grid(g.R, tiles, 32, k_cn_uniform<15>, p, g.lmax, (g.wmax - 1) / 2);
            scaldpc_host_model_smem = scratch.data();
        } else {
            if (first) run_grid((g.E + 7) / 8, tiles, 256, k_init_v2c<float>, p, g.E);
            CnGenericArgs a;
            a.exact = exact; a.first = first; a.last = last;
            a.QM = std::max(g.Q, g.QS); a.W = g.wmax; a.LMAX = g.lmax;
            run_grid(g.R, tiles, 32, k_cn_generic<float>, p, a);
        }
        run_grid((g.bvars + kVnWarps - 1) / kVnWarps, tiles, kVnWarps * 32, vn, p, last);
        if (early) {  // early_stop_round (qary.cu): row tests, freeze, count
            run_grid((g.R + 7) / 8, tiles, 256, k_qary_syndrome, p);
            running = 0;
            run_grid((tiles + 255) / 256, 1, 256, k_qary_finish, risk.data(), done.data(), slot_iters.data(), tiles, T, it,
                     &running);
            if (may_restart && it < iters && (long long)running * 8 <= T) break;
        }
    }
    g_last_work += (long long)T * it_run;
    if (verify && !early) run_grid((g.R + 7) / 8, tiles, 256, k_qary_syndrome, p);
    if (iters_out)
        for (int t = 0; t < T; ++t) iters_out[t] = early ? slot_iters[t] : iters;
    // untile
    int worst = 0;
    for (int t = 0; t < T; ++t) {
        const size_t tile = t / 32, lane = t % 32;
        for (int i = 0; i < g.N; ++i) hard[(size_t)t * g.N + i] = hard_t[(tile * g.N + i) * 32 + lane];
        if (post)
            for (int i = 0; i < g.SL; ++i) post[(size_t)t * g.SL + i] = tpost[(tile * g.SL + i) * 32 + lane];
        if (post_sum)
            for (int i = 0; i < g.SLS; ++i) post_sum[(size_t)t * g.SLS + i] = tposts[(tile * g.SLS + i) * 32 + lane];
        if (status) status[t] = st[t];
        worst = std::min(worst, st[t]);
    }
    scaldpc_host_model_smem = nullptr;
    const bool second_round = early && it_run < iters && running > 0;  // cut short: the stragglers start again, compacted
    if (verify || second_round) {  // exact-order re-decode of the flagged trials (k_risk_scan + redecode_exact, qary.cu)
        std::vector<int> list;
        for (int t = 0; t < T; ++t)
            if ((risk[t / 32] >> (t % 32)) & 1u) list.push_back(t);
        const int nf = (int)list.size();
        if (!second_round) g_last_redecoded += nf;
        if (nf) {
            std::vector<float> l2((size_t)nf * g.SL), ls2(std::max<size_t>(1, (size_t)nf * g.SLS));
            std::vector<float> p2((size_t)nf * g.SL), ps2(std::max<size_t>(1, (size_t)nf * g.SLS));
            std::vector<int8_t> h2((size_t)nf * g.N);
            for (int i = 0; i < nf; ++i) {
                memcpy(&l2[(size_t)i * g.SL], llr_in + (size_t)list[i] * g.SL, sizeof(float) * g.SL);
                if (g.SLS) memcpy(&ls2[(size_t)i * g.SLS], llr_sum_in + (size_t)list[i] * g.SLS, sizeof(float) * g.SLS);
            }
            std::vector<int> it2(nf, 0);
            const unsigned f2 = second_round ? (flags | kHmNoRestart)
                                             : ((flags | SCALDPC_CN_EXACT) & ~(SCALDPC_CN_GENERIC | SCALDPC_QARY_EARLY_STOP));
            hm_qary_decode(H, R, N, b, bsum, dv, dc, max_iter, nf, l2.data(), g.SLS ? ls2.data() : nullptr, h2.data(), p2.data(),
                           g.SLS ? ps2.data() : nullptr, nullptr, f2, nullptr, nullptr, 0, it2.data());
            if (iters_out)
                for (int i = 0; i < nf; ++i) iters_out[list[i]] = it2[i];
            for (int i = 0; i < nf; ++i) {
                memcpy(hard + (size_t)list[i] * g.N, &h2[(size_t)i * g.N], g.N);
                if (post) memcpy(post + (size_t)list[i] * g.SL, &p2[(size_t)i * g.SL], sizeof(float) * g.SL);
                if (post_sum && g.SLS)
                    memcpy(post_sum + (size_t)list[i] * g.SLS, &ps2[(size_t)i * g.SLS], sizeof(float) * g.SLS);
            }
        }
    }
    return worst;
}

extern "C" {


// Binary product-sum BP through the modelled kernels, with the host loop of scaldpc_bp_decode (bp.cu): per-trial
// freeze, polling of the running-trial count and active-trial compaction into a second buffer set (`compact` != 0).
// synd (T x m bytes), probs (n or T x n doubles); outputs decoding (T x n bytes), llr (T x n floats, may be NULL),
// iters / conv (T ints), *n_compactions.
int hostmodel_bp_decode(int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int T,
                        const uint8_t *synd, const double *probs, int probs_per_trial, int early_stop, int compact,
                        uint8_t *decoding, float *llr, int32_t *iters, int32_t *conv, int32_t *n_compactions)
{
    BpHostGraph g;
    std::string err;
    int rc = build_bp_graph(m, n, row_ptr, col, g, err);
    if (rc) return rc;
    if (max_iter == 0) max_iter = n;
    const int tiles = (T + 31) / 32;
    const size_t lanes = (size_t)tiles * 32;
    struct Set {
        std::vector<float> msg, prior_t;
        std::vector<uint32_t> synd_w, valid, done, unsat;
        std::vector<int> orig;
    } set[2];
    for (int s = 0; s < 2; ++s) {
        const size_t t = s == 0 ? tiles : (tiles + 1) / 2;
        set[s].msg.assign(std::max<size_t>(1, t * 32 * g.E), 0.0f);
        set[s].prior_t.assign(probs_per_trial ? t * 32 * n : 1, 0.0f);
        set[s].synd_w.assign(t * m, 0);
        set[s].valid.assign(t, 0);
        set[s].done.assign(t, 0);
        set[s].unsat.assign(t, 0);
        set[s].orig.assign(t * 32, -1);
    }
    std::vector<float> prior(n), post(lanes * n);
    std::vector<uint32_t> dec_w((size_t)tiles * n, 0);
    std::vector<int> vit(lanes, max_iter), vconv(lanes, 0), counts(max_iter + 2, 0), src(lanes, -1);
    for (int t = 0; t < T; ++t) {
        set[0].valid[t / 32] |= 1u << (t % 32);
        set[0].orig[t] = t;
        for (int c = 0; c < m; ++c)
            if (synd[(size_t)t * m + c] & 1) set[0].synd_w[(size_t)(t / 32) * m + c] |= 1u << (t % 32);
    }
    if (probs_per_trial) {
        for (size_t t = 0; t < lanes; ++t)
            for (int v = 0; v < n; ++v) {
                const double q = t < (size_t)T ? probs[t * n + v] : 0.5;
                set[0].prior_t[((t / 32) * n + v) * 32 + t % 32] = (float)log((1.0 - q) / q);
            }
    } else {
        for (int v = 0; v < n; ++v) prior[v] = (float)log((1.0 - probs[v]) / probs[v]);
    }
    auto view = [&](int cur) {
        BpDev p;
        p.m = m; p.n = n; p.E = g.E; p.dc_max = std::max(1, g.dc_max); p.dv_max = std::max(1, g.dv_max);
        p.per_trial_prior = probs_per_trial;
        p.row_ptr = g.row_ptr.data(); p.col = g.col.data(); p.col_ptr = g.col_ptr.data(); p.csc_off = g.csc_off.data();
        p.vn_order = g.vn_order.data();
        p.msg = set[cur].msg.data();
        p.prior = probs_per_trial ? set[cur].prior_t.data() : prior.data();
        p.synd_w = set[cur].synd_w.data(); p.dec_w = dec_w.data();
        p.post = llr ? post.data() : nullptr;
        p.done = set[cur].done.data(); p.valid = set[cur].valid.data(); p.unsat = set[cur].unsat.data();
        p.orig = set[cur].orig.data(); p.iters = vit.data(); p.conv = vconv.data(); p.counts = counts.data();
        p.nrows = nullptr;
        return p;
    };
    auto write_outputs = [&](int cur, int cur_tiles, const std::vector<uint32_t> &mask) {
        for (int tile = 0; tile < cur_tiles; ++tile)
            for (int l = 0; l < 32; ++l) {
                if (!((mask[tile] >> l) & 1u)) continue;
                const int o = set[cur].orig[tile * 32 + l];
                for (int v = 0; v < n; ++v) {
                    decoding[(size_t)o * n + v] = (dec_w[(size_t)tile * n + v] >> l) & 1u;
                    if (llr) llr[(size_t)o * n + v] = post[((size_t)tile * n + v) * 32 + l];
                }
            }
    };
    const int dmax = std::max(std::max(1, g.dc_max), std::max(1, g.dv_max));
    const int cn_wpc = 2;  // warps per CTA of the modelled check pass (any value works: items are strided)
    std::vector<float> scratch((size_t)std::max(cn_wpc * 3, 2 * 3) * dmax * 32);
    scaldpc_host_model_smem = scratch.data();
    int cur = 0, cur_tiles = tiles, compactions = 0;
    for (int it = 1; it <= max_iter; ++it) {
        const BpDev p = view(cur);
        if (it == 1) run_grid(3, 1, cn_wpc * 32, k_bp_cn<true>, p, cur_tiles);
        else run_grid(3, 1, cn_wpc * 32, k_bp_cn<false>, p, cur_tiles);
        int lo = 0;
        for (int c = 0; c <= kBpVnClasses; ++c) {
            const int hi = g.class_end[c];
            if (hi > lo) {
                switch (c) {
                case 0: run_grid(3, 1, 2 * 32, k_bp_vn<0, 1, false>, p, lo, hi, cur_tiles, -1); break;
                case 1: run_grid(3, 1, 2 * 32, k_bp_vn<1, 4, false>, p, lo, hi, cur_tiles, -1); break;
                case 2: run_grid(3, 1, 2 * 32, k_bp_vn<4, 8, false>, p, lo, hi, cur_tiles, -1); break;
                case 3: run_grid(3, 1, 2 * 32, k_bp_vn<8, 12, false>, p, lo, hi, cur_tiles, -1); break;
                case 4: run_grid(3, 1, 2 * 32, k_bp_vn<12, 16, false>, p, lo, hi, cur_tiles, -1); break;
                case 5: run_grid(3, 1, 2 * 32, k_bp_vn<16, 24, false>, p, lo, hi, cur_tiles, -1); break;
                case 6: run_grid(3, 1, 2 * 32, k_bp_vn<24, 32, false>, p, lo, hi, cur_tiles, -1); break;
                default: run_grid((hi - lo + 1) / 2, cur_tiles, 2 * 32, k_bp_vn_generic<false>, p, lo, hi, -1);
                }
            }
            lo = hi;
        }
        run_grid((m + 255) / 256, cur_tiles, 256, k_bp_syndrome, p);
        run_grid((cur_tiles + 255) / 256, 1, 256, k_bp_finish_iter, p, cur_tiles, it, early_stop, (int)(it == max_iter));
        if (early_stop && it < max_iter) {
            const int n_run = counts[it];
            if (n_run == 0) break;
            if (compact && cur_tiles >= 2 && (long long)n_run * 2 <= (long long)cur_tiles * 32) {
                write_outputs(cur, cur_tiles, set[cur].done);
                Set &a = set[cur], &b = set[cur ^ 1];
                int n_act = 0;
                const int new_tiles = (n_run + 31) / 32;
                run_grid(1, 1, 1, k_bp_compact_plan, (const uint32_t *)a.valid.data(), (const uint32_t *)a.done.data(),
                         (const int *)a.orig.data(), cur_tiles, src.data(), b.orig.data(), b.valid.data(), b.done.data(),
                         b.unsat.data(), &n_act);
                if (n_act != n_run) return -100;
                run_grid((g.E + 63) / 64, new_tiles, 256, k_bp_compact_move, (const float *)a.msg.data(), b.msg.data(),
                         (const int *)src.data(), g.E);
                if (probs_per_trial)
                    run_grid((n + 63) / 64, new_tiles, 256, k_bp_compact_move, (const float *)a.prior_t.data(),
                             b.prior_t.data(), (const int *)src.data(), n);
                run_grid((m + 7) / 8, new_tiles, 256, k_bp_compact_bits, (const uint32_t *)a.synd_w.data(),
                         b.synd_w.data(), (const int *)src.data(), m);
                cur ^= 1;
                cur_tiles = new_tiles;
                ++compactions;
            }
        }
    }
    scaldpc_host_model_smem = nullptr;
    write_outputs(cur, cur_tiles, set[cur].valid);
    for (int t = 0; t < T; ++t) {
        iters[t] = vit[t];
        conv[t] = vconv[t];
    }
    if (n_compactions) *n_compactions = compactions;
    return 0;
}


// Segmented CSR (scaldpc_bp_create_multi + scaldpc_bp_decode_multi, bp.cu): G graphs of one shape, every graph's trials in
// tiles of their own, no compaction.  row_ptr = G x (m+1), col = the G column arrays back to back, graph_of[T].
int hostmodel_bp_decode_multi(int G, int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int T,
                              const int32_t *graph_of, const uint8_t *synd, const double *probs, int early_stop,
                              uint8_t *decoding, float *llr, int32_t *iters, int32_t *conv)
{
    std::vector<BpHostGraph> gs(G);
    int Emax = 0, dc_max = 1, dv_max = 1;
    size_t coff = 0;
    for (int g = 0; g < G; ++g) {
        std::string err;
        int rc = build_bp_graph(m, n, row_ptr + (size_t)g * (m + 1), col + coff, gs[g], err);
        if (rc) return rc;
        coff += gs[g].E;
        Emax = std::max(Emax, gs[g].E);
        dc_max = std::max(dc_max, gs[g].dc_max);
        dv_max = std::max(dv_max, gs[g].dv_max);
    }
    if (max_iter == 0) max_iter = n;
    std::vector<int> a_row_ptr((size_t)G * (m + 1)), a_col((size_t)G * Emax, 0), a_col_ptr((size_t)G * (n + 1)), a_vn((size_t)G * n);
    std::vector<uint32_t> a_csc((size_t)G * Emax, 0u);
    std::vector<int> ce((size_t)G * (kBpVnClasses + 1)), cmax(kBpVnClasses + 1, 0);
    for (int g = 0; g < G; ++g) {
        const BpHostGraph &q = gs[g];
        std::copy(q.row_ptr.begin(), q.row_ptr.end(), a_row_ptr.begin() + (size_t)g * (m + 1));
        std::copy(q.col.begin(), q.col.end(), a_col.begin() + (size_t)g * Emax);
        std::copy(q.col_ptr.begin(), q.col_ptr.end(), a_col_ptr.begin() + (size_t)g * (n + 1));
        std::copy(q.vn_order.begin(), q.vn_order.end(), a_vn.begin() + (size_t)g * n);
        std::copy(q.csc_off.begin(), q.csc_off.begin() + q.E, a_csc.begin() + (size_t)g * Emax);
        for (int c = 0; c <= kBpVnClasses; ++c) {
            ce[(size_t)g * (kBpVnClasses + 1) + c] = q.class_end[c];
            cmax[c] = std::max(cmax[c], q.class_end[c] - (c ? q.class_end[c - 1] : 0));
        }
    }
    // slots: every graph's trials fill tiles of their own (trial order preserved inside a graph)
    std::vector<int> cnt(G, 0), first(G, 0);
    for (int t = 0; t < T; ++t) {
        if (graph_of[t] < 0 || graph_of[t] >= G) return SCALDPC_ERR_ARG;
        ++cnt[graph_of[t]];
    }
    int tiles = 0;
    for (int g = 0; g < G; ++g) {
        first[g] = tiles;
        tiles += (cnt[g] + 31) / 32;
    }
    std::vector<int> tile_graph(std::max(1, tiles)), orig((size_t)std::max(1, tiles) * 32, -1);
    for (int g = 0; g < G; ++g)
        for (int k = 0; k < (cnt[g] + 31) / 32; ++k) tile_graph[first[g] + k] = g;
    std::fill(cnt.begin(), cnt.end(), 0);
    for (int t = 0; t < T; ++t) orig[(size_t)first[graph_of[t]] * 32 + cnt[graph_of[t]]++] = t;
    const size_t lanes = (size_t)tiles * 32;
    std::vector<float> msg(std::max<size_t>(1, lanes * Emax), 0.0f), prior(n), post(lanes * n);
    std::vector<uint32_t> synd_w((size_t)tiles * m, 0), valid(tiles, 0), done(tiles, 0), unsat(tiles, 0), dec_w((size_t)tiles * n, 0);
    std::vector<int> vit(T, max_iter), vconv(T, 0), counts(max_iter + 2, 0);
    run_grid((tiles + 255) / 256, 1, 256, k_bp_reset_slots, valid.data(), done.data(), unsat.data(), (const int *)orig.data(),
             vit.data(), vconv.data(), tiles, max_iter);
    for (size_t s_ = 0; s_ < lanes; ++s_) {
        const int t = orig[s_];
        if (t < 0) continue;
        for (int c = 0; c < m; ++c)
            if (synd[(size_t)t * m + c] & 1) synd_w[(s_ / 32) * m + c] |= 1u << (s_ % 32);
    }
    for (int v = 0; v < n; ++v) prior[v] = (float)log((1.0 - probs[v]) / probs[v]);
    BpDev p;
    p.m = m; p.n = n; p.E = Emax; p.dc_max = dc_max; p.dv_max = dv_max; p.per_trial_prior = 0;
    p.row_ptr = a_row_ptr.data(); p.col = a_col.data(); p.col_ptr = a_col_ptr.data(); p.csc_off = a_csc.data();
    p.vn_order = a_vn.data(); p.vn_desc = nullptr;
    p.msg = msg.data(); p.prior = prior.data(); p.synd_w = synd_w.data(); p.dec_w = dec_w.data();
    p.post = llr ? post.data() : nullptr;
    p.done = done.data(); p.valid = valid.data(); p.unsat = unsat.data(); p.orig = orig.data();
    p.iters = vit.data(); p.conv = vconv.data(); p.counts = counts.data(); p.nrows = nullptr;
    p.tile_graph = tile_graph.data(); p.class_end = ce.data();
    p.g_rp = m + 1; p.g_e = Emax; p.g_cp = n + 1; p.g_n = n;
    std::vector<float> scratch((size_t)6 * std::max(dc_max, dv_max) * 32);
    scaldpc_host_model_smem = scratch.data();
    // iteration 1 as bp.cu runs it with shared priors: one table of first-pass messages per graph, formed into the bit pass
    // (k_bp_vn<.., FIRST>) instead of a check pass over the message array
    std::vector<float> first_tab((size_t)G * Emax, 0.0f);
    std::vector<int> edge_row((size_t)G * Emax, m);
    for (int g = 0; g < G; ++g)
        for (int r = 0; r < m; ++r)
            for (int e = gs[g].row_ptr[r]; e < gs[g].row_ptr[r + 1]; ++e) edge_row[(size_t)g * Emax + e] = r;
    std::vector<float> csc_tab((size_t)G * Emax, 0.0f);
    std::vector<int> csc_row((size_t)G * Emax, m), edge_csc((size_t)G * Emax, 0);
    for (int g = 0; g < G; ++g)
        for (int i = 0; i < gs[g].E; ++i) {
            const int e = gs[g].csc_edge[i];
            edge_csc[(size_t)g * Emax + e] = i;
            csc_row[(size_t)g * Emax + i] = edge_row[(size_t)g * Emax + e];
        }
    p.first_tab = first_tab.data();
    p.edge_row = edge_row.data();
    p.csc_tab = csc_tab.data();
    p.csc_row = csc_row.data();
    p.edge_csc = edge_csc.data();
    gridDim.y = G;
    for (unsigned by = 0; by < (unsigned)G; ++by)
        for (unsigned bx = 0; bx < (unsigned)((m + 127) / 128); ++bx)
            for (unsigned t = 0; t < 128; ++t) {
                blockIdx.x = bx; blockIdx.y = by; threadIdx.x = t; blockDim.x = 128;
                k_bp_first_table(p, first_tab.data());
            }
    for (int it = 1; it <= max_iter; ++it) {
        const bool first = it == 1;
        if (!first) run_grid(3, 1, 2 * 32, k_bp_cn<false>, p, tiles);
        for (int c = 0; c <= kBpVnClasses; ++c) {
            const int nv = cmax[c];
            if (nv <= 0) continue;
#define VN_(A_, B_) (first ? run_grid(3, 1, 2 * 32, k_bp_vn<A_, B_, true>, p, 0, nv, tiles, c) \
                           : run_grid(3, 1, 2 * 32, k_bp_vn<A_, B_, false>, p, 0, nv, tiles, c))
            switch (c) {
            case 0: VN_(0, 1); break;
            case 1: VN_(1, 4); break;
            case 2: VN_(4, 8); break;
            case 3: VN_(8, 12); break;
            case 4: VN_(12, 16); break;
            case 5: VN_(16, 24); break;
            case 6: VN_(24, 32); break;
            default:
                if (first) run_grid((nv + 1) / 2, tiles, 2 * 32, k_bp_vn_generic<true>, p, 0, nv, c);
                else run_grid((nv + 1) / 2, tiles, 2 * 32, k_bp_vn_generic<false>, p, 0, nv, c);
            }
#undef VN_
        }
        run_grid((m + 255) / 256, tiles, 256, k_bp_syndrome, p);
        run_grid((tiles + 255) / 256, 1, 256, k_bp_finish_iter, p, tiles, it, early_stop, (int)(it == max_iter));
        if (early_stop && it < max_iter && counts[it] == 0) break;
    }
    scaldpc_host_model_smem = nullptr;
    for (size_t s_ = 0; s_ < lanes; ++s_) {
        const int t = orig[s_];
        if (t < 0) continue;
        for (int v = 0; v < n; ++v) {
            decoding[(size_t)t * n + v] = (dec_w[(s_ / 32) * n + v] >> (s_ % 32)) & 1u;
            if (llr) llr[(size_t)t * n + v] = post[((s_ / 32) * n + v) * 32 + s_ % 32];
        }
    }
    for (int t = 0; t < T; ++t) {
        iters[t] = vit[t];
        conv[t] = vconv[t];
    }
    return 0;
}

}  // extern "C"
