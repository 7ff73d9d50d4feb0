// q-ary min-sum BP kernels for sm_100a (B200).  Hand-written CUDA; no tensor cores (the path is an irregular
// gather plus small (min,+) convolutions).
//
// Data layout in HBM ("tile" layout): trials are grouped 32 to a tile, lane = trial % 32 is the innermost
// index of every per-trial array, so a warp (= one tile) touches one 128-byte line per scalar:
//     msg     [tile][S      ][32] f32   one slot per (edge, value); edges in CSR order (check-major, column
//                                        ascending, decoder.rs:507-539); holds v2c before a check-node pass and
//                                        c2v after it (updated in place -- every slot has exactly one writer)
//     llr     [tile][bvars*Q][32] f32   channel LLRs of the B-variables        (decoder.rs:668-692)
//     llr_sum [tile][R*QS   ][32] f32   channel LLRs of the sum-variables      (decoder_special.rs:487-493)
//     post / post_sum                   final `sum` vectors, same shapes        (decoder.rs:638-642)
//     hard_t  [tile][N      ][32] i8    hard decisions                          (decoder.rs:654-657)
// All messages are stored in the CHECK's frame (already mirrored when h < 0), as the reference stores v2c
// (decoder.rs:571).  Graph indices are warp-uniform and shared by all trials (L2 / constant-cache resident).
#pragma once
#include <stdint.h>
#include <string.h>
#ifdef SCALDPC_HOST_MODEL
// tests/hostmodel/hostmodel.cpp compiles these kernels with g++ (threads run one after another) so the kernel
// logic can be checked against the oracle on a machine without a GPU.  TEST ONLY -- never part of libscaldpc.so.
#define SCALDPC_DYN_SMEM(type, name) type *name = (type *)scaldpc_host_model_smem
#else
#include <cuda_runtime.h>
#include <math_constants.h>
#define SCALDPC_DYN_SMEM(type, name) extern __shared__ type name[]
#endif
#include "../../include/scaldpc.h"

#define SC_INF (__builtin_huge_valf())

namespace scaldpc {

// Real = float everywhere on the product path; double = the fp64 verification mode (scaldpc_qary_decode_f64), which runs
// the run-time-shape kernels (k_init_v2c, k_cn_generic exact-order, k_vn) on double messages.
template <class Real>
struct QaryDevT {
    int R, N, bvars, B, BSUM, Q, QS;
    int S;    // message slots per trial
    int SL;   // bvars * Q
    int SLS;  // R * QS (0 for the generic decoder)
    const int *row_ptr;     // R+1
    const int *col;         // E
    const int8_t *sign;     // E
    const int *edge_off;    // E+1 slot offset of each edge
    const int *col_ptr;     // N+1
    const int2 *csc;        // E: {slot offset, sign} of each variable's edges, rows ascending
    Real *msg;
    const Real *llr;
    const Real *llr_sum;
    Real *post;             // may be null
    Real *post_sum;         // may be null
    int8_t *hard_t;
    int *status;            // per trial, tiles*32 entries
    // at-risk detection of the fast (forward/backward) path, last iteration only; null = off.  Bit l of risk[tile] is set
    // when trial l's decisions could differ from the reference's rounding order: a posterior whose two smallest entries are
    // closer than risk_eps, or decisions that violate a parity check.
    uint32_t *risk = nullptr;
    float risk_eps = 0.0f;
    // opt-in early termination (SCALDPC_QARY_EARLY_STOP; an extension, the reference has none: decoder_special.rs:501-504):
    // bit l of done[tile] = trial l is FROZEN -- its hard decisions / posteriors of the iteration in which every row test
    // first held are final and are not written again; a tile whose 32 bits are all set is skipped altogether.  null = off.
    const uint32_t *done = nullptr;
};
using QaryDev = QaryDevT<float>;

// +inf, min (NaN operand ignored, f32::min / f64::min) and "NaN -> +inf" for either scalar type
template <class Real>
__device__ __forceinline__ Real sc_inf() { return (Real)__builtin_huge_valf(); }
__device__ __forceinline__ float sc_min(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ double sc_min(double a, double b) { return fmin(a, b); }

// Frozen-lane helpers for the kernels (warp-uniform word, per-lane bit).
template <class Real>
__device__ __forceinline__ uint32_t done_word(const QaryDevT<Real> &p, int tile) { return p.done ? __ldg(p.done + tile) : 0u; }

// The two smallest entries of a vector (NaN = excluded = +inf).
struct Min2 {
    float m1, m2;
    __device__ __forceinline__ Min2() : m1(SC_INF), m2(SC_INF) {}
    __device__ __forceinline__ void add(float x)
    {
        const float xs = fminf(x, SC_INF);
        m2 = fminf(m2, fmaxf(m1, xs));
        m1 = fminf(m1, xs);
    }
    // true unless the runner-up is clearly (> eps) above the minimum; all-excluded vectors (inf - inf = NaN) are risky
    __device__ __forceinline__ bool risky(float eps) const { return !(m2 - m1 > eps); }
};

// Every lane of the warp calls this with its own verdict; one atomic per warp.
template <class Real>
__device__ __forceinline__ void flag_risk(const QaryDevT<Real> &p, int tile, int lane, bool risky)
{
#ifdef SCALDPC_HOST_MODEL
    if (risky) p.risk[tile] |= 1u << lane;
#else
    const unsigned m = __ballot_sync(0xffffffffu, risky);
    if (m != 0u && lane == 0) atomicOr(p.risk + tile, m);
#endif
}

__device__ __forceinline__ float min3f(float a, float b, float c)
{
#ifdef SCALDPC_HOST_MODEL
    return fminf(fminf(a, b), c);
#else
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));  // FMNMX3 on sm_100a; NaN operands ignored
    return r;
#endif
}

// NaN -> +inf, everything else unchanged (min.f32 returns the non-NaN operand): "non-finite => excluded".
__device__ __forceinline__ float sanitize(float x) { return fminf(x, SC_INF); }
__device__ __forceinline__ double sanitize(double x) { return fmin(x, (double)SC_INF); }

// Running minimum that folds two new candidates per FMNMX3.
struct MinAcc {
    float best, pend;
    int n;  // candidates seen (compile-time constant after unrolling)
    __device__ __forceinline__ MinAcc() : best(SC_INF), pend(SC_INF), n(0) {}
    __device__ __forceinline__ void add(float x)
    {
        if (n == 0) best = x;
        else if ((n & 1) == 1) pend = x;
        else best = min3f(best, pend, x);
        ++n;
    }
    __device__ __forceinline__ float get() const { return (n > 1 && (n & 1) == 0) ? fminf(best, pend) : best; }
};

#include "qary_cn_fast.cuh"
#include "qary_cn_exact.cuh"
#include "sc_logf.cuh"

// -----------------------------------------------------------------------------------------------------------
// Variable-node kernel for the B-variables (decoder.rs:634-658 / decoder_special.rs:566-590).
// One warp = (tile, variable), lane = trial.  Exactly the reference's operation order:
//   sum = channel; for each check (row order): sum = sum + perm_h(c2v)
//   per edge: prim = perm_h(sum - perm_h(c2v)); v2c = prim - prim[first strict minimum]
//   last iteration: hard = argmin_first(sum) - B, posterior = sum
// -----------------------------------------------------------------------------------------------------------
constexpr int kVnWarps = 8;

template <int Q, class Real = float>
__global__ void __launch_bounds__(kVnWarps * 32) k_vn(const QaryDevT<Real> p, const int LAST)
{
    const Real INF = sc_inf<Real>();
    const int lane = threadIdx.x & 31;
    const int var = blockIdx.x * kVnWarps + (threadIdx.x >> 5);
    const int tile = blockIdx.y;
    if (var >= p.bvars) return;
    const uint32_t frozen_w = done_word(p, tile);
    if (frozen_w == 0xffffffffu) return;  // early termination: every trial of the tile is final
    const bool frozen = (frozen_w >> lane) & 1u;
    Real *msg = p.msg + (size_t)tile * p.S * 32 + lane;
    const Real *llr = p.llr + (size_t)tile * p.SL * 32 + lane;

    Real tot[Q];
#pragma unroll
    for (int i = 0; i < Q; ++i) tot[i] = llr[(size_t)(var * Q + i) * 32];
    const int p0 = __ldg(p.col_ptr + var), p1 = __ldg(p.col_ptr + var + 1);
    for (int e = p0; e < p1; ++e) {
        const int2 es = __ldg(p.csc + e);
        const bool neg = es.y < 0;
#pragma unroll
        for (int i = 0; i < Q; ++i) tot[i] = tot[i] + msg[(size_t)(es.x + (neg ? Q - 1 - i : i)) * 32];
    }
    for (int e = p0; e < p1; ++e) {
        const int2 es = __ldg(p.csc + e);
        const bool neg = es.y < 0;
        Real prim[Q];
        Real m = INF;
#pragma unroll
        for (int i = 0; i < Q; ++i) {  // i: check frame
            const Real c = msg[(size_t)(es.x + i) * 32];
            const Real t = neg ? tot[Q - 1 - i] : tot[i];
            prim[i] = t - c;
            m = sc_min(m, prim[i]);
        }
#pragma unroll
        for (int i = 0; i < Q; ++i) msg[(size_t)(es.x + i) * 32] = sanitize(prim[i] - m);  // NaN (excluded value) travels as +inf
    }
    if (LAST) {
        Real min_val = INF;
        int arg = 0;
#pragma unroll
        for (int i = 0; i < Q; ++i)
            if (tot[i] < min_val) {
                min_val = tot[i];
                arg = i;
            }
        if (!frozen) p.hard_t[((size_t)tile * p.N + var) * 32 + lane] = (int8_t)(arg - (Q - 1) / 2);
        if (p.risk) {
            Min2 mm;
#pragma unroll
            for (int i = 0; i < Q; ++i) mm.add((float)tot[i]);
            flag_risk(p, tile, lane, !frozen && mm.risky(p.risk_eps));
        }
        if (p.post && !frozen) {
            Real *post = p.post + (size_t)tile * p.SL * 32 + lane;
#pragma unroll
            for (int i = 0; i < Q; ++i) post[(size_t)(var * Q + i) * 32] = tot[i];
        }
    }
}

// Same update for graphs whose B-variables all have degree exactly D (every Kyber H' built by make_code.py:72-94):
// the variable's whole neighbourhood -- D x Q incoming messages + Q channel LLRs -- is loaded into registers by
// D*Q + Q independent loads issued back to back (k_vn loads one edge at a time and reads every message twice), then
// updated and stored.  Operation order is k_vn's, so results are bit-identical.
template <int Q, int D>
__global__ void __launch_bounds__(kVnWarps * 32) k_vn_reg(const QaryDev p, const int LAST)
{
    const float INF = SC_INF;
    const int lane = threadIdx.x & 31;
    const int var = blockIdx.x * kVnWarps + (threadIdx.x >> 5);
    const int tile = blockIdx.y;
    if (var >= p.bvars) return;
    const uint32_t frozen_w = done_word(p, tile);
    if (frozen_w == 0xffffffffu) return;  // early termination: every trial of the tile is final
    const bool frozen = (frozen_w >> lane) & 1u;
    float *msg = p.msg + (size_t)tile * p.S * 32 + lane;
    const float *llr = p.llr + (size_t)tile * p.SL * 32 + lane + (size_t)var * Q * 32;
    const int p0 = var * D;  // regular columns: col_ptr[var] == var * D

    int2 es[D];
#pragma unroll
    for (int e = 0; e < D; ++e) es[e] = __ldg(p.csc + p0 + e);
    float tot[Q], c[D][Q];
#pragma unroll
    for (int e = 0; e < D; ++e) {
        const float *src = msg + (size_t)es[e].x * 32;
#pragma unroll
        for (int i = 0; i < Q; ++i) c[e][i] = src[i * 32];
    }
#pragma unroll
    for (int i = 0; i < Q; ++i) tot[i] = llr[i * 32];
#pragma unroll
    for (int e = 0; e < D; ++e) {
        if (es[e].y < 0) {  // warp-uniform
#pragma unroll
            for (int i = 0; i < Q; ++i) tot[i] = tot[i] + c[e][Q - 1 - i];
        } else {
#pragma unroll
            for (int i = 0; i < Q; ++i) tot[i] = tot[i] + c[e][i];
        }
    }
#pragma unroll
    for (int e = 0; e < D; ++e) {
        float prim[Q];
        float m = INF;
        if (es[e].y < 0) {
#pragma unroll
            for (int i = 0; i < Q; ++i) prim[i] = tot[Q - 1 - i] - c[e][i];
        } else {
#pragma unroll
            for (int i = 0; i < Q; ++i) prim[i] = tot[i] - c[e][i];
        }
#pragma unroll
        for (int i = 0; i < Q; ++i) m = fminf(m, prim[i]);
        float *dst = msg + (size_t)es[e].x * 32;
#pragma unroll
        for (int i = 0; i < Q; ++i) dst[i * 32] = sanitize(prim[i] - m);  // NaN (excluded value) travels as +inf
    }
    if (LAST) {
        float min_val = INF;
        int arg = 0;
#pragma unroll
        for (int i = 0; i < Q; ++i)
            if (tot[i] < min_val) {
                min_val = tot[i];
                arg = i;
            }
        if (!frozen) p.hard_t[((size_t)tile * p.N + var) * 32 + lane] = (int8_t)(arg - (Q - 1) / 2);
        if (p.risk) {
            Min2 mm;
#pragma unroll
            for (int i = 0; i < Q; ++i) mm.add(tot[i]);
            flag_risk(p, tile, lane, !frozen && mm.risky(p.risk_eps));
        }
        if (p.post && !frozen) {
            float *post = p.post + (size_t)tile * p.SL * 32 + lane;
#pragma unroll
            for (int i = 0; i < Q; ++i) post[(size_t)(var * Q + i) * 32] = tot[i];
        }
    }
}

// First-iteration initialisation of the B-edges for the runtime-shape path: v2c = perm_h(channel)
// (decoder.rs:567-573).  One warp = (tile, edge).
template <class Real = float>
__global__ void __launch_bounds__(256) k_init_v2c(const QaryDevT<Real> p, int E)
{
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int tile = blockIdx.y;
    if (e >= E) return;
    const int c = __ldg(p.col + e);
    if (c >= p.bvars) return;  // sum-variable edges are handled inside the check-node kernel
    const bool neg = __ldg(p.sign + e) < 0;
    const int off = __ldg(p.edge_off + e);
    Real *msg = p.msg + (size_t)tile * p.S * 32 + lane;
    const Real *llr = p.llr + (size_t)tile * p.SL * 32 + lane;
    for (int i = 0; i < p.Q; ++i) msg[(size_t)(off + i) * 32] = llr[(size_t)(c * p.Q + (neg ? p.Q - 1 - i : i)) * 32];
}

// -----------------------------------------------------------------------------------------------------------
// Runtime-shape check-node kernel: any row degree, any alphabets, both decoders, two algorithms.
//   exact != 0   "exact-order" DP: P_0 = {0:0}, P_{k+1}[u] = min_d fl(P_k[u-d] + a_k[d]); for every (j, d) the DP
//                is restarted from P_j with variable j pinned to d and the result is fl(M - a_j[d]).  Because
//                fl(x + c) is monotone in x this reproduces the reference's min over sequentially-rounded
//                configuration sums (decoder.rs:600-629, decoder_special.rs:539-553) BIT FOR BIT.
//   exact == 0   forward/backward as in k_cn_fast.
// One CTA = one warp = (tile, check); per-thread scratch lives in shared memory as [index][lane] (bank = lane,
// conflict-free).  Handles the fused sum-variable update like k_cn_fast.  Generic decoder (BSUM == 0): a check
// without any finite configuration raises SCALDPC_ERR_NO_CONFIG for that trial (decoder.rs:618).
// -----------------------------------------------------------------------------------------------------------
struct CnGenericArgs {
    int exact, first, last;
    int QM;    // max(Q, QS): stride of the per-variable input vectors in scratch
    int W;     // width of the DP tables (2 * sum of alphabets + 1, clamped per check at run time)
    int LMAX;  // max row degree
};

template <class Real = float>
__global__ void __launch_bounds__(32) k_cn_generic(const QaryDevT<Real> p, const CnGenericArgs g)
{
    SCALDPC_DYN_SMEM(double, smem_generic_raw);  // (double: 8-byte alignment for either scalar type)
    Real *smem = reinterpret_cast<Real *>(smem_generic_raw);
    const Real INF = sc_inf<Real>();
    const int lane = threadIdx.x;
    const int check = blockIdx.x, tile = blockIdx.y;
    const int e0 = __ldg(p.row_ptr + check), L = __ldg(p.row_ptr + check + 1) - e0;
    if (L <= 0) return;
    const uint32_t frozen_w = done_word(p, tile);
    if (frozen_w == 0xffffffffu) return;
    const bool frozen = (frozen_w >> lane) & 1u;
    Real *msg = p.msg + (size_t)tile * p.S * 32 + lane;
    const bool special = p.BSUM > 0;

    // scratch: a[L][QM], P[(L+1)][W], A[W], Bv[W]   (all x 32 lanes)
    Real *sa = smem + lane;
    Real *sP = sa + (size_t)g.LMAX * g.QM * 32;
    Real *sA = sP + (size_t)(g.LMAX + 1) * g.W * 32;
    Real *sB = sA + (size_t)g.W * 32;
#define A_(k, i) sa[((k)*g.QM + (i)) * 32]
#define P_(k, u) sP[((k)*g.W + (u)) * 32]
#define TA_(u) sA[(u)*32]
#define TB_(u) sB[(u)*32]

    // Half-width of the DP tables: a partial sum after k variables matters only if the remaining ones can bring it back
    // to 0, |u| <= min(sum of the first k alphabets, sum of the others); `tot` = the largest such bound over k is the
    // centre offset of every table (a uniform window: entries outside it belong to no valid configuration).
    int all = 0;
    for (int k = 0; k < L; ++k) all += (__ldg(p.col + e0 + k) >= p.bvars) ? p.BSUM : p.B;
    int tot = 0;
    for (int k = 0, pre = 0; k < L; ++k) {
        pre += (__ldg(p.col + e0 + k) >= p.bvars) ? p.BSUM : p.B;
        tot = max(tot, min(pre, all - pre));
    }
    const int W = 2 * tot + 1;  // <= g.W

    // ---- load inputs (sanitised copies in scratch)
    for (int k = 0; k < L; ++k) {
        const int c = __ldg(p.col + e0 + k);
        const int off = __ldg(p.edge_off + e0 + k);
        if (c >= p.bvars) {  // sum-variable: fused variable-node update
            const int vs = c - p.bvars;
            const bool neg = __ldg(p.sign + e0 + k) < 0;
            const Real *llrs = p.llr_sum + (size_t)tile * p.SLS * 32 + lane;
            Real m = INF;
            for (int i = 0; i < p.QS; ++i) {
                Real x = llrs[(size_t)(vs * p.QS + (neg ? p.QS - 1 - i : i)) * 32];
                if (!g.first) {
                    const Real cc = msg[(size_t)(off + i) * 32];
                    const Real t = x + cc;
                    x = t - cc;
                    m = sc_min(m, x);
                }
                A_(k, i) = x;
            }
            for (int i = 0; i < p.QS; ++i) A_(k, i) = sanitize(g.first ? A_(k, i) : A_(k, i) - m);
        } else {
            for (int i = 0; i < p.Q; ++i) A_(k, i) = sanitize(msg[(size_t)(off + i) * 32]);
        }
    }
#define BK_(k) ((__ldg(p.col + e0 + (k)) >= p.bvars) ? p.BSUM : p.B)

    // ---- forward table P_k[u], u offset by tot
    for (int u = 0; u < W; ++u) P_(0, u) = INF;
    P_(0, tot) = (Real)0;
    for (int k = 0; k < L; ++k) {
        const int bk = BK_(k);
        for (int u = 0; u < W; ++u) {
            Real best = INF;
            for (int i = 0; i <= 2 * bk; ++i) {
                const int up = u - (i - bk);
                if (up < 0 || up >= W) continue;
                best = sc_min(best, P_(k, up) + A_(k, i));
            }
            P_(k + 1, u) = best;
        }
    }
    if (!special && !(P_(L, tot) < INF)) {
        atomicMin(p.status + tile * 32 + lane, SCALDPC_ERR_NO_CONFIG);
    }

    if (g.exact) {
        for (int j = 0; j < L; ++j) {
            const int bj = BK_(j);
            const int offj = __ldg(p.edge_off + e0 + j);
            for (int i = 0; i <= 2 * bj; ++i) {
                const Real x = A_(j, i);
                Real r = INF;
                if (x < INF) {
                    for (int u = 0; u < W; ++u) {
                        const int up = u - (i - bj);
                        TA_(u) = (up >= 0 && up < W) ? P_(j, up) + x : INF;
                    }
                    for (int k = j + 1; k < L; ++k) {
                        const int bk = BK_(k);
                        for (int u = 0; u < W; ++u) {
                            Real best = INF;
                            for (int i2 = 0; i2 <= 2 * bk; ++i2) {
                                const int up = u - (i2 - bk);
                                if (up < 0 || up >= W) continue;
                                best = sc_min(best, TA_(up) + A_(k, i2));
                            }
                            TB_(u) = best;
                        }
                        for (int u = 0; u < W; ++u) TA_(u) = TB_(u);
                    }
                    const Real M = TA_(tot);
                    r = (M < INF) ? M - x : INF;
                }
                msg[(size_t)(offj + i) * 32] = r;
            }
        }
    } else {
        // backward rolling table: TA = G_{k+1}[u] = cost of the suffix when it must total (u - tot)
        for (int u = 0; u < W; ++u) TA_(u) = INF;
        TA_(tot) = (Real)0;
        for (int j = L - 1; j >= 0; --j) {
            const int bj = BK_(j);
            const int offj = __ldg(p.edge_off + e0 + j);
            for (int i = 0; i <= 2 * bj; ++i) {
                Real r = INF;
                if (A_(j, i) < INF) {
                    const int d = i - bj;
                    for (int u = 0; u < W; ++u) {
                        const int us = -(u - tot) - d + tot;
                        if (us < 0 || us >= W) continue;
                        r = sc_min(r, P_(j, u) + TA_(us));
                    }
                }
                msg[(size_t)(offj + i) * 32] = r;
            }
            if (j > 0) {
                for (int u = 0; u < W; ++u) {
                    Real best = INF;
                    for (int i = 0; i <= 2 * bj; ++i) {
                        const int up = u - (i - bj);
                        if (up < 0 || up >= W) continue;
                        best = sc_min(best, A_(j, i) + TA_(up));
                    }
                    TB_(u) = best;
                }
                for (int u = 0; u < W; ++u) TA_(u) = TB_(u);
            }
        }
    }

    // ---- last iteration: decision of the sum-variable
    if (g.last && special) {
        const int k = L - 1;
        const int c = __ldg(p.col + e0 + k);
        if (c >= p.bvars) {
            const int vs = c - p.bvars;
            const bool neg = __ldg(p.sign + e0 + k) < 0;
            const int off = __ldg(p.edge_off + e0 + k);
            const Real *llrs = p.llr_sum + (size_t)tile * p.SLS * 32 + lane;
            Real *post = p.post_sum ? p.post_sum + (size_t)tile * p.SLS * 32 + lane : nullptr;
            Real min_val = INF;
            int arg = 0;
            Min2 mm;
            for (int iv = 0; iv < p.QS; ++iv) {  // variable frame, ascending
                const int i = neg ? p.QS - 1 - iv : iv;
                const Real t = llrs[(size_t)(vs * p.QS + iv) * 32] + msg[(size_t)(off + i) * 32];
                if (t < min_val) {
                    min_val = t;
                    arg = iv;
                }
                mm.add((float)t);
                if (post && !frozen) post[(size_t)(vs * p.QS + iv) * 32] = t;
            }
            if (!frozen) p.hard_t[((size_t)tile * p.N + p.bvars + vs) * 32 + lane] = (int8_t)(arg - p.BSUM);
            if (p.risk) flag_risk(p, tile, lane, !frozen && mm.risky(p.risk_eps));
        }
    }
#undef A_
#undef P_
#undef TA_
#undef TB_
#undef BK_
}

// -----------------------------------------------------------------------------------------------------------
// Check node of the GENERIC decoder (decoder.rs:585-631) when every variable has the same alphabet [-B, B]
// (BSUM == 0): k_cn_generic's forward/backward algorithm with the alphabet as a template parameter -- inner loops
// over the Q values fully unrolled, tables padded by B entries of +inf on both sides so that no index is ever
// tested, every stage restricted to the partial sums that can still return to 0 (|u| <= r_k = min(sum of the first k
// alphabets, sum of the others)).  One warp = (tile, check), lane = trial, tables in shared memory as [index][lane].
//   forward   P_0 = {0: 0},  P_{k+1}[u] = min_d P_k[u - d] + a_k[d]
//   backward  G_L = {0: 0},  G_j[t] = min_d a_j[d] + G_{j+1}[t - d]          (t = total of variables j .. L-1)
//   output    c2v_j[d] = min_u P_j[u] + G_{j+1}[-u - d],  +inf where a_j[d] is not finite (decoder.rs:307-311)
// A trial whose check has no finite configuration gets SCALDPC_ERR_NO_CONFIG (decoder.rs:618).
// The reference's criterion case (benches/decoder.rs, Decoder<450,150,3,7,15,7>, 10^4 trials): k_cn_generic 202 ms per
// launch (full-range tables, one warp per SM), 30 ms with the reachable window, 4.1 ms here (u loops unrolled by 2 for
// loads in flight; P_0 and P_L are scalars, which leaves 72 KB per warp = three CTAs per SM: the pass is bound by
// shared-memory latency).
// -----------------------------------------------------------------------------------------------------------
// Register tiles of k_cn_uniform.  conv: TU outputs dst[t] = min_i (src[t - i + Q - 1] + a[i]) from ONE window of
// Q + TU - 1 consecutive table entries (src -> entry of the lowest index; entries are 32 floats apart), candidates in
// the order i = 0 .. Q-1 as the scalar code.  comb: folds TU values f[t] of P_j into out[i] = min(out[i], f[t] + g[t + i]),
// g[jj] = the entry jj BELOW gtop.
template <int Q, int TU>
__device__ __forceinline__ void cn_uniform_conv(const float *src, const float (&a)[Q], float *dst)
{
    float w[Q + TU - 1];
#pragma unroll
    for (int j = 0; j < Q + TU - 1; ++j) w[j] = src[j * 32];
#pragma unroll
    for (int t = 0; t < TU; ++t) {
        MinAcc acc;
#pragma unroll
        for (int i = 0; i < Q; ++i) acc.add(w[t - i + Q - 1] + a[i]);
        dst[t * 32] = acc.get();
    }
}

template <int Q, int TU>
__device__ __forceinline__ void cn_uniform_comb(const float *fsrc, const float *gtop, float (&out)[Q])
{
    float f[TU], g[Q + TU - 1];
#pragma unroll
    for (int t = 0; t < TU; ++t) f[t] = fsrc[t * 32];
#pragma unroll
    for (int jj = 0; jj < Q + TU - 1; ++jj) g[jj] = gtop[-jj * 32];
#pragma unroll
    for (int i = 0; i < Q; ++i) {
        if (TU == 1) {
            out[i] = fminf(out[i], f[0] + g[i]);
        } else {
#pragma unroll
            for (int t = 0; t < TU; t += 2) out[i] = min3f(out[i], f[t] + g[t + i], f[t + 1] + g[t + 1 + i]);
        }
    }
}

// Shared-memory rows (of 32 floats) of k_cn_uniform for rows of up to L variables of alphabet [-B, B], window half
// width HH: the middle variables' messages, P_1 .. P_{L-1} at their exact padded widths, two G buffers.
__host__ __device__ inline int cn_uniform_reach(int k, int L, int B, int HH)
{
    const int a = k * B, b = (L - k) * B;
    const int m = a < b ? a : b;
    return m < HH ? m : HH;
}
__host__ __device__ inline int cn_uniform_p_rows(int L, int B, int HH)
{
    int rows = 0;
    for (int k = 1; k <= L - 1; ++k) rows += 2 * (cn_uniform_reach(k + 1, L, B, HH) + B) + 1;
    return rows;
}
__host__ __device__ inline int cn_uniform_rows(int L, int Q, int HH)
{
    const int B = (Q - 1) / 2;
    return (L > 2 ? (L - 2) * Q : 0) + cn_uniform_p_rows(L, B, HH) + 2 * (2 * (HH + B) + 1);
}

// One CTA (one warp) = one (check, tile); lane = trial.  Shared memory per item: the messages of the middle variables
// (the first and the last stay in registers), P_1 .. P_{L-1} at their exact padded widths 2 (r_{k+1} + B) + 1 and two
// G buffers -- 55 KB for (L, B) = (7, 7), so FOUR items are resident per SM, one warp per scheduler.
template <int Q>
__global__ void __launch_bounds__(32) k_cn_uniform(const QaryDev p, const int LMAX, const int HH)
{
    SCALDPC_DYN_SMEM(float, smem);
    constexpr int B = (Q - 1) / 2;
    const float INF = SC_INF;
    const int lane = threadIdx.x;
    const int check = blockIdx.x, tile = blockIdx.y;
    const int e0 = __ldg(p.row_ptr + check), L = __ldg(p.row_ptr + check + 1) - e0;
    if (L <= 0) return;
    if (done_word(p, tile) == 0xffffffffu) return;
    float *msg = p.msg + (size_t)tile * p.S * 32 + lane;
    const int WP = 2 * (HH + B) + 1;  // width of a G buffer; index of partial sum u is u + HH + B
    float *sA = smem + lane;                                                     // [L-2][Q]: variables 1 .. L-2
    float *sP = sA + (size_t)(LMAX > 2 ? LMAX - 2 : 0) * Q * 32;                 // P_1 .. P_{L-1}, back to back
    float *sG = sP + (size_t)cn_uniform_p_rows(LMAX, B, HH) * 32;                // [2][WP]
#define UA_(k, i) sA[(((k)-1) * Q + (i)) * 32]
#define UG_(b, u) sG[((b)*WP + (u) + HH + B) * 32]
    auto reach = [&](int k) { return cn_uniform_reach(k, L, B, HH); };  // r_k
    auto phw = [&](int k) { return reach(k + 1) + B; };                 // half width of the stored P_k

    float afirst[Q], alast[Q];  // messages of variable 0 and variable L-1
    {
        const int off0 = __ldg(p.edge_off + e0), offl = __ldg(p.edge_off + e0 + L - 1);
#pragma unroll
        for (int i = 0; i < Q; ++i) afirst[i] = sanitize(msg[(size_t)(off0 + i) * 32]);
#pragma unroll
        for (int i = 0; i < Q; ++i) alast[i] = sanitize(msg[(size_t)(offl + i) * 32]);
    }
    for (int k = 1; k < L - 1; ++k) {
        const int off = __ldg(p.edge_off + e0 + k);
#pragma unroll
        for (int i = 0; i < Q; ++i) UA_(k, i) = sanitize(msg[(size_t)(off + i) * 32]);
    }
    const int prows = cn_uniform_p_rows(L, B, HH);
    for (int i = 0; i < prows; ++i) sP[(size_t)i * 32] = INF;
    for (int i = 0; i < 2 * WP; ++i) sG[(size_t)i * 32] = INF;

    // ---- forward: P_1 = 0 + a_0, then P_{k+1} from P_k for k = 1 .. L-2; P_L only at u = 0
    float pl0;           // P_L[0]: the cheapest complete configuration
    int poff = 0;        // first row of the current P table
    float *Pc = sP;      // its entry u = 0
    if (L == 1) {
        pl0 = 0.0f + afirst[B];
    } else {
        const int r1 = reach(1);
        Pc = sP + (size_t)(poff + phw(1)) * 32;
#pragma unroll
        for (int i = 0; i < Q; ++i)
            if (i - B >= -r1 && i - B <= r1) Pc[(i - B) * 32] = 0.0f + afirst[i];
        for (int k = 1; k < L - 1; ++k) {
            float a[Q];
#pragma unroll
            for (int i = 0; i < Q; ++i) a[i] = UA_(k, i);
            const int r = reach(k + 1);
            const int noff = poff + 2 * phw(k) + 1;
            float *Pn = sP + (size_t)(noff + phw(k + 1)) * 32;
            int u = -r;  // P_{k+1}[u] = min_i P_k[u - (i - B)] + a[i]
            for (; u + 7 <= r; u += 8) cn_uniform_conv<Q, 8>(Pc + (u - B) * 32, a, Pn + u * 32);
            if (u + 3 <= r) {
                cn_uniform_conv<Q, 4>(Pc + (u - B) * 32, a, Pn + u * 32);
                u += 4;
            }
            for (; u <= r; ++u) cn_uniform_conv<Q, 1>(Pc + (u - B) * 32, a, Pn + u * 32);
            poff = noff;
            Pc = Pn;
        }
        MinAcc acc;  // Pc = P_{L-1}
#pragma unroll
        for (int i = 0; i < Q; ++i) acc.add(Pc[-(i - B) * 32] + alast[i]);
        pl0 = acc.get();
    }
    if (!(pl0 < INF)) atomicMin(p.status + tile * 32 + lane, SCALDPC_ERR_NO_CONFIG);

    // ---- backward + combine; G_{j+1} lives in sG[cur]; Pc / poff walk back from P_{L-1} to P_1
    int cur = 0;
    UG_(0, 0) = 0.0f;
    for (int j = L - 1; j >= 0; --j) {
        float a[Q];
        if (j == L - 1) {
#pragma unroll
            for (int i = 0; i < Q; ++i) a[i] = alast[i];
        } else if (j == 0) {
#pragma unroll
            for (int i = 0; i < Q; ++i) a[i] = afirst[i];
        } else {
#pragma unroll
            for (int i = 0; i < Q; ++i) a[i] = UA_(j, i);
        }
        const int r = reach(j);
        // out[i] = min over u of P_j[u] + G_{j+1}[-u - (i - B)]: every candidate is one addition and no candidate is NaN or
        // -0 (sanitised inputs), so the minimum does not depend on how the candidates are grouped
        float out[Q];
        if (j == 0) {
#pragma unroll
            for (int i = 0; i < Q; ++i) out[i] = 0.0f + UG_(cur, -(i - B));
        } else {
#pragma unroll
            for (int i = 0; i < Q; ++i) out[i] = INF;
            int u = -r;
            for (; u + 7 <= r; u += 8) cn_uniform_comb<Q, 8>(Pc + u * 32, &UG_(cur, -u + B), out);
            if (u + 3 <= r) {
                cn_uniform_comb<Q, 4>(Pc + u * 32, &UG_(cur, -u + B), out);
                u += 4;
            }
            for (; u <= r; ++u) cn_uniform_comb<Q, 1>(Pc + u * 32, &UG_(cur, -u + B), out);
        }
        const int offj = __ldg(p.edge_off + e0 + j);
#pragma unroll
        for (int i = 0; i < Q; ++i) msg[(size_t)(offj + i) * 32] = (a[i] < INF) ? out[i] : INF;
        if (j > 0) {
            const int nxt = cur ^ 1;
            for (int i = 0; i < WP; ++i) sG[((size_t)nxt * WP + i) * 32] = INF;
            int t = -r;  // G_j[t] = min_i a[i] + G_{j+1}[t - (i - B)]
            for (; t + 7 <= r; t += 8) cn_uniform_conv<Q, 8>(&UG_(cur, t - B), a, &UG_(nxt, t));
            if (t + 3 <= r) {
                cn_uniform_conv<Q, 4>(&UG_(cur, t - B), a, &UG_(nxt, t));
                t += 4;
            }
            for (; t <= r; ++t) cn_uniform_conv<Q, 1>(&UG_(cur, t - B), a, &UG_(nxt, t));
            cur = nxt;
            if (j > 1) {
                poff -= 2 * phw(j - 1) + 1;
                Pc = sP + (size_t)(poff + phw(j - 1)) * 32;
            }
        }
    }
#undef UA_
#undef UG_
}

// -----------------------------------------------------------------------------------------------------------
// Layout conversion kernels (host row-major [T][n] <-> tile layout [tile][n][32]) with a shared-memory transpose
// so both sides are coalesced.  k_load_priors also applies into_llr (decoder.rs:668-692).
// -----------------------------------------------------------------------------------------------------------
constexpr int kTrRows = 8;  // threads.y

// into_llr of one prior row, in place (decoder.rs:668-692): f32 sum in index order (assert 1 +- 0.001, :683-684),
// NaN-ignoring maximum (:671-680), llr[i] = ln(max / p[i]) (:687) with an IEEE f32 division and glibc's logf bit for
// bit (sc_logf.cuh).  Returns 0 or the SCALDPC_ERR_* the reference panics with.
__device__ __forceinline__ int into_llr_row(float *v, int q, const ScLogfTab *ltab)
{
    float sum = 0.0f, mx = 0.0f;
    bool have = false;
    for (int i = 0; i < q; ++i) {
        const float x = v[i];
        sum += x;
        if (x == x && (!have || x > mx)) {
            mx = x;
            have = true;
        }
    }
    int err = 0;
    if (!have) err = SCALDPC_ERR_NO_MAX;
    else if (!(sum < 1.0f + 0.001f) || !(sum > 1.0f - 0.001f)) err = SCALDPC_ERR_PROB_SUM;
    for (int i = 0; i < q; ++i) v[i] = sc_logf(__fdiv_rn(mx, v[i]), ltab);
    return err;
}

// in: [T][rows*q] probabilities or LLRs; out: [tile][rows*q][32].  One CTA = one tile x `rpb` rows.
// `map` (may be null): slot t of the batch is row map[t] of `in` (re-decode of a sub-list of trials).
__global__ void __launch_bounds__(32 * kTrRows) k_load_priors(const float *__restrict__ in, float *__restrict__ out,
                                                              int T, int rows, int q, int rpb, int is_llr,
                                                              int *__restrict__ status, const int *__restrict__ map)
{
    SCALDPC_DYN_SMEM(float, sm);
    const int tile = blockIdx.y;
    const int r0 = blockIdx.x * rpb;
    const int nr = min(rpb, rows - r0);
    const int n = nr * q;             // floats per trial in this chunk
    const int pitch = (rpb * q) | 1;  // odd pitch: conflict-free column access
    const int tx = threadIdx.x, ty = threadIdx.y;
    const size_t per_trial = (size_t)rows * q;
    __shared__ ScLogfTab ltab[16];  // per-lane table index: shared memory, not the constant bank
    if (ty == 0 && tx < 16) ltab[tx] = sc_logf_tab[tx];
    // coalesced load: thread tx walks the chunk, ty walks trials
    for (int t = ty; t < 32; t += kTrRows) {
        const int trial = tile * 32 + t;
        const size_t row = trial < T ? (size_t)(map ? map[trial] : trial) : 0;
        for (int i = tx; i < n; i += 32)
            sm[t * pitch + i] = trial < T ? in[row * per_trial + (size_t)r0 * q + i] : (is_llr ? 0.0f : 1.0f / q);
    }
    __syncthreads();
    if (!is_llr) {
        // thread (tx = trial, ty = row): sum / max / ln(max / p)
        for (int r = ty; r < nr; r += kTrRows) {
            const int err = into_llr_row(sm + tx * pitch + r * q, q, ltab);
            if (err && status && tile * 32 + tx < T) atomicMin(status + tile * 32 + tx, err);
        }
        __syncthreads();
    }
    float *o = out + ((size_t)tile * per_trial + (size_t)r0 * q) * 32;
    for (int i = ty; i < n; i += kTrRows) o[(size_t)i * 32 + tx] = sm[tx * pitch + i];
}

// fp64 verification mode: probabilities (f32, row-major [T][rows*q], pydecoder.rs:53) -> LLRs in double, tile layout.
// into_llr (decoder.rs:668-692) with every operation in double: sum and maximum of the row, ln(max / p).  One warp =
// (tile, variable), lane = trial; speed is no concern here.  CUDA's log() is within 1 ulp, not correctly rounded: the
// results agree with a CPU f64 evaluation to ~1e-15 relative, not bit for bit.
__global__ void __launch_bounds__(256) k_load_priors_f64(const float *__restrict__ in, double *__restrict__ out, int T, int rows,
                                                         int q, int is_llr, int *__restrict__ status)
{
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int tile = blockIdx.y;
    if (row >= rows) return;
    const int trial = tile * 32 + lane;
    const bool live = trial < T;
    const float *v = in + ((size_t)(live ? trial : 0) * rows + row) * q;
    double *o = out + ((size_t)tile * rows + row) * q * 32 + lane;
    if (is_llr || !live) {
        for (int i = 0; i < q; ++i) o[(size_t)i * 32] = live ? (double)v[i] : 0.0;
        return;
    }
    double sum = 0.0, mx = 0.0;
    bool have_max = false;
    for (int i = 0; i < q; ++i) {
        const double x = (double)v[i];
        sum += x;
        if (x == x && (!have_max || x > mx)) {  // NotNan::new filters NaN (decoder.rs:676-680)
            mx = x;
            have_max = true;
        }
    }
    int err = 0;
    if (!have_max) err = SCALDPC_ERR_NO_MAX;
    else if (!(sum < 1.0 + 0.001) || !(sum > 1.0 - 0.001)) err = SCALDPC_ERR_PROB_SUM;
    if (err && status) atomicMin(status + trial, err);
    for (int i = 0; i < q; ++i) o[(size_t)i * 32] = log(mx / (double)v[i]);
}

// ---- priors given as ORACLE OBSERVATIONS -----------------------------------------------------
// The prior of a variable depends only on the bit pattern y the side-channel oracle returned for it
// (max_likelihood.py:87-105 s_distribution_from_hard_y; kyber.py:362-376 get_channel_probabilities), so a batch of
// priors is a small table [n_rows][q] plus one row index per (trial, variable): 2 bytes instead of 4q cross PCIe /
// are read from HBM, and into_llr runs once per table row instead of once per trial.

// into_llr of the table rows (one thread per row; rows are tiny and the kernel runs once per decode call).
__global__ void __launch_bounds__(128) k_table_llr(const float *__restrict__ tab, float *__restrict__ out,
                                                   int *__restrict__ row_status, int n_rows, int q, int is_llr)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n_rows) return;
    float *v = out + (size_t)row * q;
    for (int i = 0; i < q; ++i) v[i] = tab[(size_t)row * q + i];
    row_status[row] = is_llr ? 0 : into_llr_row(v, q, sc_logf_tab);
}

constexpr int kObsRows = 64;  // variables per CTA of k_load_observed

// obs [T][rows] u16 row indices -> out [tile][rows*q][32] channel LLRs.  One CTA = one tile x kObsRows variables:
// coalesced load of the index block into shared memory, then one warp per variable (lane = trial) copies its table
// row (L1/L2-resident) into the tile layout with coalesced 128-byte stores.  Per-trial status: the row's into_llr
// status, SCALDPC_ERR_ARG for an index >= n_rows.
__global__ void __launch_bounds__(256) k_load_observed(const uint16_t *__restrict__ obs, const float *__restrict__ tab,
                                                       const int *__restrict__ row_status, float *__restrict__ out, int T,
                                                       int rows, int q, int n_rows, int *__restrict__ status,
                                                       const int *__restrict__ map)
{
    __shared__ uint16_t so[32][kObsRows + 2];  // pitch 33 words: conflict-free column reads
    const int tile = blockIdx.y, r0 = blockIdx.x * kObsRows;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int t = ty; t < 32; t += 8) {
        const int trial = tile * 32 + t;
        const size_t row = trial < T ? (size_t)(map ? map[trial] : trial) : 0;
        for (int c = tx; c < kObsRows; c += 32)
            so[t][c] = (trial < T && r0 + c < rows) ? obs[row * rows + r0 + c] : (uint16_t)0xffff;
    }
    __syncthreads();
    const bool real = tile * 32 + tx < T;
    for (int c = ty; c < kObsRows && r0 + c < rows; c += 8) {
        const int idx = so[tx][c];
        const bool ok = real && idx < n_rows;
        if (real && status) {
            const int err = idx < n_rows ? row_status[idx] : SCALDPC_ERR_ARG;
            if (err) atomicMin(status + tile * 32 + tx, err);
        }
        const float *src = tab + (size_t)(ok ? idx : 0) * q;
        float *o = out + ((size_t)tile * rows + r0 + c) * q * 32 + tx;
        for (int i = 0; i < q; ++i) o[(size_t)i * 32] = ok ? src[i] : 0.0f;
    }
}

// tile layout -> row-major.  src [tile][n][32] -> dst [T][n]; with `map`, slot t goes to row map[t] of dst.
template <typename T_>
__global__ void __launch_bounds__(32 * kTrRows) k_untile(const T_ *__restrict__ src, T_ *__restrict__ dst, int T, int n,
                                                         const int *__restrict__ map)
{
    __shared__ T_ sm[32][33];
    const int tile = blockIdx.y;
    const int c0 = blockIdx.x * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int c = ty; c < 32; c += kTrRows)
        if (c0 + c < n) sm[c][tx] = src[((size_t)tile * n + c0 + c) * 32 + tx];
    __syncthreads();
    for (int t = ty; t < 32; t += kTrRows) {
        const int trial = tile * 32 + t;
        if (trial < T && c0 + tx < n) dst[(size_t)(map ? map[trial] : trial) * n + c0 + tx] = sm[tx][t];
    }
}

// ---- at-risk detection, part 2: the row test of the final hard decisions,
// sum_j h_ij * s_j (+ s_sum,i) == 0 for every check (SURVEY.md 7 "hard parts"; kyber.py:85-92 / :374-375 make the
// decoded sum-variable equal -ssum_i).  One warp = (tile, check), lane = trial; a violated row flags the trial.
__global__ void __launch_bounds__(256) k_qary_syndrome(const QaryDev p)
{
    const int lane = threadIdx.x & 31;
    const int check = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int tile = blockIdx.y;
    if (check >= p.R) return;
    const int e0 = __ldg(p.row_ptr + check), e1 = __ldg(p.row_ptr + check + 1);
    const int8_t *hard = p.hard_t + (size_t)tile * p.N * 32 + lane;
    int s = 0;
    for (int e = e0; e < e1; ++e) {
        const int v = hard[(size_t)__ldg(p.col + e) * 32];
        s += __ldg(p.sign + e) < 0 ? -v : v;
    }
    flag_risk(p, tile, lane, s != 0);
}

// ---- early termination: close one iteration.  risk[tile] holds this iteration's verdicts (a violated row test, or a
// posterior too close to a tie to trust); a valid, not yet frozen trial with a clean word is FROZEN now (done bit set, its
// iteration count recorded), the others keep running.  Afterwards risk[tile] = the trials still running, which is both
// the at-risk list of the verification (k_risk_scan) and the list a restart re-stages.  One thread per tile.
__global__ void __launch_bounds__(256) k_qary_finish(uint32_t *__restrict__ risk, uint32_t *__restrict__ done,
                                                     int *__restrict__ iters, int tiles, int T, int it,
                                                     int *__restrict__ running)
{
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles) return;
    const int rem = T - tile * 32;
    const uint32_t valid = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
    uint32_t d = done[tile];
    uint32_t touched = ~d & valid;  // trials that ran this iteration
    const uint32_t newly = touched & ~risk[tile];
    while (touched) {
#ifdef SCALDPC_HOST_MODEL
        const int l = __builtin_ctz(touched);
#else
        const int l = __ffs(touched) - 1;
#endif
        iters[tile * 32 + l] = it;  // last iteration this trial took part in
        touched &= touched - 1u;
    }
    d |= newly | ~valid;
    done[tile] = d;
    risk[tile] = ~d;
    const int n = __popc(~d);
#ifdef SCALDPC_HOST_MODEL
    *running += n;
#else
    if (n) atomicAdd(running, n);
#endif
}

// dst[map[i]] = src[i], i < n (map null: identity).
__global__ void __launch_bounds__(256) k_scatter_int(const int *__restrict__ src, const int *__restrict__ map,
                                                     int *__restrict__ dst, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[map ? map[i] : i] = src[i];
}

// list[i] <- map[list[i]], i < *n_dev: slot indices of a round -> trial indices of the batch.
__global__ void __launch_bounds__(256) k_compose_map(int *__restrict__ list, const int *__restrict__ map,
                                                     const int *__restrict__ n_dev)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < *n_dev) list[i] = map[list[i]];
}

#ifndef SCALDPC_HOST_MODEL
// Flag words -> ascending list of the flagged trials (< T) + their count, and the most severe per-trial status.
// One CTA of 1024 threads.  out[0] = count, out[1] = min over status[0 .. T), out[2] = gstate[0] before this call.
// With `glist` / `gstate` the flagged trials are also APPENDED, as batch-wide indices `base + t`, to the list of the
// whole call: gstate[0] = entries so far, gstate[1] = most severe status so far.  Calls are serialised by the stream.
__global__ void __launch_bounds__(1024) k_risk_scan(const uint32_t *__restrict__ risk, int tiles, int T,
                                                    const int *__restrict__ status, int *__restrict__ list,
                                                    int *__restrict__ out, int *__restrict__ glist,
                                                    int *__restrict__ gstate, int base)
{
    const int gold = gstate ? gstate[0] : 0;
    __shared__ int wsum[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int t0 = 0; t0 < tiles; t0 += 1024) {
        const int tile = t0 + tid;
        uint32_t word = (risk && tile < tiles) ? risk[tile] : 0u;
        const int rem = T - tile * 32;
        if (rem < 32) word &= rem <= 0 ? 0u : ((1u << rem) - 1u);
        const int c = __popc(word);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += v;
            }
            wsum[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        int pos = s_base + (warp ? wsum[warp - 1] : 0) + incl - c;
        while (word) {
            const int b = __ffs(word) - 1;
            list[pos] = tile * 32 + b;
            if (glist) glist[gold + pos] = base + tile * 32 + b;
            ++pos;
            word &= word - 1u;
        }
        __syncthreads();
        if (tid == 0) s_base += wsum[31];
        __syncthreads();
    }
    int w = 0;
    if (status)
        for (int i = tid; i < T; i += 1024) w = min(w, status[i]);
#pragma unroll
    for (int o = 16; o; o >>= 1) w = min(w, __shfl_xor_sync(0xffffffffu, w, o));
    if (lane == 0) wsum[warp] = w;
    __syncthreads();
    if (tid == 0) {
        for (int i = 1; i < 32; ++i) w = min(w, wsum[i]);
        out[0] = s_base;
        out[1] = w;
        out[2] = gold;
        if (gstate) {
            gstate[0] = gold + s_base;
            gstate[1] = min(gstate[1], w);
        }
    }
}

// Rows list[i] (i < scan[0]) of src [.][n] -> rows scan[2] + i of dst, as long as they fit (< cap): the priors of a
// chunk's flagged trials are put aside before the chunk's staging buffer is overwritten.  blockIdx.y strides over the
// rows, the gridDim.x CTAs of a row copy a slice each (one CTA per row was latency-bound: 20 us for three rows).
template <typename T_>
__global__ void __launch_bounds__(256) k_gather_rows(const T_ *__restrict__ src, T_ *__restrict__ dst,
                                                     const int *__restrict__ list, const int *__restrict__ scan, int n,
                                                     int cap)
{
    const int count = scan[0], off = scan[2];
    for (int i = blockIdx.y; i < count; i += gridDim.y) {
        if (off + i >= cap) return;
        const T_ *s_ = src + (size_t)list[i] * n;
        T_ *d_ = dst + (size_t)(off + i) * n;
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) d_[k] = s_[k];
    }
}
#endif

}  // namespace scaldpc
