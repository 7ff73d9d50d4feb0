// Binary product-sum BP kernels for sm_100a (B200): flooding schedule, syndrome-match early stop, per-trial
// freeze and active-trial compaction.  Replaces ldpc.bp_decoder(bp_method="product_sum") as called at
// simulate/hqc.py:694-708 and simulate/decode.py:155-171 (third-party PyPI ldpc==0.1.3, algorithm as summarised in
// SURVEY.md appendix B).  Hand-written CUDA; no tensor cores (irregular gather + scalar updates, HBM-bound).
//
// Arithmetic: upstream keeps fp64 probability RATIOS r = p1/p0 and multiplies running products.  Here:
//   * bit -> check messages travel as signed FOLDED PROBABILITIES  q = copysign(u, L),  u = 1/(1+e^|L|) = min(p0,p1)
//     in (0, 1/2], L = log(p0/p1) the fp32 LLR.  tanh(|L|/2) = 1 - 2u, so upstream's running product of
//     (2/(1+r) - 1) becomes the parity recurrence  U <- U + u - 2Uu  (U = probability that the magnitudes seen so far
//     have odd parity, always <= 1/2: only positive terms, no cancellation, exact at u = 0 and u = 1/2) plus an XOR
//     of the signs; prefix / suffix recurrences in the row's column order replace upstream's forward / backward
//     products (never total-divided-by-self);
//   * check -> bit messages travel as LLRs  c2v = +-ln((1-U)/U)  (upstream: c2b = (1-x)/(1+x)), so the bit node sums
//     where upstream multiplies (no under/overflow for heavy columns) and keeps upstream's sequential semantics
//     including the "NaN -> neutral element" resets of the running product (0 * inf in ratios == inf - inf in LLRs).
// Cost per edge and iteration: ex2 + rcp (fold) and rcp + lg2 (unfold) on the MUFU pipe and ~40 other instructions;
// the earlier phi(x) = -log tanh(x/2) formulation needed ~110 and left both passes issue-bound.
//
// Data layout in HBM ("tile" layout, shared with the q-ary decoder): 32 trials per tile, lane = trial % 32 innermost.
//     msg     [tile][E][32] f32   q(v2c) before a check pass, c2v (LLR) after it (in place; one writer per slot)
//     prior   [n] f32 (shared by all trials)  or  [tile][n][32] f32 (per-trial channel_probs)
//     synd_w  [tile][m]  u32      bit l = syndrome bit of trial l of the tile
//     dec_w   [tile][n]  u32      bit l = current hard decision of trial l      (warp ballot)
//     post    [tile][n][32] f32   posterior LLRs (only when the caller asks for them)
//     done    [tile] u32          bit l = trial l has converged (outputs frozen);  valid [tile] u32 = real trials
//     unsat   [tile] u32          bit l = some check of trial l is unsatisfied this iteration
//     orig    [tile*32] i32       original trial index of a slot (slots move when the active trials are compacted)
//     counts  [max_iter+2] i32    counts[it] = trials still running after iteration it
#pragma once
#include <stdint.h>
#ifdef SCALDPC_HOST_MODEL
#define SCALDPC_DYN_SMEM(type, name) type *name = (type *)scaldpc_host_model_smem
#else
#include <cuda_runtime.h>
#define SCALDPC_DYN_SMEM(type, name) extern __shared__ type name[]
#endif
#include "../../include/scaldpc.h"

namespace scaldpc {

#include "tma.cuh"

struct BpDev {
    int m, n, E;
    int dc_max, dv_max;
    int per_trial_prior;      // 0: prior[n]; 1: prior[tile][n][32]
    const int *row_ptr;       // m+1
    const int *col;           // E
    const int *col_ptr;       // n+1
    const uint32_t *csc_off;  // E: byte offset (edge * 128) of each variable's edges inside a tile's message block, rows ascending
    const int *vn_order;      // n: variables sorted by degree (ascending), the bit pass runs one launch per class
    const int4 *vn_desc;      // n: {var, col_ptr[var], degree, 0} of vn_order[i] in one 16-byte load (device kernels)
    float *msg;
    const float *prior;
    const uint32_t *synd_w;
    uint32_t *dec_w;
    float *post;              // may be null
    uint32_t *done;
    const uint32_t *valid;
    uint32_t *unsat;
    const int *orig;          // tiles*32
    int *iters;               // T, indexed by ORIGINAL trial
    int *conv;                // T
    int *counts;              // max_iter+2
    const int *nrows;         // T, indexed by ORIGINAL trial: active checks (a prefix of the rows) of every trial, or null
    // Segmented CSR: G graphs of one shape m x n behind one handle -- the reference draws a NEW code per
    // simulated attack (hqc.py:1056).  Every TILE belongs to one graph (its 32 lanes are that attack's decodes), so graph
    // indices stay warp-uniform.  The per-graph tables are concatenated with fixed strides; null = single graph.
    const int *tile_graph = nullptr;  // [tiles]
    int g_rp = 0, g_e = 0, g_cp = 0, g_n = 0;  // strides: m+1 (row_ptr), E = max edges (col, csc_off, first table), n+1, n
    const int *class_end = nullptr;   // [G][kBpVnClasses+1]: end of every degree class in each graph's sorted variable order
    // first iteration with shared priors: the check-to-bit messages are +-first_tab[e] (k_bp_first_table); the FIRST variant
    // of the bit pass forms them on the fly instead of reading a message array that k_bp_first_fill would have to write
    const float *first_tab = nullptr;  // [G][E]
    const int *edge_row = nullptr;     // [G][E] check of every edge
    const float *csc_tab = nullptr;    // [G][E] first_tab in CSC order (position of csc_off), written by k_bp_first_table
    const int *csc_row = nullptr;      // [G][E] check of every edge, CSC order
    const int *edge_csc = nullptr;     // [G][E] CSC position of every edge
};

constexpr int kBpVnClassCount = 7;  // == kBpVnClasses (bp_graph.h)

__device__ __forceinline__ int bp_graph(const BpDev &p, int tile) { return p.tile_graph ? __ldg(p.tile_graph + tile) : 0; }

#ifdef SCALDPC_HOST_MODEL
__device__ __forceinline__ uint32_t bp_bits(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    return u;
}
__device__ __forceinline__ float bp_float(uint32_t u)
{
    float x;
    memcpy(&x, &u, 4);
    return x;
}
__device__ __forceinline__ float bp_ex2(float x) { return exp2f(x); }
__device__ __forceinline__ float bp_lg2(float x) { return log2f(x); }
__device__ __forceinline__ float bp_rcp(float x) { return 1.0f / x; }
#else
__device__ __forceinline__ uint32_t bp_bits(float x) { return __float_as_uint(x); }
__device__ __forceinline__ float bp_float(uint32_t u) { return __uint_as_float(u); }
// Single MUFU instructions (flush-to-zero forms: no subnormal fix-up code around them).
__device__ __forceinline__ float bp_ex2(float x)
{
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float bp_lg2(float x)
{
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float bp_rcp(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
#endif
__device__ __forceinline__ bool bp_signbit(float x) { return bp_bits(x) >> 31; }

// LLR -> signed folded probability  q = copysign(1/(1+e^|L|), L): FMUL, MUFU.EX2, FADD, MUFU.RCP, LOP3.
// fold(+-inf) = +-0, fold(+-0) = +-1/2, NaN -> NaN.  e^|L| overflows for |L| > 88.7, where u = 0 ("certain"; the fp64
// upstream already rounds tanh(|L|/2) to exactly 1 for |L| > 37).
__device__ __forceinline__ float bp_fold(float L)
{
    const float u = bp_rcp(1.0f + bp_ex2(1.4426950408889634f * fabsf(L)));
    return bp_float(bp_bits(u) | (bp_bits(L) & 0x80000000u));
}
// Odd-parity probability U in [0, 1/2] -> LLR magnitude ln((1-U)/U) = ln(1/U - 1) >= 0: MUFU.RCP, FADD, MUFU.LG2, FMUL.
// unfold(0) = +inf, unfold(1/2) = 0 (+- 3e-7 absolute), NaN -> NaN.
__device__ __forceinline__ float bp_unfold(float U) { return 0.6931471805599453f * bp_lg2(bp_rcp(U) - 1.0f); }
// One more magnitude u joins a parity: U + u - 2Uu (two FFMA).
__device__ __forceinline__ float bp_parity(float U, float u) { return fmaf(u, fmaf(-2.0f, U, 1.0f), U); }

__device__ __forceinline__ float bp_prior_of(const BpDev &p, int tile, int v, int lane)
{
    return p.per_trial_prior ? p.prior[((size_t)tile * p.n + v) * 32 + lane] : __ldg(p.prior + v);
}

// Nested graphs (the decode_every loop, hqc.py:953-984, decodes on the first R_k rows of one growing H): every trial has
// its own prefix of active checks.  An inactive check sends the neutral message (LLR +0: the bit node's sums are
// unchanged, exactly as if the row were absent) and is left out of the syndrome test.
__device__ __forceinline__ int bp_row_limit(const BpDev &p, int tile, int lane)
{
    if (!p.nrows) return p.m;
    const int o = __ldg(p.orig + tile * 32 + lane);
    return o >= 0 ? __ldg(p.nrows + o) : p.m;
}

__device__ __forceinline__ bool bp_tile_done(const BpDev &p, int tile) { return __ldg(p.done + tile) == __ldg(p.valid + tile); }

// -----------------------------------------------------------------------------------------------------------
// Check-to-bit pass.  Work item = (tile, check); one warp per item, lane = trial.  A row's messages are ONE
// contiguous block of d x 128 B in the tile layout.  Persistent kernel: warps stride over the items; lane 0 fetches
// the NEXT item's block with one bulk-async copy (TMA, cp.async.bulk + mbarrier complete_tx) into the other half of
// a double buffer while the warp works on the current one, so the HBM latency hides behind ~2 k instructions of
// arithmetic (v0 issued 4 plain loads at a time from 16 warps per SM: latency-bound at 0.13 of the HBM peak).
// Per lane: P[k] = parity of the magnitudes before k (shared memory, [k][lane], conflict-free), then backwards
// c2v_k = sign * unfold(parity(P[k], suffix)), stored straight to global memory (coalesced 128 B per warp store).
// FIRST: q is computed from the channel prior (upstream's init b2c = p/(1-p)); nothing is staged.
// -----------------------------------------------------------------------------------------------------------
constexpr int kBpCnWarpsMax = 16;

template <bool FIRST>
__global__ void __launch_bounds__(kBpCnWarpsMax * 32) k_bp_cn(const BpDev p, const int tiles)
{
    SCALDPC_DYN_SMEM(float, smem);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    const int gw = warp * gridDim.x + blockIdx.x, nw = gridDim.x * wpc;  // warp-major: consecutive items on different SMs
    const int stride = p.dc_max * 32;                // floats per per-warp array
    float *wbase = smem + (size_t)warp * 3 * stride;  // stage[2], P
    float *P = wbase + 2 * stride + lane;

    const int step_t = nw / p.m, step_c = nw - step_t * p.m;  // item -> item + nw without a division per item
    struct Item {
        int tile, check, e0, d, cb;  // cb: offset of the tile's graph in the column table
    };
    // moves to the next item (stride nw) whose tile still has running trials; tile >= tiles: none left
    auto advance = [&](Item &it, bool first) {
        if (!first) {
            it.tile += step_t;
            it.check += step_c;
            if (it.check >= p.m) {
                it.check -= p.m;
                ++it.tile;
            }
        }
        while (it.tile < tiles && bp_tile_done(p, it.tile)) {
            it.tile += step_t;
            it.check += step_c;
            if (it.check >= p.m) {
                it.check -= p.m;
                ++it.tile;
            }
        }
        if (it.tile < tiles) {
            const int g = bp_graph(p, it.tile);
            const int *rp = p.row_ptr + (size_t)g * p.g_rp + it.check;
            it.e0 = __ldg(rp);
            it.d = __ldg(rp + 1) - it.e0;
            it.cb = g * p.g_e;
        }
    };
#ifndef SCALDPC_HOST_MODEL
    __shared__ uint64_t bars[kBpCnWarpsMax][2];
    uint32_t phases = 0;  // bit b = parity to wait for on bars[warp][b]
    if (!FIRST) {
        if (lane == 0) {
            sc_mbar_init(&bars[warp][0], 1);
            sc_mbar_init(&bars[warp][1], 1);
        }
        __syncwarp();
    }
    auto issue = [&](const Item &it, int b) {  // lane 0 only
        if (it.d <= 0) return;
        sc_mbar_expect_tx(&bars[warp][b], (uint32_t)it.d * 128u);
        sc_bulk_g2s(wbase + b * stride, p.msg + ((size_t)it.tile * p.E + it.e0) * 32, (uint32_t)it.d * 128u, &bars[warp][b]);
    };
#endif

    Item cur;
    cur.tile = gw / p.m;
    cur.check = gw - cur.tile * p.m;
    advance(cur, true);
    int b = 0;
#ifndef SCALDPC_HOST_MODEL
    if (!FIRST && cur.tile < tiles && lane == 0) issue(cur, 0);
#endif
    while (cur.tile < tiles) {
        Item nxt = cur;
        advance(nxt, false);
#ifndef SCALDPC_HOST_MODEL
        if (!FIRST && nxt.tile < tiles && lane == 0) {
            sc_fence_proxy_async();  // the other half was read (generic proxy) during the previous item
            issue(nxt, b ^ 1);
        }
#endif
        const int tile = cur.tile, check = cur.check, e0 = cur.e0, d = cur.d;
        float *msg = p.msg + ((size_t)tile * p.E + e0) * 32 + lane;
        if (d > 0) {
#ifdef SCALDPC_HOST_MODEL
            const float *a = msg;  // the host model reads the row in place
#else
            const float *a = wbase + b * stride + lane;
            if (!FIRST) {
                sc_mbar_wait(&bars[warp][b], (phases >> b) & 1u);
                phases ^= 1u << b;
            }
#endif
            // sign word: bit 31 = (-1)^synd * product of the signs of ALL inputs (lower bits are garbage)
            uint32_t neg = (__ldg(p.synd_w + (size_t)tile * p.m + check) >> lane) << 31;
            const uint32_t keep = check < bp_row_limit(p, tile, lane) ? 0xffffffffu : 0u;  // inactive check: +0
            float pre = 0.0f;
            if (FIRST) {
#ifdef SCALDPC_HOST_MODEL
                float *park = msg;  // host model: parked in place, re-read below by the same lane
#else
                float *park = wbase + lane;  // stage[0] is unused by the first pass
#endif
#pragma unroll 4
                for (int k = 0; k < d; ++k) {
                    const float v = bp_prior_of(p, tile, __ldg(p.col + cur.cb + e0 + k), lane);
                    const float q = bp_fold(v);
                    P[k * 32] = pre;
                    park[(size_t)k * 32] = q;
                    pre = bp_parity(pre, fabsf(q));
                    neg ^= bp_bits(v);
                }
                a = park;
            } else {
#pragma unroll 8
                for (int k = 0; k < d; ++k) {
                    const float x = a[k * 32];
                    P[k * 32] = pre;
                    pre = bp_parity(pre, fabsf(x));
                    neg ^= bp_bits(x);
                }
            }
            float suf = 0.0f;
#pragma unroll 4
            for (int k = d - 1; k >= 0; --k) {
                const float x = a[k * 32];
                const float mag = bp_unfold(bp_parity(P[k * 32], suf));  // parity of the OTHER magnitudes
                msg[(size_t)k * 32] = bp_float((bp_bits(mag) ^ ((neg ^ bp_bits(x)) & 0x80000000u)) & keep);  // sign of the OTHERS
                suf = bp_parity(suf, fabsf(x));
            }
        }
#ifndef SCALDPC_HOST_MODEL
        __syncwarp();  // every lane is done with stage[b] before lane 0 re-arms it (next iteration)
#endif
        cur = nxt;
        b ^= 1;
    }
}

// -----------------------------------------------------------------------------------------------------------
// First check pass when all trials share one prior vector (the HQC call site: channel_probs is one list,
// hqc.py:684-699).  Then upstream's init b2c = p/(1-p) is the same for every trial, and so is the magnitude of every
// first-iteration check-to-bit message; only its sign depends on the trial, through the syndrome bit.  So
//   k_bp_first_table  runs k_bp_cn<FIRST>'s arithmetic ONCE per check (one thread per check, syndrome bit 0) and
//                     leaves the signed message of every edge in tab[E]; the same device functions in the same
//                     order, so the values are bit-identical to what every lane of k_bp_cn<true> computes;
//   k_bp_first_fill   streams msg[tile][e][lane] = tab[e] with the sign flipped where the trial's syndrome bit is set:
//                     a pure write at HBM speed instead of 4 MUFU + two dependent index loads per edge and trial
//                     (k_bp_cn<true>: 2.5 ms per 4096 hqc128 trials against 0.93 ms for a later check pass).
// -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_bp_first_table(const BpDev p, float *__restrict__ tab_all)
{
    const int check = blockIdx.x * blockDim.x + threadIdx.x;
    if (check >= p.m) return;
    const int g = blockIdx.y;  // one table per graph
    const int *row_ptr = p.row_ptr + (size_t)g * p.g_rp;
    const int *col = p.col + (size_t)g * p.g_e;
    float *tab = tab_all + (size_t)g * p.g_e;
    const int e0 = __ldg(row_ptr + check), d = __ldg(row_ptr + check + 1) - e0;
    uint32_t neg = 0;  // syndrome bit 0
    float pre = 0.0f;
    for (int k = 0; k < d; ++k) {  // the prefix parities wait in the row's own table entries (round 1 recomputed them: O(d^2)
                                   // folds per row, 82 us for hqc128 -- a quarter of a single decode(syndrome) call)
        const float v = __ldg(p.prior + __ldg(col + e0 + k));
        tab[e0 + k] = pre;
        pre = bp_parity(pre, fabsf(bp_fold(v)));
        neg ^= bp_bits(v);
    }
    float suf = 0.0f;
    for (int k = d - 1; k >= 0; --k) {
        const float P = tab[e0 + k];
        const float x = bp_fold(__ldg(p.prior + __ldg(col + e0 + k)));
        const float mag = bp_unfold(bp_parity(P, suf));
        tab[e0 + k] = bp_float(bp_bits(mag) ^ ((neg ^ bp_bits(x)) & 0x80000000u));
        if (p.csc_tab) ((float *)p.csc_tab)[(size_t)g * p.g_e + __ldg(p.edge_csc + (size_t)g * p.g_e + e0 + k)] = tab[e0 + k];
        suf = bp_parity(suf, fabsf(x));
    }
}

#ifndef SCALDPC_HOST_MODEL
// One warp per (tile, edge), lane = trial.  edge_row[e] = check of edge e.
__global__ void __launch_bounds__(256) k_bp_first_fill(const BpDev p, const float *__restrict__ tab,
                                                       const int *__restrict__ edge_row, const int tiles)
{
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.y;
    if (bp_tile_done(p, tile)) return;
    const int e_lo = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 8;
    float *msg = p.msg + ((size_t)tile * p.E) * 32 + lane;
    const int limit = bp_row_limit(p, tile, lane);
    const size_t gb = (size_t)bp_graph(p, tile) * p.g_e;  // this tile's graph: its table and its edge -> row map
    tab += gb;
    edge_row += gb;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int e = e_lo + i;
        if (e < p.E) {
            const int row = __ldg(edge_row + e);
            const uint32_t w = __ldg(p.synd_w + (size_t)tile * p.m + row);
            msg[(size_t)e * 32] = row < limit ? bp_float(bp_bits(__ldg(tab + e)) ^ ((w >> lane) << 31)) : 0.0f;
        }
    }
}

#endif
// -----------------------------------------------------------------------------------------------------------
// Bit-to-check pass + decision.  Work item = (tile, variable); one warp per item, lane = trial.  Upstream's two
// sweeps over the column (rows ascending): forward running value seeded with the prior, backward running value
// seeded with the neutral element; a NaN running value is reset to the neutral element (ratio 1 == LLR 0).  Stores
// q = fold(v2c).  Decision: upstream decodes 1 when ratio >= 1, i.e. LLR <= 0.  Converged (done) lanes keep their old
// outputs.
// A variable's incoming messages are d scattered 128-byte lines.  Persistent kernel, registers only: a warp gathers
// the column of its NEXT item into registers (DMAX independent loads) before it works on the current one, so every
// resident warp always has a full column in flight (v1 loaded, computed, stored in sequence: 12.8 warps stalled on
// long_scoreboard per issue, 0.34 of the HBM peak; a cp.async/shared-memory variant hid the latency but executed
// 2.3x the instructions of the check pass and became issue-bound at 0.43).  The variables are sorted by degree and
// the pass runs one launch per degree class DMAX, so light columns do not pay for heavy ones.
// -----------------------------------------------------------------------------------------------------------
constexpr int kBpVnWarps = 8;

template <bool LEAN = false>  // LEAN: no posterior output (p.post is known to be null)
__device__ __forceinline__ void bp_vn_decide(const BpDev &p, int tile, int var, int lane, uint32_t done, float post)
{
    const bool active = !((done >> lane) & 1u);
    if (!LEAN && p.post && active) p.post[((size_t)tile * p.n + var) * 32 + lane] = post;
#ifdef SCALDPC_HOST_MODEL
    uint32_t &w = p.dec_w[(size_t)tile * p.n + var];
    if (active) w = (w & ~(1u << lane)) | ((post <= 0.0f ? 1u : 0u) << lane);
#else
    const uint32_t bits = __ballot_sync(0xffffffffu, post <= 0.0f);
    if (lane == 0) {
        // w = (w & done) | (bits & ~done) without waiting for w: two fire-and-forget reductions (the word has one
        // writer); a plain read-modify-write stalled the whole warp for an L2 round trip per item (17 % of all samples)
        uint32_t *w = p.dec_w + (size_t)tile * p.n + var;
        if (done == 0u) {
            *w = bits;
        } else {
            atomicAnd(w, done | bits);
            atomicOr(w, bits & ~done);
        }
    }
#endif
}

// An item of the bit pass and everything its phases need to know about it.
struct BpVnItem {
    int tile, j, var, p0, d;
};

// `base` + byte offset, with the base kept as an opaque 64-bit register so that the address is one 64-bit add.
__device__ __forceinline__ float *bp_at(float *base, uint32_t byte_off) { return (float *)((char *)base + byte_off); }

// The two sweeps of one variable of degree exactly D (straight-line code, no guards).  Upstream resets a NaN running value to the neutral element; a NaN can only
// appear when +inf meets -inf, and then the forward total is NaN too, so the sweeps run WITHOUT the per-step tests and a
// NaN total sends the lane through the tested version (identical results; 4 fewer instructions per edge otherwise).
template <int D, int DMAX, bool LEAN = false>
__device__ __forceinline__ void bp_vn_body(const BpDev &p, const BpVnItem &it, int lane, const float (&c)[DMAX],
                                           const float prior)
{
    const uint32_t *o = p.csc_off + it.p0;  // re-read (L1): carrying the offsets in registers costs a resident CTA per SM
    const uint32_t done = __ldg(p.done + it.tile);
    float *msg = p.msg + (size_t)it.tile * p.E * 32 + lane;
#ifndef SCALDPC_HOST_MODEL
    asm volatile("" : "+l"(msg));
#endif
    float F[D > 0 ? D : 1];
    float temp = prior;
#pragma unroll
    for (int i = 0; i < D; ++i) {
        F[i] = temp;
        temp += c[i];
    }
    float post = temp;
    if (post == post) {
        temp = 0.0f;
#pragma unroll
        for (int i = D - 1; i >= 0; --i) {
            *bp_at(msg, __ldg(o + i)) = bp_fold(F[i] + temp);
            temp += c[i];
        }
    } else {  // rare: +inf and -inf among the column's messages -- upstream's "NaN -> neutral element" semantics
        temp = prior;
#pragma unroll
        for (int i = 0; i < D; ++i) {
            F[i] = temp;
            temp += c[i];
            if (temp != temp) temp = 0.0f;
        }
        post = temp;
        temp = 0.0f;
#pragma unroll
        for (int i = D - 1; i >= 0; --i) {
            *bp_at(msg, __ldg(o + i)) = bp_fold(F[i] + temp);
            temp += c[i];
            if (temp != temp) temp = 0.0f;
        }
    }
    bp_vn_decide<LEAN>(p, it.tile, it.var, lane, done, post);
}

template <int D, int DMIN, int DMAX, bool LEAN = false>
__device__ __forceinline__ void bp_vn_dispatch(const BpDev &p, const BpVnItem &it, int lane, const float (&c)[DMAX],
                                               const float prior)
{
    if constexpr (D < DMIN) {
        bp_vn_body<DMAX, DMAX, LEAN>(p, it, lane, c, prior);  // unreachable for a correctly sorted class; keeps the code total
    } else {
        if (it.d == D)
            bp_vn_body<D, DMAX, LEAN>(p, it, lane, c, prior);
        else
            bp_vn_dispatch<D - 1, DMIN, DMAX, LEAN>(p, it, lane, c, prior);
    }
}

// Degree class (DMIN .. DMAX].  The class of degree <= 1 is instantiated as <0, 1> and also takes isolated variables.
// FIRST (iteration 1, shared priors): the incoming messages are not read from memory but formed from the first-pass table,
// c = +-first_tab[e] with the sign of the trial's syndrome bit of the edge's check (what k_bp_first_fill would have written;
// an inactive check of a nested decode sends +0) -- the first iteration then writes the message array once and never reads it.
// SHARED ("lean"): the priors are known to be shared by the trials and no posterior is wanted (the HQC call site): the
// per-trial prior's and the posterior's address arithmetic, which run-time tests only predicate off (~20 issue slots per
// item), is not generated.
template <int DMIN, int DMAX, bool FIRST, bool SHARED = false>
// Segmented CSR (p.class_end set): `cls` is the degree class, [v_lo, v_hi) = [0, largest class size over the graphs);
// every tile takes its own graph's slice of the class and skips the j beyond it.
__global__ void __launch_bounds__(kBpVnWarps * 32) k_bp_vn(const BpDev p, const int v_lo, const int v_hi, const int tiles,
                                                           const int cls)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    const int gw = blockIdx.x * wpc + warp, nw = gridDim.x * wpc;  // block-major: a CTA's warps share the descriptor / offset lines (warp-major measured slower here)
    const int nv = v_hi - v_lo;
    const int step_t = nw / nv, step_j = nw - step_t * nv;  // item -> item + nw without a division per item

    // moves to the next item (stride nw) whose tile still has running trials; tile >= tiles: none left
    auto advance = [&](BpVnItem &it, bool first) {
        bool step = !first;
        int g = 0, lo = v_lo;
        while (true) {
            if (step) {
                it.tile += step_t;
                it.j += step_j;
                if (it.j >= nv) {
                    it.j -= nv;
                    ++it.tile;
                }
            }
            step = true;
            if (it.tile >= tiles) break;
            if (bp_tile_done(p, it.tile)) continue;
            if (p.class_end) {  // this tile's graph: its slice [lo, hi) of the class
                g = bp_graph(p, it.tile);
                const int *ce = p.class_end + g * (kBpVnClassCount + 1);
                lo = cls > 0 ? __ldg(ce + cls - 1) : 0;
                if (it.j >= __ldg(ce + cls) - lo) continue;
            }
            break;
        }
        if (it.tile < tiles) {
#ifdef SCALDPC_HOST_MODEL
            it.var = __ldg(p.vn_order + (size_t)g * p.g_n + lo + it.j);
            const int *cp = p.col_ptr + (size_t)g * p.g_cp + it.var;
            const int c0 = __ldg(cp);
            it.d = __ldg(cp + 1) - c0;  // in (DMIN, DMAX] (host guarantees)
            it.p0 = g * p.g_e + c0;     // position in the (concatenated) csc_off table
#else
            // {variable, col_ptr, degree} of the sorted position in ONE 16-byte load instead of vn_order -> col_ptr x 2
            const int4 q = __ldg(p.vn_desc + (size_t)g * p.g_n + lo + it.j);
            it.var = q.x;
            it.d = q.z;
            it.p0 = g * p.g_e + q.y;
#endif
        }
    };
    auto gather = [&](const BpVnItem &it, float (&c)[DMAX]) {
        float *src = p.msg + (size_t)it.tile * p.E * 32 + lane;
#ifndef SCALDPC_HOST_MODEL
        asm volatile("" : "+l"(src));
#endif
        const uint32_t *off = p.csc_off + it.p0;
        // FIRST: the column's rows and first-pass messages sit in CSC order next to the offsets (three independent
        // sequential index loads, then ONE dependent load of the tile's syndrome word, L1-resident)
        const int *rows = p.csc_row + it.p0;
        const float *tabs = p.csc_tab + it.p0;
        const int limit = FIRST ? bp_row_limit(p, it.tile, lane) : 0;
        const uint32_t *sw = p.synd_w + (size_t)it.tile * p.m;
#pragma unroll
        for (int i = 0; i < DMAX; ++i)
            if ((DMIN > 0 && i <= DMIN) || i < it.d) {  // d > DMIN: the first are unguarded
                if (FIRST) {
                    const int row = __ldg(rows + i);
                    const uint32_t v = bp_bits(__ldg(tabs + i)) ^ ((__ldg(sw + row) >> lane) << 31);
                    c[i] = row < limit ? bp_float(v) : 0.0f;
                } else {
                    c[i] = *bp_at(src, __ldg(off + i));
                }
            }
    };

    BpVnItem cur;
    cur.tile = gw / nv;
    cur.j = gw - cur.tile * nv;
    advance(cur, true);
    float c[DMAX], cn[DMAX];
    float pr = 0.0f, prn = 0.0f;  // the variable's prior travels with its column (44 % of the stall samples sat on it)
    if (cur.tile < tiles) {
        gather(cur, c);
        pr = SHARED ? __ldg(p.prior + cur.var) : bp_prior_of(p, cur.tile, cur.var, lane);
    }
    while (cur.tile < tiles) {
        BpVnItem nxt = cur;
        advance(nxt, false);
        if (nxt.tile < tiles) {
            gather(nxt, cn);
            prn = SHARED ? __ldg(p.prior + nxt.var) : bp_prior_of(p, nxt.tile, nxt.var, lane);
        }
        bp_vn_dispatch<DMAX, (DMIN == 0 ? 0 : DMIN + 1), DMAX, SHARED>(p, cur, lane, c, pr);
        cur = nxt;
        pr = prn;
#pragma unroll
        for (int i = 0; i < DMAX; ++i) c[i] = cn[i];
    }
}

#ifndef SCALDPC_HOST_MODEL
// -----------------------------------------------------------------------------------------------------------
// The same pass (iterations >= 2) with the column gather as a software pipeline of per-lane asynchronous copies.
// ncu on k_bp_vn (profiles/r2n_ncu_bp_vn_hqc128.txt): 3.75 TB/s, long_scoreboard 10 warps per issue, 48-64 registers =
// 40-32 resident warps with ~1.3 columns in flight each -- ~30 KB in flight per SM is exactly what 3.7 TB/s needs at
// ~1.2 us of loaded latency (Little), and every byte in flight pins a destination REGISTER until it lands.  Here the
// landing zone is shared memory: lane l copies ITS 4 bytes of every line of a column with cp.async (LDGSTS, no
// destination register), kBpVnStages columns per warp are always in flight, and a column costs one LDS per line when its
// turn comes.  A lane only ever reads what it copied itself, so cp.async.wait_group is the only synchronisation -- no
// mbarrier, no __syncwarp (the per-lane 128-byte BULK copies of round 1's k_bp_vn_tma were serialised by the TMA unit).
// Arithmetic and stores are bp_vn_body's: bit-identical to k_bp_vn, which stays for iteration 1 (FIRST) and the host model.
// -----------------------------------------------------------------------------------------------------------
template <int DMAX>
struct BpVnPipe {
    static constexpr int stages = DMAX <= 12 ? 4 : (DMAX <= 16 ? 3 : 2);
    static constexpr size_t smem_bytes = (size_t)kBpVnWarps * stages * DMAX * 32 * sizeof(float);
};

__device__ __forceinline__ void sc_cp_async4(float *smem_dst, const float *gsrc)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void sc_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void sc_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int DMIN, int DMAX>
__global__ void __launch_bounds__(kBpVnWarps * 32) k_bp_vn_pipe(const BpDev p, const int v_lo, const int v_hi, const int tiles,
                                                                const int cls)
{
    constexpr int S = BpVnPipe<DMAX>::stages;
    extern __shared__ __align__(128) float sc_vn_ring[];  // [warp][stage][DMAX][32]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    const int gw = blockIdx.x * wpc + warp, nw = gridDim.x * wpc;  // block-major: a CTA's warps share the descriptor / offset lines (warp-major measured slower here)
    const int nv = v_hi - v_lo;
    const int step_t = nw / nv, step_j = nw - step_t * nv;
    float *ring = sc_vn_ring + (size_t)warp * S * DMAX * 32 + lane;

    auto advance = [&](BpVnItem &it, bool first) {  // as k_bp_vn
        bool step = !first;
        int g = 0, lo = v_lo;
        while (true) {
            if (step) {
                it.tile += step_t;
                it.j += step_j;
                if (it.j >= nv) {
                    it.j -= nv;
                    ++it.tile;
                }
            }
            step = true;
            if (it.tile >= tiles) break;
            if (bp_tile_done(p, it.tile)) continue;
            if (p.class_end) {
                g = bp_graph(p, it.tile);
                const int *ce = p.class_end + g * (kBpVnClassCount + 1);
                lo = cls > 0 ? __ldg(ce + cls - 1) : 0;
                if (it.j >= __ldg(ce + cls) - lo) continue;
            }
            break;
        }
        if (it.tile < tiles) {
            it.var = __ldg(p.vn_order + (size_t)g * p.g_n + lo + it.j);
            const int *cp = p.col_ptr + (size_t)g * p.g_cp + it.var;
            const int c0 = __ldg(cp);
            it.d = __ldg(cp + 1) - c0;
            it.p0 = g * p.g_e + c0;
        }
    };
    auto issue = [&](const BpVnItem &it, int s) {  // lane l: its 4 bytes of every line of the column -> stage s
        const float *src = p.msg + (size_t)it.tile * p.E * 32 + lane;
        const uint32_t *off = p.csc_off + it.p0;
        float *dst = ring + (size_t)s * DMAX * 32;
#pragma unroll
        for (int i = 0; i < DMAX; ++i)
            if ((DMIN > 0 && i <= DMIN) || i < it.d) sc_cp_async4(dst + i * 32, (const float *)((const char *)src + __ldg(off + i)));
    };

    BpVnItem it[S], x;
    float pr[S];
    x.tile = gw / nv;
    x.j = gw - x.tile * nv;
    advance(x, true);
#pragma unroll
    for (int s = 0; s < S; ++s) {
        it[s] = x;
        pr[s] = 0.0f;
        if (x.tile < tiles) {
            issue(x, s);
            pr[s] = bp_prior_of(p, x.tile, x.var, lane);
            advance(x, false);
        }
        sc_cp_async_commit();
    }
    while (true) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            if (it[s].tile >= tiles) return;  // items are consumed in order: nothing is pending behind an empty stage
            sc_cp_async_wait<S - 1>();
            float c[DMAX];
            const float *st = ring + (size_t)s * DMAX * 32;
#pragma unroll
            for (int i = 0; i < DMAX; ++i)
                if ((DMIN > 0 && i <= DMIN) || i < it[s].d) c[i] = st[i * 32];
            bp_vn_dispatch<DMAX, (DMIN == 0 ? 0 : DMIN + 1), DMAX>(p, it[s], lane, c, pr[s]);
            it[s] = x;  // refill the stage (its values sit in registers now)
            if (x.tile < tiles) {
                issue(x, s);
                pr[s] = bp_prior_of(p, x.tile, x.var, lane);
                advance(x, false);
            }
            sc_cp_async_commit();
        }
    }
}
#endif

// Any degree (columns heavier than the largest register class): per-lane scratch [i][lane] in shared memory,
// one CTA pass over (variable, tile) without prefetch.
// Segmented CSR: cls >= 0 -> the tile's own slice of that class (v_lo = 0, v_hi = largest slice); cls < 0 -> [v_lo, v_hi)
// of every graph's sorted order (all variables in the single-launch tail).
template <bool FIRST>
__global__ void __launch_bounds__(kBpVnWarps * 32) k_bp_vn_generic(const BpDev p, const int v_lo, const int v_hi, const int cls)
{
    constexpr bool first = FIRST;  // compile-time: a run-time test inside the gather loop keeps its loads from being batched
    SCALDPC_DYN_SMEM(float, smem);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    int idx = v_lo + blockIdx.x * wpc + warp;
    const int tile = blockIdx.y;
    if (idx >= v_hi) return;
    const uint32_t done = __ldg(p.done + tile);
    if (done == __ldg(p.valid + tile)) return;
    const int g = bp_graph(p, tile);
    if (p.class_end && cls >= 0) {
        const int *ce = p.class_end + g * (kBpVnClassCount + 1);
        const int lo = cls > 0 ? __ldg(ce + cls - 1) : 0;
        idx = lo + (idx - v_lo);
        if (idx >= __ldg(ce + cls)) return;
    }
#ifdef SCALDPC_HOST_MODEL
    const int var = __ldg(p.vn_order + (size_t)g * p.g_n + idx);
    const int *cp = p.col_ptr + (size_t)g * p.g_cp + var;
    const int p0 = g * p.g_e + __ldg(cp), d = __ldg(cp + 1) - __ldg(cp);
#else
    const int4 q = __ldg(p.vn_desc + (size_t)g * p.g_n + idx);  // {variable, col_ptr, degree}: one load
    const int var = q.x, p0 = g * p.g_e + q.y, d = q.z;
#endif
    float *C = smem + (size_t)warp * 2 * p.dv_max * 32 + lane;
    float *F = C + (size_t)p.dv_max * 32;
    float *msg = p.msg + (size_t)tile * p.E * 32 + lane;
    const uint32_t *off = p.csc_off + p0;

    float temp = bp_prior_of(p, tile, var, lane);
    const int limit = first ? bp_row_limit(p, tile, lane) : 0;
#pragma unroll 4
    for (int i = 0; i < d; ++i) {
        float c;
        if (first) {  // as k_bp_vn<.., FIRST>: +-first_tab[e], sign from the syndrome bit of the edge's check
            const int row = __ldg(p.csc_row + p0 + i);
            const uint32_t v = bp_bits(__ldg(p.csc_tab + p0 + i)) ^ ((__ldg(p.synd_w + (size_t)tile * p.m + row) >> lane) << 31);
            c = row < limit ? bp_float(v) : 0.0f;
        } else {
            c = *bp_at(msg, __ldg(off + i));
        }
        C[i * 32] = c;
        F[i * 32] = temp;
        temp += c;
        if (temp != temp) temp = 0.0f;
    }
    const float post = temp;
    temp = 0.0f;
    for (int i = d - 1; i >= 0; --i) {
        *bp_at(msg, __ldg(off + i)) = bp_fold(F[i * 32] + temp);
        temp += C[i * 32];
        if (temp != temp) temp = 0.0f;
    }
    bp_vn_decide(p, tile, var, lane, done, post);
}

// -----------------------------------------------------------------------------------------------------------
// Syndrome test, bit-parallel over the 32 trials of a tile: one thread = (tile, check) XORs the decision words of
// the check's variables and compares with the syndrome word; mismatching trials are OR-ed into unsat[tile].
// -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bp_syndrome(const BpDev p)
{
    const int tile = blockIdx.y;
    const int check = blockIdx.x * blockDim.x + threadIdx.x;
    if (bp_tile_done(p, tile)) return;
#ifndef SCALDPC_HOST_MODEL
    __shared__ int lim[32];
    if (p.nrows) {  // uniform
        if (threadIdx.x < 32) lim[threadIdx.x] = bp_row_limit(p, tile, threadIdx.x);
        __syncthreads();
    }
#endif
    uint32_t bad = 0;
    if (check < p.m) {
        const uint32_t *dec = p.dec_w + (size_t)tile * p.n;
        uint32_t par = __ldg(p.synd_w + (size_t)tile * p.m + check);
        const int g = bp_graph(p, tile);
        const int *rp = p.row_ptr + (size_t)g * p.g_rp + check;
        const int *col = p.col + (size_t)g * p.g_e;
        const int e0 = __ldg(rp), e1 = __ldg(rp + 1);
#pragma unroll 8  // eight (index, word) load pairs in flight: the rolled loop waited for every word before the next index
        for (int e = e0; e < e1; ++e) par ^= dec[__ldg(col + e)];
        bad = par;
        if (p.nrows) {  // trials for which this check is inactive cannot fail it
            uint32_t act = 0;
#ifdef SCALDPC_HOST_MODEL
            for (int l = 0; l < 32; ++l) act |= (check < bp_row_limit(p, tile, l) ? 1u : 0u) << l;
#else
#pragma unroll
            for (int l = 0; l < 32; ++l) act |= (check < lim[l] ? 1u : 0u) << l;
#endif
            bad &= act;
        }
    }
#ifdef SCALDPC_HOST_MODEL
    p.unsat[tile] |= bad;
#else
    bad = __reduce_or_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicOr(p.unsat + tile, bad);
#endif
}

// End of iteration `it`: trials whose syndrome matched are frozen (iters = it, converged = 1); counts[it] receives
// the number of trials still running.  One thread per tile.
__global__ void __launch_bounds__(256) k_bp_finish_iter(const BpDev p, int tiles, int it, int early_stop, int last)
{
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles) return;
    const uint32_t valid = p.valid[tile], done = p.done[tile];
    if (done == valid) return;
    const uint32_t ok = valid & ~done & ~p.unsat[tile];
    p.unsat[tile] = 0;
    if (early_stop || last) {
        for (int l = 0; l < 32; ++l)
            if ((ok >> l) & 1u) {
                const int o = p.orig[tile * 32 + l];
                p.iters[o] = it;
                p.conv[o] = 1;
            }
    }
    if (early_stop) {
        p.done[tile] = done | ok;
        const int running = __popc(valid & ~(done | ok));
        if (running) {
#ifdef SCALDPC_HOST_MODEL
            p.counts[it] += running;
#else
            atomicAdd(p.counts + it, running);
#endif
        }
    }
}

// One word of device memory -> page-locked host memory (UVA: the host pointer is valid on the device).  The host reads it
// after synchronising the stream.
#ifndef SCALDPC_HOST_MODEL
__global__ void k_bp_publish(const int *__restrict__ src, int *dst)
{
    *dst = *src;
    __threadfence_system();
}
#endif

// Reset of the per-decode state.  One thread per tile: valid mask from T, done = 0, identity slot map,
// iters = max_iter, conv = 0.
__global__ void __launch_bounds__(256) k_bp_reset(uint32_t *valid, uint32_t *done, uint32_t *unsat, int *orig, int *iters,
                                                  int *conv, int tiles, int T, int max_iter)
{
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles) return;
    const int left = T - tile * 32;
    valid[tile] = left >= 32 ? 0xffffffffu : ((1u << left) - 1u);
    done[tile] = 0;
    unsat[tile] = 0;
    for (int l = 0; l < 32; ++l) {
        const int t = tile * 32 + l;
        orig[t] = t < T ? t : -1;
        if (t < T) {
            iters[t] = max_iter;
            conv[t] = 0;
        }
    }
}

// The same reset when the slot -> trial map is given (segmented CSR: every graph's trials fill their own tiles, the
// last tile of a graph is padded with -1): valid = slots holding a trial.
__global__ void __launch_bounds__(256) k_bp_reset_slots(uint32_t *valid, uint32_t *done, uint32_t *unsat, const int *orig,
                                                        int *iters, int *conv, int tiles, int max_iter)
{
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles) return;
    uint32_t v = 0;
    for (int l = 0; l < 32; ++l) {
        const int t = orig[tile * 32 + l];
        if (t >= 0) {
            v |= 1u << l;
            iters[t] = max_iter;
            conv[t] = 0;
        }
    }
    valid[tile] = v;
    done[tile] = 0;
    unsat[tile] = 0;
}

// -----------------------------------------------------------------------------------------------------------
// Active-trial compaction.  When at most half of the slots still run, the running trials are packed into the
// leading lanes of fewer tiles (order preserved) so that later iterations touch only live data:
//   plan   (one CTA)   exclusive scan of the per-tile running counts -> src[new slot] = old slot (-1 = padding),
//                      new slot map / valid masks, n_act
//   move   per-trial columns of msg (E lines per tile), of the per-trial priors (n lines) and the syndrome
//          bit-planes (m words) are gathered into a second buffer set (ping-pong; the new set needs <= half the slots)
// Finished trials are written to the caller's output arrays before their slots are dropped (k_bp_unpack_* with the
// `done` masks).  dec_w / post are rewritten for running trials by every bit pass, so they are not moved.
// -----------------------------------------------------------------------------------------------------------
constexpr int kBpPlanThreads = 1024;

__global__ void __launch_bounds__(kBpPlanThreads) k_bp_compact_plan(const uint32_t *valid, const uint32_t *done,
                                                                     const int *orig, int tiles, int *src, int *orig2,
                                                                     uint32_t *valid2, uint32_t *done2, uint32_t *unsat2,
                                                                     int *n_act_out)
{
    __shared__ int part[kBpPlanThreads + 1];
    const int nt = blockDim.x, tid = threadIdx.x;
    const int per = (tiles + nt - 1) / nt;
    const int t0 = min(tiles, tid * per), t1 = min(tiles, t0 + per);
    int cnt = 0;
    for (int t = t0; t < t1; ++t) cnt += __popc(valid[t] & ~done[t]);
    part[tid] = cnt;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < nt; ++i) {
            const int c = part[i];
            part[i] = run;
            run += c;
        }
        part[nt] = run;
        *n_act_out = run;
    }
    __syncthreads();
    int pos = part[tid];
    for (int t = t0; t < t1; ++t) {
        uint32_t a = valid[t] & ~done[t];
        for (int l = 0; l < 32; ++l)
            if ((a >> l) & 1u) {
                src[pos] = t * 32 + l;
                orig2[pos] = orig[t * 32 + l];
                ++pos;
            }
    }
    const int n_act = part[nt], new_tiles = (n_act + 31) / 32;
    for (int t = tid; t < new_tiles; t += nt) {
        const int left = n_act - t * 32;
        valid2[t] = left >= 32 ? 0xffffffffu : ((1u << left) - 1u);
        done2[t] = 0;
        unsat2[t] = 0;
    }
    for (int s = n_act + tid; s < new_tiles * 32; s += nt) {
        src[s] = -1;
        orig2[s] = -1;
    }
}

// dst[(t', line, l')] = src_arr[(tile(s), line, lane(s))], s = src[t'*32 + l'].  One warp per 8 lines.
__global__ void __launch_bounds__(256) k_bp_compact_move(const float *__restrict__ from, float *__restrict__ to,
                                                         const int *__restrict__ src, int lines)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.y;
    const int l0 = (blockIdx.x * (blockDim.x >> 5) + warp) * 8;
    const int s = src[tile * 32 + lane];
    const float *f = from + ((size_t)(s >> 5) * lines) * 32 + (s & 31);
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (s >= 0 && l0 + i < lines) ? f[(size_t)(l0 + i) * 32] : 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        if (l0 + i < lines) to[((size_t)tile * lines + l0 + i) * 32 + lane] = v[i];
}

// Bit-planes: to[t'][w] bit l' = from[tile(s)][w] bit lane(s).  One warp per word.
__global__ void __launch_bounds__(256) k_bp_compact_bits(const uint32_t *__restrict__ from, uint32_t *__restrict__ to,
                                                         const int *__restrict__ src, int words)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.y;
    const int w = blockIdx.x * (blockDim.x >> 5) + warp;
    if (w >= words) return;
    const int s = src[tile * 32 + lane];
    const uint32_t bit = s >= 0 ? (from[(size_t)(s >> 5) * words + w] >> (s & 31)) & 1u : 0u;
#ifdef SCALDPC_HOST_MODEL
    uint32_t &o = to[(size_t)tile * words + w];
    o = (lane == 0 ? 0u : o) | (bit << lane);
#else
    const uint32_t word = __ballot_sync(0xffffffffu, bit != 0);
    if (lane == 0) to[(size_t)tile * words + w] = word;
#endif
}

#ifndef SCALDPC_HOST_MODEL
// ---- input / output conversion (coalesced on both sides) ----------------------------------------------------

// channel_probs (fp64, probability of a 1) -> prior LLR log((1-p)/p) in fp32 (computed in fp64: upstream's
// p/(1-p) is fp64).  Shared: out[n].  Per trial: in [T][n] -> out [tile][n][32] through a shared-memory transpose.
__global__ void __launch_bounds__(256) k_bp_prior_shared(const double *__restrict__ probs, float *__restrict__ out, int n)
{
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) {
        const double q = probs[v];
        out[v] = (float)log((1.0 - q) / q);
    }
}

// `orig` (may be null): slot -> trial map; null = identity.
__global__ void __launch_bounds__(256) k_bp_prior_tiled(const double *__restrict__ probs, float *__restrict__ out, int T, int n,
                                                        const int *__restrict__ orig)
{
    __shared__ float sm[32][33];
    const int tile = blockIdx.y, v0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int t = ty; t < 32; t += 8) {
        const int trial = orig ? orig[tile * 32 + t] : tile * 32 + t;
        float L = 0.0f;
        if (trial >= 0 && trial < T && v0 + tx < n) {
            const double q = probs[(size_t)trial * n + v0 + tx];
            L = (float)log((1.0 - q) / q);
        }
        sm[t][tx] = L;
    }
    __syncthreads();
    for (int c = ty; c < 32; c += 8)
        if (v0 + c < n) out[((size_t)tile * n + v0 + c) * 32 + tx] = sm[tx][c];
}

// syndromes [T][m] bytes -> synd_w [tile][m] words
__global__ void __launch_bounds__(256) k_bp_pack_synd(const uint8_t *__restrict__ synd, uint32_t *__restrict__ out, int T, int m,
                                                      const int *__restrict__ orig)
{
    __shared__ uint8_t sm[32][36];
    const int tile = blockIdx.y, c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int t = ty; t < 32; t += 8) {
        const int trial = orig ? orig[tile * 32 + t] : tile * 32 + t;
        sm[t][tx] = (trial >= 0 && trial < T && c0 + tx < m) ? (synd[(size_t)trial * m + c0 + tx] & 1) : 0;
    }
    __syncthreads();
    for (int c = ty; c < 32; c += 8) {
        const uint32_t w = __ballot_sync(0xffffffffu, sm[tx][c] != 0);
        if (tx == 0 && c0 + c < m) out[(size_t)tile * m + c0 + c] = w;
    }
}

// dec_w [tile][n] words -> decoding [orig trial][n] bytes, for the slots selected by mask[tile]
__global__ void __launch_bounds__(256) k_bp_unpack_dec(const uint32_t *__restrict__ dec_w, uint8_t *__restrict__ out,
                                                       const int *__restrict__ orig, const uint32_t *__restrict__ mask, int n)
{
    const int tile = blockIdx.y, v = blockIdx.x * 32 + (threadIdx.x & 31);
    const int ty = threadIdx.x >> 5;
    const uint32_t sel = mask[tile];
    if (v >= n || sel == 0) return;
    const uint32_t w = dec_w[(size_t)tile * n + v];
    for (int t = ty; t < 32; t += 8)
        if ((sel >> t) & 1u) out[(size_t)orig[tile * 32 + t] * n + v] = (w >> t) & 1u;
}

// The statistics hqc.decode derives from the decoded vector (hqc.py:709-758), one warp per trial.  H = [Hin | I_m]
// (hqc.py:680): columns < n_msg are message bits, column n_msg + i is the measurement error of check i.
// stats[t] = {unsatisfied, good_flips, bad_flips, found_bad_satisfied_checks, found_bad_unsatisfied_checks, success}:
//   good/bad flips        decoded message bits inside / outside the true support (:717-725)
//   unsatisfied           checks with syndrome bit 1 (:727-728)
//   found_bad_*           hqc.decode scores the CORRECTED vector decoded = msg ^ error (bpd.decode of a length-n word,
//                         hqc.py:703-708), so decoded[n_msg + i] = c_i ^ e_i with c_i the measured check bit and e_i the
//                         decoded measurement error (`dec` here is the error estimate e):
//                         found_bad_unsatisfied = #{c_i = 1 and not decoded} = #{c_i = 1 and e_i = 1} (:729-731),
//                         found_bad_satisfied   = #{c_i = 0 and decoded}     = #{c_i = 0 and e_i = 1} (:732-734)
//   success               support(decoded[:n_msg]) == truth (:741-749)
__global__ void __launch_bounds__(256) k_bp_score(const uint8_t *__restrict__ dec, const uint8_t *__restrict__ synd,
                                                  const int32_t *__restrict__ truth, const int32_t *__restrict__ n_checks,
                                                  int32_t *__restrict__ stats, int T, int n, int n_msg, int weight)
{
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (t >= T) return;
    const uint8_t *d = dec + (size_t)t * n;
    const uint8_t *s = synd + (size_t)t * (n - n_msg);
    int flips = 0, good = 0, unsat = 0, fbs = 0, fbu = 0;
    for (int i = lane; i < n_msg; i += 32) flips += d[i] != 0;
    for (int k = lane; k < weight; k += 32) {
        const int pos = truth[(size_t)t * weight + k];
        good += pos >= 0 && pos < n_msg && d[pos] != 0;
    }
    const int active = n_checks ? min(max(n_checks[t], 0), n - n_msg) : n - n_msg;  // nested graph: the first rows only
    for (int i = lane; i < active; i += 32) {
        const bool c = s[i] != 0, e = d[n_msg + i] != 0;
        unsat += c;
        fbu += c && e;
        fbs += !c && e;
    }
    flips = __reduce_add_sync(0xffffffffu, flips);
    good = __reduce_add_sync(0xffffffffu, good);
    unsat = __reduce_add_sync(0xffffffffu, unsat);
    fbs = __reduce_add_sync(0xffffffffu, fbs);
    fbu = __reduce_add_sync(0xffffffffu, fbu);
    if (lane == 0) {
        int32_t *o = stats + (size_t)t * 6;
        o[0] = unsat;
        o[1] = good;
        o[2] = flips - good;
        o[3] = fbs;
        o[4] = fbu;
        o[5] = (flips == good && good == weight) ? 1 : 0;
    }
}

// post [tile][n][32] -> llr [orig trial][n], for the slots selected by mask[tile]
__global__ void __launch_bounds__(256) k_bp_untile_f32(const float *__restrict__ src, float *__restrict__ dst,
                                                       const int *__restrict__ orig, const uint32_t *__restrict__ mask, int n)
{
    __shared__ float sm[32][33];
    const int tile = blockIdx.y, c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const uint32_t sel = mask[tile];
    if (sel == 0) return;
    for (int c = ty; c < 32; c += 8)
        if (c0 + c < n) sm[c][tx] = src[((size_t)tile * n + c0 + c) * 32 + tx];
    __syncthreads();
    for (int t = ty; t < 32; t += 8)
        if (((sel >> t) & 1u) && c0 + tx < n) dst[(size_t)orig[tile * 32 + t] * n + c0 + tx] = sm[tx][t];
}
#endif  // !SCALDPC_HOST_MODEL

}  // namespace scaldpc
