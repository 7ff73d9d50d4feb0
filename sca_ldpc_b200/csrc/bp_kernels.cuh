// Binary product-sum BP kernels for sm_100a (B200): flooding schedule, syndrome-match early stop, per-trial
// freeze + tile retirement.  Replaces ldpc.bp_decoder(bp_method="product_sum") as called at
// simulate/hqc.py:694-708 and simulate/decode.py:155-171 (third-party PyPI ldpc==0.1.3, algorithm as summarised in
// SURVEY.md appendix B).  Hand-written CUDA; no tensor cores (irregular gather + scalar updates, HBM-bound).
//
// Arithmetic: upstream keeps fp64 probability RATIOS r = p1/p0 and multiplies running products; here messages are
// fp32 LLRs L = log(p0/p1) = -log r, so every product becomes a sum and the check-node rule
//     c2b = (1-x)/(1+x),  x = (-1)^synd * prod_{others} (2/(1+r_k) - 1) = (-1)^synd * prod tanh(L_k/2)
// becomes, with phi(x) = -log tanh(x/2) (its own inverse),
//     |c2v| = phi( sum_{others} phi(|L_k|) ),   sign(c2v) = (-1)^synd * prod_{others} sign(L_k).
// The sums over "others" are forward/backward running sums in the row's column order (upstream's running products),
// never total-minus-self, so there is no cancellation.  The bit-node rule keeps upstream's sequential semantics
// including "NaN -> neutral element" resets of the running product (0 * inf in ratios == inf - inf in LLRs).
//
// Data layout in HBM ("tile" layout, shared with the q-ary decoder): 32 trials per tile, lane = trial % 32 innermost.
//     msg     [tile][E][32] f32   v2c before a check pass, c2v after it (in place; one writer per slot)
//     prior   [n] f32 (shared by all trials)  or  [tile][n][32] f32 (per-trial channel_probs)
//     synd_w  [tile][m]  u32      bit l = syndrome bit of trial l of the tile
//     dec_w   [tile][n]  u32      bit l = current hard decision of trial l      (warp ballot)
//     post    [tile][n][32] f32   posterior LLRs (only when the caller asks for them)
//     done    [tile] u32          bit l = trial l has converged (outputs frozen);  valid [tile] u32 = real trials
//     unsat   [tile] u32          bit l = some check of trial l is unsatisfied this iteration
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

struct BpDev {
    int m, n, E;
    int dc_max, dv_max;
    int per_trial_prior;      // 0: prior[n]; 1: prior[tile][n][32]
    const int *row_ptr;       // m+1
    const int *col;           // E
    const int *col_ptr;       // n+1
    const int *csc_edge;      // E
    float *msg;
    const float *prior;
    const uint32_t *synd_w;
    uint32_t *dec_w;
    float *post;              // may be null
    uint32_t *done;
    const uint32_t *valid;
    uint32_t *unsat;
    int *iters;               // tiles*32
    int *conv;                // tiles*32
    int *n_active;            // [1] tiles still running after the last finished iteration
};

#ifdef SCALDPC_HOST_MODEL
__device__ __forceinline__ bool bp_signbit(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    return u >> 31;
}
#else
__device__ __forceinline__ bool bp_signbit(float x) { return __float_as_uint(x) >> 31; }
#endif

// phi(x) = -log(tanh(x/2)) = log((1+e^-x)/(1-e^-x)) for x >= 0; phi(0) = +inf, phi(+inf) = 0, phi(NaN) = NaN.
// Three ranges keep the relative error at a few ulp without expm1/log1p calls.
__device__ __forceinline__ float bp_phi(float x)
{
    if (x < 0.125f) {  // -log(y) + y^2/3 - 7y^4/90, y = x/2
        const float y = 0.5f * x, y2 = y * y;
        return y2 * (0.33333334f - 0.07777778f * y2) - logf(y);
    }
    const float t = expf(-x);
    if (t < 0.03125f) {  // 2 atanh(t) = 2t (1 + t^2/3 + t^4/5 + ...)
        const float t2 = t * t;
        return 2.0f * t * (1.0f + t2 * (0.33333334f + 0.2f * t2));
    }
    return logf((1.0f + t) / (1.0f - t));  // NaN lands here and stays NaN
}

__device__ __forceinline__ float bp_prior_of(const BpDev &p, int tile, int v, int lane)
{
    return p.per_trial_prior ? p.prior[((size_t)tile * p.n + v) * 32 + lane] : __ldg(p.prior + v);
}

constexpr int kBpWarps = 4;  // warps per CTA of the check / bit kernels (each warp owns 2*dmax*128 B of scratch)

// -----------------------------------------------------------------------------------------------------------
// Check-to-bit pass.  One warp = (tile, check), lane = trial.  Scratch per lane: A[k] = phi(|L_k|) with the sign
// of L_k, P[k] = exclusive prefix sum of phi.  FIRST: v2c is the channel prior (upstream's init b2c = p/(1-p)).
// -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBpWarps * 32) k_bp_cn(const BpDev p, const int FIRST)
{
    SCALDPC_DYN_SMEM(float, smem);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int check = blockIdx.x * kBpWarps + warp;
    const int tile = blockIdx.y;
    if (check >= p.m) return;
    if (__ldg(p.done + tile) == __ldg(p.valid + tile)) return;  // every trial of the tile has converged
    const int e0 = __ldg(p.row_ptr + check), d = __ldg(p.row_ptr + check + 1) - e0;
    float *A = smem + (size_t)warp * 2 * p.dc_max * 32 + lane;
    float *P = A + (size_t)p.dc_max * 32;
    float *msg = p.msg + ((size_t)tile * p.E + e0) * 32 + lane;

    bool neg = (__ldg(p.synd_w + (size_t)tile * p.m + check) >> lane) & 1u;
    float pre = 0.0f;
#pragma unroll 4
    for (int k = 0; k < d; ++k) {
        const float v = FIRST ? bp_prior_of(p, tile, __ldg(p.col + e0 + k), lane) : msg[(size_t)k * 32];
        const float ph = bp_phi(fabsf(v));
        A[k * 32] = copysignf(ph, v);
        P[k * 32] = pre;
        pre += ph;
        neg ^= bp_signbit(v);
    }
    float suf = 0.0f;
#pragma unroll 4
    for (int k = d - 1; k >= 0; --k) {
        const float a = A[k * 32];
        const float mag = bp_phi(P[k * 32] + suf);
        msg[(size_t)k * 32] = (neg != bp_signbit(a)) ? -mag : mag;
        suf += fabsf(a);
    }
}

// -----------------------------------------------------------------------------------------------------------
// Bit-to-check pass + decision.  One warp = (tile, variable), lane = trial.  Upstream's two sweeps over the
// column (rows ascending): forward running value seeded with the prior, backward running value seeded with the
// neutral element; a NaN running value is reset to the neutral element (ratio 1 == LLR 0).
// Decision: upstream decodes 1 when ratio >= 1, i.e. LLR <= 0.  Converged (done) lanes keep their old outputs.
// -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBpWarps * 32) k_bp_vn(const BpDev p)
{
    SCALDPC_DYN_SMEM(float, smem);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int var = blockIdx.x * kBpWarps + warp;
    const int tile = blockIdx.y;
    if (var >= p.n) return;
    const uint32_t done = __ldg(p.done + tile);
    if (done == __ldg(p.valid + tile)) return;
    const int p0 = __ldg(p.col_ptr + var), d = __ldg(p.col_ptr + var + 1) - p0;
    float *C = smem + (size_t)warp * 2 * p.dv_max * 32 + lane;
    float *F = C + (size_t)p.dv_max * 32;
    float *msg = p.msg + (size_t)tile * p.E * 32 + lane;

    float temp = bp_prior_of(p, tile, var, lane);
#pragma unroll 4
    for (int i = 0; i < d; ++i) {
        const float c = msg[(size_t)__ldg(p.csc_edge + p0 + i) * 32];
        C[i * 32] = c;
        F[i * 32] = temp;
        temp += c;
        if (temp != temp) temp = 0.0f;
    }
    const float post = temp;
    temp = 0.0f;
    for (int i = d - 1; i >= 0; --i) {
        msg[(size_t)__ldg(p.csc_edge + p0 + i) * 32] = F[i * 32] + temp;
        temp += C[i * 32];
        if (temp != temp) temp = 0.0f;
    }
    const bool active = !((done >> lane) & 1u);
    if (p.post && active) p.post[((size_t)tile * p.n + var) * 32 + lane] = post;
#ifdef SCALDPC_HOST_MODEL
    uint32_t &w = p.dec_w[(size_t)tile * p.n + var];
    if (active) w = (w & ~(1u << lane)) | ((post <= 0.0f ? 1u : 0u) << lane);
#else
    const uint32_t bits = __ballot_sync(0xffffffffu, post <= 0.0f);
    if (lane == 0) {
        uint32_t *w = p.dec_w + (size_t)tile * p.n + var;
        *w = (*w & done) | (bits & ~done);
    }
#endif
}

// -----------------------------------------------------------------------------------------------------------
// Syndrome test, bit-parallel over the 32 trials of a tile: one thread = (tile, check) XORs the decision words of
// the check's variables and compares with the syndrome word; mismatching trials are OR-ed into unsat[tile].
// -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bp_syndrome(const BpDev p)
{
    const int tile = blockIdx.y;
    const int check = blockIdx.x * blockDim.x + threadIdx.x;
    if (__ldg(p.done + tile) == __ldg(p.valid + tile)) return;
    uint32_t bad = 0;
    if (check < p.m) {
        const uint32_t *dec = p.dec_w + (size_t)tile * p.n;
        uint32_t par = __ldg(p.synd_w + (size_t)tile * p.m + check);
        const int e0 = __ldg(p.row_ptr + check), e1 = __ldg(p.row_ptr + check + 1);
        for (int e = e0; e < e1; ++e) par ^= dec[__ldg(p.col + e)];
        bad = par;
    }
#ifdef SCALDPC_HOST_MODEL
    p.unsat[tile] |= bad;
#else
    bad = __reduce_or_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicOr(p.unsat + tile, bad);
#endif
}

// End of iteration `it`: trials whose syndrome matched are frozen (iters = it, converged = 1); counts the tiles
// that still have running trials.  One thread per tile.
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
                p.iters[tile * 32 + l] = it;
                p.conv[tile * 32 + l] = 1;
            }
    }
    if (early_stop) {
        p.done[tile] = done | ok;
        if ((done | ok) != valid) {
#ifdef SCALDPC_HOST_MODEL
            ++*p.n_active;
#else
            atomicAdd(p.n_active, 1);
#endif
        }
    }
}

// Reset of the per-decode state.  One thread per tile: valid mask from T, done = 0, iters = max_iter, conv = 0.
__global__ void __launch_bounds__(256) k_bp_reset(uint32_t *valid, uint32_t *done, uint32_t *unsat, int *iters, int *conv,
                                                  int tiles, int T, int max_iter)
{
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles) return;
    const int left = T - tile * 32;
    valid[tile] = left >= 32 ? 0xffffffffu : ((1u << left) - 1u);
    done[tile] = 0;
    unsat[tile] = 0;
    for (int l = 0; l < 32; ++l) {
        iters[tile * 32 + l] = max_iter;
        conv[tile * 32 + l] = 0;
    }
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

__global__ void __launch_bounds__(256) k_bp_prior_tiled(const double *__restrict__ probs, float *__restrict__ out, int T, int n)
{
    __shared__ float sm[32][33];
    const int tile = blockIdx.y, v0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int t = ty; t < 32; t += 8) {
        const int trial = tile * 32 + t;
        float L = 0.0f;
        if (trial < T && v0 + tx < n) {
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
__global__ void __launch_bounds__(256) k_bp_pack_synd(const uint8_t *__restrict__ synd, uint32_t *__restrict__ out, int T, int m)
{
    __shared__ uint8_t sm[32][36];
    const int tile = blockIdx.y, c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int t = ty; t < 32; t += 8) {
        const int trial = tile * 32 + t;
        sm[t][tx] = (trial < T && c0 + tx < m) ? (synd[(size_t)trial * m + c0 + tx] & 1) : 0;
    }
    __syncthreads();
    for (int c = ty; c < 32; c += 8) {
        const uint32_t w = __ballot_sync(0xffffffffu, sm[tx][c] != 0);
        if (tx == 0 && c0 + c < m) out[(size_t)tile * m + c0 + c] = w;
    }
}

// dec_w [tile][n] words -> decoding [T][n] bytes
__global__ void __launch_bounds__(256) k_bp_unpack_dec(const uint32_t *__restrict__ dec_w, uint8_t *__restrict__ out, int T, int n)
{
    const int tile = blockIdx.y, v = blockIdx.x * 32 + (threadIdx.x & 31);
    const int ty = threadIdx.x >> 5;
    if (v >= n) return;
    const uint32_t w = dec_w[(size_t)tile * n + v];
    for (int t = ty; t < 32; t += 8) {
        const int trial = tile * 32 + t;
        if (trial < T) out[(size_t)trial * n + v] = (w >> t) & 1u;
    }
}

// post [tile][n][32] -> llr [T][n]
__global__ void __launch_bounds__(256) k_bp_untile_f32(const float *__restrict__ src, float *__restrict__ dst, int T, int n)
{
    __shared__ float sm[32][33];
    const int tile = blockIdx.y, c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int c = ty; c < 32; c += 8)
        if (c0 + c < n) sm[c][tx] = src[((size_t)tile * n + c0 + c) * 32 + tx];
    __syncthreads();
    for (int t = ty; t < 32; t += 8) {
        const int trial = tile * 32 + t;
        if (trial < T && c0 + tx < n) dst[(size_t)trial * n + c0 + tx] = sm[tx][t];
    }
}
#endif  // !SCALDPC_HOST_MODEL

}  // namespace scaldpc
