// Exact-order check-node kernel of the special q-ary decoder (decoder_special.rs:506-563), regular rows only.
// Included by qary_kernels.cuh INSIDE namespace scaldpc, after qary_cn_fast.cuh (shares F2 / cn_conv_fwd / CnFastShape).
//
// The reference enumerates every configuration d of the check's SW B-variables, forms
//     S(d) = ((((0 + a_0[d_0]) + a_1[d_1]) + ...) + a_{SW-1}[d_{SW-1}]) + a_sum[-sum d]      sequentially in f32
// (decoder_special.rs:539-546) and emits beta_j[d_j] = min S(d) - a_j[d_j] (:548-553).  fl(x + c) and fl(x - c) are
// monotone in x, so the minimum commutes with every rounded operation:
//     beta_j[d] = fl( M_j[d] - a_j[d] ),   M_j[d] = min over the configurations with d_j = d of S
// and M_j[d] is a forward dynamic programme that never leaves the reference's order of additions:
//     F_0 = {0: 0},  F_{k+1}[u] = min_e fl(F_k[u - e] + a_k[e])                 (prefix sums of variables 0 .. k-1)
//     T_j[v] = fl(F_j[v] + a_j[d])            (variable j pinned to d; v = total of the variables before j)
//     T_{k+1}[v] = min_e fl(T_k[v - e] + a_k[e]),  k = j+1 .. SW-1              (v = total of the others so far)
//     M_j[d] = min_v fl(T_SW[v] + a_sum[-(v + d)])
// and for the sum-variable beta_sum[t] = fl( fl(F_SW[-t] + a_sum[t]) - a_sum[t] ).  Results are BIT-IDENTICAL to the
// enumeration (oracle algo 0 / 1); the cost is ~SW/2 * Q forward passes instead of three (k_cn_fast), all in
// registers with compile-time indices: one warp per (tile, check), lane = trial, the pinned value d is the only
// run-time index (one shared-memory column per lane holds the sanitised inputs for it).
//
// k_cn_generic (exact = 1) computes the same thing for any row shape with run-time loops over shared-memory tables,
// ~30x slower; this kernel is what makes the exact-order RE-DECODE of at-risk trials affordable.

constexpr int kCnExactWarps = 4;

template <int SW, int B>
struct CnExactShape {
    using Sh = CnFastShape<SW, B>;
    static constexpr size_t smem_bytes = (size_t)kCnExactWarps * Sh::NS * 32 * sizeof(float);
};

// SPLIT: the re-decode of a FEW at-risk trials (one or two tiles: a few hundred items, not enough to fill the GPU with one warp
// each) is bound by the ~15 k dependent instructions of one warp.  One CTA of NR = ceil(SW/2) warps takes an item and warp r pins
// the variables r and SW-1-r -- the pinned variable j costs SW-1-j convolutions per value, so every warp gets SW-1 of them; all
// warps read the inputs, one barrier, then each writes its own variables' slots (the messages are updated in place).
// (A first variant with SW warps per item, one pinned variable each, was slower than the plain kernel: 35 us against 27 us --
// unbalanced warps, and 251 registers x 192 threads left one CTA per SM = two waves for 256 items.)
template <int SW>
struct CnExactSplit {
    static constexpr int NR = (SW + 1) / 2;
};

template <int SW, int B, bool SPLIT = false>
__global__ void __launch_bounds__(kCnExactWarps * 32, 2) k_cn_exact(const QaryDev p, const int n_items, const int FIRST,
                                                                    const int LAST)
{
    static_assert(SW >= 2, "k_cn_exact needs at least two B-variables per check");
    static_assert(CnExactSplit<SW>::NR <= kCnExactWarps, "the split variant uses one CTA of NR warps per item");
    using Sh = CnFastShape<SW, B>;
    constexpr int Q = Sh::Q, BS = Sh::BS, QS = Sh::QS, NS = Sh::NS;
    constexpr int NP = (SW - 1) * B + 1;  // pairs of the widest table over the B-variables: 2 (SW-1) B + 1 entries
    const float INF = SC_INF;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int item = SPLIT ? (int)blockIdx.x : blockIdx.x * kCnExactWarps + warp;
    const int role = SPLIT ? warp : -1;  // SPLIT: blockDim.x == NR * 32
    if (item >= n_items) return;
    const int tile = item / p.R, check = item - tile * p.R;
    const int e0 = __ldg(p.row_ptr + check);
    const int base = __ldg(p.edge_off + e0);
    const int vs = __ldg(p.col + e0 + SW) - p.bvars;
    const bool negs = __ldg(p.sign + e0 + SW) < 0;
    float *msg = p.msg + (size_t)tile * p.S * 32 + lane;
    float *out = msg + (size_t)base * 32;  // the check's slot block
#ifdef SCALDPC_HOST_MODEL
    float sx_[NS];
#define SX(idx) sx_[(idx)]
#define SXP(idx, off) sx_[(idx) + (off)]
#else
    extern __shared__ __align__(128) float sc_cn_smem[];
    float *sx_ = sc_cn_smem + (size_t)warp * NS * 32 + lane;
#define SX(idx) sx_[(idx)*32]
#define SXP(idx, off) sx_[((idx) + (off)) * 32]
#endif

    // ---- inputs -> registers (check frame); iteration 1 takes them from the channel (decoder_special.rs:480-486)
    float a[SW][Q];
#pragma unroll
    for (int k = 0; k < SW; ++k) {
        if (FIRST) {
            const int var = __ldg(p.col + e0 + k);
            const bool neg = __ldg(p.sign + e0 + k) < 0;
            const float *src = p.llr + ((size_t)tile * p.SL + (size_t)var * Q) * 32 + lane;
#pragma unroll
            for (int i = 0; i < Q; ++i) a[k][i] = src[(neg ? Q - 1 - i : i) * 32];
        } else {
#pragma unroll
            for (int i = 0; i < Q; ++i) a[k][i] = out[(size_t)(k * Q + i) * 32];
        }
    }
    // the sum-variable: v2c = perm_h(channel) in iteration 1, else its own update from channel + previous c2v
    // (decoder_special.rs:591-604), check frame
    const float *chs = p.llr_sum + ((size_t)tile * p.SLS + (size_t)vs * QS) * 32 + lane;
    float as[QS];
#pragma unroll
    for (int i = 0; i < QS; ++i) as[i] = chs[(negs ? QS - 1 - i : i) * 32];
    if (!FIRST) {
        float m = INF;
#pragma unroll
        for (int i = 0; i < QS; ++i) {
            const float c = out[(size_t)(SW * Q + i) * 32];
            const float tot = as[i] + c;  // sum = channel + perm_h(c2v)
            as[i] = tot - c;              // prim = perm_h(sum - perm_h(c2v))
            m = fminf(m, as[i]);          // value of the first strict minimum (NaN skipped)
        }
#pragma unroll
        for (int i = 0; i < QS; ++i) as[i] = as[i] - m;
    }
#pragma unroll
    for (int k = 0; k < SW; ++k)
#pragma unroll
        for (int i = 0; i < Q; ++i) {
            a[k][i] = sanitize(a[k][i]);
            SX(k * Q + i) = a[k][i];
        }
#pragma unroll
    for (int i = 0; i < QS; ++i) SX(SW * Q + i) = sanitize(as[i]);  // `as` itself is dead from here on
#ifndef SCALDPC_HOST_MODEL
    if (SPLIT) __syncthreads();  // every warp of the item has read the slots before anyone overwrites them
#endif

    // ---- per variable j: every value d pinned in turn; Fp = F_j as pairs (F_j[v] <-> prefix total v - j*B)
    F2 Fp[NP];
#pragma unroll
    for (int j = 0; j < SW; ++j) {
        const int nF = 2 * j * B + 1;  // entries of F_j (compile-time after unrolling)
        const bool mine = !SPLIT || role == (j < SW - 1 - j ? j : SW - 1 - j);
#pragma unroll 1
        for (int d = 0; d < (mine ? Q : 0); ++d) {
            const float x = SX(j * Q + d);
            // T = F_j + a_j[d]; index v of T <-> total of the OTHER variables so far, offset (number of them) * B
            F2 Tp[NP];
            if (j == 0) {
                Tp[0] = f2_make(0.0f + x, INF);
            } else {
#pragma unroll
                for (int jj = 0; jj < (nF + 1) / 2; ++jj) Tp[jj] = f2_add_s(Fp[jj], x);
            }
            float Ts[2 * NP + Q];
            if (j == SW - 1) {
#pragma unroll
                for (int v = 0; v < nF; ++v) Ts[v] = (v & 1) ? f2_hi(Tp[v >> 1]) : f2_lo(Tp[v >> 1]);
            }
#pragma unroll
            for (int k = j + 1; k < SW; ++k) {
                const int nin = 2 * (k - 1) * B + 1;
                cn_conv_fwd<Q>(Tp, nin, a[k], Ts);
                if (k < SW - 1) {
                    Ts[nin + Q - 1] = INF;
#pragma unroll
                    for (int jj = 0; jj < (nin + Q) / 2; ++jj) Tp[jj] = f2_make(Ts[2 * jj], Ts[2 * jj + 1]);
                }
            }
            // M = min_v T[v] + a_sum[-(v' + d')]: index of a_sum = (QS - 1) - d - v
            MinAcc acc;
#pragma unroll
            for (int v = 0; v < 2 * (SW - 1) * B + 1; ++v) acc.add(Ts[v] + SXP(SW * Q + QS - 1 - d, -v));
            const float M = acc.get();
            out[(size_t)(j * Q + d) * 32] = (x < INF && M < INF) ? M - x : INF;
        }
        // F_{j+1} = F_j (+) a_j
        float Fs[2 * NP + Q];
        if (j == 0) {
#pragma unroll
            for (int i = 0; i < Q; ++i) Fs[i] = 0.0f + a[0][i];
        } else {
            cn_conv_fwd<Q>(Fp, nF, a[j], Fs);
        }
        if (j < SW - 1) {
            Fs[nF + Q - 1] = INF;
#pragma unroll
            for (int jj = 0; jj < (nF + Q) / 2; ++jj) Fp[jj] = f2_make(Fs[2 * jj], Fs[2 * jj + 1]);
        } else if (!SPLIT || role == 0) {
            // ---- the sum-variable: beta_sum[t] = (F_SW[-t] + a_sum[t]) - a_sum[t]; Fs = F_SW, QS entries
            float min_val = INF;
            int arg = negs ? QS - 1 : 0;
            float *post = (LAST && p.post_sum) ? p.post_sum + (size_t)tile * p.SLS * 32 + lane : nullptr;
#pragma unroll
            for (int i = 0; i < QS; ++i) {  // i = t + BS (check frame), F_SW index QS - 1 - i
                const float xs = SX(SW * Q + i), f = Fs[QS - 1 - i];
                const float c = (xs < INF && f < INF) ? (f + xs) - xs : INF;
                out[(size_t)(SW * Q + i) * 32] = c;
                if (LAST) {  // decision of the sum-variable (decoder_special.rs:606-608), first minimum in ITS frame
                    const int iv = negs ? QS - 1 - i : i;
                    const float tot = chs[iv * 32] + c;
                    if (negs ? (tot <= min_val && tot < INF) : (tot < min_val)) {
                        min_val = tot;
                        arg = i;
                    }
                    if (post) post[(size_t)(vs * QS + iv) * 32] = tot;
                }
            }
            if (LAST) {
                const int argv = negs ? QS - 1 - arg : arg;
                p.hard_t[((size_t)tile * p.N + p.bvars + vs) * 32 + lane] = (int8_t)(argv - BS);
            }
        }
    }
#undef SX
#undef SXP
}
