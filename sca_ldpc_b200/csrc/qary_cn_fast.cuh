// Fast check-node kernel of the special q-ary decoder (decoder_special.rs:506-563), regular rows only.
// Included by qary_kernels.cuh INSIDE namespace scaldpc.
//
// Work item = (tile of 32 trials, check); one warp per item, lane = trial.  Every check has exactly SW B-variables
// followed by its sum-variable and BSUM == SW*B, so its message slots are ONE contiguous block of
// (SW*Q + QS) x 128 B in the tile layout.  The kernel is persistent (one CTA of kCnWarps warps per SM, warps
// stride over the items) and decouples memory from arithmetic:
//   * lane 0 of a warp issues the NEXT item's inputs as bulk-async copies (TMA, cp.async.bulk.shared.global with
//     mbarrier complete_tx): the message block and the sum-variable's channel LLRs, ~10-15 KB per item;
//   * all lanes wait on the mbarrier, pull their column of the staged block into registers, and the buffer is
//     immediately re-armed for the following item, so the HBM latency of item i+1 hides behind the ~3-4 k
//     arithmetic instructions of item i (v0 issued plain loads just before use: 71 % long-scoreboard stalls);
//   * results go straight from registers to global memory (coalesced 128 B stores, fire and forget).
//
// Arithmetic: the reference's brute-force enumeration of (2B+1)^SW configurations is replaced by (min,+) convolutions
// (SURVEY.md appendix A.4).  Round 1 ran ONE forward/backward chain with the wide sum-variable table S (QS entries) at its
// end; since round 2 the chain is SPLIT around S, which needs ~12 % fewer (min,+) candidates (SW = 6, B = 3: 1562 against
// 1778) because S meets the half-width tables once instead of every prefix meeting a suffix that contains S.
// With P = SW/2 left variables a_0 .. a_{P-1} and M = SW - P right variables:
//     Fl_1 = a_0,       Fl_{k+1}[u] = min_d Fl_k[u-d] + a_k[d]                (prefix chain of the left variables)
//     Fr_1 = a_{SW-1},  Fr_{k+1}[u] = min_d Fr_k[u-d] + a_{SW-1-k}[d]         (suffix chain of the right variables)
//     TR[v] = min_u Fl_P[u] + S[-u-v]      (left half + sum-variable, given that the right variables total v)
//     TL[u] = min_v Fr_M[v] + S[-u-v]      (right half + sum-variable, given that the left variables total u)
// and each half is an ordinary backward + combine pass whose terminal table is TR (right half, variables x_k = a_{SW-1-k},
// prefixes Fr_k) or TL (left half, x_k = a_k, prefixes Fl_k):
//     G_CNT = terminal,  G_k[u] = min_d x_k[d] + G_{k+1}[u+d],   c2v_{x_k}[d] = min_u F_k[u] + G_{k+1}[u+d]
// c2v_sum[t] = F_SW[-t] is needed by two launches per decode only and is formed by carrying the LEFT chain on through the
// right variables, i.e. in the reference's own order of additions (decoder_special.rs:539-545).
// "Non-finite input => value excluded" (sanitize) and "own entry non-finite => +inf out" (decoder.rs:307-311,
// decoder_special.rs:549-553) as before.  The degree-1 sum-variable's own update (decoder_special.rs:591-609) is fused: its
// messages never travel through HBM in middle iterations, and on the last iteration its hard decision / posterior are
// emitted from here.  oracle/qary_impl.inc cn_split is this arithmetic on the CPU, bit for bit (tests/test_hostmodel.py).

// ---- packed f32x2 arithmetic (sm_100a: add.rn.f32x2 -> FADD2 with a broadcast scalar operand).  One instruction
// adds the same scalar to two neighbouring table entries, halving the FADD count of the (min,+) convolutions.
// Rounding is IEEE round-to-nearest per element, identical to the scalar FADD (and to the host model).
#ifdef SCALDPC_HOST_MODEL
struct F2 {
    float lo, hi;
};
__device__ __forceinline__ F2 f2_make(float lo, float hi) { return F2{lo, hi}; }
__device__ __forceinline__ float f2_lo(const F2 &x) { return x.lo; }
__device__ __forceinline__ float f2_hi(const F2 &x) { return x.hi; }
__device__ __forceinline__ F2 f2_add_s(const F2 &x, float s) { return F2{x.lo + s, x.hi + s}; }
#else
struct F2 {
    unsigned long long v;
};
__device__ __forceinline__ F2 f2_make(float lo, float hi)
{
    F2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float f2_lo(const F2 &x)
{
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x.v));
    return lo;
}
__device__ __forceinline__ float f2_hi(const F2 &x)
{
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x.v));
    return hi;
}
__device__ __forceinline__ F2 f2_add_s(const F2 &x, float s)
{
    F2 r;
    asm("{\n\t.reg .b64 t;\n\tmov.b64 t, {%2, %2};\n\tadd.rn.f32x2 %0, %1, t;\n\t}" : "=l"(r.v) : "l"(x.v), "f"(s));
    return r;
}
#endif

// out[o] = min_i X[o - i] + a[i], o = 0 .. nx + Q - 2 (X given as pairs, nx entries).  The sum X[j] + a[i] for even j
// also yields X[j+1] + a[i], which the next output consumes (carry).
template <int Q, int NP, int NO>
__device__ __forceinline__ void cn_conv_fwd(const F2 (&Xp)[NP], const int nx, const float (&a)[Q], float (&out)[NO])
{
    F2 carry[Q];
#pragma unroll
    for (int o = 0; o < 2 * NP + Q; ++o) {
        if (o < nx + Q - 1) {
            MinAcc acc;
#pragma unroll
            for (int i = 0; i < Q; ++i) {
                const int j = o - i;
                if (j >= 0 && j < nx) {
                    if ((j & 1) == 0) {
                        carry[i] = f2_add_s(Xp[j >> 1], a[i]);
                        acc.add(f2_lo(carry[i]));
                    } else {
                        acc.add(f2_hi(carry[i]));
                    }
                }
            }
            out[o] = acc.get();
        }
    }
}

// out[o] = min_i a[i] + X[o + i], o = 0 .. no - 1 (X given as pairs, at least no + Q - 1 entries).
template <int Q, int NP, int NO>
__device__ __forceinline__ void cn_conv_bwd(const F2 (&Xp)[NP], const int no, const float (&a)[Q], float (&out)[NO])
{
    F2 carry[Q];
#pragma unroll
    for (int o = 0; o < 2 * NP; ++o) {
        if (o < no) {
            MinAcc acc;
#pragma unroll
            for (int i = 0; i < Q; ++i) {
                const int j = o + i;
                if ((j & 1) == 0 || o == 0) carry[i] = f2_add_s(Xp[j >> 1], a[i]);
                acc.add((j & 1) ? f2_hi(carry[i]) : f2_lo(carry[i]));
            }
            out[o] = acc.get();
        }
    }
}

// out[o] = min_i F[i] + G[i + o], i = 0 .. NF-1, o = 0 .. NO-1 (F and G given as pairs; G holds at least NF + NO - 1 entries).
// One packed add serves two neighbouring outputs.
template <int NF, int NO, int NPF, int NPG>
__device__ __forceinline__ void cn_comb(const F2 (&Fp)[NPF], const F2 (&Gp)[NPG], float (&out)[NO + 1])
{
    MinAcc acc[NO];
#pragma unroll
    for (int j = 0; j < NF; ++j) {
        const float s = (j & 1) ? f2_hi(Fp[j >> 1]) : f2_lo(Fp[j >> 1]);
#pragma unroll
        for (int jj = j >> 1; jj <= (j + NO - 1) >> 1; ++jj) {
            const F2 t = f2_add_s(Gp[jj], s);
            const int i0 = 2 * jj - j, i1 = 2 * jj + 1 - j;
            if (i0 >= 0 && i0 < NO) acc[i0].add(f2_lo(t));
            if (i1 >= 0 && i1 < NO) acc[i1].add(f2_hi(t));
        }
    }
#pragma unroll
    for (int i = 0; i < NO; ++i) out[i] = acc[i].get();
}

constexpr int kCnWarps = 8;  // warps per persistent CTA; 8 x 32 x 255 registers = one SM's register file

#include "tma.cuh"

// Floats of staging buffer per lane: the check's message block followed by the sum-variable's channel LLRs.
template <int SW, int B>
struct CnFastShape {
    static constexpr int Q = 2 * B + 1, BS = SW * B, QS = 2 * BS + 1;
    static constexpr int NS = SW * Q + QS;  // message slots of one check
    static constexpr int NB = NS + QS;      // + channel LLRs of the sum-variable
    static constexpr size_t smem_bytes = (size_t)kCnWarps * NB * 32 * sizeof(float);
};

// MODE bits: which iteration of the decode this launch is.
//   kCnFirst  iteration 1: inputs are the variables' channel LLRs (k_init_v2c fused)
//   kCnLast   last iteration: emits the sum-variable's hard decision / posterior
//   kCnWsum   writes the sum-variable's c2v (F_SW) to its message slots -- only the second-to-last iteration needs to
//   kCnRsum   reads it back and forms the sum-variable's v2c as the reference does, normalise((channel + c2v) - c2v)
//             (decoder_special.rs:591-604) -- only the last iteration does
// Every other iteration takes the sum-variable's v2c straight from the channel, normalise(channel) -- formed once by
// iteration 1 and parked, sanitised and in reading order, in the sum-variable's own (otherwise idle) message slots, so a
// middle launch stages ONE contiguous block per item -- and neither computes nor moves its c2v: (a + c) - c is `a` up to one rounding, a value excluded by c2v = +inf stays
// unreachable through the B-variables (their exclusions are sticky), and the only other effect of an exclusion -- a
// different normalisation constant -- shifts all of the check's outputs uniformly, which the variable node's own
// normalisation removes from every later message.  It survives only in the posteriors, which is why the LAST
// iteration uses the reference's formula on the c2v the iteration before it stored (oracle algo 3 models exactly this).
// "Own entry non-finite => +inf out" (decoder.rs:307-311) is NOT applied to the B-variables' messages: an own entry v2c[i] is
// non-finite exactly when the variable's total[i] was, that is when its channel LLR[i] is +inf or some check finds the value
// infeasible given the other variables' exclusions -- both persist (exclusions only grow), so total[i] stays +inf whether this
// check adds +inf or a finite r[i] to it, every message derived from total[i] is excluded (+inf after sanitising) either way,
// and the finite entries of every total never see a masked term.  Decisions, posteriors and every later message are
// bit-identical to the masked computation (the oracle's algo 3 keeps the mask; tests/test_hostmodel.py and the GPU parity
// tests compare bit for bit, priors with excluded values included); 2 x SW x Q ALU instructions per item saved.
constexpr bool kCnMaskOwn = false;
constexpr int kCnFirst = 1, kCnLast = 2, kCnWsum = 4, kCnRsum = 8;

template <int SW, int B, int MODE>
__global__ void __launch_bounds__(kCnWarps * 32, 1) k_cn_fast(const QaryDev p, const int n_items)
{
    constexpr bool FIRST = MODE & kCnFirst, LAST = MODE & kCnLast, WSUM = MODE & kCnWsum, RSUM = MODE & kCnRsum;
    constexpr bool HOME = !FIRST && !RSUM;  // the sum-variable's v2c is the table iteration 1 parked in its message slots
    static_assert(!(FIRST && RSUM), "the first iteration has no previous c2v");
    static_assert(SW >= 2, "k_cn_fast needs at least two B-variables per check");
    using Sh = CnFastShape<SW, B>;
    constexpr int Q = Sh::Q, BS = Sh::BS, QS = Sh::QS, NS = Sh::NS;
    constexpr int P = SW / 2, M = SW - P;  // left / right variables of the split chain (see the header comment)
    const float INF = SC_INF;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // warp-major numbering: consecutive items go to different CTAs, so a batch with fewer items than resident warps (one tile:
    // the reference-shaped single decode) spreads over the SMs one warp per scheduler instead of filling a few SMs
    const int gw = warp * gridDim.x + blockIdx.x, nw = gridDim.x * kCnWarps;
#ifdef SCALDPC_HOST_MODEL
    float hbuf[Sh::NB];  // the host model stages per thread
#define SC_BUF(idx) hbuf[(idx)]
#define SC_STAGE(dst_slot, src, nslots) \
    for (int s_ = 0; s_ < (nslots); ++s_) hbuf[(dst_slot) + s_] = (src)[(size_t)s_ * 32 + lane]
#else
    extern __shared__ __align__(128) float sc_cn_smem[];
    __shared__ uint64_t bars[kCnWarps];
    float *bufbase = sc_cn_smem + (size_t)warp * Sh::NB * 32;
    const float *buf = bufbase + lane;
    uint64_t *bar = &bars[warp];
    uint32_t phase = 0;
    if (lane == 0) sc_mbar_init(bar, 1);
    __syncwarp();
#define SC_BUF(idx) buf[(idx)*32]
#define SC_STAGE(dst_slot, src, nslots) sc_bulk_g2s(bufbase + (dst_slot)*32, (src), (nslots)*128u, bar)
#endif

    // Stage the inputs of `item` (device: lane 0 only, asynchronous; host model: every thread, synchronous).
    auto issue = [&](int item) {
        const int tile = item / p.R, check = item - tile * p.R;
        if (done_word(p, tile) == 0xffffffffu) return;  // early termination: the item is skipped, nothing is staged
        const int e0 = __ldg(p.row_ptr + check);
#ifndef SCALDPC_HOST_MODEL
        sc_mbar_expect_tx(bar, (uint32_t)((FIRST ? SW * Q + QS : (RSUM ? NS + QS : NS)) * 128));
#endif
        if (FIRST) {  // v2c = perm_h(channel) (decoder_special.rs:480-486): gather the variables' channel LLRs
            for (int k = 0; k < SW; ++k) {
                const int var = __ldg(p.col + e0 + k);
                SC_STAGE(k * Q, p.llr + ((size_t)tile * p.SL + (size_t)var * Q) * 32, Q);
            }
        } else {
            SC_STAGE(0, p.msg + ((size_t)tile * p.S + __ldg(p.edge_off + e0)) * 32, NS);  // HOME: the table; RSUM: c2v_sum
        }
        if (FIRST || RSUM) {  // the sum-variable's channel LLRs
            const int vs = __ldg(p.col + e0 + SW) - p.bvars;
            SC_STAGE(NS, p.llr_sum + ((size_t)tile * p.SLS + (size_t)vs * QS) * 32, QS);
        }
    };

#ifndef SCALDPC_HOST_MODEL
    if (gw < n_items && lane == 0) issue(gw);
#endif
    for (int item = gw; item < n_items; item += nw) {
        const int tile = item / p.R, check = item - tile * p.R;
        const uint32_t frozen_w = done_word(p, tile);
        if (frozen_w == 0xffffffffu) {  // every trial of the tile is final: `issue` staged nothing for this item either
#ifndef SCALDPC_HOST_MODEL
            if (item + nw < n_items && lane == 0) issue(item + nw);
#endif
            continue;
        }
        const bool frozen = (frozen_w >> lane) & 1u;
        const int e0 = __ldg(p.row_ptr + check);
        const int base = __ldg(p.edge_off + e0);
        const int vs = __ldg(p.col + e0 + SW) - p.bvars;
        const bool negs = __ldg(p.sign + e0 + SW) < 0;
        const int sbase = base + SW * Q;
        float *msg = p.msg + (size_t)tile * p.S * 32 + lane;
#ifdef SCALDPC_HOST_MODEL
        issue(item);
#else
        sc_mbar_wait(bar, phase);
        phase ^= 1u;
#endif
        // ---- staged inputs -> registers (messages are already in the check frame after the first iteration)
        float a[SW][Q];
#pragma unroll
        for (int k = 0; k < SW; ++k) {
            const bool neg = FIRST && (__ldg(p.sign + e0 + k) < 0);
#pragma unroll
            for (int i = 0; i < Q; ++i) a[k][i] = FIRST ? SC_BUF(k * Q + (neg ? Q - 1 - i : i)) : SC_BUF(k * Q + i);
        }
        // the sum-variable: v2c from channel + previous c2v (decoder_special.rs:591-604), check frame
        float as[QS];
        if (HOME) {  // normalise(channel), sanitised, check frame, REVERSED: iteration 1 left it in the sum-variable's idle slots
#pragma unroll
            for (int i = 0; i < QS; ++i) as[i] = SC_BUF(SW * Q + QS - 1 - i);
        } else {
#pragma unroll
            for (int i = 0; i < QS; ++i) as[i] = SC_BUF(NS + i);
            if (negs) {  // warp-uniform; never taken for H = [H' | I]
#pragma unroll
                for (int i = 0; i < QS / 2; ++i) {
                    const float t = as[i];
                    as[i] = as[QS - 1 - i];
                    as[QS - 1 - i] = t;
                }
            }
        }
        if (RSUM) {
            float m = INF;
#pragma unroll
            for (int i = 0; i < QS; ++i) {
                const float c = SC_BUF(SW * Q + i);
                const float tot = as[i] + c;  // sum = channel + perm_h(c2v)
                as[i] = tot - c;              // prim = perm_h(sum - perm_h(c2v))
                m = fminf(m, as[i]);          // value of the first strict minimum (NaN skipped)
            }
#pragma unroll
            for (int i = 0; i < QS; ++i) as[i] = as[i] - m;
        }
        if (FIRST && !LAST && !WSUM) {
            // Every later iteration but the last takes the sum-variable's v2c straight from the channel, normalise(channel)
            // (value of the first strict minimum, NaN skipped, subtracted; then sanitised): the same table every time, so it is
            // formed ONCE, here, and parked in the sum-variable's message slots, which nothing else uses until the
            // second-to-last iteration overwrites them with c2v_sum.  Stored reversed = in the order the (min,+) passes read it.
            MinAcc macc;
#pragma unroll
            for (int i = 0; i < QS; ++i) macc.add(as[i]);
            const float m = macc.get();
            float *home = p.msg + ((size_t)tile * p.S + sbase) * 32 + lane;
#pragma unroll
            for (int i = 0; i < QS; ++i) home[(size_t)(QS - 1 - i) * 32] = sanitize(as[i] - m);
        }
#ifndef SCALDPC_HOST_MODEL
        if (!LAST) {  // every lane holds its inputs: re-arm the buffer with the next item
            __syncwarp();
            if (item + nw < n_items && lane == 0) {
                sc_fence_proxy_async();
                issue(item + nw);
            }
        }
#endif
        if (FIRST) {  // channel LLRs may hold NaN; the variable-node kernels store their messages sanitised
#pragma unroll
            for (int k = 0; k < SW; ++k)
#pragma unroll
                for (int i = 0; i < Q; ++i) a[k][i] = sanitize(a[k][i]);
        }
        if (!HOME) {
#pragma unroll
            for (int i = 0; i < QS; ++i) as[i] = sanitize(as[i]);
        }

        // ---- prefix chain of the left variables and suffix chain of the right ones, as pairs:
        //      Fl[k] = a_0 (+) .. (+) a_{k-1},  Fr[k] = a_{SW-1} (+) .. (+) a_{SW-k};  entry j <-> u = j - k*B, j = 0 .. 2kB
        F2 Fl[P + 1][P * B + 1], Fr[M + 1][M * B + 1];
#pragma unroll
        for (int jj = 0; jj < (Q + 1) / 2; ++jj) {
            Fl[1][jj] = f2_make(a[0][2 * jj], 2 * jj + 1 < Q ? a[0][2 * jj + 1] : INF);
            Fr[1][jj] = f2_make(a[SW - 1][2 * jj], 2 * jj + 1 < Q ? a[SW - 1][2 * jj + 1] : INF);
        }
#pragma unroll
        for (int k = 1; k < P; ++k) {
            float Fs[2 * P * B + 2];
            cn_conv_fwd<Q>(Fl[k], 2 * k * B + 1, a[k], Fs);
            Fs[2 * (k + 1) * B + 1] = INF;
#pragma unroll
            for (int jj = 0; jj < (k + 1) * B + 1; ++jj) Fl[k + 1][jj] = f2_make(Fs[2 * jj], Fs[2 * jj + 1]);
        }
#pragma unroll
        for (int k = 1; k < M; ++k) {
            float Fs[2 * M * B + 2];
            cn_conv_fwd<Q>(Fr[k], 2 * k * B + 1, a[SW - 1 - k], Fs);
            Fs[2 * (k + 1) * B + 1] = INF;
#pragma unroll
            for (int jj = 0; jj < (k + 1) * B + 1; ++jj) Fr[k + 1][jj] = f2_make(Fs[2 * jj], Fs[2 * jj + 1]);
        }
        float *out = msg + (size_t)base * 32;  // the check's slot block: constant offsets from here on
        float cs[QS];
        if (LAST || WSUM) {  // c2v of the sum-variable: F_SW[-t].  The left chain is carried on through the right variables IN THE
                             // REFERENCE'S ORDER of additions (((a_0 + a_1) + ..) + a_{SW-1}, decoder_special.rs:539-545), which keeps
                             // the sum-variables' posteriors within an ulp or two of the reference's; two launches per decode pay for it.
            F2 Fc[BS + 1];
#pragma unroll
            for (int jj = 0; jj < P * B + 1; ++jj) Fc[jj] = Fl[P][jj];
            float Fs[QS + 1];
#pragma unroll
            for (int k = P; k < SW; ++k) {
                cn_conv_fwd<Q>(Fc, 2 * k * B + 1, a[k], Fs);
                Fs[2 * (k + 1) * B + 1] = INF;
                if (k + 1 < SW) {
#pragma unroll
                    for (int jj = 0; jj < (k + 1) * B + 1; ++jj) Fc[jj] = f2_make(Fs[2 * jj], Fs[2 * jj + 1]);
                }
            }
#pragma unroll
            for (int i = 0; i < QS; ++i) {  // i = t + BS, F_SW index QS - 1 - i
                cs[i] = (as[i] == INF) ? INF : Fs[QS - 1 - i];
                if (WSUM) out[(size_t)(SW * Q + i) * 32] = cs[i];
            }
        }
        if (LAST) {  // decision of the sum-variable (decoder_special.rs:606-608): first minimum in the VARIABLE's frame
            float min_val = INF;
            int arg = negs ? QS - 1 : 0;
            float *post = p.post_sum ? p.post_sum + (size_t)tile * p.SLS * 32 + lane : nullptr;
            Min2 mm;
#pragma unroll
            for (int i = 0; i < QS; ++i) {
                const int iv = negs ? QS - 1 - i : i;
                const float tot = SC_BUF(NS + iv) + cs[i];
                if (negs ? (tot <= min_val && tot < INF) : (tot < min_val)) {
                    min_val = tot;
                    arg = i;
                }
                mm.add(tot);
                if (post && !frozen) post[(size_t)(vs * QS + iv) * 32] = tot;
            }
            const int argv = negs ? QS - 1 - arg : arg;
            if (!frozen) p.hard_t[((size_t)tile * p.N + p.bvars + vs) * 32 + lane] = (int8_t)(argv - BS);
            if (p.risk) flag_risk(p, tile, lane, !frozen && mm.risky(p.risk_eps));
#ifndef SCALDPC_HOST_MODEL
            __syncwarp();
            if (item + nw < n_items && lane == 0) {
                sc_fence_proxy_async();
                issue(item + nw);
            }
#endif
        }

        // ---- the two terminals.  Sp[j] = a_sum[-(j - BS)] as pairs (cost of the sum-variable given the B-variables' total);
        //      TR[j] = min_i Fl[P][i] + Sp[i + j]  (left half + sum-variable, given the right variables' total j - M*B)
        //      TL[j] = min_i Fr[M][i] + Sp[i + j]  (right half + sum-variable, given the left variables' total j - P*B)
        float TR[2 * M * B + 2], TL[2 * P * B + 2];
        {
            F2 Sp[BS + 1];
#pragma unroll
            for (int jj = 0; jj < BS + 1; ++jj)
                Sp[jj] = f2_make(as[QS - 1 - 2 * jj], 2 * jj + 1 < QS ? as[QS - 2 - 2 * jj] : INF);
            cn_comb<2 * P * B + 1, 2 * M * B + 1>(Fl[P], Sp, TR);
            cn_comb<2 * M * B + 1, 2 * P * B + 1>(Fr[M], Sp, TL);
            TR[2 * M * B + 1] = INF;
            TL[2 * P * B + 1] = INF;
        }

        // ---- each half: backward + combine with its terminal.  Gp = G_{k+1} as pairs, G_{k+1}[j] <-> u = j - (k+1)*B
        //      right half: x_k = a_{SW-1-k}, prefixes Fr[k];  left half: x_k = a_k, prefixes Fl[k]
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const int CNT = side ? P : M;  // the right half first (its prefix tables die first)
            F2 Gp[(P > M ? P : M) * B + 1];
#pragma unroll
            for (int jj = 0; jj < CNT * B + 1; ++jj)
                Gp[jj] = side ? f2_make(TL[2 * jj], TL[2 * jj + 1]) : f2_make(TR[2 * jj], TR[2 * jj + 1]);
#pragma unroll
            for (int k = CNT - 1; k >= 0; --k) {
                const int var = side ? k : SW - 1 - k;
                const int nk = 2 * k * B + 1;
                // c2v_{x_k}[i] = min_j F_k[j] + G_{k+1}[j + i]
                float r[Q];
                if (k == 0) {
#pragma unroll
                    for (int i = 0; i < Q; ++i) r[i] = (i & 1) ? f2_hi(Gp[i >> 1]) : f2_lo(Gp[i >> 1]);
                } else {
                    MinAcc acc[Q];
#pragma unroll
                    for (int j = 0; j < nk; ++j) {
                        const F2 fp = side ? Fl[k < P ? k : 0][j >> 1] : Fr[k < M ? k : 0][j >> 1];
                        const float s = (j & 1) ? f2_hi(fp) : f2_lo(fp);
#pragma unroll
                        for (int jj = j >> 1; jj <= (j + Q - 1) >> 1; ++jj) {
                            const F2 t = f2_add_s(Gp[jj], s);
                            const int i0 = 2 * jj - j, i1 = 2 * jj + 1 - j;
                            if (i0 >= 0 && i0 < Q) acc[i0].add(f2_lo(t));
                            if (i1 >= 0 && i1 < Q) acc[i1].add(f2_hi(t));
                        }
                    }
#pragma unroll
                    for (int i = 0; i < Q; ++i) r[i] = acc[i].get();
                }
#pragma unroll
                for (int i = 0; i < Q; ++i) out[(size_t)(var * Q + i) * 32] = kCnMaskOwn ? ((a[var][i] == INF) ? INF : r[i]) : r[i];
                if (k > 0) {  // G_k[o] = min_i x_k[i] + G_{k+1}[o + i]
                    float Gs[2 * (P > M ? P : M) * B + 2];
                    cn_conv_bwd<Q>(Gp, nk, a[var], Gs);
                    Gs[nk] = INF;
#pragma unroll
                    for (int jj = 0; jj < k * B + 1; ++jj) Gp[jj] = f2_make(Gs[2 * jj], Gs[2 * jj + 1]);
                }
            }
        }
    }
#undef SC_BUF
#undef SC_STAGE
}
