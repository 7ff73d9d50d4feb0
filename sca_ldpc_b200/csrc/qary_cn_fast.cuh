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
// Arithmetic (unchanged from v0): the reference's brute-force enumeration of (2B+1)^SW configurations is replaced
// by a forward / backward (min,+) convolution:
//     F_1 = a_0,  F_{k+1}[u] = min_d F_k[u-d] + a_k[d]                      (prefix sums, |u| <= kB)
//     G_SW[u] = a_sum[-u],  G_k[u] = min_d a_k[d] + G_{k+1}[u+d]            (cost of the suffix given prefix u)
//     c2v_k[d] = min_u F_k[u] + G_{k+1}[u+d],   c2v_sum[t] = F_SW[-t]
// with "non-finite input => value excluded" (sanitize) and "own entry non-finite => +inf out"
// (decoder.rs:307-311, decoder_special.rs:549-553).  The degree-1 sum-variable's own update
// (decoder_special.rs:591-609) is fused: its v2c is recomputed from the channel LLR and the previous c2v, so it
// never travels through HBM, and on the last iteration its hard decision / posterior are emitted from here.

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

template <int SW, int B, bool FIRST, bool LAST>
__global__ void __launch_bounds__(kCnWarps * 32, 1) k_cn_fast(const QaryDev p, const int n_items)
{
    using Sh = CnFastShape<SW, B>;
    constexpr int Q = Sh::Q, BS = Sh::BS, QS = Sh::QS, NS = Sh::NS;
    const float INF = SC_INF;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * kCnWarps + warp, nw = gridDim.x * kCnWarps;
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
        const int e0 = __ldg(p.row_ptr + check);
#ifndef SCALDPC_HOST_MODEL
        sc_mbar_expect_tx(bar, (uint32_t)((FIRST ? SW * Q + QS : Sh::NB) * 128));
#endif
        if (FIRST) {  // v2c = perm_h(channel) (decoder_special.rs:480-486): gather the variables' channel LLRs
            for (int k = 0; k < SW; ++k) {
                const int var = __ldg(p.col + e0 + k);
                SC_STAGE(k * Q, p.llr + ((size_t)tile * p.SL + (size_t)var * Q) * 32, Q);
            }
        } else {
            SC_STAGE(0, p.msg + ((size_t)tile * p.S + __ldg(p.edge_off + e0)) * 32, NS);
        }
        const int vs = __ldg(p.col + e0 + SW) - p.bvars;
        SC_STAGE(NS, p.llr_sum + ((size_t)tile * p.SLS + (size_t)vs * QS) * 32, QS);
    };

#ifndef SCALDPC_HOST_MODEL
    if (gw < n_items && lane == 0) issue(gw);
#endif
    for (int item = gw; item < n_items; item += nw) {
        const int tile = item / p.R, check = item - tile * p.R;
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
        // ---- staged inputs -> registers
        float a[SW][Q];
#pragma unroll
        for (int k = 0; k < SW; ++k) {
            const bool neg = FIRST && (__ldg(p.sign + e0 + k) < 0);  // later iterations: already in the check frame
#pragma unroll
            for (int i = 0; i < Q; ++i) a[k][i] = SC_BUF(k * Q + (neg ? Q - 1 - i : i));
        }
        // the sum-variable: v2c from channel + previous c2v (decoder_special.rs:591-604), check frame
        float as[QS];
#pragma unroll
        for (int i = 0; i < QS; ++i) as[i] = SC_BUF(NS + (negs ? QS - 1 - i : i));
        if (!FIRST) {
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
#ifndef SCALDPC_HOST_MODEL
        if (!LAST) {  // every lane holds its inputs: re-arm the buffer with the next item
            __syncwarp();
            if (item + nw < n_items && lane == 0) {
                sc_fence_proxy_async();
                issue(item + nw);
            }
        }
#endif
#pragma unroll
        for (int k = 0; k < SW; ++k)
#pragma unroll
            for (int i = 0; i < Q; ++i) a[k][i] = sanitize(a[k][i]);
#pragma unroll
        for (int i = 0; i < QS; ++i) as[i] = sanitize(as[i]);

        // ---- forward: F[k][u + k*B], k = 1..SW-1
        float F[SW][2 * (SW - 1) * B + 1];
#pragma unroll
        for (int i = 0; i < Q; ++i) F[1 % SW][i] = a[0][i];
#pragma unroll
        for (int k = 1; k < SW - 1; ++k) {
#pragma unroll
            for (int u = -(k + 1) * B; u <= (k + 1) * B; ++u) {
                MinAcc acc;
#pragma unroll
                for (int d = -B; d <= B; ++d) {
                    const int up = u - d;
                    if (up >= -k * B && up <= k * B) acc.add(F[k][up + k * B] + a[k][d + B]);
                }
                F[k + 1][u + (k + 1) * B] = acc.get();
            }
        }
        // ---- c2v of the sum-variable: F_SW[-t]
        float cs[QS];
#pragma unroll
        for (int t = -BS; t <= BS; ++t) {
            const int u = -t;
            MinAcc acc;
            if (SW == 1) {
                if (u >= -B && u <= B) acc.add(a[0][u + B]);
            } else {
                constexpr int k = SW - 1;
#pragma unroll
                for (int d = -B; d <= B; ++d) {
                    const int up = u - d;
                    if (up >= -k * B && up <= k * B) acc.add(F[k][up + k * B] + a[k][d + B]);
                }
            }
            const float r = acc.get();
            cs[t + BS] = (as[t + BS] == INF) ? INF : r;
            msg[(size_t)(sbase + t + BS) * 32] = cs[t + BS];
        }
        if (LAST) {  // decision of the sum-variable (decoder_special.rs:606-608): first minimum in the VARIABLE's frame
            float min_val = INF;
            int arg = negs ? QS - 1 : 0;
            float *post = p.post_sum ? p.post_sum + (size_t)tile * p.SLS * 32 + lane : nullptr;
#pragma unroll
            for (int i = 0; i < QS; ++i) {
                const int iv = negs ? QS - 1 - i : i;
                const float tot = SC_BUF(NS + iv) + cs[i];
                if (negs ? (tot <= min_val && tot < INF) : (tot < min_val)) {
                    min_val = tot;
                    arg = i;
                }
                if (post) post[(size_t)(vs * QS + iv) * 32] = tot;
            }
            const int argv = negs ? QS - 1 - arg : arg;
            p.hard_t[((size_t)tile * p.N + p.bvars + vs) * 32 + lane] = (int8_t)(argv - BS);
#ifndef SCALDPC_HOST_MODEL
            __syncwarp();
            if (item + nw < n_items && lane == 0) {
                sc_fence_proxy_async();
                issue(item + nw);
            }
#endif
        }

        // ---- backward + combine.  Gc = G_{k+1}, indexed [u + (k+1)*B]
        float Gc[QS], Gn[QS];
#pragma unroll
        for (int i = 0; i < QS; ++i) Gc[i] = as[QS - 1 - i];  // G_SW[u] = a_sum[-u]
#pragma unroll
        for (int k = SW - 1; k >= 0; --k) {
#pragma unroll
            for (int d = -B; d <= B; ++d) {  // c2v_k[d] = min_u F_k[u] + G_{k+1}[u + d]
                MinAcc acc;
                if (k == 0) {
                    acc.add(Gc[d + B]);
                } else {
#pragma unroll
                    for (int u = -k * B; u <= k * B; ++u) acc.add(F[k][u + k * B] + Gc[u + d + (k + 1) * B]);
                }
                const float r = acc.get();
                msg[(size_t)(base + k * Q + d + B) * 32] = (a[k][d + B] == INF) ? INF : r;
            }
            if (k > 0) {
#pragma unroll
                for (int u = -k * B; u <= k * B; ++u) {
                    MinAcc acc;
#pragma unroll
                    for (int d = -B; d <= B; ++d) acc.add(a[k][d + B] + Gc[u + d + (k + 1) * B]);
                    Gn[u + k * B] = acc.get();
                }
#pragma unroll
                for (int i = 0; i < QS; ++i) Gc[i] = Gn[i];
            }
        }
    }
#undef SC_BUF
#undef SC_STAGE
}
