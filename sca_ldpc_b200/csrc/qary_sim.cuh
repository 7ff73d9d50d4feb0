// On-device Kyber trial synthesis: the Monte-Carlo loop around the decoder, without the host.
//
// Restates, per trial, what the reference composes in Python:
//   simulate/kyber.py:35-47        centred-binomial(eta) secret coefficients
//   simulate/kyber.py:85-92        ssum_i = sum_j H[i][j] * s_j
//   simulate/max_likelihood.py:9-18   SimpleOracle: every code bit is reported correctly with probability `accuracy`
//   simulate/max_likelihood.py:87-105, kyber.py:362-376   posterior given the observed pattern == row `y` of the tables
//   success = decoded[:bvars] == s, and decoded[bvars:] == -ssum  (the composition of SURVEY.md 3.2; the decoder
//   construction it ends in is kyber.py:396-402)
// The reference draws from os.urandom / `random`; a reproducible, partition-independent stream replaces them:
// Philox4x32-10 keyed by `seed`, counter = (trial_lo, trial_hi, variable index, stream << 16 | draw >> 2), draw k of a
// variable = word (k & 3).  stream 0 (secret coefficient j): draw 0 = secret bits (low eta bits minus next eta bits),
// draws 1..nbits = one uniform u32 per code bit; stream 1 (check sum i): draws 0..nbits_sum-1.  Code bit k is
// flipped iff u32 >= floor(accuracy * 2^32).  oracle/kyber_sim.py is the numpy restatement the tests compare with.
#pragma once
#include <cstdint>

#include "qary_kernels.cuh"

namespace scaldpc {

struct Philox4 {
    uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                          uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1;
        c3 = (uint32_t)p0;
        c0 = n0;
        c2 = n2;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return Philox4{c0, c1, c2, c3};
}

struct SimRng {  // draws of one variable of one trial, four at a time
    uint32_t t_lo, t_hi, index, stream, k0, k1;
    Philox4 cur;
    int block;
    __device__ __forceinline__ uint32_t draw(int k)
    {
        if ((k >> 2) != block) {
            block = k >> 2;
            cur = philox4x32_10(t_lo, t_hi, index, (stream << 16) | (uint32_t)block, k0, k1);
        }
        const int w = k & 3;
        return w == 0 ? cur.x : w == 1 ? cur.y : w == 2 ? cur.z : cur.w;
    }
};

struct SimParams {
    uint64_t seed, first_trial;
    uint64_t flip_thr;  // flip iff (u64)u32 >= flip_thr
    int eta, nbits, nbits_sum;
};

__device__ __forceinline__ uint32_t sim_observe(SimRng &g, uint32_t word, int nbits, int first_draw, uint64_t thr)
{
    for (int k = 0; k < nbits; ++k)
        if ((uint64_t)g.draw(first_draw + k) >= thr) word ^= 1u << k;
    return word;
}

// one thread per (trial, secret coefficient): secret [T][bvars] int8, obs [T][bvars] uint16
__global__ void __launch_bounds__(256) k_sim_secret(const SimParams sp, const uint16_t *__restrict__ code, int8_t *__restrict__ secret,
                                                    uint16_t *__restrict__ obs, const int T, const int bvars, const int B)
{
    const size_t n = (size_t)T * bvars;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint64_t trial = sp.first_trial + i / bvars;
        const int j = (int)(i % bvars);
        SimRng g{(uint32_t)trial, (uint32_t)(trial >> 32), (uint32_t)j, 0u, (uint32_t)sp.seed, (uint32_t)(sp.seed >> 32), {}, -1};
        const uint32_t bits = g.draw(0), m = (1u << sp.eta) - 1u;
        const int s = __popc(bits & m) - __popc((bits >> sp.eta) & m);  // kyber.py:35-47
        secret[i] = (int8_t)s;
        obs[i] = (uint16_t)sim_observe(g, code[s + B], sp.nbits, 1, sp.flip_thr);
    }
}

// one thread per (trial, check): ssum [T][R] int16 = H' s (kyber.py:85-92), obs_sum [T][R] uint16
__global__ void __launch_bounds__(256) k_sim_sum(const SimParams sp, const QaryDev p, const uint16_t *__restrict__ code_sum,
                                                 const int8_t *__restrict__ secret, int16_t *__restrict__ ssum,
                                                 uint16_t *__restrict__ obs_sum, int *__restrict__ bad, const int T)
{
    const size_t n = (size_t)T * p.R;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t t = i / p.R;
        const uint64_t trial = sp.first_trial + t;
        const int r = (int)(i % p.R);
        const int8_t *s = secret + t * p.bvars;
        int acc = 0;
        for (int e = p.row_ptr[r]; e < p.row_ptr[r + 1]; ++e) {
            const int c = p.col[e];
            if (c < p.bvars) acc += (int)p.sign[e] * (int)s[c];
        }
        ssum[i] = (int16_t)acc;
        if (acc < -p.BSUM || acc > p.BSUM) {  // cannot happen when BSUM >= (row weight) * eta
            *bad = 1;
            acc = 0;
        }
        SimRng g{(uint32_t)trial, (uint32_t)(trial >> 32), (uint32_t)r, 1u, (uint32_t)sp.seed, (uint32_t)(sp.seed >> 32), {}, -1};
        obs_sum[i] = (uint16_t)sim_observe(g, code_sum[acc + p.BSUM], sp.nbits_sum, 0, sp.flip_thr);
    }
}

// one warp per trial: counters {trials, successes (secret and sum variables right), trials whose secret is right,
// wrong secret coefficients, wrong sum variables}
__global__ void __launch_bounds__(256) k_sim_count(const int8_t *__restrict__ hard, const int8_t *__restrict__ secret,
                                                   const int16_t *__restrict__ ssum, unsigned long long *__restrict__ counters,
                                                   const int T, const int N, const int bvars)
{
    const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    unsigned long long ok = 0, okb = 0, eb = 0, es = 0, nt = 0;
    for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < T; t += gridDim.x * wpb) {
        const int8_t *h = hard + (size_t)t * N;
        int wb = 0, ws = 0;
        for (int j = lane; j < bvars; j += 32) wb += h[j] != secret[(size_t)t * bvars + j];
        for (int i = lane; i < N - bvars; i += 32) ws += (int)h[bvars + i] != -(int)ssum[(size_t)t * (N - bvars) + i];
        wb = __reduce_add_sync(0xffffffffu, wb);
        ws = __reduce_add_sync(0xffffffffu, ws);
        ++nt;
        ok += (wb | ws) == 0;
        okb += wb == 0;
        eb += wb;
        es += ws;
    }
    if (lane == 0 && nt) {
        atomicAdd(counters + 0, nt);
        atomicAdd(counters + 1, ok);
        atomicAdd(counters + 2, okb);
        atomicAdd(counters + 3, eb);
        atomicAdd(counters + 4, es);
    }
}

}  // namespace scaldpc
