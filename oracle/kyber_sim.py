"""TEST INFRASTRUCTURE -- CPU restatement (numpy) of the on-device Kyber trial generator.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product never does.

What is restated and where it comes from:
  simulate/kyber.py:35-47        secret coefficients ~ centred binomial(eta)            (secret_from_bits)
  simulate/kyber.py:85-92        ssum_i = sum_j H[i][j] * s_j                           (generate)
  simulate/max_likelihood.py:9-18   SimpleOracle: a bit is reported correctly w.p. accuracy  (observe)
The reference takes its randomness from os.urandom / `random`, so there is nothing of the reference's to reproduce bit
for bit; the random stream is this framework's own definition (sca_ldpc_b200/csrc/qary_sim.cuh) and this file is its
second, independent statement: Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11;
pinned by the Random123 known-answer vectors in tests/test_kyber_sim.py), key = seed, counter = (trial_lo, trial_hi,
variable index, stream << 16 | draw >> 2), draw k = word k & 3.  Stream 0 = secret coefficient j (draw 0: secret bits,
draws 1..nbits: one uniform per code bit), stream 1 = check sum i (draws 0..nbits_sum-1).  A bit is flipped iff
u32 >= floor(accuracy * 2^32).  Parity of the DISTRIBUTIONS with the reference's sampling is tested statistically.
"""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
S32 = np.uint64(32)


def philox4x32_10(counter, key):
    """counter (..., 4) uint32-valued, key (2,) -> (..., 4) uint32.  Ten rounds, key bumped between rounds."""
    c = [np.asarray(counter[..., i], dtype=np.uint64) for i in range(4)]
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        c = [(p1 >> S32) ^ c[1] ^ np.uint64(k0), p1 & MASK, (p0 >> S32) ^ c[3] ^ np.uint64(k1), p0 & MASK]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return np.stack(c, axis=-1).astype(np.uint32)


def draws(seed, trials, index, stream, n):
    """(len(trials), len(index), n) uint32: draws 0..n-1 of every (trial, variable) of a stream."""
    trials = np.asarray(trials, dtype=np.uint64)
    index = np.asarray(index, dtype=np.uint64)
    nb = (n + 3) // 4
    ctr = np.empty((len(trials), len(index), nb, 4), dtype=np.uint64)
    ctr[..., 0] = (trials & MASK)[:, None, None]
    ctr[..., 1] = (trials >> S32)[:, None, None]
    ctr[..., 2] = index[None, :, None]
    ctr[..., 3] = (np.uint64(stream) << np.uint64(16)) | np.arange(nb, dtype=np.uint64)[None, None, :]
    key = (int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF)
    out = philox4x32_10(ctr, key)
    return out.reshape(len(trials), len(index), nb * 4)[..., :n]


def flip_threshold(accuracy: float) -> int:
    return 1 << 32 if accuracy >= 1.0 else int(accuracy * 4294967296.0)


def observe(code_words, u, thr):
    """code_words (...,) ints, u (..., nbits) uint32 uniforms: bit k flipped iff u[..., k] >= thr."""
    flips = (u.astype(np.uint64) >= np.uint64(thr)).astype(np.int64)
    return (np.asarray(code_words, dtype=np.int64) ^ (flips << np.arange(u.shape[-1])).sum(-1)).astype(np.uint16)


def secret_from_bits(bits, eta):
    b = np.asarray(bits, dtype=np.uint64)
    pop = lambda x: sum(((x >> np.uint64(k)) & np.uint64(1)).astype(np.int64) for k in range(eta))  # noqa: E731
    return (pop(b) - pop(b >> np.uint64(eta))).astype(np.int8)


def generate(H, T, seed, first_trial, code, nbits, code_sum, nbits_sum, accuracy, eta, B, BSUM):
    """-> secret (T, bvars) int8, ssum (T, R) int16, obs (T, bvars) uint16, obs_sum (T, R) uint16."""
    H = np.asarray(H)
    R, N = H.shape
    bvars = N - R
    trials = np.uint64(first_trial) + np.arange(T, dtype=np.uint64)
    thr = flip_threshold(float(accuracy))
    d = draws(seed, trials, np.arange(bvars), 0, 1 + nbits)
    secret = secret_from_bits(d[..., 0], eta)
    obs = observe(np.asarray(code)[secret.astype(np.int64) + B], d[..., 1:], thr)
    ssum = (secret.astype(np.int64) @ H[:, :bvars].astype(np.int64).T).astype(np.int16)
    ds = draws(seed, trials, np.arange(R), 1, nbits_sum)
    obs_sum = observe(np.asarray(code_sum)[ssum.astype(np.int64) + BSUM], ds, thr)
    return secret, ssum, obs, obs_sum
