"""Synthetic inputs for the decoder hot path: parity-check matrices and oracle-derived priors.

Host-side (numpy) input producers for the workloads of BASELINE.json / SURVEY.md 8(d).  They restate, vectorised
and with SEEDED randomness, what the reference composes from
  simulate/make_code.py:50-94     negacyclic +-1 q-ary QC blocks + identity            (qc_parity_check)
  simulate/kyber.py:35-47,60-64   centred-binomial secrets and their prior             (sample_secrets, cbd_prior)
  simulate/kyber.py:85-92         ssum_i = sum_j H[i][j] * s_j                         (KyberWorkload.generate)
  simulate/max_likelihood.py:41-105  Bayes posterior of a value given noisy oracle bits (posterior_table)
  simulate/kyber.py:362-376       prior packing, sum distribution REVERSED             (KyberWorkload.generate)
  simulate/hqc.py:661-708, make_code.py:220-244, distance_spectrum.py:47-87           (HqcWorkload)
The posterior of a variable depends only on the observed bit pattern y, so it is a table lookup
[2^nbits, Q]; trial generation is: secret -> code word -> bit flips w.p. 1-accuracy -> table row.
Nothing here runs on the decode path.
"""
from __future__ import annotations

from dataclasses import dataclass
from math import comb

import numpy as np

KYBER = {  # name -> (eta of the secret, number of 256-coefficient blocks)     kyber.py:8-24
    "kyber512": (3, 2),
    "kyber768": (2, 3),
    "kyber1024": (2, 4),
}
BLOCK_LEN = 256


def negacyclic_block(block_len: int, column_weight: int, rng: np.random.RandomState) -> np.ndarray:
    """One +-1 negacyclic block (make_code.py:50-68): `column_weight` distinct start columns c0 drawn with
    rng.randint(0, block_len-1); entry (i, (c0+i) mod L) = s0 * (-1)^floor((c0+i)/L), s0 = +1 iff c0 == 0."""
    starts = []
    while len(starts) < column_weight:
        i = int(rng.randint(0, block_len - 1))
        if i not in starts:
            starts.append(i)
    block = np.zeros((block_len, block_len), dtype=np.int8)
    rows = np.arange(block_len)
    for c0 in starts:
        pos = c0 + rows
        sign = np.where(pos >= block_len, -1, 1) * (1 if c0 == 0 else -1)
        block[rows, pos % block_len] = sign
    return block


def qc_parity_check(block_len: int, sum_weight: int, num_blocks: int, seed, check_blocks: int = 1) -> np.ndarray:
    """H = [H' | I] with H' a check_blocks x num_blocks grid of negacyclic blocks (make_code.py:72-94).
    `seed` is an int or a RandomState; the same seed gives the reference's matrix."""
    if sum_weight % num_blocks:
        raise NotImplementedError("sum_weight must be a multiple of num_blocks (make_code.py:79-90)")
    rng = seed if isinstance(seed, np.random.RandomState) else np.random.RandomState(seed)
    cw = sum_weight // num_blocks
    grid = [[negacyclic_block(block_len, cw, rng) for _ in range(num_blocks)] for _ in range(check_blocks)]
    m = np.block(grid)
    return np.concatenate((m, np.identity(block_len * check_blocks, dtype=np.int8)), axis=1).astype(np.int8)


def cbd_prior(eta: int, sum_weight: int = 1) -> np.ndarray:
    """Distribution of a sum of `sum_weight` centred-binomial(eta) coefficients over [-B, B], B = sum_weight*eta
    (kyber.py:60-64)."""
    B = sum_weight * eta
    n = 2 * B
    return np.array([comb(n, s + B) for s in range(-B, B + 1)], dtype=np.float64) / 2.0 ** n


def sample_secrets(rs: np.random.RandomState, shape, eta: int) -> np.ndarray:
    """Centred binomial(eta) coefficients (kyber.py:35-47, with a seeded generator instead of os.urandom)."""
    bits = rs.randint(0, 2, size=tuple(shape) + (2 * eta,), dtype=np.int8)
    return (bits[..., :eta].sum(-1) - bits[..., eta:].sum(-1)).astype(np.int8)


def random_coding(num_values: int, nbits: int, rs: np.random.RandomState) -> np.ndarray:
    """Seeded binary code table (num_values x nbits), code words distinct while 2^nbits allows it, value 0's
    neighbours spread first (the role of kyber.py:113-266's pattern databases for shapes they do not cover)."""
    words = rs.permutation(2 ** nbits)
    idx = np.arange(num_values) % (2 ** nbits)
    codes = words[idx]
    return ((codes[:, None] >> np.arange(nbits)[None, :]) & 1).astype(np.int8)


def posterior_table(coding: np.ndarray, prior: np.ndarray, accuracy: float) -> np.ndarray:
    """table[y, s] = Pr[S = s | Y = y] for a SimpleOracle of the given accuracy
    (max_likelihood.py:9-18, 41-105).  y is the integer with bit k = observed bit k.  float64."""
    coding = np.asarray(coding)
    nv, nbits = coding.shape
    ys = (np.arange(2 ** nbits)[:, None] >> np.arange(nbits)[None, :]) & 1          # [Y, nbits]
    agree = (ys[:, None, :] == coding[None, :, :]).sum(-1)                           # [Y, nv]
    like = accuracy ** agree * (1.0 - accuracy) ** (nbits - agree)
    joint = like * np.asarray(prior, dtype=np.float64)[None, :]
    z = joint.sum(1, keepdims=True)
    with np.errstate(invalid="ignore", divide="ignore"):
        t = np.where(z > 0, joint / z, 1.0 / nv)  # an impossible observation (accuracy 1) never occurs
    return t


@dataclass
class KyberBatch:
    priors: np.ndarray       # (T, bvars, Q)   float32   channel_output      (kyber.py:368)
    priors_sum: np.ndarray   # (T, R, QS)      float32   channel_output_sum  (kyber.py:369, reversed :374-375)
    secret: np.ndarray       # (T, bvars)      int8
    ssum: np.ndarray         # (T, R)          int16


@dataclass
class KyberObserved:
    obs: np.ndarray          # (T, bvars)      uint16    observed bit pattern of every secret coefficient
    obs_sum: np.ndarray      # (T, R)          uint16    observed bit pattern of every check sum
    secret: np.ndarray       # (T, bvars)      int8
    ssum: np.ndarray         # (T, R)          int16


class KyberWorkload:
    """One BASELINE.json Kyber configuration: H, code tables, posterior tables; generate(T, seed) draws trials."""

    def __init__(self, variant="kyber768", sum_weight=6, check_blocks=2, accuracy=0.95, nbits=3, nbits_sum=4,
                 seed=0, coding=None, coding_sum=None):
        self.variant = variant
        self.eta, self.num_blocks = KYBER[variant]
        self.sum_weight, self.check_blocks, self.accuracy = sum_weight, check_blocks, float(accuracy)
        self.H = qc_parity_check(BLOCK_LEN, sum_weight, self.num_blocks, seed, check_blocks)
        self.R, self.N = self.H.shape
        self.bvars = self.N - self.R
        self.B, self.BSUM = self.eta, self.eta * sum_weight
        self.Q, self.QS = 2 * self.B + 1, 2 * self.BSUM + 1
        rs = np.random.RandomState((int(seed) * 7919 + 17) % (2 ** 31))
        self.coding = np.asarray(coding, np.int8) if coding is not None else random_coding(self.Q, nbits, rs)
        self.coding_sum = (np.asarray(coding_sum, np.int8) if coding_sum is not None
                           else random_coding(self.QS, nbits_sum, rs))
        self.table = posterior_table(self.coding, cbd_prior(self.eta), self.accuracy).astype(np.float32)
        self.table_sum = posterior_table(self.coding_sum, cbd_prior(self.eta, sum_weight), self.accuracy)
        self.table_sum = self.table_sum[:, ::-1].astype(np.float32).copy()  # kyber.py:374-375: reversed
        self.Hs = self.H[:, : self.bvars].astype(np.int16)

    @property
    def decoder_kwargs(self):
        return dict(B=self.B, BSUM=self.BSUM, DC=self.sum_weight + 1)

    @property
    def simulate_kwargs(self):
        """Arguments of QaryDecoder.simulate (on-device trial synthesis): tables, code words (bit k of the word =
        oracle bit k, the integer _observe builds), bit counts, accuracy, eta."""
        word = lambda c: (c.astype(np.int64) << np.arange(c.shape[1])).sum(-1).astype(np.uint16)  # noqa: E731
        return dict(table=self.table, table_sum=self.table_sum, code=word(self.coding), code_sum=word(self.coding_sum),
                    nbits=self.coding.shape[1], nbits_sum=self.coding_sum.shape[1], accuracy=self.accuracy, eta=self.eta)

    def _observe(self, rs, values, coding, offset):
        bits = coding[values.astype(np.int64) + offset]                           # (..., nbits)
        flips = rs.random_sample(bits.shape) >= self.accuracy
        y = bits ^ flips
        return (y.astype(np.int64) << np.arange(bits.shape[-1])).sum(-1)

    def generate(self, T: int, seed=0) -> KyberBatch:
        rs = np.random.RandomState(seed)
        s = sample_secrets(rs, (T, self.bvars), self.eta)
        ssum = (s.astype(np.int16) @ self.Hs.T).astype(np.int16)                   # kyber.py:85-92
        y = self._observe(rs, s, self.coding, self.B)
        ysum = self._observe(rs, ssum, self.coding_sum, self.BSUM)
        return KyberBatch(self.table[y], self.table_sum[ysum], s, ssum)

    def generate_observed(self, T: int, seed=0) -> "KyberObserved":
        """The same trials as generate(T, seed), as oracle observations: row indices into self.table / self.table_sum
        instead of the expanded prior vectors (priors == table[obs], priors_sum == table_sum[obs_sum])."""
        rs = np.random.RandomState(seed)
        s = sample_secrets(rs, (T, self.bvars), self.eta)
        ssum = (s.astype(np.int16) @ self.Hs.T).astype(np.int16)
        y = self._observe(rs, s, self.coding, self.B)
        ysum = self._observe(rs, ssum, self.coding_sum, self.BSUM)
        return KyberObserved(y.astype(np.uint16), ysum.astype(np.uint16), s, ssum)

    @staticmethod
    def success(batch, hard: np.ndarray) -> np.ndarray:
        """Per-trial success: decoded secret == s and decoded sum-variables == -ssum."""
        bv = batch.secret.shape[1]
        return (hard[:, :bv] == batch.secret).all(1) & (hard[:, bv:].astype(np.int16) == -batch.ssum).all(1)


# -----------------------------------------------------------------------------------------------------------------
# The reference's own micro-benchmark input (simulate_rs/benches/decoder.rs:38-62, "medium decoder"): the generic
# single-alphabet Decoder<450,150,3,7,15,7>, all-zero message, ONE corrupted symbol (0.1 on 0, 0.9 on +7).
# -----------------------------------------------------------------------------------------------------------------
@dataclass
class GenericBatch:
    priors: np.ndarray       # (T, N, Q) float32
    priors_sum: None
    secret: np.ndarray       # (T, N) int8: the transmitted word (all zero)


class DecoderBenchWorkload:
    """H = benches/parity_check_150_450.txt (passed in as a dense int8 matrix), B = 7; trial t corrupts symbol
    1 + (t mod (N-1))' -- the criterion input is trial 0."""

    def __init__(self, H, B=7, seed=0):
        self.H = np.ascontiguousarray(H, dtype=np.int8)
        self.R, self.N = self.H.shape
        self.bvars, self.B, self.BSUM = self.N, B, 0
        self.Q, self.QS = 2 * B + 1, 0
        self.sum_weight, self.check_blocks, self.accuracy = int((self.H != 0).sum(1).max()) - 1, 0, 0.9  # DC = sum_weight + 1

    @property
    def decoder_kwargs(self):
        return dict(B=self.B, DV=3, DC=7)

    def generate(self, T: int, seed=0) -> GenericBatch:
        pri = np.zeros((T, self.N, self.Q), np.float32)
        pri[:, :, self.B] = 1.0
        pos = (1 + (np.arange(T) + seed) % (self.N - 1)) if T > 1 else np.array([1])
        pri[np.arange(T), pos, self.B] = 0.1
        pri[np.arange(T), pos, 2 * self.B] = 0.9
        return GenericBatch(pri, None, np.zeros((T, self.N), np.int8))

    @staticmethod
    def success(batch, hard):
        return (hard == batch.secret).all(1)


# -----------------------------------------------------------------------------------------------------------------
# HQC (binary) workload: shared sparse graph, T independent noise realisations   
# -----------------------------------------------------------------------------------------------------------------
HQC = {  # N, omega          dependencies/liboqs-rs-bindings/src/hqc.rs:346-352, 378-384, 410-416
    "hqc128": (17669, 66),
    "hqc192": (35851, 100),
    "hqc256": (57637, 131),
}


def sparse_first_row(N: int, W: int, rs: np.random.RandomState) -> np.ndarray:
    """Support of a random weight-W vector whose cyclic distance spectrum has multiplicity <= 1
    (the property distance_spectrum.py:47-87 `gen_array_ds_multiplicity(N, W, 1, rng)` enforces; a seeded
    rejection sampler, not the reference's stream)."""
    supp, dists = [], set()
    while len(supp) < W:
        c = int(rs.randint(0, N))
        if c in supp:
            continue
        new = set()
        ok = True
        for x in supp:
            d = min((c - x) % N, (x - c) % N)
            if d in dists or d in new:
                ok = False
                break
            new.add(d)
        if ok:
            supp.append(c)
            dists |= new
    return np.array(sorted(supp), dtype=np.int64)


@dataclass
class HqcBatch:
    syndromes: np.ndarray   # (T, R) uint8  measured (noisy) check bits = syndrome of msg = [0]*N ++ checks
    y: np.ndarray           # (omega,) secret support shared by the batch  or (T, omega)


class HqcWorkload:
    """H = [Hin | I_R] with Hin = R rows of the N x N circulant of a sparse first row (hqc.py:680, 885-908),
    built directly as CSR (never the dense circulant of make_code.py:242-243)."""

    def __init__(self, variant="hqc128", W=50, R=3000, rho=0.9942, seed=0):
        self.N, self.omega = HQC[variant] if isinstance(variant, str) else variant
        self.W, self.R, self.rho = W, R, rho
        rs = np.random.RandomState(seed)
        self.supp = sparse_first_row(self.N, W, rs)                 # support of the circulant's first COLUMN r1
        self.bits = np.sort(rs.choice(self.N, R, replace=False))    # which rows of Hgen are measured
        # circulant(c)[i, j] = c[(i - j) mod N]  =>  row b has ones at columns (b - r) mod N, r in supp
        cols = (self.bits[:, None] - self.supp[None, :]) % self.N
        cols = np.sort(cols, axis=1)
        ident = self.N + np.arange(R)[:, None]
        self.col = np.concatenate((cols, ident), axis=1).astype(np.int32).reshape(-1)
        self.row_ptr = (np.arange(R + 1) * (W + 1)).astype(np.int32)
        self.m, self.n = R, self.N + R
        self.channel_probs = np.concatenate((np.full(self.N, self.omega / self.N), np.full(R, 1.0 - rho)))
        self._cols = cols

    def generate(self, T: int, seed=0, per_trial_secret=True):
        rs = np.random.RandomState(seed)
        synd = np.zeros((T, self.R), np.uint8)
        ys = np.zeros((T, self.omega), np.int64)
        for t in range(T):
            y = np.sort(rs.choice(self.N, self.omega, replace=False)) if (per_trial_secret or t == 0) else ys[0]
            ys[t] = y
            e = np.zeros(self.N, np.uint8)
            e[y] = 1
            true_checks = e[self._cols].sum(1) & 1                  # (y * r1)[b] mod 2, hqc.py:1062-1075
            noise = (rs.random_sample(self.R) >= self.rho).astype(np.uint8)
            synd[t] = true_checks ^ noise
        return HqcBatch(synd, ys)

    def success(self, batch: HqcBatch, decoding: np.ndarray) -> np.ndarray:
        """support(decoded[:N]) == y (hqc.py:717-749)."""
        truth = np.zeros((decoding.shape[0], self.N), np.uint8)
        np.put_along_axis(truth, batch.y, 1, axis=1)
        return (decoding[:, : self.N] == truth).all(1)
