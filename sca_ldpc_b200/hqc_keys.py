"""HQC key files -> the secret supports (x, y) the attack simulation decodes for  (SURVEY.md 8f.4).

Reference interface restated (paths relative to the reference root):
  simulate-with-python/simulate/hqc.py:74-93        read_or_generate_keypair: the key file is pickle((pk_bytes, sk_bytes))
  simulate-with-python/simulate/hqc.py:1043-1046    y_sparse = HQC.secrets_from_key(priv)[1]
  dependencies/liboqs-rs-bindings/src/hqc.rs:187-202  secrets_from_key -> OQS_KEM_hqc_*_secret_key_from_string(x, y, pk, sk)
  dependencies/liboqs-rs-bindings/src/hqc.rs:346-352, 378-384, 410-416   PARAM_N / PARAM_OMEGA of the three sets
The C side (liboqs fork, PQCLEAN_HQCRMRS{128,192,256}_CLEAN_*) is NOT in the reference tree (submodule missing), so what
is restated here is the published HQC round-3 (2020-10, "hqc-rmrs") algorithm those symbols name:
  sk = sk_seed[40] || pk,   pk = pk_seed[40] || s                                   (parsing.c)
  seedexpander = the NIST AES-256 seed expander (rng.c): key = seed[0:32], counter = seed[32:40] || maxlen(be32) || ctr(be32)
  x = vect_set_random_fixed_weight(sk_expander, omega), y = ..._by_coordinates(sk_expander, omega)   (vector.c: 3 bytes per
      candidate, rejection above floor(2^24 / n) * n, reduction mod n, duplicates re-drawn)
  h = vect_set_random(pk_expander)  (n bits, little-endian bytes);  public key equation  s = x + h * y  in GF(2)[X]/(X^n - 1)
"Unverifiable against liboqs on this box" -- but NOT unverified: `secrets_from_key(..., verify=True)` checks the public-key
equation on the key itself, which a wrong expander, byte order, sampling rule or x/y order cannot satisfy (2^-n chance).
Everything here is host-side and runs once per key; nothing is on the decode path.
"""
from __future__ import annotations

import io
import pickle

import numpy as np

PARAMS = {  # name -> (n, omega)
    "hqc128": (17669, 66),
    "hqc192": (35851, 100),
    "hqc256": (57637, 131),
}
SEED_BYTES = 40
SEEDEXPANDER_MAX_LENGTH = 4294967295

# ---------------------------------------------------------------------------------------------------------------------
# AES-256 block encryption (FIPS-197), table-free and tiny: ~150 blocks per key are needed, speed is irrelevant.
# ---------------------------------------------------------------------------------------------------------------------


def _xtime(a: int) -> int:
    a <<= 1
    return (a ^ 0x11B) & 0xFF if a & 0x100 else a


def _make_sbox():
    # multiplicative inverse in GF(2^8) followed by the affine map
    exp, log = [0] * 510, [0] * 256
    x = 1
    for i in range(255):
        exp[i] = x
        log[x] = i
        x ^= _xtime(x)  # multiply by 3
    for i in range(255, 510):
        exp[i] = exp[i - 255]
    sbox = [0] * 256
    for a in range(256):
        inv = 0 if a == 0 else exp[255 - log[a]]
        s = inv
        for _ in range(4):
            inv = ((inv << 1) | (inv >> 7)) & 0xFF
            s ^= inv
        sbox[a] = s ^ 0x63
    return sbox


_SBOX = _make_sbox()


class AES256:
    """Encrypt-only AES-256 (the seed expander uses ECB encryption of a counter block)."""

    def __init__(self, key: bytes):
        if len(key) != 32:
            raise ValueError("AES-256 needs a 32-byte key")
        w = [list(key[4 * i: 4 * i + 4]) for i in range(8)]
        rcon = 1
        for i in range(8, 60):
            t = list(w[i - 1])
            if i % 8 == 0:
                t = [_SBOX[t[1]] ^ rcon, _SBOX[t[2]], _SBOX[t[3]], _SBOX[t[0]]]
                rcon = _xtime(rcon)
            elif i % 8 == 4:
                t = [_SBOX[b] for b in t]
            w.append([a ^ b for a, b in zip(w[i - 8], t)])
        self._rk = [sum(w[4 * r: 4 * r + 4], []) for r in range(15)]

    def encrypt_block(self, block: bytes) -> bytes:
        s = [b ^ k for b, k in zip(block, self._rk[0])]
        for r in range(1, 15):
            s = [_SBOX[b] for b in s]
            s = [s[(i + 4 * (i % 4)) % 16] for i in range(16)]  # ShiftRows (column-major state)
            if r != 14:
                t = []
                for c in range(4):
                    a = s[4 * c: 4 * c + 4]
                    x = a[0] ^ a[1] ^ a[2] ^ a[3]
                    t += [a[i] ^ x ^ _xtime(a[i] ^ a[(i + 1) % 4]) for i in range(4)]
                s = t
            s = [b ^ k for b, k in zip(s, self._rk[r])]
        return bytes(s)


class SeedExpander:
    """The NIST API seed expander (rng.c seedexpander_init / seedexpander): AES-256-ECB over a 16-byte counter block
    diversifier[8] || maxlen (big-endian 32 bit) || block counter (big-endian 32 bit, starts at 0)."""

    def __init__(self, seed: bytes, maxlen: int = SEEDEXPANDER_MAX_LENGTH):
        if len(seed) != SEED_BYTES:
            raise ValueError("seed must be 40 bytes (32 key + 8 diversifier)")
        self._aes = AES256(seed[:32])
        self._ctr = bytearray(seed[32:40] + maxlen.to_bytes(4, "big") + b"\0\0\0\0")
        self._buf, self._pos = b"\0" * 16, 16
        self.calls = 0  # seedexpander() invocations: the quantity HQC's timing side channel leaks (hqc.py:96-107)

    def read(self, n: int) -> bytes:
        self.calls += 1
        out = bytearray()
        while n > 0:
            if self._pos == 16:
                self._buf = self._aes.encrypt_block(bytes(self._ctr))
                self._pos = 0
                for i in range(15, 11, -1):  # increment the 32-bit block counter
                    self._ctr[i] = (self._ctr[i] + 1) & 0xFF
                    if self._ctr[i]:
                        break
            take = min(n, 16 - self._pos)
            out += self._buf[self._pos: self._pos + take]
            self._pos += take
            n -= take
        return bytes(out)


def fixed_weight_support(exp: SeedExpander, n: int, weight: int) -> list[int]:
    """vect_set_random_fixed_weight[_by_coordinates] (hqc-rmrs vector.c): `weight` distinct positions in draw order."""
    threshold = (2 ** 24 // n) * n
    nbytes = 3 * weight
    rnd, j = exp.read(nbytes), 0
    out: list[int] = []
    while len(out) < weight:
        while True:
            if j == nbytes:
                rnd, j = exp.read(nbytes), 0
            v = (rnd[j] << 16) | (rnd[j + 1] << 8) | rnd[j + 2]
            j += 3
            if v < threshold:
                break
        v %= n
        if v not in out:
            out.append(v)
    return out


def _bits_le(data: bytes, n: int) -> np.ndarray:
    """n bits of a little-endian packed vector (load8_arr / store8_arr: bit i = byte[i / 8] >> (i % 8))."""
    return np.unpackbits(np.frombuffer(data, np.uint8), bitorder="little")[:n].astype(np.uint8)


def param_set_of(pk: bytes, sk: bytes) -> str:
    for name, (n, _) in PARAMS.items():
        if len(pk) == SEED_BYTES + (n + 7) // 8 and len(sk) == SEED_BYTES + len(pk):
            return name
    raise ValueError(f"not an HQC round-3 key pair: public key {len(pk)} bytes, secret key {len(sk)} bytes")


def public_key_check(pk: bytes, x_support, y_support, n: int) -> bool:
    """s == x + h * y in GF(2)[X] / (X^n - 1), with h expanded from the public seed and s read from the public key."""
    h = _bits_le(SeedExpander(pk[:SEED_BYTES]).read((n + 7) // 8), n)
    s = _bits_le(pk[SEED_BYTES:], n)
    acc = np.zeros(n, np.uint8)
    acc[np.asarray(x_support, np.int64)] ^= 1
    for pos in y_support:  # X^pos * h = h rotated by pos
        acc ^= np.roll(h, int(pos))
    return bool(np.array_equal(acc, s))


def secrets_from_key(sk: bytes, pk: bytes | None = None, verify: bool = True):
    """(x_support, y_support) as sorted int64 arrays -- what HQC.secrets_from_key(priv) returns in sparse form
    (hqc.py:1043-1046 uses [1], the support of y).  verify=True checks the public-key equation and raises ValueError if it
    does not hold (wrong file, or a different HQC revision than round-3 hqc-rmrs)."""
    sk = bytes(sk)
    pk_in = sk[SEED_BYTES:]
    if pk is not None and bytes(pk) != pk_in:
        raise ValueError("the public key stored inside the secret key differs from the one given")
    name = param_set_of(pk_in, sk)
    n, omega = PARAMS[name]
    exp = SeedExpander(sk[:SEED_BYTES])
    x = fixed_weight_support(exp, n, omega)
    y = fixed_weight_support(exp, n, omega)
    if verify and not public_key_check(pk_in, x, y, n):
        raise ValueError("public-key equation s = x + h*y does not hold: not a round-3 hqc-rmrs key")
    return np.sort(np.asarray(x, np.int64)), np.sort(np.asarray(y, np.int64))


class _NoGlobals(pickle.Unpickler):
    def find_class(self, module, name):  # the key file is two bytes objects in a tuple: nothing needs importing
        raise pickle.UnpicklingError(f"key file references {module}.{name}: refusing to unpickle code")


def load_keyfile(path: str):
    """hqc.py:74-93: pickle((public_key, secret_key)).  Returns (pk bytes, sk bytes); never imports anything the file names."""
    with open(path, "rb") as f:
        key = _NoGlobals(io.BytesIO(f.read())).load()
    if not (isinstance(key, (tuple, list)) and len(key) == 2):
        raise ValueError("key file does not hold a (public key, secret key) pair")
    pk, sk = (bytes(bytearray(k)) for k in key)
    param_set_of(pk, sk)
    return pk, sk
