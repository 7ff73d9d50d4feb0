"""Drop-in for the reference's `simulate_rs` extension module (decoder classes only).

The reference registers one pyo3 class per compile-time shape (simulate_rs/src/lib.rs:32-75) and its Python side
finds them BY NAME: `getattr(simulate_rs, f"DecoderN{n}R{r}V{v}C{c}B{B}")` (simulate/decode.py:227-229) or
`from simulate_rs import DecoderN1280R512SW6` (simulate/kyber.py:396-402).  Here the classes are synthesised on
first lookup (PEP 562 module __getattr__), so every shape exists:

    DecoderN{n}R{r}V{v}C{c}B{b}     generic single-alphabet decoder, Q = 2b+1        (pydecoder.rs:13-71)
    DecoderN{n}R{r}SW{sw}           special decoder; ONLY the names the reference itself uses (lib.rs:54-75 and the
                                    commented example kyber.py:385-392), all Kyber-768: B = 2, BSUM = 2*sw, DC = sw+1
    DecoderN{n}R{r}SW{sw}B{b}       special decoder with the alphabet spelled out: BSUM = b*sw (Kyber-512: B3, ...)
    make_special_decoder_class(...) explicit factory for anything else

Use as `import sca_ldpc_b200.simulate_rs as simulate_rs` or install it under the reference's name with
`sca_ldpc_b200.install_as_simulate_rs()`.  The Hqc128/192/256 crypto-oracle classes (simulate_rs/src/hqc.rs)
are out of scope.
"""
from __future__ import annotations

import re

from .qary import QaryDecoder

_GENERIC = re.compile(r"DecoderN(\d+)R(\d+)V(\d+)C(\d+)B(\d+)")
_SPECIAL = re.compile(r"DecoderN(\d+)R(\d+)SW(\d+)")
_SPECIAL_B = re.compile(r"DecoderN(\d+)R(\d+)SW(\d+)B(\d+)")
_classes = {}

# lib.rs:32-75 -- the shapes the reference itself registers (DV of the special classes is not in the name)
REGISTERED = {
    "DecoderN6R3V3C4B7": dict(N=6, R=3, DV=3, DC=4, B=7),
    "DecoderN450R150V3C7B1": dict(N=450, R=150, DV=3, DC=7, B=1),
    "DecoderN1024R256SW6": dict(N=1024, R=256, DV=2, DC=7, B=2, BSUM=12),
    "DecoderN1280R512SW6": dict(N=1280, R=512, DV=4, DC=7, B=2, BSUM=12),
    # kyber.py:385-392 (commented example, "decoders used during testing"); same alphabets, sum_weight 3
    "DecoderN1024R256SW3": dict(N=1024, R=256, DV=0, DC=4, B=2, BSUM=6),
}


def _make(name, N, R, B, BSUM=0, DV=0, DC=0):
    doc = (f"{name}(py_parity_check: ndarray[int8, ({R}, {N})], iterations: int)\n\n"
           f"B200 q-ary min-sum decoder, N={N}, R={R}, DV={DV or 'any'}, DC={DC or 'any'}, B={B}"
           + (f", BSUM={BSUM}" if BSUM else ""))

    def __init__(self, py_parity_check, iterations, device=0):
        QaryDecoder.__init__(self, py_parity_check, iterations, B=B, BSUM=BSUM, DV=DV, DC=DC, shape=(R, N),
                             device=device)

    return type(name, (QaryDecoder,), {"__init__": __init__, "__doc__": doc, "__module__": __name__})


def make_decoder_class(N, R, DV, DC, B, name=None):
    """register_py_decoder_class!(m <= Name {N, R, DV, DC, B})  (pydecoder.rs:11-71)"""
    name = name or f"DecoderN{N}R{R}V{DV}C{DC}B{B}"
    return _classes.setdefault(name, _make(name, N, R, B, 0, DV, DC))


def make_special_decoder_class(N, R, DC, B, BSUM, DV=0, name=None):
    """register_py_decoder_special_class!(m <= Name {N, R, DV, DC, B, BSUM})  (pydecoder.rs:85-150)"""
    name = name or f"DecoderN{N}R{R}SW{DC - 1}B{B}"
    return _classes.setdefault(name, _make(name, N, R, B, BSUM, DV, DC))


def __getattr__(name):
    if name in _classes:
        return _classes[name]
    if name in REGISTERED:
        p = REGISTERED[name]
        if "BSUM" in p:
            return make_special_decoder_class(p["N"], p["R"], p["DC"], p["B"], p["BSUM"], p["DV"], name=name)
        return make_decoder_class(p["N"], p["R"], p["DV"], p["DC"], p["B"], name=name)
    m = _GENERIC.fullmatch(name)
    if m:
        n, r, v, c, b = map(int, m.groups())
        return make_decoder_class(n, r, v, c, b, name=name)
    m = _SPECIAL_B.fullmatch(name)
    if m:
        n, r, sw, b = map(int, m.groups())
        return make_special_decoder_class(n, r, sw + 1, b, b * sw, 0, name=name)
    if _SPECIAL.fullmatch(name):
        # the name does not say which alphabets the class has (B, BSUM are const generics of the registration,
        # lib.rs:54-75): guessing B = 2 would silently crop the priors of e.g. a Kyber-512 (B = 3) decoder
        raise AttributeError(f"module 'simulate_rs' has no attribute '{name}': not a shape the reference registers; "
                             f"spell the alphabet out ('{name}B<b>') or use make_special_decoder_class(N, R, DC, B, BSUM)")
    raise AttributeError(f"module 'simulate_rs' has no attribute '{name}'")
