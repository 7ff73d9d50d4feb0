"""sca_ldpc_b200 -- B200-native (sm_100a CUDA behind a C ABI) belief-propagation decoders for SCA-LDPC.

    from sca_ldpc_b200 import simulate_rs            # DecoderN{n}R{r}V{v}C{c}B{b} / DecoderN{n}R{r}SW{sw} classes
    from sca_ldpc_b200.ldpc import bp_decoder        # ldpc.bp_decoder-compatible binary product-sum BP
    from sca_ldpc_b200 import QaryDecoder            # explicit-parameter q-ary decoder (any B / BSUM)

Importing the package does not load the CUDA library; the first decoder construction does, and fails loudly
if sca_ldpc_b200/libscaldpc.so has not been built (python -m sca_ldpc_b200._build).
"""
from __future__ import annotations

import sys

from . import _lib, simulate_rs
from ._lib import ScaldpcError
from .qary import QaryDecoder

__all__ = ["QaryDecoder", "ScaldpcError", "simulate_rs", "install_as_simulate_rs", "install_as_ldpc", "build"]
__version__ = "0.1.0"


def build(force: bool = False) -> str:
    from . import _build

    return _build.build(force=force)


def install_as_simulate_rs():
    """Make `import simulate_rs` resolve to this package's decoder classes (reference: simulate/decode.py:227,
    simulate/kyber.py:396)."""
    sys.modules["simulate_rs"] = simulate_rs
    return simulate_rs


def install_as_ldpc():
    """Make `from ldpc import bp_decoder` resolve to the B200 decoder (reference: simulate/hqc.py:24,
    simulate/decode.py:2)."""
    from . import ldpc

    sys.modules["ldpc"] = ldpc
    sys.modules["ldpc.codes"] = ldpc.codes
    return ldpc
