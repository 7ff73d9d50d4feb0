"""Test-time shim: import the reference's *Python* helpers (simulate.kyber, simulate.make_code, ...) from
/root/reference when it is present (it is in the build container, NOT on the GPU box).

Used only by tests/golden/make_golden.py (fixture generation) and by the tests that compare this repository's host-side
restatements with the reference's own Python functions (tests/test_workloads.py, tests/test_oracle_golden.py,
tests/test_hostmodel_bp.py); all of them skip those comparisons when /root/reference is absent.  Nothing here is on the
product path.

The reference modules import three things that are not installed here:
  * coloredlogs            (simulate/utils.py:7)            -> empty stub
  * simulate_rs            (simulate/hqc.py:21)             -> stub exposing Hqc128/192/256.params() and whatever
                                                               decoder module the caller passes in
  * ldpc                   (simulate/hqc.py:24, decode.py:2) -> stub whose bp_decoder is the class passed in
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import numpy as np

REF_ROOT = "/root/reference"
REF_PY = os.path.join(REF_ROOT, "simulate-with-python")

# dependencies/liboqs-rs-bindings/src/hqc.rs:346-352, 378-384, 410-416 (PARAM_* constants)
HQC_PARAMS = {
    "Hqc128": {"N": 17669, "N1": 46, "N2": 384, "N1N2": 17664, "OMEGA": 66},
    "Hqc192": {"N": 35851, "N1": 56, "N2": 640, "N1N2": 35840, "OMEGA": 100},
    "Hqc256": {"N": 57637, "N1": 90, "N2": 640, "N1N2": 57600, "OMEGA": 131},
}


def available() -> bool:
    return os.path.isdir(os.path.join(REF_PY, "simulate"))


def _hqc_stub(name):
    class _Hqc:
        @staticmethod
        def params(what):
            return HQC_PARAMS[name][what]

        @staticmethod
        def name():
            return name

    _Hqc.__name__ = name
    return _Hqc


def rep_code(n):
    """ldpc.codes.rep_code: (n-1) x n repetition-code parity-check matrix (adjacent pairs)."""
    H = np.zeros((n - 1, n), dtype=int)
    for i in range(n - 1):
        H[i, i] = 1
        H[i, i + 1] = 1
    return H


def install(bp_decoder_cls=None, decoder_module=None):
    """Install the stubs and return the imported reference package `simulate`.

    bp_decoder_cls   class used as ldpc.bp_decoder
    decoder_module   module (or object) whose attributes are looked up as simulate_rs.Decoder* classes
    """
    if not available():
        raise RuntimeError("/root/reference is not present")
    sys.modules.setdefault("coloredlogs", types.ModuleType("coloredlogs"))

    rs = types.ModuleType("simulate_rs")
    for name in HQC_PARAMS:
        setattr(rs, name, _hqc_stub(name))
    if decoder_module is not None:
        rs.__getattr__ = lambda attr: getattr(decoder_module, attr)  # PEP 562
    sys.modules["simulate_rs"] = rs

    ldpc = types.ModuleType("ldpc")
    ldpc.bp_decoder = bp_decoder_cls
    codes = types.ModuleType("ldpc.codes")
    codes.rep_code = rep_code
    code_util = types.ModuleType("ldpc.code_util")
    code_util.get_code_parameters = lambda H: {}
    ldpc.codes, ldpc.code_util = codes, code_util
    sys.modules["ldpc"] = ldpc
    sys.modules["ldpc.codes"] = codes
    sys.modules["ldpc.code_util"] = code_util

    if REF_PY not in sys.path:
        sys.path.insert(0, REF_PY)
    for m in [k for k in sys.modules if k == "simulate" or k.startswith("simulate.")]:
        del sys.modules[m]
    pkg = importlib.import_module("simulate")
    for sub in ("utils", "make_code", "distance_spectrum", "max_likelihood", "kyber", "decode", "hqc"):
        importlib.import_module(f"simulate.{sub}")
    return pkg
