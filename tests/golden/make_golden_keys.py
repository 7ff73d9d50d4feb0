#!/usr/bin/env python
"""Fixture of the HQC key files the reference ships (test-hqc{128,192,256}.key, written by simulate/hqc.py:74-93 with a
liboqs keypair): the raw (pk, sk) bytes and the secret supports sca_ldpc_b200.hqc_keys derives from them, kept only
after the public-key equation s = x + h*y has been checked on each key.  Run in the build container (reads
/root/reference); the GPU box only sees the committed tests/golden/hqc_keys.npz.

    python tests/golden/make_golden_keys.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from sca_ldpc_b200 import hqc_keys as K  # noqa: E402

REF = "/root/reference"


def main():
    out = {}
    for name in ("hqc128", "hqc192", "hqc256"):
        pk, sk = K.load_keyfile(os.path.join(REF, f"test-{name}.key"))
        x, y = K.secrets_from_key(sk, pk, verify=True)  # raises unless s == x + h*y
        n, omega = K.PARAMS[name]
        assert len(x) == omega and len(y) == omega and x.max() < n and y.max() < n
        out[f"{name}_pk"] = np.frombuffer(pk, np.uint8)
        out[f"{name}_sk"] = np.frombuffer(sk, np.uint8)
        out[f"{name}_x"] = x
        out[f"{name}_y"] = y
        print(name, "public-key equation holds; y[:8] =", y[:8].tolist())
    np.savez_compressed(os.path.join(HERE, "hqc_keys.npz"), **out)


if __name__ == "__main__":
    main()
