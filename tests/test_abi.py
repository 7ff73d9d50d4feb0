"""The C-ABI library loads and exports every symbol include/scaldpc.h declares; argument validation that needs no
GPU (graph build runs before the device is touched); creation fails loudly without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import util
from sca_ldpc_b200 import _build, _lib


@pytest.fixture(scope="module")
def lib():
    _build.build()
    return _lib.lib()


def _declared():
    text = open(os.path.join(util.ROOT, "include", "scaldpc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(scaldpc_[a-z0-9_]+)\s*\(", text)))


def test_exports_every_declared_symbol(lib):
    names = _declared()
    assert len(names) >= 16
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/scaldpc.h but not exported"
    assert sorted(_lib.SYMBOLS) == names


def test_version_and_device_count(lib):
    assert "sm_100a" in _lib.version()
    assert _lib.device_count() >= 0


def _create(lib, H, b, bsum, dv=0, dc=0):
    H = np.ascontiguousarray(H, np.int8)
    h = C.c_void_p()
    rc = lib.scaldpc_qary_create(H.ctypes.data, H.shape[0], H.shape[1], b, bsum, dv, dc, 3, 0, C.byref(h))
    msg = lib.scaldpc_last_error().decode()
    if h:
        lib.scaldpc_qary_destroy(h)
    return rc, msg


def test_qary_create_validation(lib):
    """decoder_special.rs:388-395, :513-522 and decoder.rs:465-473 construction panics -> status codes."""
    H = np.zeros((2, 5), np.int8)
    H[0, [0, 1, 3]] = 1
    H[1, [1, 2, 4]] = -1
    rc, msg = _create(lib, H, 2, 5)
    assert rc == _lib.ERR_STRUCTURE and "multiple of B" in msg
    rc, msg = _create(lib, H, 2, 4, dc=2)
    assert rc == _lib.ERR_DEGREE
    rc, msg = _create(lib, H, 2, 2)
    assert rc == _lib.ERR_RANGE
    Hbad = H.copy()
    Hbad[0, 4] = 1  # two sum-variables in one row
    rc, msg = _create(lib, Hbad, 2, 4)
    assert rc in (_lib.ERR_STRUCTURE, _lib.ERR_DEGREE)
    rc, msg = _create(lib, H, 0, 0)
    assert rc == _lib.ERR_ARG
    rc, msg = _create(lib, H[:, :3], 1, 0, dv=1)
    assert rc == _lib.ERR_DEGREE
    # a valid graph gets as far as the device
    rc, msg = _create(lib, H, 2, 4)
    if _lib.device_count() == 0:
        assert rc == _lib.ERR_CUDA and "no CPU path" in msg
    else:
        assert rc == 0


def test_bp_create_validation(lib):
    h = C.c_void_p()
    row_ptr = np.array([0, 2, 4], np.int32)
    col = np.array([0, 1, 1, 5], np.int32)
    rc = lib.scaldpc_bp_create(2, 3, row_ptr.ctypes.data, col.ctypes.data, 10, 0, C.byref(h))
    assert rc == _lib.ERR_ARG and "out of range" in lib.scaldpc_last_error().decode()
    col = np.array([1, 0, 1, 2], np.int32)
    rc = lib.scaldpc_bp_create(2, 3, row_ptr.ctypes.data, col.ctypes.data, 10, 0, C.byref(h))
    assert rc == _lib.ERR_ARG and "ascending" in lib.scaldpc_last_error().decode()


def test_python_shims_fail_loudly_without_gpu():
    if _lib.device_count() > 0:
        pytest.skip("a CUDA device is present")
    from sca_ldpc_b200 import simulate_rs
    from sca_ldpc_b200.ldpc import bp_decoder

    z = util.load("rust_unit_tests.npz")
    with pytest.raises(_lib.ScaldpcError) as e:
        simulate_rs.DecoderN6R3V3C4B7(z["H6"], 10)
    assert e.value.code == _lib.ERR_CUDA
    with pytest.raises(_lib.ScaldpcError):
        bp_decoder(np.array([[1, 1, 0], [0, 1, 1]]), error_rate=0.1, max_iter=3)


def test_simulate_rs_class_lookup():
    """Name-based class lookup of simulate/decode.py:227-229 and simulate/kyber.py:396-402."""
    from sca_ldpc_b200 import simulate_rs

    n, r, v, c, B = 450, 150, 3, 7, 1
    cls = getattr(simulate_rs, f"DecoderN{n}R{r}V{v}C{c}B{B}")
    assert cls.__name__ == "DecoderN450R150V3C7B1" and cls is simulate_rs.DecoderN450R150V3C7B1
    assert simulate_rs.DecoderN1280R512SW6.__name__ == "DecoderN1280R512SW6"
    with pytest.raises(AttributeError):
        simulate_rs.NoSuchDecoder
    # special-decoder names do not carry the alphabets: only the reference's own shapes resolve by the bare name
    # (lib.rs:54-75, kyber.py:385-392); other shapes spell B out or go through the factory
    assert simulate_rs.DecoderN1024R256SW3.__name__ == "DecoderN1024R256SW3"
    with pytest.raises(AttributeError, match="make_special_decoder_class"):
        simulate_rs.DecoderN768R256SW6
    assert simulate_rs.DecoderN768R256SW6B3 is simulate_rs.make_special_decoder_class(768, 256, 7, 3, 18, name="DecoderN768R256SW6B3")
    with pytest.raises(TypeError):  # pydecoder.rs:24: PyReadonlyArray2<i8>
        simulate_rs.DecoderN6R3V3C4B7(np.zeros((3, 6), np.int64), 10)
    with pytest.raises(IndexError):  # smaller than the class's const shape (pydecoder.rs:31)
        simulate_rs.DecoderN6R3V3C4B7(np.zeros((2, 6), np.int8), 10)


def test_qary_create_csr_validation(lib):
    """scaldpc_qary_create_csr: the sparse constructor validates its CSR before it touches the device
    and then applies the same construction checks as the dense one (decoder_special.rs:388-395, decoder.rs:465-473)."""
    def create(rp, col, val, R, N, b, bsum, dc=0):
        rp, col, val = np.asarray(rp, np.int32), np.asarray(col, np.int32), np.asarray(val, np.int8)
        h = C.c_void_p()
        rc = lib.scaldpc_qary_create_csr(R, N, rp.ctypes.data, col.ctypes.data, val.ctypes.data, b, bsum, 0, dc, 3, 0, C.byref(h))
        msg = lib.scaldpc_last_error().decode()
        if h:
            lib.scaldpc_qary_destroy(h)
        return rc, msg

    # H = [[1, 1, 0 | 1, 0], [0, -1, 1 | 0, 1]]
    rp, col, val = [0, 3, 6], [0, 1, 3, 1, 2, 4], [1, 1, 1, -1, 1, 1]
    rc, msg = create(rp, col, val, 2, 5, 2, 4)
    assert rc in (0, _lib.ERR_CUDA), msg              # a valid graph: only the missing device can stop it
    rc, msg = create(rp, [1, 0, 3, 1, 2, 4], val, 2, 5, 2, 4)
    assert rc == _lib.ERR_ARG and "ascending" in msg
    rc, msg = create(rp, [0, 1, 5, 1, 2, 4], val, 2, 5, 2, 4)
    assert rc == _lib.ERR_ARG and "out of range" in msg
    rc, msg = create(rp, col, [1, 0, 1, -1, 1, 1], 2, 5, 2, 4)
    assert rc == _lib.ERR_ARG and "zero" in msg
    rc, msg = create(rp, col, val, 2, 5, 2, 5)
    assert rc == _lib.ERR_STRUCTURE and "multiple of B" in msg
    rc, msg = create(rp, col, val, 2, 5, 2, 4, dc=2)
    assert rc == _lib.ERR_DEGREE
    rc, msg = create(rp, col, val, 2, 5, 2, 2)
    assert rc == _lib.ERR_RANGE
