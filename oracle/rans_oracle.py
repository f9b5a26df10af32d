"""TEST INFRASTRUCTURE -- ctypes front end of oracle/rans_oracle.c (the CPU oracle coder).

Mirrors the surface of the reference's pybind module MLCodec_rans / MLCodec_CXX
(reference: DCVC-DC/src/cpp/py_rans/py_rans.cpp:227-243, ops/ops.cpp:84-91) so tests read
like calls into the reference.  Never imported by dcvc_b200/.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(HERE, "_build", "librans_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(HERE, "rans_oracle.c")
    hdr = os.path.join(HERE, "third_party", "rans64.h")
    if not force and os.path.isfile(_SO) and \
            os.path.getmtime(_SO) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-shared", "-fPIC", "-DNDEBUG",
                           "-I", HERE, src, "-o", _SO, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        vp, ci, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_long
        L.oracle_pmf_to_quantized_cdf.argtypes = [vp, ci, ci, vp]
        L.oracle_pmf_to_quantized_cdf.restype = ci
        L.oracle_enc_new.argtypes = [ci]
        L.oracle_enc_new.restype = vp
        L.oracle_enc_reset.argtypes = [vp]
        L.oracle_enc_free.argtypes = [vp]
        L.oracle_enc_push.argtypes = [vp, vp, vp, ci, vp, ci, vp, vp]
        L.oracle_enc_flush.argtypes = [vp]
        L.oracle_enc_flush.restype = cl
        L.oracle_enc_stream.argtypes = [vp]
        L.oracle_enc_stream.restype = vp
        L.oracle_dec_new.argtypes = []
        L.oracle_dec_new.restype = vp
        L.oracle_dec_free.argtypes = [vp]
        L.oracle_dec_set_stream.argtypes = [vp, vp, cl]
        L.oracle_dec_set_stream.restype = ci
        L.oracle_dec_decode.argtypes = [vp, vp, ci, vp, ci, vp, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def pmf_to_quantized_cdf(pmf, precision=16):
    pmf = np.ascontiguousarray(np.asarray(pmf, dtype=np.float32))
    cdf = np.zeros(pmf.size + 1, dtype=np.uint32)
    rc = lib().oracle_pmf_to_quantized_cdf(_p(pmf), pmf.size, precision, _p(cdf))
    if rc != 0:
        raise ValueError(f"pmf_to_quantized_cdf failed ({rc})")
    return cdf


def _tables(cdfs, sizes, offsets):
    cdfs = np.ascontiguousarray(cdfs, dtype=np.int32)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32).reshape(-1)
    offsets = np.ascontiguousarray(offsets, dtype=np.int32).reshape(-1)
    assert cdfs.ndim == 2 and cdfs.shape[0] == sizes.size == offsets.size
    return cdfs, sizes, offsets


class RansEncoder:
    """Same call sequence as MLCodec_rans.RansEncoder (multiThread only changes threading)."""

    def __init__(self, multi_thread=False, stream_part=1):
        self._h = lib().oracle_enc_new(int(stream_part))
        if not self._h:
            raise ValueError("stream_part must be in 1..16")
        self._len = 0

    def __del__(self):
        if getattr(self, "_h", None):
            try:
                lib().oracle_enc_free(self._h)
            except Exception:
                pass
            self._h = None

    def reset(self):
        lib().oracle_enc_reset(self._h)

    def encode_with_indexes(self, symbols, indexes, cdfs, cdfs_sizes, offsets):
        s = np.ascontiguousarray(symbols, dtype=np.int16).reshape(-1)
        i = np.ascontiguousarray(indexes, dtype=np.int16).reshape(-1)
        assert s.size == i.size
        c, z, o = _tables(cdfs, cdfs_sizes, offsets)
        lib().oracle_enc_push(self._h, _p(s), _p(i), s.size, _p(c), c.shape[1], _p(z), _p(o))

    def flush(self):
        self._len = lib().oracle_enc_flush(self._h)

    def get_encoded_stream(self):
        ptr = lib().oracle_enc_stream(self._h)
        return np.frombuffer(ctypes.string_at(ptr, self._len), dtype=np.uint8).copy()


class RansDecoder:
    def __init__(self, stream_part=1):
        self._h = lib().oracle_dec_new()
        self._parts = stream_part

    def __del__(self):
        if getattr(self, "_h", None):
            try:
                lib().oracle_dec_free(self._h)
            except Exception:
                pass
            self._h = None

    def set_stream(self, encoded):
        e = np.ascontiguousarray(np.frombuffer(bytes(encoded), dtype=np.uint8))
        rc = lib().oracle_dec_set_stream(self._h, _p(e), e.size)
        if rc < 0:
            raise ValueError(f"malformed stream ({rc})")

    def decode_stream(self, indexes, cdfs, cdfs_sizes, offsets):
        i = np.ascontiguousarray(indexes, dtype=np.int16).reshape(-1)
        c, z, o = _tables(cdfs, cdfs_sizes, offsets)
        out = np.zeros(i.size, dtype=np.int16)
        lib().oracle_dec_decode(self._h, _p(i), i.size, _p(c), c.shape[1], _p(z), _p(o), _p(out))
        return out
