"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module (see oracle/oracle.h).  Parity is unpinned by the reference (a README with
no code); the restatement follows SURVEY.md §8 rows a3/a4/a6/a9.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

OP_COPY, OP_SCALE, OP_ADD, OP_TRIAD, OP_VECTORADD, OP_TRIAD_UNFUSED = 0, 1, 2, 3, 4, 5
SEED_Y_OFFSET = 1 << 32   # b200v_stream fills y from seed + this (include/b200validate.h)

_f32p = ctypes.POINTER(ctypes.c_float)
_u32p = ctypes.POINTER(ctypes.c_uint32)


def build(force=False):
    """Compile liboracle.so with oracle/Makefile (gcc only; no GPU needed)."""
    src_m = max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("oracle.c", "oracle.h"))
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < src_m:
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        L.oracle_fill_rand.argtypes = [_f32p, _f32p, ctypes.c_size_t]
        L.oracle_fill_hash.argtypes = [_f32p, ctypes.c_size_t, ctypes.c_uint64]
        L.oracle_fill_hash_mt.argtypes = [_f32p, ctypes.c_size_t, ctypes.c_uint64, ctypes.c_int]
        L.oracle_splitmix64.argtypes = [ctypes.c_uint64]
        L.oracle_splitmix64.restype = ctypes.c_uint64
        L.oracle_vadd.argtypes = [_f32p, _f32p, _f32p, ctypes.c_size_t]
        L.oracle_vadd_mt.argtypes = [_f32p, _f32p, _f32p, ctypes.c_size_t, ctypes.c_int]
        L.oracle_verify_tolerance.argtypes = [_f32p, _f32p, _f32p, ctypes.c_size_t]
        L.oracle_verify_tolerance.restype = ctypes.c_longlong
        L.oracle_verify_tolerance_mt.argtypes = [_f32p, _f32p, _f32p, ctypes.c_size_t, ctypes.c_int]
        L.oracle_verify_tolerance_mt.restype = ctypes.c_longlong
        L.oracle_stream.argtypes = [ctypes.c_int, _f32p, _f32p, _f32p, ctypes.c_float, ctypes.c_size_t]
        L.oracle_stream_mt.argtypes = [ctypes.c_int, _f32p, _f32p, _f32p, ctypes.c_float,
                                       ctypes.c_size_t, ctypes.c_int]
        L.oracle_fnv1a64_words.argtypes = [_u32p, ctypes.c_size_t]
        L.oracle_fnv1a64_words.restype = ctypes.c_uint64
        L.oracle_digest.argtypes = [_u32p, ctypes.c_size_t]
        L.oracle_digest.restype = ctypes.c_uint64
        L.oracle_digest_from.argtypes = [_u32p, ctypes.c_size_t, ctypes.c_uint64]
        L.oracle_digest_from.restype = ctypes.c_uint64
        L.oracle_allreduce_expected.argtypes = [ctypes.c_size_t, ctypes.c_int]
        L.oracle_allreduce_expected.restype = ctypes.c_float
        _lib = L
    return _lib


def _p(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_f32p)


def _words(a):
    assert a.dtype.itemsize == 4 and a.flags["C_CONTIGUOUS"]
    return a.view(np.uint32).ctypes.data_as(_u32p)


def fill_rand(n):
    """(A, B) as the sample's main() initialises them (glibc rand(), default seed)."""
    A = np.empty(n, np.float32)
    B = np.empty(n, np.float32)
    lib().oracle_fill_rand(_p(A), _p(B), n)
    return A, B


def fill_hash(n, seed, threads=1):
    X = np.empty(n, np.float32)
    if threads > 1:
        lib().oracle_fill_hash_mt(_p(X), n, seed & (2**64 - 1), threads)
    else:
        lib().oracle_fill_hash(_p(X), n, seed & (2**64 - 1))
    return X


def vadd(A, B, threads=1, out=None):
    C = np.empty_like(A) if out is None else out
    if threads > 1:
        lib().oracle_vadd_mt(_p(A), _p(B), _p(C), A.size, threads)
    else:
        lib().oracle_vadd(_p(A), _p(B), _p(C), A.size)
    return C


def verify_tolerance(A, B, C, threads=1):
    """First index failing the sample's 1e-5 check, or -1."""
    if threads > 1:
        return int(lib().oracle_verify_tolerance_mt(_p(A), _p(B), _p(C), A.size, threads))
    return int(lib().oracle_verify_tolerance(_p(A), _p(B), _p(C), A.size))


def stream(op, x, y=None, s=0.0, threads=1, out=None):
    dst = np.empty_like(x) if out is None else out
    yp = _p(y) if y is not None else ctypes.cast(None, _f32p)
    if threads > 1:
        lib().oracle_stream_mt(op, _p(dst), _p(x), yp, s, x.size, threads)
    else:
        lib().oracle_stream(op, _p(dst), _p(x), yp, s, x.size)
    return dst


def fnv1a64_words(a):
    return int(lib().oracle_fnv1a64_words(_words(a), a.size))


def digest(a):
    return int(lib().oracle_digest(_words(a), a.size))


def digest_from(a, first_index):
    """Digest of a piece of a larger array; the pieces' digests add up modulo 2^64."""
    return int(lib().oracle_digest_from(_words(a), a.size, first_index))


def allreduce_input(n, rank):
    """What rank `rank` contributes to the probe's all-reduce: float((i*(2r+1) + r) & 15)."""
    i = np.arange(n, dtype=np.int64)
    return ((i * (2 * rank + 1) + rank) & 15).astype(np.float32)


def allreduce_expected(n, nranks):
    """Closed form of the probe's reduced buffer: sum over r of float((i*(2r+1) + r) & 15)."""
    i = np.arange(n, dtype=np.int64)
    out = np.zeros(n, np.float32)
    for r in range(nranks):
        out += ((i * (2 * r + 1) + r) & 15).astype(np.float32)
    return out


# ---- numpy restatements, used to cross-check the C oracle in the CPU tests -------------------

def np_splitmix64(x):
    x = (np.asarray(x, np.uint64) + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def np_fill_hash(n, seed):
    with np.errstate(over="ignore"):
        i = np.arange(n, dtype=np.uint64) + np.uint64(seed & (2**64 - 1))
        top = (np_splitmix64(i) >> np.uint64(40)).astype(np.uint32)
    return top.astype(np.float32) * np.float32(2.0 ** -24)


def np_digest(a):
    w = a.view(np.uint32).astype(np.uint64)
    with np.errstate(over="ignore"):
        z = w + np.uint64(0x9E3779B97F4A7C15) * (np.arange(a.size, dtype=np.uint64) + np.uint64(1))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        return int(np.add.reduce(z, dtype=np.uint64))
