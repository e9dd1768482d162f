"""ctypes binding of libb200validate.so — the same C ABI a cgo binding would use
(include/b200validate.h; INTEGRATION.md shows the Go side).

The library is the product: there is no Python or CPU fallback.  Loading fails loudly when the
shared object has not been built (run `python -c "import __graft_entry__ as g; g.build()"`), and
every compute call raises B200VError when no CUDA device is usable.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200validate.so")
HEADER_PATH = os.path.normpath(os.path.join(_HERE, "..", "include", "b200validate.h"))

OP_COPY, OP_SCALE, OP_ADD, OP_TRIAD, OP_VECTORADD, OP_TRIAD_UNFUSED = range(6)
OP_NAMES = {OP_COPY: "copy", OP_SCALE: "scale", OP_ADD: "add", OP_TRIAD: "triad",
            OP_VECTORADD: "vectoradd", OP_TRIAD_UNFUSED: "triad_unfused"}
OP_BYTES_PER_ELEM = {OP_COPY: 8, OP_SCALE: 8, OP_ADD: 12, OP_TRIAD: 12, OP_VECTORADD: 12,
                     OP_TRIAD_UNFUSED: 12}

E_BADARG, E_MISMATCH, E_NCCL, E_NODEVICE, E_NONCCL, E_NOMEM = -1, -2, -3, -4, -5, -6
NCCL_UNIQUE_ID_BYTES = 128
SEED_Y_OFFSET = 1 << 32


class Info(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char * 256), ("cc_major", ctypes.c_int), ("cc_minor", ctypes.c_int),
                ("sm_count", ctypes.c_int), ("mem_clock_khz", ctypes.c_int),
                ("mem_bus_bits", ctypes.c_int), ("sm_clock_khz", ctypes.c_int),
                ("l2_bytes", ctypes.c_uint64), ("hbm_bytes", ctypes.c_uint64)]


class Timing(ctypes.Structure):
    _fields_ = [("h2d_ms", ctypes.c_float), ("kernel_ms", ctypes.c_float), ("d2h_ms", ctypes.c_float),
                ("total_ms", ctypes.c_float), ("grid", ctypes.c_int), ("block", ctypes.c_int),
                ("chunks", ctypes.c_int), ("kernel_launches", ctypes.c_int)]


class StreamResult(ctypes.Structure):
    _fields_ = [("best_ms", ctypes.c_float), ("median_ms", ctypes.c_float), ("mean_ms", ctypes.c_float),
                ("best_gbs", ctypes.c_double), ("median_gbs", ctypes.c_double),
                ("bytes_per_launch", ctypes.c_uint64), ("digest", ctypes.c_uint64),
                ("mismatches", ctypes.c_uint64), ("first_mismatch", ctypes.c_uint64),
                ("l2_resident", ctypes.c_int), ("grid", ctypes.c_int), ("block", ctypes.c_int),
                ("kernel_launches", ctypes.c_int)]


class ArRow(ctypes.Structure):
    _fields_ = [("bytes", ctypes.c_uint64), ("best_ms", ctypes.c_float), ("median_ms", ctypes.c_float),
                ("algbw_gbs", ctypes.c_double), ("busbw_gbs", ctypes.c_double),
                ("mismatches", ctypes.c_uint64), ("nccl_result", ctypes.c_int), ("algo", ctypes.c_int)]


AR_ALGO_NAMES = {0: "nccl", 1: "multimem", 2: "p2p", 3: "push", 4: "oneshot", 5: "hybrid"}


class B200VError(RuntimeError):
    def __init__(self, code, what):
        self.code = code
        super().__init__("%s failed: %d (%s)" % (what, code, strerror(code)))


_vp, _sz, _i, _f, _u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_float, ctypes.c_uint64
_P = ctypes.POINTER

# name -> (restype, argtypes); mirrors include/b200validate.h declaration by declaration
SIGNATURES = {
    "b200v_version": (None, [_P(_i), _P(_i)]),
    "b200v_strerror": (ctypes.c_char_p, [_i]),
    "b200v_device_count": (_i, [_P(_i)]),
    "b200v_device_info": (_i, [_i, _P(Info)]),
    "b200v_host_alloc": (_i, [_P(_vp), _sz]),
    "b200v_host_free": (_i, [_vp]),
    "b200v_host_alloc_ex": (_i, [_P(_vp), _sz, _i]),
    "b200v_vectoradd": (_i, [_i, _vp, _vp, _vp, _sz, _P(Timing)]),
    "b200v_vectoradd_noalloc": (_i, [_i, _vp, _vp, _vp, _sz, _vp]),
    "b200v_stream_host": (_i, [_i, _i, _vp, _vp, _vp, _f, _sz, _P(Timing)]),
    "b200v_vectoradd_device": (_i, [_i, _vp, _vp, _vp, _sz, _vp]),
    "b200v_stream_device": (_i, [_i, _i, _vp, _vp, _vp, _f, _sz, _vp]),
    "b200v_kernel_name": (_i, [_i, _vp, _vp, _vp, _sz, ctypes.c_char_p, _sz]),
    "b200v_vectoradd_device_sampleshape": (_i, [_i, _vp, _vp, _vp, _sz, _vp]),
    "b200v_fill_device": (_i, [_i, _vp, _sz, _u64, _vp]),
    "b200v_digest_device": (_i, [_i, _vp, _sz, _P(_u64), _vp]),
    "b200v_check_device": (_i, [_i, _i, _vp, _vp, _vp, _f, _sz, _P(_u64), _P(_u64), _vp]),
    "b200v_stream": (_i, [_i, _i, _sz, _f, _u64, _i, _i, _P(StreamResult)]),
    "b200v_allreduce_probe": (_i, [_i, _P(_i), _sz, _sz, _i, _i, _P(ArRow), _i, _P(_i)]),
    "b200v_nccl_unique_id": (_i, [ctypes.c_char_p]),
    "b200v_allreduce_probe_rank": (_i, [_i, _i, _i, ctypes.c_char_p, _sz, _sz, _i, _i, _P(ArRow), _i, _P(_i)]),
    "b200v_nccl_version": (_i, [_P(_i)]),
    "b200v_last_nccl_error": (_i, []),
    "b200v_nvls_allreduce_probe": (_i, [_i, _P(_i), _sz, _sz, _i, _i, _P(ArRow), _i, _P(_i)]),
    "b200v_last_driver_error": (_i, []),
    "b200v_nvls_algo_for": (_i, [_i, _sz]),
    "b200v_nvls_allreduce_probe_rank": (_i, [_i, _i, _i, ctypes.c_char_p, _sz, _sz, _i, _i, _P(ArRow), _i, _P(_i)]),
    "b200v_ar_comm_create": (_i, [_i, _i, _i, ctypes.c_char_p, _P(_vp)]),
    "b200v_ar_comm_allreduce": (_i, [_vp, _vp, _sz, _vp]),
    "b200v_ar_comm_destroy": (_i, [_vp]),
    "b200v_nvls_comm_create": (_i, [_i, _i, _i, ctypes.c_char_p, _sz, _P(_vp)]),
    "b200v_nvls_comm_allreduce": (_i, [_vp, _vp, _sz, _vp]),
    "b200v_nvls_comm_check": (_i, [_vp]),
    "b200v_nvls_comm_destroy": (_i, [_vp]),
    "b200v_shutdown": (_i, []),
}

_lib = None


def lib():
    """The loaded shared library.  Raises OSError when it has not been built: no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError("%s is missing: build it with `make -C %s lib` (needs nvcc); this package "
                          "has no CPU fallback" % (LIB_PATH, _HERE))
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def strerror(code):
    return lib().b200v_strerror(code).decode()


def _check(code, what, allow=()):
    if code != 0 and code not in allow:
        raise B200VError(code, what)
    return code


def _ptr(a):
    """Address of a numpy array, torch tensor, ctypes buffer or raw int."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    if hasattr(a, "ctypes"):
        return a.ctypes.data
    return ctypes.addressof(a)


def version():
    a, b = _i(), _i()
    lib().b200v_version(ctypes.byref(a), ctypes.byref(b))
    return a.value, b.value


def device_count():
    n = _i()
    rc = lib().b200v_device_count(ctypes.byref(n))
    if rc == E_NODEVICE:
        return 0
    _check(rc, "b200v_device_count")
    return n.value


def device_info(dev=0):
    info = Info()
    _check(lib().b200v_device_info(dev, ctypes.byref(info)), "b200v_device_info")
    return info


HOST_WRITE_ONLY = 1


def host_alloc(nbytes, flags=0):
    p = _vp()
    _check(lib().b200v_host_alloc_ex(ctypes.byref(p), nbytes, flags), "b200v_host_alloc_ex")
    return p.value


def host_free(ptr):
    _check(lib().b200v_host_free(ptr), "b200v_host_free")


class PinnedArray:
    """n elements of pinned host memory from b200v_host_alloc, viewed as a numpy array (.array).
    Freed by close() or when the object is collected; the array must not outlive it."""

    def __init__(self, n, dtype="float32", write_only=False):
        import numpy as np
        dt = np.dtype(dtype)
        self.nbytes = max(1, n * dt.itemsize)
        self.ptr = host_alloc(self.nbytes, HOST_WRITE_ONLY if write_only else 0)
        buf = (ctypes.c_char * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dt, count=n)

    def close(self):
        if self.ptr:
            self.array = None
            host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def vectoradd(hA, hB, hC, n=None, dev=0):
    """hC = hA + hB + 0.0f through the host-buffer entry point.  Returns the Timing struct."""
    n = hA.size if n is None else n
    t = Timing()
    _check(lib().b200v_vectoradd(dev, _ptr(hA), _ptr(hB), _ptr(hC), n, ctypes.byref(t)), "b200v_vectoradd")
    return t


NOALLOC_MAX_ELEMS = 1 << 20
E_TOOLARGE = -7


def vectoradd_noalloc(hA, hB, hC, n=None, dev=0, stream=None):
    """hC = hA + hB + 0.0f through b200v_vectoradd_noalloc: no device allocation, n capped at
    NOALLOC_MAX_ELEMS (B200VError with code E_TOOLARGE beyond it)."""
    n = hA.size if n is None else n
    _check(lib().b200v_vectoradd_noalloc(dev, _ptr(hA), _ptr(hB), _ptr(hC), n, stream),
           "b200v_vectoradd_noalloc")


def stream_host(op, hDst, hX, hY=None, s=0.0, n=None, dev=0):
    n = hX.size if n is None else n
    t = Timing()
    _check(lib().b200v_stream_host(dev, op, _ptr(hDst), _ptr(hX), _ptr(hY), s, n, ctypes.byref(t)),
           "b200v_stream_host")
    return t


def vectoradd_device(dA, dB, dC, n, dev=0, stream=None):
    _check(lib().b200v_vectoradd_device(dev, _ptr(dA), _ptr(dB), _ptr(dC), n, stream), "b200v_vectoradd_device")


def vectoradd_device_sampleshape(dA, dB, dC, n, dev=0, stream=None):
    _check(lib().b200v_vectoradd_device_sampleshape(dev, _ptr(dA), _ptr(dB), _ptr(dC), n, stream),
           "b200v_vectoradd_device_sampleshape")


def kernel_name(op, dDst, dX, dY, n):
    """Name of the kernel the library dispatches for this op and these device pointers."""
    buf = ctypes.create_string_buffer(256)
    _check(lib().b200v_kernel_name(op, _ptr(dDst), _ptr(dX), _ptr(dY), n, buf, len(buf)), "b200v_kernel_name")
    return buf.value.decode()


def stream_device(op, dDst, dX, dY, s, n, dev=0, stream=None):
    _check(lib().b200v_stream_device(dev, op, _ptr(dDst), _ptr(dX), _ptr(dY), s, n, stream),
           "b200v_stream_device")


def fill_device(dX, n, seed, dev=0, stream=None):
    _check(lib().b200v_fill_device(dev, _ptr(dX), n, seed & (2**64 - 1), stream), "b200v_fill_device")


def digest_device(dWords, n, dev=0, stream=None):
    out = _u64()
    _check(lib().b200v_digest_device(dev, _ptr(dWords), n, ctypes.byref(out), stream), "b200v_digest_device")
    return out.value


def check_device(op, dDst, dX, dY, s, n, dev=0, stream=None):
    """(mismatches, first_mismatch) of dDst against op recomputed element by element."""
    bad, first = _u64(), _u64()
    _check(lib().b200v_check_device(dev, op, _ptr(dDst), _ptr(dX), _ptr(dY), s, n, ctypes.byref(bad),
                                    ctypes.byref(first), stream), "b200v_check_device", allow=(E_MISMATCH,))
    return bad.value, first.value


def stream(op, n, s=3.0, seed=0xB200, warmup=5, iters=20, dev=0):
    r = StreamResult()
    _check(lib().b200v_stream(dev, op, n, s, seed, warmup, iters, ctypes.byref(r)), "b200v_stream",
           allow=(E_MISMATCH,))
    return r


def _rows(rows, n):
    return [rows[i] for i in range(n)]


def allreduce_probe(devs, min_bytes=1024, max_bytes=1 << 30, warmup=5, iters=20):
    _settle_nccl()
    devs = list(devs)
    arr = (_i * len(devs))(*devs)
    rows = (ArRow * 64)()
    n = _i()
    _check(lib().b200v_allreduce_probe(len(devs), arr, min_bytes, max_bytes, warmup, iters, rows, 64,
                                       ctypes.byref(n)), "b200v_allreduce_probe")
    return _rows(rows, n.value)


E_UNSUPPORTED, E_DRIVER, E_TIMEOUT = -8, -9, -10


def nvls_algo_for(nranks, nbytes):
    """Name of the kernel the hand-written all-reduce picks (host logic only, no GPU needed)."""
    return AR_ALGO_NAMES[_check(lib().b200v_nvls_algo_for(nranks, nbytes), "b200v_nvls_algo_for", allow=(1, 2, 3, 4, 5))]


def nvls_allreduce_probe(devs, min_bytes=1024, max_bytes=1 << 30, warmup=5, iters=20):
    """The hand-written NVLink-multicast all-reduce probe; same rows as allreduce_probe."""
    devs = list(devs)
    arr = (_i * len(devs))(*devs)
    rows = (ArRow * 64)()
    n = _i()
    _check(lib().b200v_nvls_allreduce_probe(len(devs), arr, min_bytes, max_bytes, warmup, iters, rows, 64,
                                            ctypes.byref(n)), "b200v_nvls_allreduce_probe")
    return _rows(rows, n.value)


def nvls_allreduce_probe_rank(dev, rank, nranks, rendezvous, min_bytes=1024, max_bytes=1 << 30, warmup=5,
                              iters=20):
    """One rank of the hand-written all-reduce probe, one process per GPU (collective call)."""
    rows = (ArRow * 64)()
    n = _i()
    _check(lib().b200v_nvls_allreduce_probe_rank(dev, rank, nranks, rendezvous.encode(), min_bytes, max_bytes,
                                                 warmup, iters, rows, 64, ctypes.byref(n)),
           "b200v_nvls_allreduce_probe_rank")
    return _rows(rows, n.value)


def _settle_nccl():
    """A process gets ONE libnccl.so.2: whichever is mapped first.  The library dlopen()s the
    system's on first use; torch brings a newer one and fails to import if the older is already
    there.  In a Python process (tests, bench.py) let torch map its copy first; the library then
    finds it already loaded and uses it.  No torch, no problem: the system's is used."""
    try:
        import torch  # noqa: F401
    except Exception:
        pass


def nccl_unique_id():
    _settle_nccl()
    buf = ctypes.create_string_buffer(NCCL_UNIQUE_ID_BYTES)
    _check(lib().b200v_nccl_unique_id(buf), "b200v_nccl_unique_id")
    return buf.raw


def allreduce_probe_rank(dev, rank, nranks, uid, min_bytes=1024, max_bytes=1 << 30, warmup=5, iters=20):
    assert len(uid) == NCCL_UNIQUE_ID_BYTES
    _settle_nccl()
    rows = (ArRow * 64)()
    n = _i()
    _check(lib().b200v_allreduce_probe_rank(dev, rank, nranks, uid, min_bytes, max_bytes, warmup, iters,
                                            rows, 64, ctypes.byref(n)), "b200v_allreduce_probe_rank")
    return _rows(rows, n.value)


class ArComm:
    """One rank of an NCCL communicator that all-reduces caller-owned device buffers
    (b200v_ar_comm_*).  Collective: every rank constructs it, with the id rank 0 made."""

    def __init__(self, dev, rank, nranks, uid):
        assert len(uid) == NCCL_UNIQUE_ID_BYTES
        _settle_nccl()
        h = _vp()
        _check(lib().b200v_ar_comm_create(dev, rank, nranks, uid, ctypes.byref(h)), "b200v_ar_comm_create")
        self.h = h

    def allreduce(self, dBuf, count, stream=None):
        _check(lib().b200v_ar_comm_allreduce(self.h, _ptr(dBuf), count, stream), "b200v_ar_comm_allreduce")

    def close(self):
        if self.h:
            _check(lib().b200v_ar_comm_destroy(self.h), "b200v_ar_comm_destroy")
            self.h = None


class NvlsComm:
    """One rank of the hand-written NVLink all-reduce on caller-owned device buffers
    (b200v_nvls_comm_*).  Collective; `rendezvous` names the abstract Unix socket the ranks of this
    host meet on."""

    def __init__(self, dev, rank, nranks, rendezvous, max_bytes):
        h = _vp()
        _check(lib().b200v_nvls_comm_create(dev, rank, nranks, rendezvous.encode(), max_bytes, ctypes.byref(h)),
               "b200v_nvls_comm_create")
        self.h = h

    def allreduce(self, dBuf, count, stream=None):
        _check(lib().b200v_nvls_comm_allreduce(self.h, _ptr(dBuf), count, stream), "b200v_nvls_comm_allreduce")

    def check(self):
        """Synchronises the device; raises if a kernel gave up waiting for a peer."""
        _check(lib().b200v_nvls_comm_check(self.h), "b200v_nvls_comm_check")

    def close(self):
        if self.h:
            _check(lib().b200v_nvls_comm_destroy(self.h), "b200v_nvls_comm_destroy")
            self.h = None


def nccl_version():
    _settle_nccl()
    v = _i()
    _check(lib().b200v_nccl_version(ctypes.byref(v)), "b200v_nccl_version")
    return v.value


def shutdown():
    _check(lib().b200v_shutdown(), "b200v_shutdown")
