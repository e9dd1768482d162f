"""GPU tests of how callers actually reach the host-buffer entry points: from several OS threads
(the cgo model: one goroutine locked to a thread per device), with host pointers that are not
nicely aligned, and repeatedly with changing sizes (the slot ring is reused and regrown)."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_unaligned_host_pointers(capi, oracle, gpu):
    n = 100003
    base_a = oracle.fill_hash(n + 3, 31)
    base_b = oracle.fill_hash(n + 3, 32)
    out = np.full(n + 3, np.nan, np.float32)
    for off in (0, 1, 2, 3):                       # 4-byte aligned only
        a, b, c = base_a[off:off + n], base_b[off:off + n], out[off:off + n]
        out[:] = np.nan
        capi.vectoradd(a, b, c)
        assert c.tobytes() == oracle.vadd(a.copy(), b.copy()).tobytes()
        assert np.isnan(out[:off]).all() and np.isnan(out[off + n:]).all()


def test_ring_is_reused_and_regrown_across_calls(capi, oracle, gpu):
    for n in (1000, (1 << 22) + 5, 17, (1 << 24) + (1 << 23) + 11, 50000, 3 * (1 << 20) + 1):
        a = oracle.fill_hash(n, n, threads=4)
        b = oracle.fill_hash(n, n + 1, threads=4)
        c = np.zeros(n, np.float32)
        t = capi.vectoradd(a, b, c)
        assert c.tobytes() == oracle.vadd(a, b, threads=4).tobytes(), n
        assert t.kernel_launches == t.chunks >= 1


def test_concurrent_calls_from_threads(capi, oracle, gpu):
    """Two threads per visible GPU call b200v_vectoradd at once: calls on one device serialise on
    the library's per-device lock, calls on different devices run in parallel; all are exact."""
    n = (1 << 22) + 12345
    a = oracle.fill_hash(n, 41, threads=4)
    b = oracle.fill_hash(n, 42, threads=4)
    want = oracle.vadd(a, b, threads=4).tobytes()
    results, errors = {}, []

    def work(dev, k):
        try:
            out = np.zeros(n, np.float32)
            for _ in range(3):
                capi.vectoradd(a, b, out, dev=dev)
            results[(dev, k)] = out.tobytes() == want
        except Exception as e:  # surfaced below
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(dev, k)) for dev in range(min(gpu, 4)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert len(results) == len(threads) and all(results.values())
