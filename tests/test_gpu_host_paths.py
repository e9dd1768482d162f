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


def test_repeated_calls_do_not_leak_device_memory(capi, gpu):
    """Entry points that allocate per call (stream rows, digest, check, the probes) give it all
    back; the host path keeps only its bounded slot ring."""
    import torch
    n = (1 << 20) + 3
    x = torch.rand(n, device="cuda")
    a = np.ones(n, np.float32)
    c = np.zeros(n, np.float32)

    def one_round():
        capi.stream(capi.OP_TRIAD, n, warmup=1, iters=2)
        capi.digest_device(x.data_ptr(), n)
        capi.check_device(capi.OP_COPY, x.data_ptr(), x.data_ptr(), None, 0.0, n)
        capi.allreduce_probe([0], min_bytes=1 << 10, max_bytes=1 << 12, warmup=1, iters=2)
        capi.vectoradd(a, a, c)
        m = capi.NOALLOC_MAX_ELEMS
        capi.vectoradd_noalloc(a[:m], a[:m], c[:m])

    for _ in range(3):
        one_round()                                   # warm: ring, staging, NCCL, lazy module loads
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(25):
        one_round()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    # a leaked per-call buffer would cost >= 100 MiB here; allow NCCL's own small internal growth
    assert free0 - free1 < (64 << 20), "device memory shrank by %d bytes over 25 rounds" % (free0 - free1)


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


@pytest.mark.parametrize("policy", ["pinned", "wc", "thp"])
def test_host_alloc_policies_give_the_same_bytes(capi, oracle, gpu, policy, monkeypatch):
    """b200v_host_alloc_ex under each B200V_HOST_POLICY: inputs flagged write-only (they may come
    back as write-combined memory, which the CPU only writes here), the output readable, the result
    identical to the oracle's, and every block released by b200v_host_free."""
    monkeypatch.setenv("B200V_HOST_POLICY", policy)
    n = (1 << 22) + 77
    a = oracle.fill_hash(n, 0x51, threads=4)
    b = oracle.fill_hash(n, 0x52, threads=4)
    ha, hb = capi.PinnedArray(n, write_only=True), capi.PinnedArray(n, write_only=True)
    hc = capi.PinnedArray(n)
    try:
        ha.array[:] = a                              # CPU writes only
        hb.array[:] = b
        hc.array[:] = np.nan
        t = capi.vectoradd(ha.array, hb.array, hc.array, n)
        assert hc.array.tobytes() == oracle.vadd(a, b, threads=4).tobytes()
        assert t.chunks >= 1 and t.kernel_launches == t.chunks
    finally:
        for h in (ha, hb, hc):
            h.close()
