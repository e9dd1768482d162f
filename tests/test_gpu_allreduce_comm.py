"""GPU parity tests of the all-reduce on CALLER-OWNED device buffers (b200v_ar_comm_* over NCCL,
b200v_nvls_comm_* over the hand-written NVLink kernels): the test fills each rank's buffer, the
library reduces it in place, the test reads it back and compares it byte for byte with the CPU
oracle's closed form — nothing inside the library judges the result.

On a one-GPU lease the NCCL communicator has one rank (the sum is the input), which still runs
the real communicator set-up and ncclAllReduce on the caller's buffer; the hand-written kernels
need a multicast object, which the driver refuses to make for a single device, so they are
covered from two GPUs up (and by bench.py at N = 2, 4, 8, which checks them against the same
oracle on every rank).  With two or more GPUs the ranks run as host threads, one per GPU.
"""
import os
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZES = (1003, 4096, (1 << 18) + 5, 1 << 22)          # floats: ragged, tiny, ragged mid, 16 MiB


def _nranks(gpu):
    return 1 if gpu < 2 else (2 if gpu < 4 else (4 if gpu < 8 else 8))


def _run_ranks(nranks, fn):
    """fn(rank) on one host thread per rank; re-raises the first failure."""
    errs = [None] * nranks

    def wrap(r):
        try:
            fn(r)
        except BaseException as e:      # noqa: BLE001 - reported below
            errs[r] = e
    th = [threading.Thread(target=wrap, args=(r,)) for r in range(nranks)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for e in errs:
        if e is not None:
            raise e


def _reduce_and_compare(make_comm, nranks, oracle, sizes, repeats=1):
    import torch
    got = {}

    def rank_fn(r):
        torch.cuda.set_device(r)
        comm = make_comm(r)
        try:
            for n in sizes:
                buf = torch.from_numpy(oracle.allreduce_input(n, r)).cuda(r)
                guard = torch.full((n + 64,), -77.0, device="cuda:%d" % r)
                guard[:n] = buf
                st = torch.cuda.current_stream(r).cuda_stream
                for _ in range(repeats):                     # back to back, nothing in between
                    comm.allreduce(guard.data_ptr(), n, stream=st)
                torch.cuda.synchronize(r)
                if hasattr(comm, "check"):
                    comm.check()
                host = guard.cpu().numpy()
                assert (host[n:] == -77.0).all(), "wrote past the end at n=%d" % n
                got[(r, n)] = host[:n].copy()
        finally:
            comm.close()
    _run_ranks(nranks, rank_fn)
    for n in sizes:
        want = oracle.allreduce_expected(n, nranks) * np.float32(nranks ** (repeats - 1))
        for r in range(nranks):
            assert got[(r, n)].tobytes() == want.tobytes(), "rank %d n=%d repeats=%d" % (r, n, repeats)


def test_nccl_comm_allreduce_matches_the_oracle(capi, gpu, oracle):
    nranks = _nranks(gpu)
    uid = capi.nccl_unique_id()
    _reduce_and_compare(lambda r: capi.ArComm(r, r, nranks, uid), nranks, oracle, SIZES)
    uid = capi.nccl_unique_id()
    _reduce_and_compare(lambda r: capi.ArComm(r, r, nranks, uid), nranks, oracle, SIZES[:3], repeats=3)


@pytest.mark.parametrize("algo", ["auto", "multimem", "oneshot", "p2p", "push", "hybrid"])
def test_nvls_comm_allreduce_matches_the_oracle(capi, gpu, oracle, algo, monkeypatch):
    nranks = _nranks(gpu)
    if nranks < 2:
        pytest.skip("needs at least two GPUs (no multicast object for one device)")
    if algo == "auto":
        monkeypatch.delenv("B200V_NVLS_ALGO", raising=False)
    else:
        monkeypatch.setenv("B200V_NVLS_ALGO", algo)
    name = "b200v-comm-test-%d-%s" % (os.getpid(), algo)
    max_bytes = 4 * max(SIZES) + 64

    def make(r):
        try:
            return capi.NvlsComm(r, r, nranks, name, max_bytes)
        except capi.B200VError as e:
            if e.code == capi.E_UNSUPPORTED:
                pytest.skip("no multicast support on this box (%d rank(s))" % nranks)
            raise
    _reduce_and_compare(make, nranks, oracle, SIZES)
    name += "-b2b"
    _reduce_and_compare(make, nranks, oracle, SIZES[:3], repeats=3)


@pytest.mark.parametrize("algo", ["push", "p2p"])
def test_peer_kernels_add_in_rank_order(capi, gpu, algo, monkeypatch):
    """Arbitrary floats, where the order of additions shows: the two peer kernels promise the
    sequential sum over ranks 0..n-1 (one owner per slice adds in rank order and broadcasts), so
    every rank must hold exactly the bytes of that fp32 sum."""
    import torch
    nranks = _nranks(gpu)
    if nranks < 2:
        pytest.skip("needs at least two GPUs")
    monkeypatch.setenv("B200V_NVLS_ALGO", algo)
    n = (1 << 20) + 37
    rng = np.random.default_rng(2024)
    inputs = [(rng.standard_normal(n) * 10.0 ** rng.integers(-3, 4, n)).astype(np.float32) for _ in range(nranks)]
    want = inputs[0].copy()
    for r in range(1, nranks):
        want = want + inputs[r]                          # fp32, rank order
    got = [None] * nranks
    name = "b200v-order-%d-%s" % (os.getpid(), algo)

    def rank_fn(r):
        torch.cuda.set_device(r)
        try:
            comm = capi.NvlsComm(r, r, nranks, name, 4 * n + 64)
        except capi.B200VError as e:
            if e.code == capi.E_UNSUPPORTED:
                return
            raise
        try:
            buf = torch.from_numpy(inputs[r]).cuda(r)
            comm.allreduce(buf.data_ptr(), n, stream=torch.cuda.current_stream(r).cuda_stream)
            torch.cuda.synchronize(r)
            comm.check()
            got[r] = buf.cpu().numpy()
        finally:
            comm.close()
    _run_ranks(nranks, rank_fn)
    if any(g is None for g in got):
        pytest.skip("no multicast support on this box")
    for r in range(nranks):
        assert got[r].tobytes() == want.tobytes(), "rank %d" % r


def test_a_rank_that_never_arrives_times_out_instead_of_hanging(capi, gpu, monkeypatch):
    """A health probe must not wedge a GPU.  Rank 0 starts an all-reduce that rank 1 never joins:
    its kernel gives up after the bounded spin (200 ms here), the stream drains, and
    b200v_nvls_comm_check reports B200V_E_TIMEOUT."""
    import time

    import torch
    if gpu < 2:
        pytest.skip("needs at least two GPUs")
    monkeypatch.setenv("B200V_NVLS_TIMEOUT_MS", "200")
    comms = [None, None]

    def make(r):
        torch.cuda.set_device(r)
        try:
            comms[r] = capi.NvlsComm(r, r, 2, "b200v-comm-timeout-%d" % os.getpid(), 1 << 20)
        except capi.B200VError as e:
            if e.code != capi.E_UNSUPPORTED:
                raise
    _run_ranks(2, make)
    if comms[0] is None or comms[1] is None:
        pytest.skip("no multicast support on this box")
    try:
        torch.cuda.set_device(0)
        buf = torch.ones(4096, device="cuda:0")
        t0 = time.time()
        comms[0].allreduce(buf.data_ptr(), buf.numel())      # rank 1 stays away
        with pytest.raises(capi.B200VError) as ei:
            comms[0].check()
        assert ei.value.code == capi.E_TIMEOUT
        assert time.time() - t0 < 10.0                       # bounded: nowhere near a hang
        torch.cuda.synchronize(0)                            # the GPU is still usable
        assert float(torch.ones(8, device="cuda:0").sum()) == 8.0
    finally:
        _run_ranks(2, lambda r: comms[r].close())


def test_nvls_comm_rejects_what_it_cannot_hold(capi, gpu):
    import torch
    with pytest.raises(capi.B200VError) as ei:
        capi.NvlsComm(0, 0, 1, "b200v-comm-one-%d" % os.getpid(), 1 << 20)
    assert ei.value.code == capi.E_BADARG                    # one rank is not a collective
    if gpu < 2:
        return
    comms = [None, None]

    def rank_fn(r):
        torch.cuda.set_device(r)
        try:
            comms[r] = capi.NvlsComm(r, r, 2, "b200v-comm-small-%d" % os.getpid(), 1 << 20)
        except capi.B200VError as e:
            if e.code != capi.E_UNSUPPORTED:
                raise
    _run_ranks(2, rank_fn)
    if comms[0] is None or comms[1] is None:
        pytest.skip("no multicast support on this box")
    try:
        buf = torch.zeros(4 << 20, device="cuda:0")
        with pytest.raises(capi.B200VError) as ei:
            comms[0].allreduce(buf.data_ptr(), buf.numel())
        assert ei.value.code == capi.E_TOOLARGE
        comms[0].allreduce(buf.data_ptr(), 0)                # empty: a no-op
        with pytest.raises(capi.B200VError) as ei:
            comms[0].allreduce(None, 16)
        assert ei.value.code == capi.E_BADARG
    finally:
        _run_ranks(2, lambda r: comms[r].close())
