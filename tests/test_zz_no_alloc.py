"""GPU test of b200v_vectoradd_noalloc, the vectorAdd entry point for callers that cannot allocate
device memory (a health probe re-running the validation on a GPU whose HBM belongs to a tenant's
job: cudaMalloc may fail there, and cudaMalloc/cudaFree would stall the tenant's streams).

The entry point stages through a fixed array inside the library instead of allocating, so what
is checked is exactly what such a caller relies on: called through the C ABI it returns the same
bytes as the regular entry point b200v_vectoradd and as the CPU oracle (bit-exact, the bar every
parity test here uses), it still does so on a second call that finds the previous call's data in
the shared array, it leaves the GPU's free memory exactly where it was once the first call has
put the array in place, and it refuses sizes beyond its fixed capacity without touching the
output.

This file sorts last on purpose (pytest collects files alphabetically), so a run that stops at
the first failure has already run every other test.  It does not depend on them: it passes alone
in a fresh process too.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(n, seed):
    rng = np.random.default_rng(seed)
    a = rng.random(n, dtype=np.float32)
    b = rng.random(n, dtype=np.float32)
    # sprinkle the cases where "+ 0.0f" and denormal handling are visible
    if n >= 8:
        a[:4] = np.array([-0.0, 1e-45, -1.5, 3.0e38], np.float32)
        b[:4] = np.array([-0.0, 1e-45, 1.5, 3.0e38], np.float32)
    return a, b


def test_noalloc_matches_regular_entry_point_on_two_calls_in_a_row(capi, gpu, oracle):
    cap = capi.NOALLOC_MAX_ELEMS
    # Call 1 fills the whole staging array; call 2 is shorter, ragged and has different data, so
    # any element it picked up from call 1's leftovers would differ from the regular entry point.
    for n, seed in ((cap, 101), (50000 + 3, 202)):
        a, b = _inputs(n, seed)
        want = np.full(n, np.nan, np.float32)
        capi.vectoradd(a, b, want)                      # existing entry point
        got = np.full(n, np.nan, np.float32)
        capi.vectoradd_noalloc(a, b, got)               # new entry point, default stream
        assert got.tobytes() == want.tobytes(), "n=%d" % n
        assert got.tobytes() == oracle.vadd(a, b).tobytes(), "n=%d vs the CPU oracle" % n
        assert not np.isnan(got).any()


def test_noalloc_leaves_free_device_memory_unchanged_after_the_first_call(capi, gpu):
    """The promise a probe on a tenant's GPU relies on: after the first call on a GPU (which loads
    the module and with it the fixed array) further calls reserve and release nothing.  Free
    memory as the driver reports it is the same before and after, at both ends of the size
    range."""
    import torch
    n = capi.NOALLOC_MAX_ELEMS
    a, b = _inputs(n, 31)
    out = np.empty(n, np.float32)
    capi.vectoradd_noalloc(a, b, out)                   # set-up call
    torch.cuda.synchronize(0)
    free0, _total = torch.cuda.mem_get_info(0)
    for m in (n, 1000, n):
        capi.vectoradd_noalloc(a[:m], b[:m], out[:m])
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 == free0, "free device memory moved by %d bytes" % (free0 - free1)


def test_noalloc_on_a_second_gpu(capi, gpu, oracle):
    """Every GPU has its own copy of the array and its own one-time address lookup."""
    if gpu < 2:
        pytest.skip("needs a second GPU; the single-GPU lease covers device 0 above")
    n = 300001
    a, b = _inputs(n, 77)
    want = oracle.vadd(a, b).tobytes()
    for dev in (1, 0, 1):
        got = np.full(n, np.nan, np.float32)
        capi.vectoradd_noalloc(a, b, got, dev=dev)
        assert got.tobytes() == want, "dev=%d" % dev


def test_noalloc_on_a_caller_stream_and_tiny_sizes(capi, gpu):
    import torch
    st = torch.cuda.Stream(device=0)    # explicit: the library leaves whichever GPU it used last current
    for n in (0, 1, 7, 2048, 2049):
        a, b = _inputs(n, 7 + n)
        want = np.zeros(n, np.float32)
        capi.vectoradd(a, b, want)
        got = np.full(n, -1.0, np.float32)
        capi.vectoradd_noalloc(a, b, got, stream=st.cuda_stream)
        assert got.tobytes() == want.tobytes(), "n=%d" % n


def test_noalloc_refuses_sizes_beyond_its_capacity(capi, gpu):
    n = capi.NOALLOC_MAX_ELEMS + 1
    a, b = _inputs(n, 5)
    out = np.full(n, -7.0, np.float32)
    with pytest.raises(capi.B200VError) as ei:
        capi.vectoradd_noalloc(a, b, out)
    assert ei.value.code == capi.E_TOOLARGE
    assert "fixed capacity" in str(ei.value)
    assert (out == -7.0).all()                          # nothing was computed or copied
    with pytest.raises(capi.B200VError) as ei:
        capi.vectoradd_noalloc(a, None, out, n=16)
    assert ei.value.code == capi.E_BADARG
