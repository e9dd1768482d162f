"""Property-based GPU parity (hypothesis): random sizes, ops, scalars, host offsets and device
alignments — every draw must match the CPU oracle byte for byte, through both the host-buffer and
the device-pointer entry points."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu

OPS = ["copy", "scale", "add", "triad", "vectoradd", "triad_unfused"]
SIZES = st.one_of(st.integers(0, 70), st.integers(2040, 2060), st.integers(8180, 8200),
                  st.integers(0, 3 << 20), st.sampled_from([1 << 20, (1 << 22) + 1, 3 * (1 << 22) + 7]))
SCALARS = st.sampled_from([3.0, -1.0, 0.5, 1e-3, -7.25, 1e20, 0.0])


def make_inputs(n, off, seed, oracle):
    x = oracle.fill_hash(n + 8, seed)
    y = oracle.fill_hash(n + 8, seed + oracle.SEED_Y_OFFSET)
    x[: min(8, x.size)] = np.array([-0.0, 0.0, 1e-45, -1e-45, 3e38, -3e38, 1.0, -1.0], np.float32)[: min(8, x.size)]
    return x[off:off + n], y[off:off + n]


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(n=SIZES, name=st.sampled_from(OPS), s=SCALARS, off=st.integers(0, 7), seed=st.integers(0, 2**32))
def test_host_entry_point_matches_oracle(capi, oracle, gpu, n, name, s, off, seed):
    x, y = make_inputs(n, off, seed, oracle)
    out = np.full(n + 2, np.nan, np.float32)
    capi.stream_host(getattr(capi, "OP_" + name.upper()), out[1:1 + n], x, y, s, n=n)
    want = oracle.stream(getattr(oracle, "OP_" + name.upper()), x.copy(), y.copy(), s)
    assert out[1:1 + n].tobytes() == want.tobytes()
    assert np.isnan(out[0]) and np.isnan(out[-1])           # nothing written outside the range


@settings(max_examples=30, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(n=st.integers(1, 300000), name=st.sampled_from(OPS), s=SCALARS, od=st.integers(0, 9), ox=st.integers(0, 9),
       oy=st.integers(0, 9), seed=st.integers(0, 2**32))
def test_device_entry_point_matches_oracle(capi, oracle, gpu, n, name, s, od, ox, oy, seed):
    import torch
    hx, hy = make_inputs(n + 16, 0, seed, oracle)
    dx = torch.from_numpy(hx.copy()).cuda()
    dy = torch.from_numpy(hy.copy()).cuda()
    dd = torch.full((n + 16,), float("nan"), device="cuda")
    capi.stream_device(getattr(capi, "OP_" + name.upper()), dd[od:].data_ptr(), dx[ox:].data_ptr(),
                       dy[oy:].data_ptr(), s, n)
    torch.cuda.synchronize()
    want = oracle.stream(getattr(oracle, "OP_" + name.upper()), hx[ox:ox + n].copy(), hy[oy:oy + n].copy(), s)
    assert dd[od:od + n].cpu().numpy().tobytes() == want.tobytes()
    assert bool(torch.isnan(dd[:od]).all()) and bool(torch.isnan(dd[od + n:]).all())
