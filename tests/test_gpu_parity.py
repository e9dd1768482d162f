"""GPU parity tests: every kernel, through the C ABI, bit for bit against the CPU oracle.

Bar (BASELINE.json north_star): bit-exact — IEEE fp32 elementwise arithmetic is deterministic, so
the comparison is on raw bytes with no tolerance.  The sample's own 1e-5 verification loop is run
as well where the sample would run it.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 0xB200
SIZES = [1, 3, 4, 5, 7, 8, 9, 255, 256, 257, 2047, 2048, 2049, 50000, (1 << 20) + 1]
OPS = ["copy", "scale", "add", "triad", "vectoradd", "triad_unfused"]


def op_id(mod, name):
    return getattr(mod, "OP_" + name.upper())


def same_bytes(a, b):
    return a.tobytes() == b.tobytes()


@pytest.fixture(scope="module")
def torch_cuda(gpu):
    import torch
    assert torch.cuda.is_available()
    return torch


# ---- cfg0: the sample itself ---------------------------------------------------------------------

def test_cfg0_sample_vectoradd(capi, oracle, golden, gpu):
    n = 50000
    A, B = oracle.fill_rand(n)
    C = np.zeros(n, np.float32)
    t = capi.vectoradd(A, B, C)
    assert same_bytes(C, oracle.vadd(A, B))
    assert oracle.verify_tolerance(A, B, C) == -1            # the sample's own check
    assert "%016x" % oracle.fnv1a64_words(C) == golden["sample"]["C_fnv1a64_words"]
    assert t.kernel_launches == 1 and t.chunks == 1 and t.grid >= 1 and t.block >= 32
    assert t.total_ms > 0


@pytest.mark.parametrize("n", [0] + SIZES)
def test_vectoradd_host_sizes(capi, oracle, gpu, n):
    A = oracle.fill_hash(n, SEED)
    B = oracle.fill_hash(n, SEED + oracle.SEED_Y_OFFSET)
    C = np.full(n, np.nan, np.float32)
    capi.vectoradd(A, B, C)
    assert same_bytes(C, oracle.vadd(A, B))


def test_vectoradd_host_multichunk_pipeline(capi, oracle, gpu):
    """More chunks than ring slots, ragged last chunk, pageable and pinned host memory."""
    n = 5 * (1 << 22) + 12345
    A = oracle.fill_hash(n, 1, threads=8)
    B = oracle.fill_hash(n, 2, threads=8)
    want = oracle.vadd(A, B, threads=8)
    C = np.zeros(n, np.float32)
    t = capi.vectoradd(A, B, C)
    assert t.chunks > 3 and t.kernel_launches == t.chunks          # more chunks than ring slots
    assert same_bytes(C, want)
    pa, pb, pc = capi.PinnedArray(n), capi.PinnedArray(n), capi.PinnedArray(n)
    pa.array[:] = A
    pb.array[:] = B
    pc.array[:] = 0
    capi.vectoradd(pa.array, pb.array, pc.array)
    assert same_bytes(pc.array, want)
    for p in (pa, pb, pc):
        p.close()


@pytest.mark.parametrize("name", OPS)
@pytest.mark.parametrize("n", [1, 5, 257, 4099, (1 << 20) + 1])
def test_stream_host_ops(capi, oracle, golden, gpu, name, n):
    x = oracle.fill_hash(n, SEED)
    y = oracle.fill_hash(n, SEED + oracle.SEED_Y_OFFSET)
    d = np.full(n, np.nan, np.float32)
    capi.stream_host(op_id(capi, name), d, x, y, 3.0)
    assert same_bytes(d, oracle.stream(op_id(oracle, name), x, y, 3.0))
    assert "%016x" % oracle.digest(d) == golden["ops"]["sizes"][str(n)][name]["digest"]


def test_special_values(capi, oracle, gpu):
    """Signed zeros, denormals (no flush-to-zero), infinities, cancellation, huge values."""
    f = np.float32
    A = np.array([-0.0, -0.0, 0.0, 1.5, np.inf, -np.inf, 1e-45, -1e-45, 1e-39, 3.4e38, 3.4e38, 1.0,
                  2 ** -126, 16777216.0, -1e-38, 0.1], f)
    B = np.array([-0.0, 0.0, -0.0, -1.5, 1.0, 1.0, 1e-45, 1e-45, 1e-39, 3.4e38, -3.4e38, 2 ** -24,
                  -2 ** -127, 1.0, 1e-38, 0.2], f)
    for name in OPS:
        for s in (3.0, -0.5, 1e-30):
            d = np.zeros(A.size, f)
            capi.stream_host(op_id(capi, name), d, A, B, s)
            assert same_bytes(d, oracle.stream(op_id(oracle, name), A, B, s)), (name, s)
    # NaN in -> NaN out (payload bits are not part of the contract: x86 and the GPU differ there)
    d = np.zeros(2, f)
    capi.vectoradd(np.array([np.nan, np.inf], f), np.array([1.0, -np.inf], f), d)
    assert np.isnan(d).all()


# ---- device-pointer entry points --------------------------------------------------------------------

@pytest.mark.parametrize("name", OPS)
def test_stream_device_alignments_and_inplace(capi, oracle, torch_cuda, name):
    """Vector path (32- and 16-byte aligned), scalar path (4-byte aligned offsets) and in-place."""
    torch = torch_cuda
    n = 100003
    pad = 16
    hx = oracle.fill_hash(n + pad, 5)
    hy = oracle.fill_hash(n + pad, 6)
    dx = torch.from_numpy(hx).cuda()
    dy = torch.from_numpy(hy).cuda()
    op_c, op_o = op_id(capi, name), op_id(oracle, name)
    for off_d, off_x, off_y in [(0, 0, 0), (4, 4, 4), (8, 0, 4), (1, 1, 1), (0, 3, 2), (2, 0, 0), (5, 7, 9)]:
        dd = torch.full((n + pad,), float("nan"), device="cuda")
        capi.stream_device(op_c, dd[off_d:].data_ptr(), dx[off_x:].data_ptr(), dy[off_y:].data_ptr(),
                           3.0, n, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        got = dd[off_d:off_d + n].cpu().numpy()
        want = oracle.stream(op_o, hx[off_x:off_x + n].copy(), hy[off_y:off_y + n].copy(), 3.0)
        assert same_bytes(got, want), (name, off_d, off_x, off_y)
        # nothing outside [off_d, off_d + n) was written
        rest = torch.cat([dd[:off_d], dd[off_d + n:]])
        assert bool(torch.isnan(rest).all())
    # in place: dst aliases x
    d2 = dx.clone()
    capi.stream_device(op_c, d2.data_ptr(), d2.data_ptr(), dy.data_ptr(), 3.0, n,
                       stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert same_bytes(d2[:n].cpu().numpy(), oracle.stream(op_o, hx[:n].copy(), hy[:n].copy(), 3.0))


def test_sampleshape_kernel_equals_tuned_kernel(capi, oracle, torch_cuda):
    torch = torch_cuda
    n = (1 << 22) + 77
    a = torch.empty(n, device="cuda")
    b = torch.empty(n, device="cuda")
    capi.fill_device(a.data_ptr(), n, 21)
    capi.fill_device(b.data_ptr(), n, 22)
    c1 = torch.zeros(n, device="cuda")
    c2 = torch.ones(n, device="cuda")
    capi.vectoradd_device(a.data_ptr(), b.data_ptr(), c1.data_ptr(), n)
    capi.vectoradd_device_sampleshape(a.data_ptr(), b.data_ptr(), c2.data_ptr(), n)
    torch.cuda.synchronize()
    assert bool((c1.view(torch.int32) == c2.view(torch.int32)).all())
    assert same_bytes(c1.cpu().numpy(), oracle.vadd(oracle.fill_hash(n, 21), oracle.fill_hash(n, 22)))


@pytest.mark.parametrize("n", [0, 1, 1000, (1 << 20) + 3])
def test_fill_and_digest_device(capi, oracle, torch_cuda, n):
    torch = torch_cuda
    x = torch.zeros(max(n, 1), device="cuda")
    capi.fill_device(x.data_ptr(), n, SEED)
    torch.cuda.synchronize()
    host = x[:n].cpu().numpy()
    assert same_bytes(host, oracle.fill_hash(n, SEED))
    assert capi.digest_device(x.data_ptr(), n) == oracle.digest(host)


def test_check_device_fault_injection(capi, oracle, torch_cuda):
    """The device-side verification loop passes on a good result and pins the first bad element
    when one is corrupted (the failure path of SURVEY.md §5)."""
    torch = torch_cuda
    n = 300001
    a = torch.empty(n, device="cuda")
    b = torch.empty(n, device="cuda")
    c = torch.empty(n, device="cuda")
    capi.fill_device(a.data_ptr(), n, 1)
    capi.fill_device(b.data_ptr(), n, 2)
    capi.vectoradd_device(a.data_ptr(), b.data_ptr(), c.data_ptr(), n)
    assert capi.check_device(capi.OP_VECTORADD, c.data_ptr(), a.data_ptr(), b.data_ptr(), 0.0, n) == (0, 2**64 - 1)
    c[200000] += 1.0
    c[123456] = -c[123456]
    c[299999] += 0.5
    assert capi.check_device(capi.OP_VECTORADD, c.data_ptr(), a.data_ptr(), b.data_ptr(), 0.0, n) == (3, 123456)


@pytest.mark.parametrize("name", OPS)
def test_stream_row(capi, oracle, gpu, name):
    n = (1 << 21) + 5
    r = capi.stream(op_id(capi, name), n, s=3.0, seed=SEED, warmup=2, iters=5)
    x = oracle.fill_hash(n, SEED, threads=4)
    y = oracle.fill_hash(n, SEED + oracle.SEED_Y_OFFSET, threads=4)
    assert r.mismatches == 0
    assert r.digest == oracle.digest(oracle.stream(op_id(oracle, name), x, y, 3.0))
    assert r.bytes_per_launch == n * capi.OP_BYTES_PER_ELEM[op_id(capi, name)]
    # small arrays are timed as batches of back-to-back launches: a multiple of the 5 samples
    assert r.kernel_launches >= 5 and r.kernel_launches % 5 == 0
    assert r.best_ms > 0 and r.best_ms <= r.median_ms
    assert r.l2_resident == 1


# ---- every compiled kernel variant (vector 128/256-bit, one-shot grids, TMA bulk) -------------------

VARIANT_SCRIPT = r"""
import sys
sys.path.insert(0, %(root)r)
import numpy as np
import __graft_entry__ as e
capi = e.load_package().capi
o = e.load_oracle()
ok = True
for n in (4, 1023, 4096, 4097, 50000, (1 << 22) + 13):
    x = o.fill_hash(n, 3); y = o.fill_hash(n, 4)
    for name in ("copy", "scale", "add", "triad", "vectoradd", "triad_unfused"):
        d = np.zeros(n, np.float32)
        capi.stream_host(getattr(capi, "OP_" + name.upper()), d, x, y, 3.0)
        if d.tobytes() != o.stream(getattr(o, "OP_" + name.upper()), x, y, 3.0).tobytes():
            print("MISMATCH", name, n); ok = False
print("VARIANT_OK" if ok else "VARIANT_BAD")
"""


@pytest.mark.parametrize("variant", ["t8u1b1024", "t4u2b1024", "v8u1b256-oneshot", "clc8u1b1024",
                                     "v8u2b256-persist8", "v4u4b256-persist8", "bulk2048s3b128"])
def test_kernel_variants(gpu, variant):
    env = dict(os.environ, B200V_EW_VARIANT=variant)
    out = subprocess.run([sys.executable, "-c", VARIANT_SCRIPT % {"root": ROOT}], env=env,
                         capture_output=True, text=True, timeout=600)
    assert "VARIANT_OK" in out.stdout, out.stdout + out.stderr


# ---- cfg1: full size, plus size-independent properties ------------------------------------------------

def test_cfg1_full_size_bit_exact(capi, oracle, torch_cuda):
    """2^28 elements per array (1 GiB each) on the device; inputs generated on the device by the
    counter-based generator, regenerated by the oracle on the host; whole output compared."""
    torch = torch_cuda
    n = 1 << 28
    a = torch.empty(n, device="cuda")
    b = torch.empty(n, device="cuda")
    c = torch.empty(n, device="cuda")
    capi.fill_device(a.data_ptr(), n, SEED)
    capi.fill_device(b.data_ptr(), n, SEED + capi.SEED_Y_OFFSET)
    capi.vectoradd_device(a.data_ptr(), b.data_ptr(), c.data_ptr(), n)
    torch.cuda.synchronize()
    threads = os.cpu_count() or 1
    hA = oracle.fill_hash(n, SEED, threads=threads)
    hB = oracle.fill_hash(n, SEED + oracle.SEED_Y_OFFSET, threads=threads)
    want = oracle.vadd(hA, hB, threads=threads)
    got = c.cpu().numpy()
    assert same_bytes(got, want)
    assert oracle.verify_tolerance(hA, hB, got, threads=threads) == -1
    dev_digest = capi.digest_device(c.data_ptr(), n)
    assert dev_digest == oracle.digest(want)
    # properties that need no oracle: commutativity, device re-check, triad(s=1) == add on exact sums
    c2 = torch.empty(n, device="cuda")
    capi.vectoradd_device(b.data_ptr(), a.data_ptr(), c2.data_ptr(), n)
    assert capi.digest_device(c2.data_ptr(), n) == dev_digest
    assert capi.check_device(capi.OP_VECTORADD, c.data_ptr(), a.data_ptr(), b.data_ptr(), 0.0, n)[0] == 0
    capi.stream_device(capi.OP_TRIAD, c2.data_ptr(), a.data_ptr(), b.data_ptr(), 1.0, n)
    assert capi.digest_device(c2.data_ptr(), n) == dev_digest      # fma(1, b, a) == a + b, inputs >= 0
    capi.stream_device(capi.OP_COPY, c2.data_ptr(), c.data_ptr(), None, 0.0, n)
    assert capi.digest_device(c2.data_ptr(), n) == dev_digest


def test_cfg1_host_path_full_size(capi, oracle, gpu):
    """The reference-facing call at full size: host buffers in, host buffer out, 64 chunks."""
    n = 1 << 28
    threads = os.cpu_count() or 1
    pa, pb, pc = capi.PinnedArray(n), capi.PinnedArray(n), capi.PinnedArray(n)
    lib = oracle.lib()
    import ctypes
    f32p = ctypes.POINTER(ctypes.c_float)
    lib.oracle_fill_hash_mt(pa.array.ctypes.data_as(f32p), n, SEED, threads)
    lib.oracle_fill_hash_mt(pb.array.ctypes.data_as(f32p), n, SEED + oracle.SEED_Y_OFFSET, threads)
    t = capi.vectoradd(pa.array, pb.array, pc.array)
    assert t.chunks >= 16 and t.kernel_launches == t.chunks
    want = oracle.vadd(pa.array, pb.array, threads=threads)
    assert same_bytes(pc.array, want)
    for p in (pa, pb, pc):
        p.close()


# ---- argument checking ------------------------------------------------------------------------------

def test_bad_arguments(capi, gpu):
    A = np.ones(8, np.float32)
    with pytest.raises(capi.B200VError) as ei:
        capi.vectoradd(A, A, A, dev=gpu + 5)
    assert ei.value.code == capi.E_BADARG
    with pytest.raises(capi.B200VError) as ei:
        capi.vectoradd(A, None, A)
    assert ei.value.code == capi.E_BADARG
    with pytest.raises(capi.B200VError) as ei:
        capi.stream_host(17, A, A, A)
    assert ei.value.code == capi.E_BADARG
    with pytest.raises(capi.B200VError) as ei:
        capi.stream(capi.OP_COPY, 0)
    assert ei.value.code == capi.E_BADARG
    info = capi.device_info(0)
    assert info.sm_count > 0 and info.l2_bytes > 0 and info.hbm_bytes > 0 and info.cc_major >= 10


# ---- NVLink probe and the real executables ---------------------------------------------------------

def test_allreduce_probe_single_device(capi, gpu):
    """One rank: the all-reduce is the identity, which still exercises communicator set-up, the
    fill/check kernels and the row bookkeeping on a 1-GPU box."""
    assert capi.nccl_version() >= 22000
    rows = capi.allreduce_probe([0], min_bytes=1 << 10, max_bytes=1 << 22, warmup=1, iters=3)
    assert [r.bytes for r in rows] == [1 << k for k in range(10, 23)]
    assert all(r.mismatches == 0 and r.nccl_result == 0 and r.best_ms > 0 for r in rows)
    assert all(r.busbw_gbs == 0.0 for r in rows)            # 2(n-1)/n = 0 at n = 1


def test_allreduce_probe_all_devices(capi, gpu):
    if gpu < 2:
        pytest.skip("needs at least two GPUs")
    rows = capi.allreduce_probe(list(range(gpu)), min_bytes=1 << 10, max_bytes=1 << 26, warmup=2, iters=5)
    assert len(rows) == 17
    assert all(r.mismatches == 0 and r.nccl_result == 0 for r in rows)   # bit-exact vs closed form
    top = rows[-1]
    assert abs(top.busbw_gbs - top.algbw_gbs * 2 * (gpu - 1) / gpu) < 1e-6
    assert top.busbw_gbs > 50.0


def _run_bin(name, tmp_path, **env):
    exe = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "bin", name)
    e = dict(os.environ, OUTPUT_DIR=str(tmp_path))
    e.update({k: str(v) for k, v in env.items()})
    p = subprocess.run([exe], env=e, capture_output=True, text=True, timeout=600)
    return p.returncode, p.stdout, p.stderr


def test_vectoradd_executable_on_gpu(gpu, tmp_path):
    rc, out, err = _run_bin("vectorAdd", tmp_path)
    assert rc == 0, err
    lines = out.splitlines()
    assert lines[0] == "[Vector addition of 50000 elements]" and lines[-2:] == ["Test PASSED", "Done"]
    assert "CUDA kernel launch with 7 blocks of 1024 threads" in lines


def test_validator_executable_fanout_on_gpu(gpu, tmp_path):
    """The drop-in entrypoint as kubelet runs it: every visible GPU validated, status files
    written; then a corrupted element makes it exit 1 and leave no cuda-ready."""
    rc, out, err = _run_bin("nvidia-validator", tmp_path, COMPONENT="cuda")
    assert rc == 0, err
    want = ["cuda-gpu%d-ready" % i for i in range(gpu)] + ["cuda-ready"]
    assert sorted(f for f in os.listdir(tmp_path)) == sorted(want)
    assert out.count("Test PASSED") == gpu
    rc, out, err = _run_bin("nvidia-validator", tmp_path, COMPONENT="cuda", B200V_INJECT_FAULT=31337)
    assert rc == 1 and "Result verification failed at element 31337!" in err
    assert "cuda-ready" not in os.listdir(tmp_path)
    rc, out, err = _run_bin("nvidia-validator", tmp_path, COMPONENT="cuda", B200V_ELEMS=(1 << 24) + 3)
    assert rc == 0 and "cuda-ready" in os.listdir(tmp_path)
