"""GPU parity at the largest configured size and past both 32-bit boundaries: 2^32 + 7 elements
per array (16 GiB each; BASELINE.json configs[2] tops out at 16 GiB per array).  The sample's
kernel indexes with `int`; every kernel here indexes with size_t, and this is the test that
would catch a 2^31 or 2^32 wrap anywhere in the launch arithmetic.

The oracle cannot hold 48 GiB comfortably, so it works in 2^26-element pieces: inputs come from
the counter-based generator (piece k starts at seed + k*2^26), and the digest is a sum over
elements, so the pieces' digests add up to the digest the device computes over the whole array."""
import concurrent.futures
import os

import pytest

pytestmark = pytest.mark.gpu

SEED = 0xB200
N = (1 << 32) + 7
PIECE = 1 << 26


def test_vectoradd_past_the_32_bit_boundaries(capi, oracle, gpu):
    import torch
    free, _total = torch.cuda.mem_get_info(0)
    if free < 3 * 4 * N + (2 << 30):
        pytest.skip("needs 50 GiB of free device memory")
    a = torch.empty(N, dtype=torch.float32, device="cuda")
    b = torch.empty(N, dtype=torch.float32, device="cuda")
    c = torch.empty(N, dtype=torch.float32, device="cuda")
    capi.fill_device(a.data_ptr(), N, SEED)
    capi.fill_device(b.data_ptr(), N, SEED + capi.SEED_Y_OFFSET)
    capi.vectoradd_device(a.data_ptr(), b.data_ptr(), c.data_ptr(), N)
    torch.cuda.synchronize()

    def piece(k):
        lo = k * PIECE
        n = min(PIECE, N - lo)
        x = oracle.fill_hash(n, SEED + lo)
        y = oracle.fill_hash(n, SEED + oracle.SEED_Y_OFFSET + lo)
        z = oracle.vadd(x, y)
        return oracle.digest_from(x, lo), oracle.digest_from(z, lo)

    pieces = (N + PIECE - 1) // PIECE
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as pool:
        parts = list(pool.map(piece, range(pieces)))
    want_a = sum(p[0] for p in parts) % 2**64
    want_c = sum(p[1] for p in parts) % 2**64

    assert capi.digest_device(a.data_ptr(), N) == want_a          # the generator, past 2^32
    assert capi.digest_device(c.data_ptr(), N) == want_c          # the kernel, past 2^32
    assert capi.check_device(capi.OP_VECTORADD, c.data_ptr(), a.data_ptr(), b.data_ptr(), 0.0, N)[0] == 0
    # the last elements (indices >= 2^32) against the oracle, bytes and all
    tail = 7 + 1024
    x = oracle.fill_hash(tail, SEED + N - tail)
    y = oracle.fill_hash(tail, SEED + oracle.SEED_Y_OFFSET + N - tail)
    assert c[N - tail:].cpu().numpy().tobytes() == oracle.vadd(x, y).tobytes()
    # the other ops reuse the same launch arithmetic: triad with s = 1 equals the add bit for bit
    capi.stream_device(capi.OP_TRIAD, c.data_ptr(), a.data_ptr(), b.data_ptr(), 1.0, N)
    assert capi.digest_device(c.data_ptr(), N) == want_c
