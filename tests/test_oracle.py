"""CPU tests of the oracle: pins against the committed golden vectors, against SURVEY.md §8(c)'s
recorded values, and against independent restatements (numpy, and a pure-Python glibc rand())."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return [int(v) for v in np.asarray(a, np.float32).view(np.uint32)]


def glibc_rand_stream(count, seed=1):
    """Independent restatement of glibc's default rand() (TYPE_3 additive feedback generator,
    degree 31, separation 3; published algorithm of random_r.c): does not call libc."""
    r = [0] * (344 + count)
    r[0] = seed
    for i in range(1, 31):
        hi, lo = divmod(r[i - 1], 127773)
        word = 16807 * lo - 2836 * hi
        if word < 0:
            word += 2147483647
        r[i] = word
    for i in range(31, 34):
        r[i] = r[i - 31]
    for i in range(34, 344 + count):
        r[i] = (r[i - 31] + r[i - 3]) & 0xFFFFFFFF
    return [v >> 1 for v in r[344:]]


def test_sample_inputs_match_independent_glibc_rand(oracle):
    n = 2000
    A, B = oracle.fill_rand(n)
    stream = glibc_rand_stream(2 * n)
    assert stream[0] == 1804289383          # the well-known first rand() of an unseeded glibc program
    rmax = np.float32(2147483647)
    # `rand() / (float)RAND_MAX`: int converted to float, divided in float
    wantA = (np.array(stream[0::2], dtype=np.int64).astype(np.float32) / rmax).astype(np.float32)
    wantB = (np.array(stream[1::2], dtype=np.int64).astype(np.float32) / rmax).astype(np.float32)
    assert bits(A) == bits(wantA)
    assert bits(B) == bits(wantB)


def test_sample_pins_from_survey(oracle):
    """SURVEY.md §8(c) / BASELINE.md §5 value pins (the survey's FNV figure could not be
    reproduced from its own stated fold and is superseded by golden.json)."""
    A, B = oracle.fill_rand(50000)
    C = oracle.vadd(A, B)
    assert "%.9g" % A[0] == "0.840187728"
    assert "%.9g" % B[0] == "0.394382924"
    assert "%.9g" % C[0] == "1.23457062"
    assert "%.9g" % A[49999] == "0.664940238"
    assert oracle.verify_tolerance(A, B, C) == -1


def test_sample_golden(oracle, golden):
    g = golden["sample"]
    A, B = oracle.fill_rand(g["n"])
    C = oracle.vadd(A, B)
    assert bits(A[:8]) == g["A_head_bits"] and bits(A[-4:]) == g["A_tail_bits"]
    assert bits(B[:8]) == g["B_head_bits"] and bits(B[-4:]) == g["B_tail_bits"]
    assert bits(C[:8]) == g["C_head_bits"] and bits(C[-4:]) == g["C_tail_bits"]
    assert "%016x" % oracle.fnv1a64_words(C) == g["C_fnv1a64_words"]
    assert "%016x" % oracle.fnv1a64_words(A) == g["A_fnv1a64_words"]
    assert "%016x" % oracle.fnv1a64_words(B) == g["B_fnv1a64_words"]
    assert "%016x" % oracle.digest(C) == g["C_digest"]
    assert oracle.verify_tolerance(A, B, C) == g["tolerance_first_bad"] == -1


def test_fill_rand_is_repeatable(oracle):
    a1, b1 = oracle.fill_rand(1000)
    a2, b2 = oracle.fill_rand(1000)
    assert a1.tobytes() == a2.tobytes() and b1.tobytes() == b2.tobytes()


def test_vadd_is_plain_ieee_add(oracle):
    A, B = oracle.fill_rand(50000)
    C = oracle.vadd(A, B)
    assert C.tobytes() == (A + B).astype(np.float32).tobytes()


def test_vadd_plus_zero_semantics(oracle):
    """The sample's `+ 0.0f` maps a -0.0f sum to +0.0f and leaves everything else alone."""
    A = np.array([-0.0, -0.0, 0.0, 1.5, -1.5, np.inf, 1e-45, -1e-45], np.float32)
    B = np.array([-0.0, 0.0, -0.0, -1.5, 1.5, 1.0, 1e-45, 1e-45], np.float32)
    C = oracle.vadd(A, B)
    assert bits(C) == bits(np.array([0.0, 0.0, 0.0, 0.0, 0.0, np.inf, 3e-45, 0.0], np.float32))
    # plain ADD keeps the negative zero
    D = oracle.stream(oracle.OP_ADD, A, B)
    assert bits(D[:1]) == [0x80000000]


def test_verify_tolerance_reports_first_bad_index(oracle):
    A, B = oracle.fill_rand(4096)
    C = oracle.vadd(A, B)
    assert oracle.verify_tolerance(A, B, C) == -1
    C[1234] += 1e-3
    C[3000] += 1.0
    assert oracle.verify_tolerance(A, B, C) == 1234
    assert oracle.verify_tolerance(A, B, C, threads=4) == 1234
    C2 = oracle.vadd(A, B)
    C2[7] = np.float32(C2[7]) + np.float32(5e-6)      # inside the sample's 1e-5 tolerance
    assert oracle.verify_tolerance(A, B, C2) == -1


def test_fill_hash_matches_numpy_and_golden(oracle, golden):
    g = golden["hash_fill"]
    for n in (0, 1, 7, 1000):
        assert oracle.fill_hash(n, g["seed"]).tobytes() == oracle.np_fill_hash(n, g["seed"]).tobytes()
    assert bits(oracle.fill_hash(8, g["seed"])) == g["head_bits"]
    assert bits(oracle.fill_hash(8, g["seed"] + oracle.SEED_Y_OFFSET)) == g["head_bits_seed_y"]
    x = oracle.fill_hash(100000, g["seed"])
    assert x.min() >= 0.0 and x.max() < 1.0
    assert oracle.fill_hash(100000, g["seed"], threads=3).tobytes() == x.tobytes()


def test_splitmix64_known_answers(oracle):
    # first outputs of the published splitmix64 with state 0 (Vigna's reference values)
    lib = oracle.lib()
    assert lib.oracle_splitmix64(0) == 0xE220A8397B1DCDAF
    assert lib.oracle_splitmix64(0x9E3779B97F4A7C15) == 0x6E789E6AA1B965F4


@pytest.mark.parametrize("size", [1, 5, 257, 4099, (1 << 20) + 1])
def test_ops_golden(oracle, golden, size):
    g = golden["ops"]
    x = oracle.fill_hash(size, g["seed_x"])
    y = oracle.fill_hash(size, g["seed_y"])
    s = np.float32(g["scalar"])
    want = {
        "copy": x,
        "scale": (s * x).astype(np.float32),
        "add": (x + y).astype(np.float32),
        "vectoradd": (x + y).astype(np.float32),          # inputs are >= 0: + 0.0f is invisible
        "triad_unfused": (x + (s * y).astype(np.float32)).astype(np.float32),
    }
    ids = {"copy": oracle.OP_COPY, "scale": oracle.OP_SCALE, "add": oracle.OP_ADD,
           "triad": oracle.OP_TRIAD, "vectoradd": oracle.OP_VECTORADD,
           "triad_unfused": oracle.OP_TRIAD_UNFUSED}
    for name, op in ids.items():
        d = oracle.stream(op, x, y, float(s))
        row = g["sizes"][str(size)][name]
        assert "%016x" % oracle.digest(d) == row["digest"], name
        assert "%016x" % oracle.fnv1a64_words(d) == row["fnv1a64_words"], name
        assert bits(d[:4]) == row["head_bits"], name
        if name in want:
            assert d.tobytes() == want[name].tobytes(), name
        assert oracle.stream(op, x, y, float(s), threads=3).tobytes() == d.tobytes(), name


def test_triad_is_single_rounding(oracle):
    """fmaf(s, y, x) equals the float64 evaluation rounded once (exact for fp32 operands: the
    product of two fp32 has 48 significant bits and the float64 sum of it with an fp32 of similar
    magnitude is exact here); the unfused form differs on some elements."""
    x = oracle.fill_hash(100000, 1)
    y = oracle.fill_hash(100000, 2)
    fused = oracle.stream(oracle.OP_TRIAD, x, y, 3.0)
    ref = (x.astype(np.float64) + 3.0 * y.astype(np.float64)).astype(np.float32)
    assert fused.tobytes() == ref.tobytes()
    unfused = oracle.stream(oracle.OP_TRIAD_UNFUSED, x, y, 3.0)
    assert (unfused != fused).any()


def test_digest_matches_numpy_and_is_position_sensitive(oracle):
    x = oracle.fill_hash(5000, 7)
    assert oracle.digest(x) == oracle.np_digest(x)
    y = x.copy()
    y[[10, 11]] = y[[11, 10]]
    assert oracle.digest(y) != oracle.digest(x)
    assert oracle.digest(x[:0]) == 0
    # sum structure: digest of a concatenation is not the sum of digests (index enters the mix)
    assert (oracle.digest(x[:2500]) + oracle.digest(x[2500:])) % 2**64 != oracle.digest(x)


def test_fnv_known_answer(oracle):
    # FNV-1a-64 of the single 32-bit word 0 folded as one unit: (basis ^ 0) * prime
    z = np.zeros(1, np.uint32)
    assert oracle.fnv1a64_words(z) == (0xcbf29ce484222325 * 0x100000001b3) % 2**64
    assert oracle.fnv1a64_words(z[:0]) == 0xcbf29ce484222325


def test_allreduce_closed_form(oracle, golden):
    for nr, want in golden["allreduce"].items():
        got = oracle.allreduce_expected(16, int(nr))
        assert bits(got) == want
        lib = oracle.lib()
        assert [lib.oracle_allreduce_expected(i, int(nr)) for i in range(16)] == list(got)
    # the pattern must depend on i at every rank count (the reason (i+r)&7 was not used)
    for nr in (2, 4, 8):
        assert len(set(oracle.allreduce_expected(64, nr).tolist())) > 2


def test_digest_of_pieces_adds_up(oracle):
    x = oracle.fill_hash(10007, 3)
    for cut in (0, 1, 4000, 10006, 10007):
        assert (oracle.digest_from(x[:cut], 0) + oracle.digest_from(x[cut:], cut)) % 2**64 == oracle.digest(x)
    assert oracle.digest_from(x, 5) != oracle.digest(x)          # the index is part of the digest


def test_threaded_vadd_matches_scalar(oracle):
    A = oracle.fill_hash(100003, 11)
    B = oracle.fill_hash(100003, 12)
    assert oracle.vadd(A, B, threads=5).tobytes() == oracle.vadd(A, B).tobytes()


def test_committed_sweep_digests_are_what_the_oracle_computes(oracle):
    """tests/golden/sweep_digests.json (bench.py's checker for the 2^30 and 2^32 rows of the STREAM
    sweep) against the live oracle at the sizes that take a moment on the CPU, through the same
    piecewise routine bench.py uses — so the committed file, the generator script and the bench's
    own digest arithmetic cannot drift apart."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    with open(os.path.join(ROOT, "tests", "golden", "sweep_digests.json")) as f:
        golden = json.load(f)
    assert golden["seed"] == bench.SEED and golden["scalar"] == 3.0
    assert {int(k) for k in golden["log2_n"]} >= set(bench.SWEEP_LOG2)
    ops = [oracle.OP_COPY, oracle.OP_SCALE, oracle.OP_ADD, oracle.OP_TRIAD]
    for lg in (18, 22, 24, 25):                      # below, at and above the 2^24-element piece size
        if str(lg) not in golden["log2_n"]:
            continue
        got = bench.oracle_sweep_digests(oracle, lg, ops, 3.0)
        want = [int(golden["log2_n"][str(lg)][name], 16) for name in ("copy", "scale", "add", "triad")]
        assert got == want, lg
    # the digest is position-sensitive: the same values one element later do not collide
    x = oracle.fill_hash(1000, bench.SEED)
    assert oracle.digest_from(x, 0) != oracle.digest_from(x, 1)
