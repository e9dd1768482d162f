"""Regenerates tests/golden/golden.json from the CPU oracle.

The reference repository has no code and no test vectors on this path, so these
are SELF-DERIVED regression pins of the restated algorithm, not vectors lifted from the reference:
  * the sample's glibc-rand() inputs and the vectorAdd output at n = 50 000 (values that SURVEY.md
    §8c / BASELINE.md §5 also record),
  * the counter-based generator, every elementwise op and both digests at ragged sizes.
tests/test_oracle.py additionally re-derives the rand() stream with an independent pure-Python
restatement of glibc's TYPE_3 generator, so the pins do not depend on the libc of the machine
that ran this script.

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as o  # noqa: E402

SEED = 0xB200
SCALAR = 3.0


def bits(a):
    return [int(v) for v in np.asarray(a, np.float32).view(np.uint32)]


def main():
    out = {"note": "self-derived pins of the CPU oracle; the reference ships no vectors (SURVEY.md §8c)"}
    n = 50000
    A, B = o.fill_rand(n)
    C = o.vadd(A, B)
    out["sample"] = {
        "n": n,
        "A_head_bits": bits(A[:8]), "B_head_bits": bits(B[:8]), "C_head_bits": bits(C[:8]),
        "A_tail_bits": bits(A[-4:]), "B_tail_bits": bits(B[-4:]), "C_tail_bits": bits(C[-4:]),
        "A0": "%.9g" % A[0], "B0": "%.9g" % B[0], "C0": "%.9g" % C[0], "A_last": "%.9g" % A[-1],
        "C_fnv1a64_words": "%016x" % o.fnv1a64_words(C),
        "A_fnv1a64_words": "%016x" % o.fnv1a64_words(A),
        "B_fnv1a64_words": "%016x" % o.fnv1a64_words(B),
        "C_digest": "%016x" % o.digest(C),
        "tolerance_first_bad": o.verify_tolerance(A, B, C),
    }
    out["hash_fill"] = {"seed": SEED, "head_bits": bits(o.fill_hash(8, SEED)),
                        "head_bits_seed_y": bits(o.fill_hash(8, SEED + o.SEED_Y_OFFSET))}
    ops = {}
    for size in (1, 5, 257, 4099, (1 << 20) + 1):
        x = o.fill_hash(size, SEED)
        y = o.fill_hash(size, SEED + o.SEED_Y_OFFSET)
        row = {}
        for op, name in ((o.OP_COPY, "copy"), (o.OP_SCALE, "scale"), (o.OP_ADD, "add"),
                         (o.OP_TRIAD, "triad"), (o.OP_VECTORADD, "vectoradd"),
                         (o.OP_TRIAD_UNFUSED, "triad_unfused")):
            d = o.stream(op, x, y, SCALAR)
            row[name] = {"digest": "%016x" % o.digest(d), "fnv1a64_words": "%016x" % o.fnv1a64_words(d),
                         "head_bits": bits(d[:4])}
        ops[str(size)] = row
    out["ops"] = {"seed_x": SEED, "seed_y": SEED + o.SEED_Y_OFFSET, "scalar": SCALAR, "sizes": ops}
    out["allreduce"] = {str(nr): bits(o.allreduce_expected(16, nr)) for nr in (1, 2, 4, 8)}
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", os.path.join(HERE, "golden.json"))


if __name__ == "__main__":
    main()
