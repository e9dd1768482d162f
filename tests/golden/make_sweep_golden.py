"""Regenerates tests/golden/sweep_digests.json: the b200v digest of the output of every STREAM op
(copy, scale, add, triad; scalar 3.0; seed 0xB200 — what b200v_stream and `B200V_MODE=stream` use)
at the sizes bench.py's `stream_sweep` reports, 2^18 ... 2^32 elements per array.

The digest is a sum over elements, so the arrays are produced and folded in 2^24-element pieces
(piece k starts at seed + k*2^24) and never exist whole: 2^32 elements need no 16 GiB buffer.
SELF-DERIVED from the CPU oracle (the reference ships no vectors); bench.py compares
the device's digests with these, and with the live oracle at the sizes that take it under a second.

    python tests/golden/make_sweep_golden.py        (minutes on 8 cores)
"""
import concurrent.futures
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as o  # noqa: E402

SEED = 0xB200
SCALAR = 3.0
PIECE = 1 << 24
LOG2_SIZES = (18, 20, 22, 24, 26, 28, 30, 32)
OPS = ((o.OP_COPY, "copy"), (o.OP_SCALE, "scale"), (o.OP_ADD, "add"), (o.OP_TRIAD, "triad"))


def piece_digests(lo, n):
    x = o.fill_hash(n, SEED + lo)
    y = o.fill_hash(n, SEED + o.SEED_Y_OFFSET + lo)
    return [o.digest_from(o.stream(op, x, y, SCALAR), lo) for op, _name in OPS]


def main():
    top = 1 << max(LOG2_SIZES)
    pieces = [(lo, min(PIECE, top - lo)) for lo in range(0, top, PIECE)]
    small = [lg for lg in LOG2_SIZES if (1 << lg) < PIECE]
    out = {"seed": SEED, "scalar": SCALAR, "note": "self-derived from the CPU oracle, piecewise", "log2_n": {}}
    for lg in small:
        out["log2_n"][str(lg)] = {name: "%016x" % d for (_op, name), d in zip(OPS, piece_digests(0, 1 << lg))}
    # prefix sums over the pieces give every power-of-two size >= PIECE in one pass
    acc = [0, 0, 0, 0]
    done = 0
    with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as pool:
        for (lo, n), ds in zip(pieces, pool.map(lambda p: piece_digests(*p), pieces)):
            acc = [(a + d) % 2**64 for a, d in zip(acc, ds)]
            done = lo + n
            if done & (done - 1) == 0 and done.bit_length() - 1 in LOG2_SIZES:
                out["log2_n"][str(done.bit_length() - 1)] = {name: "%016x" % a for (_op, name), a in zip(OPS, acc)}
                print("2^%d done" % (done.bit_length() - 1), flush=True)
    with open(os.path.join(HERE, "sweep_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote sweep_digests.json")


if __name__ == "__main__":
    main()
