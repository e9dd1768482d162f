"""GPU test of the hand-written NVLink-multicast all-reduce probe: bit-exact
against the closed form at every size, and in agreement with the NCCL probe's row format.
Needs at least two GPUs with multicast (NVSwitch) support; skipped otherwise."""
import pytest

pytestmark = pytest.mark.gpu


def test_nvls_allreduce_probe_bit_exact(capi, gpu):
    if gpu < 2:
        pytest.skip("needs at least two GPUs")
    try:
        rows = capi.nvls_allreduce_probe(list(range(gpu)), min_bytes=1 << 10, max_bytes=1 << 26, warmup=2, iters=5)
    except capi.B200VError as e:
        if e.code == capi.E_UNSUPPORTED:
            pytest.skip("no multicast support on this box")
        raise
    assert [r.bytes for r in rows] == [1 << k for k in range(10, 27)]
    assert all(r.mismatches == 0 for r in rows)              # exact sum, every element, every GPU
    top = rows[-1]
    assert top.best_ms > 0 and top.best_ms <= top.median_ms
    assert abs(top.busbw_gbs - top.algbw_gbs * 2 * (gpu - 1) / gpu) < 1e-6
    assert top.busbw_gbs > 50.0


def test_nvls_allreduce_ragged_slices_and_both_algorithms(capi, gpu, monkeypatch):
    """Sizes whose float4 count is not a multiple of the rank count (63 * 2^k vectors), so the last
    GPU's slice is short or empty, on the multimem kernel and on the peer-to-peer kernel."""
    if gpu < 2:
        pytest.skip("needs at least two GPUs")
    ranks = 2 if gpu < 4 else (4 if gpu < 8 else 8)
    for algo in ("multimem", "p2p", "push", "oneshot", "hybrid"):
        monkeypatch.setenv("B200V_NVLS_ALGO", algo)
        try:
            rows = capi.nvls_allreduce_probe(list(range(ranks)), min_bytes=63 * 16, max_bytes=63 * 16 << 12,
                                             warmup=1, iters=3)
        except capi.B200VError as e:
            if e.code == capi.E_UNSUPPORTED:
                pytest.skip("no multicast support on this box")
            raise
        assert [r.bytes for r in rows] == [63 * 16 << k for k in range(13)]
        assert all(r.mismatches == 0 for r in rows), algo
    # a tiny message: fewer vectors than ranks, most slices empty
    rows = capi.nvls_allreduce_probe(list(range(ranks)), min_bytes=16, max_bytes=64, warmup=1, iters=2)
    assert [r.bytes for r in rows] == [16, 32, 64] and all(r.mismatches == 0 for r in rows)


def _rank_worker(rank, nranks, name, out_dir):
    import json
    import os

    import __graft_entry__ as entry
    c = entry.load_package().capi
    try:
        rows = c.nvls_allreduce_probe_rank(rank, rank, nranks, name, min_bytes=1 << 12, max_bytes=1 << 24,
                                           warmup=1, iters=3)
        res = {"rows": [[int(r.bytes), int(r.mismatches), float(r.best_ms), float(r.busbw_gbs)] for r in rows]}
    except c.B200VError as e:
        res = {"error": e.code}
    with open(os.path.join(out_dir, "r%d.json" % rank), "w") as f:
        json.dump(res, f)


def test_nvls_probe_one_process_per_gpu(capi, gpu, tmp_path):
    """The per-rank form: every process drives one GPU, rank 0 hands the multicast object to the
    others as a file descriptor over the rendezvous socket."""
    if gpu < 2:
        pytest.skip("needs at least two GPUs")
    import json
    import os

    import torch.multiprocessing as mp
    nranks = 2 if gpu < 4 else (4 if gpu < 8 else 8)
    name = "b200v-nvls-test-%d" % os.getpid()
    mp.spawn(_rank_worker, args=(nranks, name, str(tmp_path)), nprocs=nranks, join=True)
    results = [json.load(open(tmp_path / ("r%d.json" % r))) for r in range(nranks)]
    if any(r.get("error") == capi.E_UNSUPPORTED for r in results):
        pytest.skip("no multicast support on this box")
    assert all("rows" in r for r in results), results
    for r in results:
        assert [row[0] for row in r["rows"]] == [1 << k for k in range(12, 25)]
        assert all(row[1] == 0 for row in r["rows"])          # bit-exact on every rank
        assert r["rows"][-1][3] > 20.0


def test_nvls_probe_argument_checks(capi, gpu):
    with pytest.raises(capi.B200VError) as ei:
        capi.nvls_allreduce_probe([0], min_bytes=1024, max_bytes=2048)
    assert ei.value.code == capi.E_BADARG                    # one GPU is not a collective
    with pytest.raises(capi.B200VError) as ei:
        capi.nvls_allreduce_probe([0, 0], min_bytes=1024, max_bytes=2048)
    assert ei.value.code == capi.E_BADARG                    # two ranks on one GPU would wait on each other
    with pytest.raises(capi.B200VError) as ei:
        capi.nvls_allreduce_probe([0, gpu + 3], min_bytes=1024, max_bytes=2048)
    assert ei.value.code == capi.E_BADARG
    with pytest.raises(capi.B200VError) as ei:
        capi.nvls_allreduce_probe([0, 0], min_bytes=8, max_bytes=2048)
    assert ei.value.code == capi.E_BADARG                    # sizes are multiples of 16 bytes
