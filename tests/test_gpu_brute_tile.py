"""The tensor-core leaf tile (csrc/brute_tile.cu: TMA loads, tcgen05.mma kind::tf32 into TMEM, tcgen05.ld) as a
conservative filter: on the same inputs as the exact CUDA-core sweep it must never drop a true candidate, and it
must not let many more through. Sizes that are not multiples of the tile sizes exercise the out-of-bounds handling."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nq,np_,k", [(4096, 2048, 100), (1000, 700, 32), (129, 31, 10)])
def test_tensor_core_filter_is_conservative(nq, np_, k):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
    from tools.brute_tile_bench import run
    r = run(nq, np_, k, reps=1, sample=nq)
    assert r["true_candidates_dropped_by_filter"] == 0, r
    assert r["candidates_filter"] >= r["candidates_exact"], r
    assert r["filter_inflation"] < 1.5, r
    assert r["cuda_counts_vs_torch_max_abs_diff"] <= 8, r        # torch.cdist rounds differently at the threshold
