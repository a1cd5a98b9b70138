"""Synthetic-input generators (frodo-slam_b200/synth.py): the graph shapes bench.py and the
tests are quoted on, and the guard against asking for more loop closures than exist."""
import pytest

from frodo_slam_b200 import synth


@pytest.mark.parametrize("cfg,K,undirected", [("pair", 2, 1), ("window8", 8, 12), ("calib32", 32, 64),
                                               ("loop64", 64, 128), ("loops:8", 8, 16),
                                               ("loops:128", 128, 256)])
def test_graph_edges_shapes(cfg, K, undirected):
    k, und = synth.graph_edges(cfg)
    assert k == K and len(und) == undirected
    assert len(set(und)) == len(und)                 # no duplicate pairs
    assert all(0 <= i < j < K for i, j in und)       # i < j, in range
    assert {(i, i + 1) for i in range(K - 1)} <= set(und) or cfg == "pair"  # the chain is there


def test_graph_edges_is_deterministic():
    assert synth.graph_edges("loop64") == synth.graph_edges("loop64")


def test_impossible_loop_count_is_rejected():
    # "loops:4" asks for 5 closures with a gap >= 2 among 4 keyframes: only 3 pairs exist.  The
    # sampling loop used to spin for ever on this.
    with pytest.raises(ValueError):
        synth.graph_edges("loops:4")


def test_keyframe_graph_fields_cpu():
    K, und = synth.graph_edges("window8")
    g = synth.keyframe_graph(K, 12, 16, und, seed=1)
    E, N = 2 * len(und), 12 * 16
    assert g["Xs"].shape == (K, N, 3) and g["Cs"].shape == (K, N, 1)
    assert g["ii"].shape == (E,) and g["idx_ii2jj"].shape == (E, N)
    assert g["valid_match"].shape == (E, N, 1) and g["Q"].shape == (E, N, 1)
    assert int(g["idx_ii2jj"].min()) >= 0 and int(g["idx_ii2jj"].max()) < N
    assert 0.5 < float(g["valid_match"].float().mean()) < 0.95  # ~80 % of the matches are kept
