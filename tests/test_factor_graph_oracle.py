"""Pins oracle/factor_graph_oracle.py against tests/golden/factor_graph.npz, the output of the reference's own
FactorGraph.add_factors / prep_two_way_edges / get_unique_kf_idx / constrain_points_to_ray python
(tests/golden/make_golden_factor_graph.py).  CPU only; bit-exact."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import factor_graph_oracle as fgo  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden", "factor_graph.npz")
MATCH_KEYS = ("idx_i2j", "idx_j2i", "valid_j", "valid_i", "Qii", "Qjj", "Qji", "Qij")


def q_close(a, b, name):
    """The fixture ran the reference on the CPU, where torch.sqrt goes through MKL's VML: <= 1 ulp but not
    correctly rounded (about 0.5 % of the values differ from IEEE sqrt).  The reference's production path is
    CUDA (sqrt.rn.f32, correctly rounded) — that is what the oracle and the kernel reproduce, and what
    tests/test_gpu_factor_graph.py compares bit-for-bit against torch on the GPU."""
    np.testing.assert_array_max_ulp(a, b, maxulp=1)
    assert (a != b).mean() < 0.02, name


def replay(g, fn):
    """feeds the fixture's batches to fn(t, ii, jj, matches, min_match_frac, is_reloc) -> ret"""
    for t in range(int(g["num_batches"])):
        m = tuple(g[f"b{t}_{k}"] for k in MATCH_KEYS)
        ret = fn(t, g[f"b{t}_ii"], g[f"b{t}_jj"], m, float(g[f"b{t}_min_match_frac"]), bool(g[f"b{t}_is_reloc"]))
        assert bool(ret) == bool(g[f"b{t}_ret"]), f"batch {t}: return value"


def test_add_factors_tail_matches_reference_python():
    g = np.load(GOLD)
    st = fgo.GraphState(int(g["n"]))

    def step(t, ii, jj, m, mmf, reloc):
        r = fgo.add_factors_tail(st, ii, jj, m, 1.5, mmf, reloc)
        assert st.ii.shape[0] == int(g[f"b{t}_num_edges_after"])
        return r

    replay(g, step)
    for name in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
        a, b = getattr(st, name), g[f"state_{name}"]
        assert a.shape == b.shape and a.dtype == b.dtype, name
        if name.startswith("Q_"):
            q_close(a, b, name)
        else:
            assert np.array_equal(a, b), name
    assert np.array_equal(fgo.get_unique_kf_idx(st), g["unique_kf_idx"])
    for name, t in zip(("ii", "jj", "idx", "valid", "Q"), fgo.prep_two_way_edges(st)):
        if name == "Q":
            q_close(t, g["twoway_Q"], name)
        else:
            assert np.array_equal(t, g[f"twoway_{name}"]), name


def test_oracle_q_is_the_ieee_value():
    """the oracle's float32 mul + sqrt equals the correctly rounded result (what sqrt.rn.f32 gives on the GPU)"""
    g = np.load(GOLD)
    for t in range(int(g["num_batches"])):
        Qii, idx, Qji = g[f"b{t}_Qii"], g[f"b{t}_idx_i2j"], g[f"b{t}_Qji"]
        prod = Qii[np.arange(idx.shape[0])[:, None], idx] * Qji  # float32 product
        exact = np.sqrt(prod.astype(np.float64)).astype(np.float32)  # innocuous double rounding: 53 >= 2*24+2
        assert np.array_equal(fgo.edge_q(Qii, idx, Qji), exact)


def test_fixture_covers_the_branches():
    g = np.load(GOLD)
    rets = [bool(g[f"b{t}_ret"]) for t in range(int(g["num_batches"]))]
    after = [int(g[f"b{t}_num_edges_after"]) for t in range(int(g["num_batches"]))]
    sizes = [len(g[f"b{t}_ii"]) for t in range(int(g["num_batches"]))]
    assert False in rets and True in rets                      # the is_reloc early return is exercised
    assert any(a - b < s for a, b, s in zip(after, [0] + after[:-1], sizes))  # some pair was rejected


def test_interleave_is_a_permutation_of_the_reference_order():
    g = np.load(GOLD)
    two = tuple(g[f"twoway_{k}"] for k in ("ii", "jj", "idx", "valid", "Q"))
    inter = fgo.interleave_two_way(two)
    E = two[0].shape[0] // 2
    for a, b in zip(two, inter):
        assert np.array_equal(b[0::2], a[:E]) and np.array_equal(b[1::2], a[E:])
    # both directions of a pair are adjacent and mirror each other
    assert np.array_equal(inter[0][0::2], inter[1][1::2]) and np.array_equal(inter[1][0::2], inter[0][1::2])


def test_constrain_points_to_ray_matches_reference_python():
    g = np.load(GOLD)
    out = fgo.constrain_points_to_ray((int(g["ray_h"]), int(g["ray_w"])), g["ray_Xs"], g["ray_K"])
    assert out.dtype == np.float32 and np.array_equal(out, g["ray_out"])


@pytest.mark.parametrize("shape", [(2, 6, 8, 3), (1, 48, 3)])
def test_constrain_points_to_ray_shapes(shape):
    rng = np.random.default_rng(3)
    X = rng.normal(size=shape).astype(np.float32)
    K = np.array([[9.0, 0, 3.5], [0, 8.0, 2.5], [0, 0, 1]], np.float32)
    out = fgo.constrain_points_to_ray((6, 8), X, K)
    assert out.shape == X.shape
    assert np.array_equal(out[..., 2], X[..., 2])
