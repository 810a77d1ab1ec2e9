"""e2SEC table dissimilarity: eSEC_analyser.getDissimilarityMatrix / similarity_manipulations.

CPU: the oracle restatement against the vectors the reference's own functions produced (oracle/gen_dissim_golden.py).
GPU: e2sec_dissimilarity (through the drop-in module and the C ABI) against the oracle and those vectors.
Integer results (triples differing in 1 / 2 / 3 components) are bit-exact; D is a float64 sum the reference takes with
numpy's pairwise summation over the per-triple terms, the device forms it from the three integers: tolerance 1e-13 absolute
on values in [0, 1] (written below as TOL).
"""
import os

import numpy as np
import pytest

from oracle import dissim_oracle as orc

TOL = 1e-13
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dissim_reference.npz")
TN = ["U", "T", "N", "A"]
SSR = ["U", "AB", "BE", "AR", "ART", "ABT", "BET", "IN", "SA", "TO", "BO", "BW", "F", "BA", "R", "L"]
DSR = ["U", "MT", "MA", "HT", "FMT", "S", "Q", "GC", "MC"]


def golden_case(name):
    g = np.load(GOLD)
    rows, nc, codes = int(g[name + "_rows"]), g[name + "_ncols"], g[name + "_codes"]
    tabs, p = [], 0
    for n in nc:
        tabs.append(codes[p:p + rows * n].reshape(rows, n))
        p += rows * n
    return g, tabs


def random_tables(seed, opal, col_counts, stickiness=0.7):
    rng = np.random.default_rng(seed)
    vocab = [TN] * opal + [SSR] * opal + [DSR] * opal
    out = []
    for n in col_counts:
        col = [v[rng.integers(len(v))] for v in vocab]
        cols = []
        for _ in range(int(n)):
            col = [c if rng.random() < stickiness else v[rng.integers(len(v))] for c, v in zip(col, vocab)]
            cols.append(list(col))
        out.append(np.array(cols, dtype="<U5").T.copy())
    return out


# ---- CPU: the oracle is pinned to the reference's numbers ------------------------------------------------------------------
@pytest.mark.parametrize("name", ["esec30", "e2sec24"])
def test_oracle_equals_reference_vectors(name):
    g, tabs = golden_case(name)
    D, counts = orc.dissimilarity_matrix(tabs)
    assert np.array_equal(D, g[name + "_D"])                  # bit-exact, same summation order as the reference
    assert np.array_equal(counts, counts.transpose(1, 0, 2))
    if name == "e2sec24":
        sim = np.array([[orc.similarity(a, b) for b in tabs] for a in tabs])
        assert np.array_equal(sim, g[name + "_similarity"])


def test_oracle_properties():
    _, tabs = golden_case("esec30")
    D, counts = orc.dissimilarity_matrix(tabs)
    assert np.all(np.diag(D) == 0.0)
    assert D[1, 10] == 0.0 and not counts[1, 10].any()       # table 10 is a copy of table 1
    assert D[5, 11] == 0.0                                    # table 11 = table 5 + its last column once more
    assert np.all(D >= 0.0) and np.all(D <= 1.0)


# ---- GPU ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["esec30", "e2sec24"])
def test_gpu_equals_reference_vectors(name):
    from manipulationlibrary_b200 import eSEC_analyser as dropin
    from manipulationlibrary_b200.engine import Engine
    g, tabs = golden_case(name)
    D = dropin.getDissimilarityMatrix(tabs)
    assert D.shape == g[name + "_D"].shape
    assert np.max(np.abs(D - g[name + "_D"])) <= TOL
    eng = Engine(0)
    D2, counts = eng.dissimilarity(tabs, divisor=10.0, want_counts=True)
    _, oc = orc.dissimilarity_matrix(tabs)
    assert np.array_equal(counts, oc)                         # the integers are exact
    assert np.array_equal(D, D2)
    if name == "e2sec24":
        ref = g[name + "_similarity"]
        for i in range(len(tabs)):
            for j in range(len(tabs)):
                assert abs(dropin.similarity_manipulations(tabs[i], tabs[j]) - ref[i, j]) <= 100 * TOL
    eng.close()


@pytest.mark.gpu
def test_gpu_dict_input_and_str_dtype():
    """the reference is called with a dict keyed 0..n-1 of unicode arrays (classification_e2sec)"""
    from manipulationlibrary_b200 import eSEC_analyser as dropin
    tabs = random_tables(5, 10, [3, 9, 1, 17])
    D = dropin.getDissimilarityMatrix({k: t for k, t in enumerate(tabs)})
    ref, _ = orc.dissimilarity_matrix(tabs)
    assert np.max(np.abs(D - ref)) <= TOL


@pytest.mark.gpu
def test_gpu_equals_oracle_random_ragged():
    from manipulationlibrary_b200.engine import Engine
    eng = Engine(0)
    for seed, opal, cols in [(1, 10, [1, 1, 2, 40, 33, 7, 64, 5, 12]), (2, 8, [6] * 5 + [23] * 4), (3, 1, [1, 2, 3]),
                             (4, 10, list(range(1, 70, 3)))]:
        tabs = random_tables(seed, opal, cols)
        D, counts = eng.dissimilarity(tabs, divisor=10.0, want_counts=True)
        ref, oc = orc.dissimilarity_matrix(tabs)
        assert np.array_equal(counts, oc)
        assert np.max(np.abs(D - ref)) <= TOL
        assert np.array_equal(D, D.T) and np.all(np.diag(D) == 0.0)
    eng.close()


@pytest.mark.gpu
def test_gpu_trailing_whitespace_and_padding_semantics():
    """chararray comparison ignores trailing whitespace (:877); a table equals itself continued by its last column"""
    from manipulationlibrary_b200.engine import Engine
    eng = Engine(0)
    t = random_tables(9, 10, [6])[0]
    spaced = np.char.add(t, " ").astype("<U5")
    longer = np.column_stack([t, t[:, -1], t[:, -1]])
    D, counts = eng.dissimilarity([t, spaced, longer], want_counts=True)
    assert not D.any() and not counts.any()
    ref, _ = orc.dissimilarity_matrix([t, spaced, longer])
    assert not ref.any()
    eng.close()


@pytest.mark.gpu
def test_gpu_paper_size_properties():
    """120 tables (the paper's setting) and a larger set: symmetry, zero diagonal, range, counts consistent with D, and a
    sample of pairs against the oracle"""
    from manipulationlibrary_b200.engine import Engine
    eng = Engine(0)
    for M, seed in [(120, 11), (1000, 12)]:
        rng = np.random.default_rng(seed)
        tabs = random_tables(seed, 10, rng.integers(4, 31, size=M))
        D, counts = eng.dissimilarity(tabs, want_counts=True)
        assert np.array_equal(D, D.T) and np.all(np.diag(D) == 0.0)
        assert np.all(D >= 0.0) and np.all(D <= 1.0)
        n = np.array([t.shape[1] for t in tabs])
        nmax = np.maximum(n[:, None], n[None, :])
        assert np.all(counts.sum(axis=2) <= nmax * 10)
        a1, a2 = np.sqrt(1.0) / np.sqrt(3.0), np.sqrt(2.0) / np.sqrt(3.0)
        rebuilt = (1.0 / (nmax * 10.0)) * ((counts[..., 0] * a1 + counts[..., 1] * a2) + counts[..., 2])
        assert np.array_equal(D, rebuilt)
        tr = [orc.triples(t) for t in tabs]
        for i, j in rng.integers(0, M, size=(60, 2)):
            oc, _, _ = orc.pair_counts(tr[i], tr[j])
            assert np.array_equal(counts[i, j], oc)
            assert abs(D[i, j] - orc.pair_score(tr[i], tr[j], 10)) <= TOL
    eng.close()


@pytest.mark.gpu
def test_gpu_argument_errors():
    from manipulationlibrary_b200.engine import Engine
    from manipulationlibrary_b200._native import E2SecError
    eng = Engine(0)
    t = random_tables(1, 10, [4])[0]
    with pytest.raises(E2SecError):
        eng.dissimilarity([t, t[:, :0]])                       # no last column to repeat (KeyError in the reference)
    with pytest.raises(ValueError):
        eng.dissimilarity([t, t[:29]])                         # ragged row counts
    with pytest.raises(E2SecError):
        eng.dissimilarity([t[:29], t[:29]])                    # rows not a multiple of 3
    eng.close()
