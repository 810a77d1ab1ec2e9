"""CPU restatement of the e2SEC dissimilarity of eSEC_analyser.py -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and the CPU-baseline legs of the bench tools may import this module; the product path
(manipulationlibrary_b200.eSEC_analyser -> libe2sec.so) never does.

Pinned: tests/test_dissimilarity.py checks it against tests/golden/dissim_reference.npz, which oracle/gen_dissim_golden.py
produced by executing the reference's own function definitions.
"""
import numpy as np


def _codes(table):
    """The triples live in a chararray of itemsize 7 (eSEC_analyser.py:877): entries are cut to 7 bytes and compared with
    trailing whitespace removed."""
    t = np.asarray(table)
    if t.dtype.kind == "S":
        t = np.char.decode(t, "ascii")
    t = t.astype(str)
    return np.char.rstrip(np.array([[s[:7] for s in row] for row in t], dtype=object).astype(str))


def triples(table):
    """_make_triple (:866-889) / _make_triple_one_manipualtion (:891-913): per column, opal rows of
    (row i, row i + opal, row i + 2 opal).  Returns an array [n_cols][opal][3]."""
    t = _codes(table)
    opal = int(t.shape[0] / 3)
    return np.stack([t[0:opal].T, t[opal:2 * opal].T, t[2 * opal:3 * opal].T], axis=2)


def pad_pair(a, b):
    """_compare_triple_matrics (:915-944): the table with fewer columns is continued by its own last column."""
    n = max(len(a), len(b))
    if len(a) < n:
        a = np.concatenate([a, np.repeat(a[-1:], n - len(a), axis=0)])
    if len(b) < n:
        b = np.concatenate([b, np.repeat(b[-1:], n - len(b), axis=0)])
    return a, b


def pair_counts(ta, tb):
    """number of triples that differ in exactly 1, 2, 3 components, and the padded column count"""
    a, b = pad_pair(ta, tb)
    l = (a != b).sum(axis=2)                                  # L1 + L2 + L3 of :831-836
    return np.array([(l == 1).sum(), (l == 2).sum(), (l == 3).sum()], np.uint32), len(a), l


def pair_score(ta, tb, divisor):
    """:837-838 (divisor 10) and :859-860 (divisor 8), with the reference's own expression for the sum"""
    _, n, l = pair_counts(ta, tb)
    # the reference's L arrays are transposed (n, opal) arrays, i.e. column-major (opal, n): np.sum walks them column by column
    d = (np.sqrt(np.asfortranarray(l.T))) / (np.sqrt(3))
    return (1 / (n * divisor)) * np.sum(d)


def dissimilarity_matrix(tables, divisor=10):
    """getDissimilarityMatrix (:782-792) -> _calc_D_shaped (:819-841).  Returns (D, counts)."""
    tr = [triples(t) for t in tables]
    M = len(tr)
    D = np.zeros((M, M))
    counts = np.zeros((M, M, 3), np.uint32)
    for i in range(M):
        for j in range(M):
            counts[i, j], _, _ = pair_counts(tr[i], tr[j])
            D[i, j] = pair_score(tr[i], tr[j], divisor)
    return D, counts


def similarity(table_1, table_2):
    """similarity_manipulations (:843-864)"""
    D = pair_score(triples(table_1), triples(table_2), 8)
    return (1 - D) * 100
