"""Generates tests/golden/dissim_reference.npz by running the REFERENCE's own functions.

TEST INFRASTRUCTURE, generation time only (needs /root/reference; never runs on the GPU box).

eSEC_analyser.py cannot be imported here (tabula, seaborn are absent), and none of the functions on this path need those
modules: the function definitions getDissimilarityMatrix, _calc_D_shaped, _make_triple, _make_triple_one_manipualtion,
_compare_triple_matrics and similarity_manipulations are taken out of the reference file's syntax tree and executed,
unmodified, in a namespace that holds numpy and copy. Nothing of the reference is written into this repository except the
numbers it computes.

    python -m oracle.gen_dissim_golden
"""
import ast
import copy
import os

import numpy as np

REF = "/root/reference/eSEC_analyser.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "dissim_reference.npz")
WANTED = {"getDissimilarityMatrix", "_calc_D_shaped", "_make_triple", "_make_triple_one_manipualtion",
          "_compare_triple_matrics", "similarity_manipulations"}

# relation codes the e2SEC tables hold after esec_to_e2sec (upper case): T/N, SSR and DSR vocabularies of the paper
TN = ["U", "T", "N", "A"]
SSR = ["U", "AB", "BE", "AR", "ART", "ABT", "BET", "IN", "SA", "TO", "BO", "BW", "F", "BA", "R", "L"]
DSR = ["U", "MT", "MA", "HT", "FMT", "S", "Q", "GC", "MC"]


def reference_namespace():
    tree = ast.parse(open(REF).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in WANTED]
    assert {n.name for n in keep} == WANTED
    ns = {"np": np, "copy": copy}
    exec(compile(ast.Module(body=keep, type_ignores=[]), REF, "exec"), ns)
    return ns


def random_table(rng, opal, n_cols, stickiness=0.7):
    """rows = 3 * opal; successive columns change a few entries, like event columns do"""
    vocab = [TN] * opal + [SSR] * opal + [DSR] * opal
    col = [v[rng.integers(len(v))] for v in vocab]
    cols = []
    for _ in range(n_cols):
        col = [c if rng.random() < stickiness else v[rng.integers(len(v))] for c, v in zip(col, vocab)]
        cols.append(list(col))
    return np.array(cols, dtype="<U5").T.copy()


def make_cases():
    rng = np.random.default_rng(20240611)
    cases = {}
    # getDissimilarityMatrix: 10-row blocks (30 rows), ragged column counts, duplicates and a one-column table
    tabs = [random_table(rng, 10, int(n)) for n in [7, 12, 12, 1, 19, 5, 30, 9, 14, 3]]
    tabs.append(tabs[1].copy())                                   # identical to table 1: D = 0 off the diagonal too
    tabs.append(np.column_stack([tabs[5], tabs[5][:, -1]]))       # table 5 with its last column repeated: D = 0 against 5
    cases["esec30"] = tabs
    # the e2SEC shape (24 rows, opal = 8) through the same function, which still divides by 10 (:838)
    cases["e2sec24"] = [random_table(rng, 8, int(n)) for n in [6, 11, 4, 16, 8, 8]]
    return cases


def pack(tabs):
    rows = tabs[0].shape[0]
    n_cols = np.array([t.shape[1] for t in tabs], np.int32)
    flat = np.concatenate([np.asarray(t, dtype="S6").reshape(-1) for t in tabs])
    return rows, n_cols, flat


def main():
    ns = reference_namespace()
    out = {}
    for name, tabs in make_cases().items():
        D = ns["getDissimilarityMatrix"](tabs)
        rows, n_cols, flat = pack(tabs)
        out[name + "_rows"] = np.int32(rows)
        out[name + "_ncols"] = n_cols
        out[name + "_codes"] = flat
        out[name + "_D"] = np.asarray(D, np.float64)
        if name == "e2sec24":
            M = len(tabs)
            sim = np.zeros((M, M))
            for i in range(M):
                for j in range(M):
                    sim[i, j] = ns["similarity_manipulations"](tabs[i], tabs[j])
            out[name + "_similarity"] = sim
    np.savez_compressed(OUT, **out)
    print("wrote", os.path.normpath(OUT), {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
