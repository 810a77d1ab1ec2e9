"""Measurement of the e2SEC table dissimilarity -- a side bench, bench.py stays the headline.

    python tools/bench_dissim.py [--tables 120] [--steps 20] [--warmup 3] [--cpu-pairs 400]

One step = getDissimilarityMatrix over `--tables` synthetic 30-row eSEC tables (4..30 columns), i.e. tables^2 table pairs.
Prints one JSON line: `value` = table pairs/s from the device time of the two kernels (tables already uploaded), `e2e` = the
same through the C-ABI call with host buffers (upload, kernels, download of D), `cpu_baseline` = the oracle port timed on a
bounded sample of pairs.  The kernels are integer compares on data that lives in L2 (1.2 kB per table); the figure reported as
`roofline` is issue-bound work: 64-bit compares per second against nothing but itself (no HBM or tensor roofline applies).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

TN = ["U", "T", "N", "A"]
SSR = ["U", "AB", "BE", "AR", "ART", "ABT", "BET", "IN", "SA", "TO", "BO", "BW", "F", "BA", "R", "L"]
DSR = ["U", "MT", "MA", "HT", "FMT", "S", "Q", "GC", "MC"]


def tables(M, seed=0, opal=10):
    rng = np.random.default_rng(seed)
    vocab = [np.array(v) for v in [TN] * opal + [SSR] * opal + [DSR] * opal]
    out = []
    for n in rng.integers(4, 31, size=M):
        cols = np.empty((3 * opal, int(n)), dtype="<U5")
        for r, v in enumerate(vocab):
            keep = rng.random(int(n)) < 0.7
            draw = v[rng.integers(len(v), size=int(n))]
            for c in range(int(n)):
                cols[r, c] = cols[r, c - 1] if (c and keep[c]) else draw[c]
        out.append(cols)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tables", type=int, default=120)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-pairs", type=int, default=400)
    a = ap.parse_args()
    from manipulationlibrary_b200.engine import Engine
    eng = Engine(0)
    tabs = tables(a.tables)
    M = len(tabs)
    n = np.array([t.shape[1] for t in tabs])
    compares = int(np.maximum(n[:, None], n[None, :])[np.triu_indices(M)].sum()) * 10 * 3     # 64-bit compares per step
    for _ in range(a.warmup):
        eng.dissimilarity(tabs)
    dev_ms, wall = [], []
    for _ in range(a.steps):
        t0 = time.perf_counter()
        D = eng.dissimilarity(tabs)
        wall.append(time.perf_counter() - t0)
        dev_ms.append(eng.dissimilarity_ms())
    dev = float(np.mean(dev_ms)) * 1e-3
    e2e = float(np.mean(wall))
    line = {"metric": "e2SEC table pairs compared per second (getDissimilarityMatrix)", "value": M * M / dev,
            "unit": "pairs/s", "n_gpus": 1, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dev * 1e3,
            "higher_is_better": True, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "%d eSEC tables, 30 rows, 4-30 columns" % M, "pairs_per_step": M * M},
            "e2e": {"value": M * M / e2e, "unit": "pairs/s", "ms_per_step": e2e * 1e3,
                    "note": "Python packing of the tables + C-ABI call with host buffers"},
            "gpu_launches": 2 * a.steps,
            "roofline": {"bound": "issue", "achieved": compares / dev / 1e9, "unit": "G compares/s", "peak": None, "frac": None,
                         "note": "tables are L2-resident; upper triangle only, both halves written"}}
    if a.cpu_pairs:
        from oracle import dissim_oracle as orc
        tr = [orc.triples(t) for t in tabs]
        rng = np.random.default_rng(1)
        pairs = rng.integers(0, M, size=(a.cpu_pairs, 2))
        t0 = time.perf_counter()
        for i, j in pairs:
            orc.pair_score(tr[i], tr[j], 10)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": a.cpu_pairs / dt, "unit": "pairs/s", "cores": 1, "kind": "port",
                                "sample": "%d random table pairs, triples prebuilt" % a.cpu_pairs}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
