"""Experiment: two contexts (two streams) on one GPU, each working through its own block of the stream, against one
context doing the same number of frames -- do the HBM-bound segmentation of one batch and the issue-bound radius / contact
kernels of another overlap?   python tools/exp_two_streams.py [--batch 400] [--steps 10]"""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from manipulationlibrary_b200 import _native as nat
from manipulationlibrary_b200.engine import Engine
from manipulationlibrary_b200.synthetic import SyntheticStream


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=400)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--pool", type=int, default=1200)
    a = ap.parse_args()
    B, K = a.batch, a.steps
    stream = SyntheticStream(bench.WIDTH, bench.HEIGHT, bench.N_OBJECTS, config=bench.STREAM_CONFIG)
    hx, hl = bench.generate_pool(stream, 0, a.pool)
    dev = torch.device("cuda", 0)
    dx, dl = torch.from_numpy(hx).to(dev), torch.from_numpy(hl).to(dev)
    p, keep = nat.make_params(**stream.params(filter_mode="closure", filter_seed=1, radius_mode="bounded"))
    rec_w = nat.record_doubles(int(p.n_entities))
    for n_ctx in (1, 2, 3):
        engs = [Engine(0) for _ in range(n_ctx)]
        for e in engs:
            e.configure(p, B, stream.N, keep)
        recs = [torch.empty((B, rec_w), dtype=torch.float64).pin_memory().numpy() for _ in engs]
        ids = np.arange(B, dtype=np.int32)

        def run(steps):
            for i in range(steps):
                for c, e in enumerate(engs):
                    s = ((i * n_ctx + c) * B) % a.pool
                    e.process_batch_device_async(dx[s].data_ptr(), dl[s].data_ptr(), B, ids, stream.N, records=recs[c])
            for e in engs:
                e.sync()

        run(3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run(K)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print("contexts %d: %.0f frames/s (%.3f ms per batch of %d)" % (n_ctx, n_ctx * K * B / dt, dt * 1e3 / (n_ctx * K), B), flush=True)
        for e in engs:
            e.close()


if __name__ == "__main__":
    main()
