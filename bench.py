#!/usr/bin/env python
"""e2SEC extraction throughput (BASELINE.json metric) on synthetic 640x480 / 10-object streams.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path (phase 1 on the GPU for a batch of `--batch` frames of the stream; the sequential
phase-2 resolver runs over all records at the end of the timed region, after the one all-gather when N > 1).
`value` = frames/s with the frames already resident in HBM; `e2e` = the same through the C ABI with pinned HOST
buffers (H2D of every frame and D2H of the records inside the timed region).  `roofline` is for the dominant
stage of the step; `cpu_baseline` is the CPU oracle (a port of the reference's per-frame pass) on the host cores.
--impl reference times that CPU path alone, frame-parallel over all host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "e2sec_frames_per_sec"
UNIT = "frames/s"
WIDTH, HEIGHT, N_OBJECTS, STREAM_CONFIG = 640, 480, 10, 2
# dram__bytes_read.sum + dram__bytes_write.sum per frame of the dominant kernel of each stage, from the ncu --set full capture
# of this workload (profiles/r1_v16_ncu_full_summary.csv: 400-frame launches): seg_stream_kernel + seg_emit_kernel,
# cell_filter_kernel, contact_work_kernel
NCU_TRAFFIC_PER_FRAME = {"segment": 8.334e6, "radius": 2.454e6, "pairs": 0.822e6}
NCU_STREAM_KERNEL_PER_FRAME = 5.770e6      # seg_stream_kernel alone: 4.922 MB read (the input, once) + 0.848 MB written
BYTES_PER_POINT = 16                     # 12 B xyz float32 + 4 B int32 label, read once


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=400, help="frames per step and per GPU (larger batches amortise kernel tails: 100 -> 72k, 400 -> 88k frames/s)")
    ap.add_argument("--pool", type=int, default=1200, help="distinct frames resident per GPU (cycled)")
    ap.add_argument("--filter-mode", default="closure", choices=["closure", "reference_seeded"])
    ap.add_argument("--radius-mode", default="bounded", choices=["bounded", "exact"],
                    help="bounded: cell-table bounds of np.median(pdist) with exact fallback; exact: always the O(n^2) selection")
    ap.add_argument("--cpu-frames", type=int, default=4, help="frames of the cpu_baseline sample (0 = skip)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--contexts", type=int, default=3,
                    help="contexts (streams) per GPU: batch i runs on context i %% C with its one-frame halo in front, so the "
                         "HBM-bound segmentation of one batch overlaps the issue-bound kernels of another (1: one stream)")
    ap.add_argument("--e2e-labels", default="u8", choices=["u8", "i32"],
                    help="host label images of the e2e leg: uint8 (e2sec_process_batch_labels8, 13 B/point over PCIe) or int32")
    return ap.parse_args()


def workload_name():
    return "synthetic %dx%d organised cloud stream, %d labelled objects" % (WIDTH, HEIGHT, N_OBJECTS)


# ---- CPU oracle legs (the only places bench.py touches oracle/) ------------------------------------------------------
def _oracle_for(stream, mode):
    from oracle import esec_oracle as eo
    return eo.E2SecOracle(stream.ground_label, stream.hand_label, None, stream.thresh, stream.translation, stream.roll,
                          stream.ground, stream.total_unique_labels, filter_mode=mode, seed=20261104)


def cpu_baseline_single(stream, first, n_frames, mode):
    """The oracle's sequential per-frame pass on one core: frames/s over n_frames consecutive frames."""
    orc = _oracle_for(stream, mode)
    frames = [stream.frame(first + i) for i in range(n_frames)]
    t0 = time.perf_counter()
    for i, (xyz, lab) in enumerate(frames):
        orc.process_frame(first + i, xyz, lab.reshape(stream.H, stream.W))
    dt = time.perf_counter() - t0
    return n_frames / dt, dt


def _ref_worker(conn, wid, first, mode):
    from manipulationlibrary_b200.synthetic import SyntheticStream
    stream = SyntheticStream(WIDTH, HEIGHT, N_OBJECTS, config=STREAM_CONFIG)
    orc = _oracle_for(stream, mode)
    nxt = None
    while True:
        msg = conn.recv()
        if msg[0] == "prep":
            nxt = (msg[1], stream.frame(msg[1]))
            conn.send("ok")
        elif msg[0] == "run":
            f, (xyz, lab) = nxt
            orc.process_frame(f, xyz, lab.reshape(stream.H, stream.W))
            conn.send("ok")
        else:
            return


def run_reference(args):
    """The reference's CPU implementation of the path (oracle port), one frame per host core per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count() or 1
    try:                                          # one worker holds ~0.5 GB of pdist output at a time
        import psutil
        cores = max(1, min(cores, int(psutil.virtual_memory().available / 2 ** 30 / 1.5)))
    except Exception:
        pass
    cores = min(cores, 64)
    ctx = mp.get_context("fork")
    span = args.steps + args.warmup + 1
    workers = []
    for w in range(cores):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_ref_worker, args=(b, w, 30 + w * span, args.filter_mode), daemon=True)
        p.start()
        workers.append((p, a, 30 + w * span))

    def step(i, timed):
        for _, c, first in workers:
            c.send(("prep", first + i))
        for _, c, _ in workers:
            c.recv()
        t0 = time.perf_counter()
        for _, c, _ in workers:
            c.send(("run",))
        for _, c, _ in workers:
            c.recv()
        return time.perf_counter() - t0

    for i in range(args.warmup):
        step(i, False)
    total = 0.0
    for i in range(args.steps):
        total += step(args.warmup + i, True)
    for p, c, _ in workers:
        c.send(("stop",))
        p.join(timeout=5)
    frames = cores * args.steps
    value = frames / total
    sample = "%d frames per step (one per host core), %d timed steps" % (cores, args.steps)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(), "filter_mode": args.filter_mode, "frames_per_step": cores},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


# ---- clocks ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, uuid):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", uuid, "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if t < t0 or t > t1 + 0.1:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- the B200 arm ----------------------------------------------------------------------------------------------------
def generate_pool(stream, first, count):
    from concurrent.futures import ThreadPoolExecutor
    xyz = np.empty((count, stream.N, 3), np.float32)
    lab = np.empty((count, stream.N), np.int32)

    def one(i):
        xyz[i], lab[i] = stream.frame(first + i)

    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
        list(ex.map(one, range(count)))
    return xyz, lab


def run_b200(args):
    import torch
    import torch.distributed as dist
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine, Resolver
    from manipulationlibrary_b200.synthetic import SyntheticStream

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    stream = SyntheticStream(WIDTH, HEIGHT, N_OBJECTS, config=STREAM_CONFIG)
    B, K, W = args.batch, args.steps, args.warmup
    N = stream.N
    pool = max(B, min(args.pool, (K + W) * B))
    pool = (pool // B) * B
    rank_first = rank * (K + W) * B               # every rank owns its block of the stream (weak scaling)

    # cpu_baseline before the CUDA context exists on this rank (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and args.cpu_frames > 0:
        v, dt = cpu_baseline_single(stream, 30, args.cpu_frames, args.filter_mode)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "%d consecutive frames of the same stream, oracle.esec_oracle.E2SecOracle.process_frame, %.1f s" % (args.cpu_frames, dt)}

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    saved_stdout = None
    if world > 1:
        # NCCL prints its version banner on stdout when the communicator comes up: keep stdout for the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    hx, hl = generate_pool(stream, rank_first, pool)
    # one more frame in front: a copy of the pool's last frame, so that batch s can be handed over together with the frame
    # before it (its halo) as one contiguous block even where the pool wraps
    hx = np.concatenate([hx[-1:], hx])
    hl = np.concatenate([hl[-1:], hl])
    hx_t = torch.from_numpy(hx).pin_memory()
    hl_t = torch.from_numpy(hl).pin_memory()
    lab8 = args.e2e_labels == "u8" and not args.no_e2e and int(hl.max()) < 256 and int(hl.min()) >= 0
    hl8_t = torch.from_numpy(hl.astype(np.uint8)).pin_memory() if lab8 else None     # what a uint8 label reader hands over
    dx = hx_t.to(dev, non_blocking=True)
    dl = hl_t.to(dev, non_blocking=True)
    torch.cuda.synchronize()

    p, keep = nat.make_params(**stream.params(filter_mode=args.filter_mode, filter_seed=20261104,
                                               radius_mode=args.radius_mode))
    C = max(1, args.contexts)
    engs = [Engine(local) for _ in range(C)]
    for e in engs:
        e.configure(p, B + 1, N, keep)
    eng = engs[0]
    ests = [torch.cuda.ExternalStream(e.stream, device=dev) for e in engs]
    est = ests[0]
    rec_w = nat.record_doubles(int(p.n_entities))
    total_frames = (K + W) * B
    # batch i writes rows [i, 1:]; row [i, 0] receives the record of its halo frame when the batch carries one
    rec_host = torch.empty((K + W, B + 1, rec_w), dtype=torch.float64).pin_memory()
    rec3 = rec_host.numpy()
    ids_all = np.arange(rank_first, rank_first + total_frames, dtype=np.int32)
    KEPT = float(nat.FLAG_HAND_LABEL | nat.FLAG_HAND_POINTS)
    import threading
    from concurrent.futures import ThreadPoolExecutor
    follower = ThreadPoolExecutor(max_workers=1)    # phase 2 follows the GPU on its own thread, in frame order
    locks = [threading.Lock() for _ in engs]        # reset + queue on a context is one unit (the follower may redo a batch)
    halo_of = {}                               # batch -> True when it was queued with its halo frame in front

    def inputs(on_device, first_frame):
        """pointers to the frame with local index first_frame (pool position + 1 because of the leading copy)"""
        s = first_frame % pool + 1
        if on_device:
            return dx[s].data_ptr(), dl[s].data_ptr()
        return hx_t[s].data_ptr(), (hl8_t if lab8 else hl_t)[s].data_ptr()

    def halo_inputs(on_device, batch_first):
        """the block that starts one frame before batch_first: pool position batch_first % pool of the extended arrays"""
        s = batch_first % pool
        if on_device:
            return dx[s].data_ptr(), dl[s].data_ptr()
        return hx_t[s].data_ptr(), (hl8_t if lab8 else hl_t)[s].data_ptr()

    def queue(e, px, pl, n, ids, out, on_device):
        if on_device:
            e.process_batch_device_async(px, pl, n, ids, N, records=out)
        else:
            e.process_batch_device_async(px, pl, n, ids, N, records=out, on_device=False, labels8=lab8)

    def step(i, on_device, seq_first):
        """Queues batch i (frames i*B .. (i+1)*B-1 of this rank's block) on context i % C.  With one context the batches of
        a sequence follow each other on its stream and the context carries the previous kept frame itself.  With several,
        every batch is its own little sequence on its context: the frame before it goes in front as the halo (its record
        lands in row 0 and is dropped) -- the same one-frame halo the frame sharding across GPUs uses."""
        c = i % C
        e = engs[c]
        ids = ids_all[i * B: (i + 1) * B]
        with locks[c]:
            if C == 1 or i == seq_first:
                if C > 1:
                    e.reset_sequence()
                px, pl = inputs(on_device, i * B)
                queue(e, px, pl, B, ids, rec3[i, 1:], on_device)
                halo_of[i] = False
            else:
                e.reset_sequence()
                px, pl = halo_inputs(on_device, i * B)
                queue(e, px, pl, B + 1, ids_all[i * B - 1: (i + 1) * B], rec3[i], on_device)
                halo_of[i] = True
        return c

    def settle(i, on_device, seq_first):
        """batch i is on the host.  If its halo frame turned out not to be a kept frame (no hand in it), the previous kept
        frame lies further back: redo the batch behind a walk back to it (sharding.run_shard's rule).  Rare, synchronous."""
        if not halo_of.get(i) or rec3[i, 0, 1] == KEPT:
            return
        e = engs[i % C]
        with locks[i % C]:
            e.sync()
            j = i * B - 2
            one = np.empty((1, rec_w), np.float64)
            while True:
                e.reset_sequence()
                if j < seq_first * B:
                    break
                px, pl = inputs(on_device, j)
                queue(e, px, pl, 1, ids_all[j: j + 1], one, on_device)
                e.sync()
                if one[0, 1] == KEPT:
                    break
                j -= 1
            px, pl = inputs(on_device, i * B)
            queue(e, px, pl, B, ids_all[i * B: (i + 1) * B], rec3[i, 1:], on_device)
            e.sync()

    def sync_all():
        for e in engs:
            e.sync()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: the records of batch i are all-gathered (NCCL, its own stream) while the GPUs work on batch i+1, and rank 0
    # resolves the gathered batch a few steps behind ON THE LAUNCH THREAD: that wait is what keeps the ranks loosely in step
    # (with free-running launches a rank's all-gather kernel sits on its SMs waiting for the slowest rank: 729 k instead of
    # 766 k frames/s at 8 GPUs) -- every shard is its own sequence here, so the shards' resolvers run
    # side by side (one C call each, the GIL is released).  No collective on the data path itself.
    if world > 1:
        gst = torch.cuda.Stream(device=dev)
        gbuf = [torch.empty((B, rec_w), dtype=torch.float64, device=dev) for _ in range(2)]
        gfull = [torch.empty((world * B, rec_w), dtype=torch.float64, device=dev) for _ in range(2)]
        ghost = torch.empty((max(K, W, 1), world * B, rec_w), dtype=torch.float64).pin_memory()
        ghost_np = ghost.numpy()
        tpool = ThreadPoolExecutor(max_workers=min(world, os.cpu_count() or 1)) if rank == 0 else None

    def pipelined_gather(on_device, first_step, n_steps):
        resolvers = [Resolver(p, keep) for _ in range(world)] if rank == 0 else None
        gdone, LAG = [], int(os.environ.get("E2SEC_BENCH_LAG", 2))      # measured at 2 GPUs: lag 1: 202.8 k, 2: 206.3 k, 3: 202.5 k, 6: 191.5 k, 9: 185.1 k frames/s

        def feed(j):
            if rank == 0:
                list(tpool.map(lambda r: resolvers[r].feed(ghost_np[j, r * B: (r + 1) * B]), range(world)))

        for i in range(n_steps):
            c = step(first_step + i, on_device, first_step)
            with torch.cuda.stream(ests[c]):
                e = torch.cuda.Event()
                e.record()
            with torch.cuda.stream(gst):
                gst.wait_event(e)                      # the batch's records are in the pinned host buffer
                gbuf[i & 1].copy_(rec_host[first_step + i, 1:], non_blocking=True)
                dist.all_gather_into_tensor(gfull[i & 1], gbuf[i & 1])
                if rank == 0:
                    ghost[i].copy_(gfull[i & 1], non_blocking=True)
                ge = torch.cuda.Event()
                ge.record()
            gdone.append(ge)
            if i >= LAG:                               # batches of slack: a rank that runs late does not stall the others
                gdone[i - LAG].synchronize()
                feed(i - LAG)
        sync_all()
        for i in range(n_steps):                       # every halo frame of the synthetic stream is a kept frame; say so if not
            if halo_of.get(first_step + i) and rec3[first_step + i, 0, 1] != KEPT:
                raise SystemExit("bench.py: batch %d starts behind a frame without a hand; run with --contexts 1" % (first_step + i))
        for j in range(max(n_steps - LAG, 0), n_steps):
            gdone[j].synchronize()
            feed(j)
        cols = 0
        if rank == 0:
            cols = sum(r.n_columns for r in resolvers)
            for r in resolvers:
                r.close()
        return cols

    def timed(on_device, first_step, n_steps):
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in engs]
        for e in engs:
            e.reset_sequence()              # the timed block is its own sequence (the resolver starts at its first frame)
        barrier()
        l0 = sum(e.launch_count for e in engs)
        t0 = time.perf_counter()
        with torch.cuda.stream(est):
            ev0.record()                    # every stream is idle here (barrier); the last stream to finish ends the region
        if world == 1:
            # one GPU: phase 2 (host, sequential) follows the GPU on its own thread instead of after it, so that the launch
            # thread keeps C batches in flight
            res = Resolver(p, keep)
            futs = []

            def follow(i, ev):
                ev.synchronize()                       # batch i is on the host; the later ones keep the GPU busy meanwhile
                settle(i, on_device, first_step)
                res.feed(rec3[i, 1:])

            for i in range(n_steps):
                c = step(first_step + i, on_device, first_step)
                with torch.cuda.stream(ests[c]):
                    e = torch.cuda.Event()
                    e.record()
                futs.append(follower.submit(follow, first_step + i, e))
                if i >= 2 * C:
                    futs[i - 2 * C].result()           # bounds how far the launches run ahead of phase 2
            for f in futs:
                f.result()
            sync_all()                                 # raises device-side errors of the queued batches
            cols = res.n_columns
            res.close()
        else:
            cols = pipelined_gather(on_device, first_step, n_steps)
        for c in range(C):
            with torch.cuda.stream(ests[c]):
                ev1[c].record()
        barrier()
        t1 = time.perf_counter()
        ms = max(max(ev0.elapsed_time(e) for e in ev1), 0.0)
        wall = (t1 - t0) * 1e3
        t = torch.tensor([ms, wall], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), sum(e.launch_count for e in engs) - l0, cols, (t0, t1)

    for e in engs:
        e.reset_sequence()
    if world > 1:
        pipelined_gather(True, 0, W)        # also brings up the NCCL communicator outside the timed region
    else:
        for i in range(W):
            step(i, True, 0)
    sync_all()
    if saved_stdout is not None:            # the warm-up has run the first collectives: the banner is out
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    uuid = str(torch.cuda.get_device_properties(dev).uuid)
    sampler = ClockSampler(uuid if uuid.startswith("GPU-") else "GPU-" + uuid) if rank == 0 else None
    time.sleep(0.3 if sampler else 0.0)
    ms, wall_ms, launches, cols, (t0, t1) = timed(True, W, K)
    clocks = sampler.stop(t0, t1) if sampler else None
    # the event pair brackets device work AND the host resolver between launches; wall is reported beside it
    frames = K * B * world
    value = frames / (ms * 1e-3)
    # how many filtered entities of the timed frames needed the exact O(n^2) path (radius field not NaN)
    Kent = int(p.n_entities)
    ent = rec3[W: W + K, 1:, nat.REC_HEADER: nat.REC_HEADER + nat.REC_ENTITY * Kent].reshape(-1, Kent, nat.REC_ENTITY)
    n_filtered = int((ent[:, :, 2] >= 0).sum())
    n_exact = int(((ent[:, :, 2] >= 0) & ~np.isnan(ent[:, :, 3])).sum())
    # stage split and the streaming kernel's own time: one more batch on one context with nothing else on the GPU
    # (inside the timed region the stages of different batches overlap, their event pairs would count each other's work)
    # -- three batches back to back, the last one's events are read (steady state, like the batches of the timed region)
    eng.reset_sequence()
    for i in range(W, W + 3):
        px, pl = inputs(True, i * B)
        queue(eng, px, pl, B, ids_all[i * B: (i + 1) * B], rec3[0, 1:], True)  # rows of a warm-up batch, no longer needed
    eng.sync()
    stages = eng.stage_ms()

    e2e = None
    if not args.no_e2e:
        for e in engs:
            e.reset_sequence()
        for i in range(C):
            step(i, False, 0)
        sync_all()
        ems, ewall, _, _, _ = timed(False, W, K)
        h2d_frames = B + (1 if C > 1 else 0)             # the halo frame travels with its batch
        e2e = {"value": frames / (ewall * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": h2d_frames * N * (13 if lab8 else BYTES_PER_POINT),
               "d2h_bytes_per_step": h2d_frames * rec_w * 8, "host_labels": "uint8" if lab8 else "int32",
               "timing": "wall clock around the C-ABI calls, pinned host buffers"}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        which = "measured" if "hbm_gbs" in peaks else "fallback"
        dom = max(stages, key=lambda k: stages[k])
        dom_ms = stages[dom]
        alg = B * N * BYTES_PER_POINT
        achieved = alg / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None
        step_gbs = frames / world * N * BYTES_PER_POINT / (ms * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(), "frames_per_step_per_gpu": B, "frames_timed": frames,
                       "pool_frames_per_gpu": pool, "filter_mode": args.filter_mode, "radius_mode": args.radius_mode,
                       "filtered_entities": n_filtered, "exact_path_entities": n_exact,
                       "l2": "inputs larger than L2 (one step reads %d MB)" % (alg // 2 ** 20),
                       "parallelism": ("frame-sharded x%d, records all-gathered per batch behind the GPU work" % world) if world > 1
                       else "one GPU, host resolver following on its own thread",
                       "contexts_per_gpu": C},
            "clocks": clocks, "gpu_launches": launches, "wall_ms_per_step": wall_ms / K, "esec_columns": cols,
            "stages_ms_solo_step": stages, "stages_ms_last_step": stages,
            "roofline": {"bound": "hbm", "kernel": dom + " stage of one step", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                         "traffic": (NCU_TRAFFIC_PER_FRAME[dom] * B) if (dom in NCU_TRAFFIC_PER_FRAME and N == 307200) else None,
                         "peak_source": which, "whole_step_gbs_per_gpu": step_gbs, "whole_step_frac": step_gbs / peak,
                         # the streaming kernel (label filter + compaction) is the HBM-bound one; the other two stages are
                         # bound by O(cells^2) / O(n_a n_b) arithmetic on L2-resident data
                         # the one kernel that streams the input (label filter + compaction, seg_stream.cuh), timed alone with
                         # its own event pair: this is the HBM-bound kernel of the path
                         "streaming_kernel": ({"name": "seg_stream_kernel", "ms": eng.seg_stream_ms,
                                               "algorithmic_gbs": alg / (eng.seg_stream_ms * 1e-3) / 1e9,
                                               "frac": alg / (eng.seg_stream_ms * 1e-3) / 1e9 / peak,
                                               "dram_gbs": (NCU_STREAM_KERNEL_PER_FRAME * B / (eng.seg_stream_ms * 1e-3) / 1e9)
                                               if N == 307200 else None}
                                              if getattr(eng, "seg_stream_ms", 0) > 0 else None),
                         "stages": {k: {"ms": stages[k], "algorithmic_gbs": alg / (stages[k] * 1e-3) / 1e9,
                                        "frac": alg / (stages[k] * 1e-3) / 1e9 / peak,
                                        "dram_gbs": (NCU_TRAFFIC_PER_FRAME[k] * B / (stages[k] * 1e-3) / 1e9) if N == 307200 else None}
                                    for k in ("segment", "radius", "pairs") if stages.get(k, 0) > 0}},
        }
        if e2e:
            out["e2e"] = e2e
        if cpu:
            out["cpu_baseline"] = cpu
        print(json.dumps(out), flush=True)
    for e in engs:
        e.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
