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
# BASELINE.json configs[1], [3], [4]: workload -> stream shape and the default step size
CONFIGS = {
    2: dict(width=640, height=480, objects=10, batch=400, pool=1200, cpu_frames=4, near=False,
            name="synthetic 640x480 organised cloud stream, 10 labelled objects"),
    4: dict(width=1280, height=720, objects=20, batch=120, pool=360, cpu_frames=1, near=True,
            name="synthetic 1280x720 stream, 20 objects (190 pairs), neighbours within +-5 mm of thresh"),
    5: dict(width=2000, height=1000, objects=30, batch=40, pool=120, cpu_frames=1, near=False,
            name="synthetic 2M-point clouds (2000x1000), 30 objects (435 pairs)"),
}
# dram__bytes_read.sum + dram__bytes_write.sum per frame of each stage's kernels: written by tools/ncu_traffic.py from an
# `ncu --set full` capture of this very command (profiles/ncu_traffic.json names the capture); absent -> traffic is null
NCU_TRAFFIC_FILE = os.path.join(ROOT, "profiles", "ncu_traffic.json")
BYTES_PER_POINT = 16                     # 12 B xyz float32 + 4 B int32 label, read once


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS), help="BASELINE.json workload: 2 = 640x480 / 10 objects "
                    "(the metric's configuration), 4 = 1280x720 / 20 objects, 5 = 2M points / 30 objects")
    ap.add_argument("--batch", type=int, default=0, help="frames per step and per GPU (0: the config's default; larger batches amortise kernel tails)")
    ap.add_argument("--pool", type=int, default=0, help="distinct frames resident per GPU, cycled (0: the config's default)")
    ap.add_argument("--blocks", type=int, default=30, help="the K-step block is repeated this often; value = median block (>= 1 s of timed work)")
    ap.add_argument("--e2e-blocks", type=int, default=5)
    ap.add_argument("--phase2", default="device", choices=["device", "host"],
                    help="device: relation classifiers as a kernel behind every batch (phase2.cuh); host: the C++ resolver (one GPU only)")
    ap.add_argument("--filter-mode", default="closure", choices=["closure", "reference_seeded"])
    ap.add_argument("--radius-mode", default="bounded", choices=["bounded", "exact"],
                    help="bounded: cell-table bounds of np.median(pdist) with exact fallback; exact: always the O(n^2) selection")
    ap.add_argument("--cpu-frames", type=int, default=-1, help="frames of the cpu_baseline sample (0 = skip, -1: the config's default)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--contexts", type=int, default=3,
                    help="contexts (streams) per GPU: batch i runs on context i %% C with its one-frame halo in front, so the "
                         "HBM-bound segmentation of one batch overlaps the issue-bound kernels of another (1: one stream)")
    ap.add_argument("--e2e-labels", default="u8", choices=["u8", "i32"],
                    help="host label images of the e2e leg: uint8 (e2sec_process_batch_labels8, 13 B/point over PCIe) or int32")
    return ap.parse_args()


def make_stream(config):
    from manipulationlibrary_b200.synthetic import SyntheticStream
    c = CONFIGS[config]
    return SyntheticStream(c["width"], c["height"], c["objects"], config=config, near_thresh=c["near"])


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


def _ref_worker(conn, wid, first, mode, config):
    stream = make_stream(config)
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
        per_worker_gb = 1.5 * (CONFIGS[args.config]["width"] * CONFIGS[args.config]["height"]) / 307200.0
        cores = max(1, min(cores, int(psutil.virtual_memory().available / 2 ** 30 / per_worker_gb)))
    except Exception:
        pass
    cores = min(cores, 64)
    ctx = mp.get_context("fork")
    span = args.steps + args.warmup + 1
    workers = []
    for w in range(cores):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_ref_worker, args=(b, w, 30 + w * span, args.filter_mode, args.config), daemon=True)
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
        "config": {"workload": CONFIGS[args.config]["name"], "filter_mode": args.filter_mode, "frames_per_step": cores},
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


def ncu_traffic(config):
    try:
        return json.load(open(NCU_TRAFFIC_FILE)).get(str(config))
    except Exception:
        return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine, Phase2, Resolver, columns_table

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    if args.phase2 == "host" and world > 1:
        raise SystemExit("bench.py: --phase2 host is the one-GPU comparison mode; several GPUs shard ONE sequence, which needs the device classifier")
    cfg = CONFIGS[args.config]
    stream = make_stream(args.config)
    B, K, W = (args.batch or cfg["batch"]), args.steps, max(args.warmup, 1)
    N = stream.N
    pool = args.pool or cfg["pool"]
    pool = max(B, (pool // B) * B)
    seq = K * B                                   # frames of a timed block per GPU; all ranks' blocks are ONE sequence
    cpu_frames = cfg["cpu_frames"] if args.cpu_frames < 0 else args.cpu_frames

    # cpu_baseline before the CUDA context exists on this rank (rank 0, N = 1 only)
    cpu, oracle_table = None, None
    if rank == 0 and world == 1 and cpu_frames > 0:
        orc = _oracle_for(stream, args.filter_mode)
        frames = [stream.frame(30 + i) for i in range(cpu_frames)]
        t0 = time.perf_counter()
        for i, (xyz, lab) in enumerate(frames):
            orc.process_frame(30 + i, xyz, lab.reshape(stream.H, stream.W))
        dt = time.perf_counter() - t0
        cpu = {"value": cpu_frames / dt, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "%d consecutive frames of the same stream, oracle.esec_oracle.E2SecOracle.process_frame, %.1f s" % (cpu_frames, dt)}
        oracle_table = (orc.table, list(orc.final_columns), frames)

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_near_gpu(torch, local)
    saved_stdout = None
    if world > 1:
        # NCCL prints its version banner on stdout when the communicator comes up: keep stdout for the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    # One sequence of world * K batches of B frames, dealt round-robin: global batch g runs on rank g % world as that rank's
    # step g // world, behind a one-frame halo (the last frame of batch g - 1).  Rank r holds `pool` distinct stream frames
    # [r * pool, (r + 1) * pool) as pool / B blocks of B frames; step j of a block of the run has the content of pool block
    # j % (pool / B).  Every block is stored with its halo frame in front -- the last frame of the pool block the
    # predecessor batch uses on ITS rank -- so a batch and its halo are one contiguous region.
    nb = pool // B

    def pred_block(r, b):
        """(rank, pool block) of the batch in front of rank r's pool block b"""
        return (r - 1, b) if r > 0 else (world - 1, (b - 1) % nb)

    hx = np.empty((nb, B + 1, N, 3), np.float32)
    hl = np.empty((nb, B + 1, N), np.int32)
    gx_, gl_ = generate_pool(stream, rank * pool, pool)
    hx[:, 1:] = gx_.reshape(nb, B, N, 3)
    hl[:, 1:] = gl_.reshape(nb, B, N)
    del gx_, gl_
    for b_ in range(nb):
        pr, pb = pred_block(rank, b_)
        hx[b_, 0], hl[b_, 0] = stream.frame(pr * pool + pb * B + B - 1)
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
    est = torch.cuda.ExternalStream(eng.stream, device=dev)
    ests = [torch.cuda.ExternalStream(e.stream, device=dev) for e in engs]
    rec_w = nat.record_doubles(int(p.n_entities))
    p2 = Phase2(eng, max_columns=seq + 8)
    # e2e / host phase 2 / verification: the records of a block on the host, batch i in rows [i, 1:] (row [i, 0]: its halo frame)
    rec_host = torch.empty((K, B + 1, rec_w), dtype=torch.float64).pin_memory()
    rec3 = rec_host.numpy()
    serial = [0]                                   # blocks run so far (tickets of the ring never repeat)
    if world > 1:
        # every rank's phase-2 mailbox is opened by the rank in front of it as peer memory (CUDA IPC, NVLink): the state of
        # the sequence then travels from kernel to kernel without the host
        mine = torch.from_numpy(p2.mailbox_handle()).to(dev)
        handles = torch.empty((world, nat.IPC_HANDLE_BYTES), dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(handles, mine)
        p2.connect(handles[(rank + 1) % world].cpu().numpy())
    from concurrent.futures import ThreadPoolExecutor
    follower = ThreadPoolExecutor(max_workers=1)

    def ids_of(j, halo):
        """frame ids (= global frame indices of the sequence) of this rank's step j, with its halo frame's in front"""
        g = j * world + rank
        return np.arange(g * B - (1 if halo else 0), (g + 1) * B, dtype=np.int32)

    def inputs(on_device, j, halo):
        blk, o = j % nb, (0 if halo else 1)
        if on_device:
            return dx[blk, o].data_ptr(), dl[blk, o].data_ptr()
        return hx_t[blk, o].data_ptr(), (hl8_t if lab8 else hl_t)[blk, o].data_ptr()

    def queue(e, px, pl, n, ids, out, on_device):
        if on_device:
            e.process_batch_device_async(px, pl, n, ids, N, records=out)
        else:
            e.process_batch_device_async(px, pl, n, ids, N, records=out, on_device=False, labels8=lab8)

    def step(j, on_device, want_host):
        """Phase 1 of this rank's step j on context j % C.  Every batch is its own little sequence on its context: the frame
        before it goes in front as the halo (its record is dropped) -- except the very first batch of the sequence, and
        except with one context on one GPU, where the batches simply follow each other on the stream."""
        e = engs[j % C]
        g = j * world + rank
        halo = g > 0 and not (C == 1 and world == 1)
        if halo or g == 0:
            e.reset_sequence()
        px, pl = inputs(on_device, j, halo)
        out = (rec3[j] if halo else rec3[j, 1:]) if want_host else None
        queue(e, px, pl, B + (1 if halo else 0), ids_of(j, halo), out, on_device)
        return e, (1 if halo else 0)

    def phase2_step(j, e, skip):
        """The relation classifiers of step j behind everything before it in the sequence.  With several GPUs the kernel
        itself waits for the state (512 bytes) in this rank's mailbox, where the kernel of the batch in front -- on the
        previous rank of the ring -- has put it over NVLink, and posts its own state into the next rank's mailbox."""
        g = j * world + rank
        if world == 1:
            p2.run_engine_batch(e, B + skip, skip=skip)
            return
        base = serial[0] * world * K + 1
        p2.run_ring_engine_batch(e, B + skip, skip, base + g if g > 0 else 0, base + g + 1 if g < world * K - 1 else 0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cap_cols = seq + 8
    exp_t = torch.zeros(16 + 36 * cap_cols, dtype=torch.uint8, device=dev)
    exp_all = [torch.empty_like(exp_t) for _ in range(world)] if (world > 1 and rank == 0) else None

    def gather_columns(out):
        """The one collective: the emitted columns (32 code bytes + frame id each) of every rank on rank 0, in frame order."""
        n = p2.sync()                                  # raises phase-2 errors (records out of order, too many columns)
        if world == 1:
            if out is not None:
                out["columns"], out["frames"] = p2.fetch()
            return n
        p2.export(exp_t.data_ptr(), cap_cols)
        dist.gather(exp_t, exp_all, dst=0)
        if rank != 0:
            return 0
        host = torch.stack(exp_all).cpu().numpy()
        parts = [Phase2.parse_export(host[r], cap_cols) for r in range(world)]
        cols = np.concatenate([c for c, _ in parts])
        fr = np.concatenate([f for _, f in parts])
        order = np.argsort(fr, kind="stable")
        if out is not None:
            out["columns"], out["frames"] = np.ascontiguousarray(cols[order]), fr[order]
        return len(fr)

    def block(on_device, want_host=False, out=None):
        """One timed block = one sequence of world * K * B frames.  -> (device ms, wall ms, launches, columns)"""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dev_p2 = args.phase2 == "device"
        if dev_p2:
            p2.reset()
        barrier()
        l0 = sum(e.launch_count for e in engs)
        t0 = time.perf_counter()
        with torch.cuda.stream(est):
            ev0.record()                    # every stream is idle here (barrier)
        if dev_p2:
            for j in range(K):
                e, skip = step(j, on_device, want_host)
                phase2_step(j, e, skip)
            serial[0] += 1
            for e in engs:
                e.sync()                                   # raises device-side errors of the queued batches
            cols = gather_columns(out)
        else:
            # the C++ resolver follows the GPU on its own thread, in frame order
            res = Resolver(p, keep)
            futs = []

            def follow(i, ev):
                ev.synchronize()
                res.feed(rec3[i, 1:])

            for i in range(K):
                step(i, on_device, True)
                with torch.cuda.stream(ests[i % C]):
                    e = torch.cuda.Event()
                    e.record()
                futs.append(follower.submit(follow, i, e))
                if i >= 2 * C:
                    futs[i - 2 * C].result()
            for f in futs:
                f.result()
            for e in engs:
                e.sync()
            cols = res.n_columns
            if out is not None:
                out["table"] = res.table()
            res.close()
        torch.cuda.synchronize()
        ev1.record()
        ev1.synchronize()
        t1 = time.perf_counter()
        t = torch.tensor([ev0.elapsed_time(ev1), (t1 - t0) * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), sum(e.launch_count for e in engs) - l0, cols, (t0, t1)

    for _ in range(max(1, (W + K - 1) // K)):        # warm-up: whole blocks (>= W steps); also brings up the NCCL communicator
        block(True)
    if saved_stdout is not None:                    # the warm-up has run the first collectives: the banner is out
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    uuid = str(torch.cuda.get_device_properties(dev).uuid)
    debug = bool(os.environ.get("E2SEC_BENCH_DEBUG"))
    sampler = ClockSampler(uuid if uuid.startswith("GPU-") else "GPU-" + uuid) if (rank == 0 or debug) else None
    time.sleep(0.3 if sampler else 0.0)
    runs = [block(True) for _ in range(max(1, args.blocks))]
    if os.environ.get("E2SEC_BENCH_DEBUG"):            # per-rank view of the timed blocks (stderr): own device time, clocks
        own = []
        for _ in range(5):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t_ = time.perf_counter()
            for j in range(K):
                step(j, True, False)
            for e in engs:
                e.sync()
            own.append((time.perf_counter() - t_) * 1e3)
        try:
            q = subprocess.run(["nvidia-smi", "-i", str(local), "--query-gpu=clocks.sm,power.draw,temperature.gpu,clocks_event_reasons.active",
                                "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
        except Exception:
            q = "?"
        print("[rank %d] phase 1 only, K steps, no ring: %s ms; idle-after-run smi: %s" % (rank, ["%.2f" % v for v in own], q), file=sys.stderr, flush=True)
    clocks = sampler.stop(runs[0][4][0], runs[-1][4][1]) if sampler else None
    if debug:
        print("[rank %d] clocks under load: %s; block ms (max over ranks): median %.2f" % (rank, clocks, statistics.median(r[0] for r in runs)),
              file=sys.stderr, flush=True)
    ms_all = sorted(r[0] for r in runs)
    ms = statistics.median(ms_all)
    wall_ms = statistics.median(r[1] for r in runs)
    launches, cols = runs[-1][2], runs[-1][3]
    frames = K * B * world
    value = frames / (ms * 1e-3)

    # ---- checks outside the timed region (rank 0 decides; a failed check ends the run loudly) --------------------------
    checks = {}
    got = {}
    block(True, want_host=True, out=got)
    ent_w = nat.REC_ENTITY * int(p.n_entities)
    ent = rec3[:, 1:, nat.REC_HEADER: nat.REC_HEADER + ent_w].reshape(-1, int(p.n_entities), nat.REC_ENTITY)
    n_filtered = int((ent[:, :, 2] >= 0).sum())
    n_exact = int(((ent[:, :, 2] >= 0) & ~np.isnan(ent[:, :, 3])).sum())
    if rank == 0 and world == 1:
        # the device classifier against the C++ resolver on the very records of that block
        res = Resolver(p, keep)
        for i in range(K):
            res.feed(rec3[i, 1:])
        want_t, want_fc = res.table()
        res.close()
        got_t, got_fc = columns_table(got["columns"], got["frames"], 3) if args.phase2 == "device" else got["table"]
        ok = got_t.shape == want_t.shape and bool((got_t == want_t).all()) and list(got_fc) == list(want_fc)
        checks["phase2_equals_host_resolver"] = ok
        if not ok:
            raise SystemExit("bench.py: device phase 2 and the host resolver disagree on the block's records")
        if oracle_table is not None:
            # the cpu_baseline frames through the GPU: same raw table as the oracle produced
            otab, ofc, oframes = oracle_table
            n = len(oframes)
            xs = np.stack([f[0] for f in oframes])
            ls = np.stack([f[1].reshape(-1) for f in oframes]).astype(np.int32)
            eng.reset_sequence()
            r = eng.process_batch(xs, ls, np.arange(30, 30 + n, dtype=np.int32))
            res = Resolver(p, keep)
            res.feed(r)
            t, fc = res.table()
            res.close()
            ok = (otab is None and t.shape[1] == 0) or (otab is not None and t.shape == otab.shape and
                                                         bool((t.astype("U5") == otab).all()) and list(fc) == list(ofc))
            checks["cpu_baseline_frames_equal_oracle"] = ok
            if not ok:
                raise SystemExit("bench.py: the GPU path and the oracle disagree on the cpu_baseline frames")

    # stage split and the streaming kernel's own time: batches on one context with nothing else on the GPU (inside the
    # timed region the stages of different batches overlap) -- three back to back, the last one's events are read
    eng.reset_sequence()
    for i in range(3):
        px, pl = inputs(True, i, False)
        queue(eng, px, pl, B, ids_of(i, False), None, True)
    eng.sync()
    stages = eng.stage_ms()

    e2e = None
    if not args.no_e2e:
        block(False, want_host=True)
        eruns = [block(False, want_host=True) for _ in range(max(1, args.e2e_blocks))]
        ewall = statistics.median(r[1] for r in eruns)
        h2d_frames = B + (1 if C > 1 else 0)             # the halo frame travels with its batch
        e2e = {"value": frames / (ewall * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": h2d_frames * N * (13 if lab8 else BYTES_PER_POINT),
               "d2h_bytes_per_step": h2d_frames * rec_w * 8, "host_labels": "uint8" if lab8 else "int32",
               "blocks": len(eruns), "best": frames / (min(r[1] for r in eruns) * 1e-3),
               "timing": "wall clock around the C-ABI calls of a block (median block), pinned host buffers, records back to the host",
               "host_numa": numa}      # rank 0's binding (bind_near_gpu): the pinned pools sit on the GPU's own NUMA node

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        which = "measured" if "hbm_gbs" in peaks else "fallback"
        dom = max(("segment", "radius", "pairs"), key=lambda k: stages.get(k, 0.0))
        dom_ms = stages[dom]
        alg = B * N * BYTES_PER_POINT
        achieved = alg / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None
        step_gbs = frames / world * N * BYTES_PER_POINT / (ms * 1e-3) / 1e9
        tr = ncu_traffic(args.config) or {}
        tr_stage = tr.get("stage_bytes_per_frame", {})
        tr_stream = tr.get("stream_kernel_bytes_per_frame")
        sk_ms = getattr(eng, "seg_stream_ms", 0)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": cfg["name"], "baseline_config": args.config, "frames_per_step_per_gpu": B,
                       "frames_timed": frames, "pool_frames_per_gpu": pool, "filter_mode": args.filter_mode,
                       "radius_mode": args.radius_mode, "filtered_entities": n_filtered, "exact_path_entities": n_exact,
                       "l2": "inputs larger than L2 (one step reads %d MB)" % (alg // 2 ** 20),
                       "parallelism": ("one sequence of %d frames, frame-sharded x%d: blocks of %d frames dealt round-robin, each behind a "
                                       "one-frame halo; the phase-2 state (512 B) goes from kernel to kernel around the ring of GPUs "
                                       "through peer memory (NVLink), one gather of the emitted columns at the end" % (frames, world, B))
                       if world > 1 else "one GPU",
                       "phase2": args.phase2, "contexts_per_gpu": C},
            "timed_blocks": {"blocks": len(runs), "steps_per_block": K, "median_ms": ms, "best_ms": ms_all[0],
                             "worst_ms": ms_all[-1], "best_value": frames / (ms_all[0] * 1e-3),
                             "total_timed_s": sum(ms_all) * 1e-3},
            "clocks": clocks, "gpu_launches": launches, "wall_ms_per_step": wall_ms / K, "esec_columns": cols,
            "checks": checks, "stages_ms_solo_step": stages,
            "roofline": {"bound": "hbm", "kernel": dom + " stage of one step", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                         "traffic": (tr_stage[dom] * B) if dom in tr_stage else None, "traffic_source": tr.get("capture"),
                         "peak_source": which, "whole_step_gbs_per_gpu": step_gbs, "whole_step_frac": step_gbs / peak,
                         # the one kernel that streams the input (label filter + compaction, seg_stream.cuh), timed alone with
                         # its own event pair: this is the HBM-bound kernel of the path; the entity-level stages work on the
                         # stored points out of L2
                         "streaming_kernel": ({"name": "seg_stream_kernel", "ms": sk_ms,
                                               "algorithmic_gbs": alg / (sk_ms * 1e-3) / 1e9,
                                               "frac": alg / (sk_ms * 1e-3) / 1e9 / peak,
                                               "dram_gbs": (tr_stream * B / (sk_ms * 1e-3) / 1e9) if tr_stream else None}
                                              if sk_ms > 0 else None),
                         "stages": {k: {"ms": stages[k], "algorithmic_gbs": alg / (stages[k] * 1e-3) / 1e9,
                                        "frac": alg / (stages[k] * 1e-3) / 1e9 / peak,
                                        "dram_gbs": (tr_stage[k] * B / (stages[k] * 1e-3) / 1e9) if k in tr_stage else None}
                                    for k in ("segment", "radius", "pairs") if stages.get(k, 0) > 0}},
        }
        if e2e:
            out["e2e"] = e2e
        if cpu:
            out["cpu_baseline"] = cpu
        print(json.dumps(out), flush=True)
    p2.close()
    for e in engs:
        e.close()
    if world > 1:
        dist.destroy_process_group()


def bind_near_gpu(torch, local):
    """Keep this rank on the cores of the NUMA node its GPU hangs off, before the pinned host pools are allocated: the
    pages of a pinned buffer land on the node of the thread that allocates them, and with eight ranks feeding eight PCIe
    links at once a pool on the far socket sends every byte over the inter-socket link first (the e2e figure at 8 GPUs).
    -> {"node": n, "cpus": count} or None when the box does not say (no sysfs entry, one node, or a cpuset without those cores)."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as fh:
            node = int(fh.read())
        if node < 0:
            return None
        cpus = set()
        with open("/sys/devices/system/node/node%d/cpulist" % node) as fh:
            for part in fh.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        mine = os.sched_getaffinity(0) & cpus
        if not mine:
            return None
        os.sched_setaffinity(0, mine)
        return {"node": node, "cpus": len(mine), "gpu": bdf}
    except Exception:
        return None


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
