"""Frame sharding of the e2SEC path across the GPUs of one box.

The reference's frame loop (manipulation_analyser.py:2271-2281) is sequential only through tiny state (o1/o2/o3
identity, sticky absent flags, previous centres, last column: globals :32-59).  Phase 1 (everything that touches
points) is per frame and needs just the clouds of the previous KEPT frame for the stale-cloud distances of
:1562-1658, so:
    * rank r takes the contiguous block  [F*r/W, F*(r+1)/W)  of the processed frames;
    * before its block it walks backwards from the block start until it has processed one kept frame (hand label
      present and filtered hand non-empty) -- the one-frame halo -- and drops those records;
    * one all-gather of the fixed-size records (torch.distributed: NCCL on the GPUs, gloo in the CPU tests)
      replaces any per-frame communication; rank 0 then runs the sequential resolver over all records.
No collective sits on the data path.
"""
import numpy as np

from . import _native as nat


def shard_range(n_frames, rank, world):
    return (n_frames * rank) // world, (n_frames * (rank + 1)) // world


def run_shard(engine, params, keepalive, source, frame_ids, rank, world, batch):
    """Phase 1 of this rank's block.  source.fetch(ids) -> (xyz (F,N,3) f32, labels (F,N) i32).  -> records."""
    frame_ids = np.asarray(frame_ids, np.int32)
    lo, hi = shard_range(len(frame_ids), rank, world)
    K = int(params.n_entities)
    out = np.empty((hi - lo, nat.record_doubles(K)), np.float64)
    if engine.max_frames is None or engine.max_frames < batch or engine._keep is None or engine._keep[0] is not params:
        xyz0, _ = source.fetch(frame_ids[:1])
        engine.configure(params, batch, xyz0.shape[1], keepalive)
    engine.reset_sequence()
    j = lo - 1
    while j >= 0:                                   # halo: the last kept frame before the block
        xyz, lab = source.fetch(frame_ids[j: j + 1])
        rec = engine.process_batch(xyz, lab, frame_ids[j: j + 1])
        if int(rec[0, 1]) == (nat.FLAG_HAND_LABEL | nat.FLAG_HAND_POINTS):
            break
        engine.reset_sequence()
        j -= 1
    for s in range(lo, hi, batch):
        ids = frame_ids[s: min(s + batch, hi)]
        xyz, lab = source.fetch(ids)
        out[s - lo: s - lo + len(ids)] = engine.process_batch(xyz, lab, ids)
    return out


def gather_records(local, n_frames, world, device=None):
    """All ranks' records in frame order on every rank.  One collective; shards may differ by one frame."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return np.asarray(local)
    per = max(shard_range(n_frames, r, world)[1] - shard_range(n_frames, r, world)[0] for r in range(world))
    width = local.shape[1]
    buf = torch.zeros((per, width), dtype=torch.float64, device=device)
    if len(local):
        buf[: len(local)] = torch.from_numpy(np.ascontiguousarray(local)).to(buf.device)
    full = torch.empty((world * per, width), dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(full, buf) if dist.get_backend() == "nccl" else \
        dist.all_gather(list(full.view(world, per, width).unbind(0)), buf)
    full = full.view(world, per, width).cpu().numpy()
    parts = []
    for r in range(world):
        a, b = shard_range(n_frames, r, world)
        parts.append(full[r, : b - a])
    return np.concatenate(parts)


# ---- one sequence dealt over the GPUs batch by batch (bench.py) ------------------------------------------------
def batch_owner(g, world):
    """Global batch g of the sequence runs on rank g % world as that rank's step g // world, behind a one-frame halo."""
    return g % world, g // world


def ring_tickets(g, n_batches, serial=0):
    """(wait, post) tickets of batch g for e2sec_phase2_run_ring: the kernel of batch g waits for the state the kernel of
    batch g - 1 posted, and posts its own for batch g + 1; 0 = nothing to wait for / post.  `serial` numbers the sequences
    run so far so that tickets never repeat."""
    base = serial * n_batches + 1
    return (base + g if g > 0 else 0), (base + g + 1 if g < n_batches - 1 else 0)


def gather_columns(columns, frames, world, device=None):
    """The one collective of the sharded run: every rank's emitted columns ((n, 32) uint8 relation codes + (n,) int32 frame
    ids) on rank 0, merged into frame order.  Other ranks get (None, None)."""
    import torch
    import torch.distributed as dist
    columns = np.ascontiguousarray(columns, np.uint8).reshape(-1, nat.PHASE2_COLUMN_BYTES)
    frames = np.ascontiguousarray(frames, np.int32).reshape(-1)
    if world == 1:
        order = np.argsort(frames, kind="stable")
        return columns[order], frames[order]
    rank = dist.get_rank()
    cnt = torch.tensor([len(frames)], dtype=torch.int64, device=device)
    counts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(counts, cnt)
    counts = [int(c.item()) for c in counts]
    width = nat.PHASE2_COLUMN_BYTES + 4
    pad = torch.zeros((max(max(counts), 1), width), dtype=torch.uint8, device=device)
    if len(frames):
        packed = np.concatenate([columns, frames.view(np.uint8).reshape(-1, 4)], axis=1)
        pad[: len(frames)] = torch.from_numpy(packed).to(pad.device)
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    if rank != 0:
        return None, None
    allp = np.concatenate([parts[r][: counts[r]].cpu().numpy() for r in range(world)])
    fr = np.ascontiguousarray(allp[:, nat.PHASE2_COLUMN_BYTES:]).view(np.int32).reshape(-1)
    order = np.argsort(fr, kind="stable")
    return np.ascontiguousarray(allp[order, : nat.PHASE2_COLUMN_BYTES]), fr[order]
