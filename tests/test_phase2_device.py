"""Phase 2 on the device (csrc/phase2.cuh) and the host resolver against the oracle and the reference's known answers:
  * the scripted scenes of tests/golden/relation_kats.npz (outputs of the reference's own analyse_maniac_manipulation) through
    CUDA records -> device phase 2, whole and cut into uneven pieces (the state carried between calls);
  * >= 96 frames (four hand cycles) of a small synthetic stream: CUDA records -> host resolver == oracle raw table,
    final_columns and o1/o2/o3; device phase 2 == host resolver;
  * a sequence longer than one 1024-frame round of the kernel; a sequence cut into shards whose state travels as a blob."""
import numpy as np
import pytest

from oracle import esec_oracle as eo
from tests.helpers import load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kats():
    z = load_golden("relation_kats")
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="module")
def engine():
    from manipulationlibrary_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()


def _dev(records):
    import torch
    return torch.from_numpy(np.ascontiguousarray(records)).cuda()


def _phase2_table(engine, records, cuts=None):
    """records (host) -> device phase 2 in the pieces `cuts` gives -> (table, frame ids, state)."""
    from manipulationlibrary_b200.engine import Phase2
    t = _dev(records)
    width = records.shape[1] * 8
    p2 = Phase2(engine, max_columns=max(16, len(records) + 1))
    edges = [0] + list(cuts or []) + [len(records)]
    for a, b in zip(edges[:-1], edges[1:]):
        if b > a:
            p2.run(t.data_ptr() + a * width, b - a)
    table, fc = p2.table()
    state = p2.state()
    p2.close()
    return table, fc, state


def test_device_phase2_matches_reference_sequences(kats, engine):
    from oracle.gen_relation_kats import scene_from_golden
    from tests.test_relation_kats import check_table, scene_params
    rng = np.random.default_rng(5)
    for k in range(int(kats["b_count"])):
        sc, exp = scene_from_golden(kats, k)
        p, keep = scene_params(sc, exp)
        F = len(sc.frames)
        engine.configure(p, F, 640 * 480, keep)
        xyz = np.empty((F, 640 * 480, 3), np.float32)
        lab = np.empty((F, 640 * 480), np.int32)
        for i in range(F):
            x, l = sc.dense(i)
            xyz[i], lab[i] = x, l.reshape(-1)
        rec = engine.process_batch(xyz, lab, np.arange(F, dtype=np.int32) * 3)
        for cuts in (None, sorted(rng.choice(np.arange(1, F), size=min(5, F - 1), replace=False).tolist())):
            table, fc, state = _phase2_table(engine, rec, cuts)
            if table.shape[1] == 0:
                table = np.full((10 * sc.relations, 1), b"-", dtype="S5")
            check_table(table, fc, state, sc, exp)


def _stream_records(engine, s, ids, batch=32):
    from manipulationlibrary_b200 import _native as nat
    p, keep = nat.make_params(**s.params())
    engine.configure(p, batch, s.N, keep)
    engine.reset_sequence()
    out = []
    for a in range(0, len(ids), batch):
        xyz, lab = s.fetch(ids[a: a + batch])
        out.append(engine.process_batch(xyz, lab, ids[a: a + batch]))
    return p, keep, np.concatenate(out)


def test_synthetic_cycles_resolver_and_phase2_equal_oracle(engine):
    from manipulationlibrary_b200.engine import Resolver
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=320, height=240, n_objects=4, config=99, cycle=24)
    ids = np.arange(0, 100, dtype=np.int32)
    p, keep, rec = _stream_records(engine, s, ids)
    orc = eo.E2SecOracle(s.ground_label, s.hand_label, None, s.thresh, s.translation, s.roll, s.ground,
                         s.total_unique_labels, filter_mode="closure")
    for f in ids:
        xyz, lab = s.frame(int(f))
        orc.process_frame(int(f), xyz, lab.reshape(s.H, s.W))
    want = orc.table
    assert want is not None and want.shape[1] >= 8, "the stream should produce a real table"
    r = Resolver(p, keep)
    r.feed(rec)
    table, fc = r.table()
    assert table.shape == want.shape and (table.astype("U5") == want).all()
    assert fc == list(orc.final_columns)
    labels, absent = r.state()
    assert labels == [orc.olabel[k] if orc.count[k] else -1 for k in "123"]
    assert absent == [orc.absent[k] for k in "123"]
    assert orc.count["1"] and orc.count["2"], "the scripted hand should bind at least o1 and o2"
    for cuts in (None, [1, 2, 37, 64]):
        t2, fc2, st2 = _phase2_table(engine, rec, cuts)
        assert t2.shape == table.shape and (t2 == table).all() and fc2 == fc and st2 == (labels, absent)
    r.close()


def test_phase2_rounds_and_shard_states(engine):
    """More frames than one round of the kernel (1024), and the same records cut into three shards whose state travels
    as the opaque blob (host and device copies): identical columns."""
    import torch
    from manipulationlibrary_b200.engine import Phase2, Resolver, columns_table
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=160, height=120, n_objects=3, config=98, cycle=16)
    ids = np.arange(0, 96, dtype=np.int32)
    p, keep, rec96 = _stream_records(engine, s, ids)
    # 12 repetitions with fresh frame ids: a periodic 1152-frame sequence (the cross tables of each repetition's first
    # frame are made consistent by re-running the join frames below)
    reps = 12
    rec = np.concatenate([rec96] * reps)
    rec[:, 0] = np.arange(len(rec))
    kept = (rec[:, 1].astype(int) & 3) == 3
    prev = -1
    for i in range(len(rec)):                       # previous-kept-frame ids consistent with the new numbering
        if kept[i]:
            rec[i, 3] = prev
            prev = i
        else:
            rec[i, 3] = prev
    r = Resolver(p, keep)
    r.feed(rec)
    table, fc = r.table()
    assert table.shape[1] > reps
    t2, fc2, st2 = _phase2_table(engine, rec)
    assert (t2 == table).all() and fc2 == fc and st2 == r.state()
    # three shards, the state handed on through a host blob and through device memory
    t = _dev(rec)
    width = rec.shape[1] * 8
    edges = [0, 300, 1030, len(rec)]
    cols, frames = [], []
    blob = None
    dev_blob = torch.zeros(nat.PHASE2_STATE_BYTES, dtype=torch.uint8, device="cuda")
    for j in range(3):
        p2 = Phase2(engine, max_columns=2048)
        if j == 1:
            p2.state_set(blob)
        if j == 2:
            p2.state_set(dev_blob.data_ptr())
        p2.run(t.data_ptr() + edges[j] * width, edges[j + 1] - edges[j])
        c, f = p2.fetch()
        cols.append(c)
        frames.append(f)
        blob = p2.state_get()
        p2.state_get(dev_blob.data_ptr())
        if j == 2:
            assert p2.state() == r.state()
        p2.close()
    t3, fc3 = columns_table(np.concatenate(cols), np.concatenate(frames), 3)
    assert (t3 == table).all() and fc3 == fc
    r.close()


def test_phase2_reports_out_of_order_records(engine):
    from manipulationlibrary_b200.engine import Phase2
    from manipulationlibrary_b200._native import E2SecError
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=160, height=120, n_objects=3, config=98, cycle=16)
    ids = np.arange(0, 12, dtype=np.int32)
    p, keep, rec = _stream_records(engine, s, ids)
    t = _dev(rec[4:])                                # starts mid-sequence without its halo
    p2 = Phase2(engine, max_columns=64)
    p2.run(t.data_ptr(), len(rec) - 4)
    with pytest.raises(E2SecError):
        p2.sync()
    p2.close()


def test_phase2_ring_through_mailboxes(engine):
    """The state handed from kernel to kernel through mailboxes (e2sec_phase2_run_ring): pieces of one sequence dealt in
    turn to three Phase2 objects whose kernels wait for / post tickets on the device -- the columns, merged by frame id,
    equal the single object's.  (Across processes the mailbox is opened through its CUDA IPC handle; bench.py does that.)"""
    from manipulationlibrary_b200.engine import Phase2, Resolver, columns_table
    from manipulationlibrary_b200.synthetic import SyntheticStream
    s = SyntheticStream(width=160, height=120, n_objects=3, config=98, cycle=16)
    ids = np.arange(0, 96, dtype=np.int32)
    p, keep, rec = _stream_records(engine, s, ids)
    r = Resolver(p, keep)
    r.feed(rec)
    table, fc = r.table()
    t = _dev(rec)
    width = rec.shape[1] * 8
    ring = [Phase2(engine, max_columns=128) for _ in range(3)]
    for i, q in enumerate(ring):
        q.connect_local(ring[(i + 1) % 3])
    edges = list(range(0, 96, 8)) + [96]
    n = len(edges) - 1
    for g in range(n):
        ring[g % 3].run_ring(t.data_ptr() + edges[g] * width, edges[g + 1] - edges[g], 100 + g if g else 0, 101 + g if g < n - 1 else 0)
    cols, frames = [], []
    for q in ring:
        c, f = q.fetch()
        cols.append(c)
        frames.append(f)
    cols, frames = np.concatenate(cols), np.concatenate(frames)
    order = np.argsort(frames, kind="stable")
    t3, fc3 = columns_table(cols[order], frames[order], 3)
    assert (t3 == table).all() and fc3 == fc
    assert ring[(n - 1) % 3].state() == r.state()
    for q in ring:
        q.close()
    r.close()
