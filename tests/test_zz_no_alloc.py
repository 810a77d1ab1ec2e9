"""e2sec_icp_correspondences_static (include/e2sec.h): the ICP correspondence step with its scratch in a file-scope
__device__ array of the library -- no cudaMalloc / cudaFree inside the call.  Called through the C ABI with device
pointers and compared with e2sec_icp_correspondences (the allocating entry point, host pointers) on two calls in a row with
different sizes, so that what the first call left in the shared scratch cannot leak into the second.  The tolerance is the
one of the frame-0 tests for this operator's own kernels: the results are bit-identical (same kernels, same order).

Collected last on purpose; needs nothing from the tests before it."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _transform(rng, angle, shift):
    ax = rng.normal(size=3)
    ax /= np.linalg.norm(ax)
    Kx = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    T = np.eye(4)
    T[:3, :3] = np.eye(3) + np.sin(angle) * Kx + (1 - np.cos(angle)) * (Kx @ Kx)
    T[:3, 3] = shift
    return np.ascontiguousarray(T)


def _case(rng, ns, nt):
    tgt = np.ascontiguousarray(rng.uniform(-0.5, 0.5, (nt, 3)))
    src = np.ascontiguousarray(tgt[rng.integers(0, nt, ns)] + rng.normal(0, 0.01, (ns, 3)))
    src[: ns // 7] += 3.0                                   # points with no target within reach
    return src, tgt, _transform(rng, 0.02, rng.normal(0, 0.005, 3))


def test_static_scratch_entry_point_equals_allocating_one():
    import torch
    from manipulationlibrary_b200 import _native as nat
    from manipulationlibrary_b200.engine import Engine
    lib, ptr = nat.lib, nat.ptr
    eng = Engine(0)
    try:
        rng = np.random.default_rng(20261105)
        cases = [_case(rng, 6000, 4500), _case(rng, 1777, 9000)]       # the second is smaller: stale scratch beyond it
        want = []
        for src, tgt, T in cases:
            st = np.zeros(17)
            rc = lib.e2sec_icp_correspondences(eng._h, ptr(src), len(src), ptr(tgt), len(tgt), ptr(T), 0.05, ptr(st))
            assert rc == 0, lib.e2sec_last_error(eng._h).decode()
            assert 0 < st[0] < len(src), "the case should have matched and unmatched points"
            want.append(st)
        dev = [(torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()) for s, t, _ in cases]
        out = torch.full((2, 17), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        for k, ((src, tgt, T), (ds, dt)) in enumerate(zip(cases, dev)):                    # two calls in a row
            rc = lib.e2sec_icp_correspondences_static(eng._h, C.c_void_p(ds.data_ptr()), len(src), C.c_void_p(dt.data_ptr()), len(tgt),
                                                      ptr(T), 0.05, C.c_void_p(out[k].data_ptr()))
            assert rc == 0, lib.e2sec_last_error(eng._h).decode()
        free_after = torch.cuda.mem_get_info()[0]
        got = out.cpu().numpy()
        for k in range(2):
            assert np.array_equal(got[k], want[k]), (k, got[k], want[k])
        assert free_after == free_before, "the static entry point must not change the device's free memory"
        # beyond the scratch: a clear error, nothing launched
        n_big = nat.STATIC_ICP_MAX_SOURCE + 1
        rc = lib.e2sec_icp_correspondences_static(eng._h, C.c_void_p(dev[0][0].data_ptr()), n_big, C.c_void_p(dev[0][1].data_ptr()), 4500,
                                                  ptr(cases[0][2]), 0.05, C.c_void_p(out[0].data_ptr()))
        assert rc == -3                                                 # E2SEC_ERR_CAPACITY
        msg = lib.e2sec_last_error(eng._h).decode()
        assert "E2SEC_STATIC_ICP_MAX_SOURCE" in msg and str(n_big) in msg
        # and the entry point still works afterwards, with the first case again
        rc = lib.e2sec_icp_correspondences_static(eng._h, C.c_void_p(dev[0][0].data_ptr()), 6000, C.c_void_p(dev[0][1].data_ptr()), 4500,
                                                  ptr(cases[0][2]), 0.05, C.c_void_p(out[1].data_ptr()))
        assert rc == 0 and np.array_equal(out[1].cpu().numpy(), want[0])
    finally:
        eng.close()
