"""b200lsap_solve_batch_noalloc_f32 (include/b200lsap.h): the batched solve without any allocation, its scratch in a
file-scope device array.  Collected last on purpose.  Checked through the C-ABI
against the existing entry point b200lsap_solve_batch_device_f32: the batched kernels place the reference's own bids,
so the tolerance of every batched test applies — exact equality — on two calls in a row (the array keeps the last
call's contents)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(lib, t, batch, rows, cols, layout, out, status, stream=0, opts=None):
    return lib.b200lsap_solve_batch_noalloc_f32(t.device.index, C.c_void_p(stream), C.c_void_p(t.data_ptr()), batch, rows,
                                                cols, layout, opts, C.c_void_p(out.data_ptr()),
                                                C.c_void_p(status.data_ptr()) if status is not None else None)


@pytest.mark.parametrize("rows,cols,batch", [(32, 48, 1300), (20, 44, 257), (9, 13, 300), (32, 48, 40000)])
def test_noalloc_batch_equals_the_existing_entry_point_twice_in_a_row(handle, oracle_lib, rows, cols, batch):
    import torch
    import auctionalgorithm_b200 as ab
    lib = ab.lib()
    t = torch.empty((batch, rows, cols), dtype=torch.float32, device="cuda:0")
    handle.generate_device(t.data_ptr(), batch, rows, cols, cols, ab.ROW_MAJOR, seed=5, problem0=11)
    want = torch.empty((batch, rows), dtype=torch.int32, device="cuda:0")
    handle.solve_batch_device(t.data_ptr(), batch, rows, cols, want.data_ptr())
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):                      # the second call finds the first call's contents in the scratch array
        got = torch.full((batch, rows), -9, dtype=torch.int32, device="cuda:0")
        status = torch.full((1,), -9, dtype=torch.int32, device="cuda:0")
        rc = _call(lib, t, batch, rows, cols, ab.ROW_MAJOR, got, status, stream)
        assert rc == ab.OK, lib.b200lsap_noalloc_last_error()
        torch.cuda.synchronize()
        assert int(status.item()) == 0
        assert torch.equal(got, want), f"call {call}"
    # against the CPU oracle too (a sample), so that both entry points cannot be wrong together
    idx = np.unique(np.concatenate([np.arange(min(batch, 50)), np.arange(max(0, batch - 50), batch)]))
    a = np.stack([oracle_lib.generate(rows, cols, 5, problem=11 + int(b)) for b in idx])
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0 and np.array_equal(got.cpu().numpy()[idx], ref)


def test_noalloc_column_major_uses_the_static_staging_area_and_rejects_what_does_not_fit(handle):
    import torch
    import auctionalgorithm_b200 as ab
    from auctionalgorithm_b200 import _capi
    lib = ab.lib()
    rows, cols = 32, 48
    batch = _capi.NOALLOC_MAX_ELEMS // (rows * cols)            # 1365: the largest column-major batch
    t = torch.empty((batch, rows, cols), dtype=torch.float32, device="cuda:0")
    handle.generate_device(t.data_ptr(), batch, rows, cols, cols, ab.ROW_MAJOR, seed=2, problem0=0)
    want = torch.empty((batch, rows), dtype=torch.int32, device="cuda:0")
    handle.solve_batch_device(t.data_ptr(), batch, rows, cols, want.data_ptr())
    t_cm = t.transpose(1, 2).contiguous()                       # every problem column-major (Eigen's default)
    torch.cuda.synchronize()
    for call in range(2):
        got = torch.full((batch, rows), -9, dtype=torch.int32, device="cuda:0")
        status = torch.zeros(1, dtype=torch.int32, device="cuda:0")
        assert _call(lib, t_cm, batch, rows, cols, ab.COL_MAJOR, got, status) == ab.OK, lib.b200lsap_noalloc_last_error()
        torch.cuda.synchronize()
        assert torch.equal(got, want) and int(status.item()) == 0
    # one problem more does not fit the staging area: refused, nothing launched
    big = torch.zeros((batch + 1, cols, rows), dtype=torch.float32, device="cuda:0")
    out = torch.full((batch + 1, rows), -9, dtype=torch.int32, device="cuda:0")
    assert _call(lib, big, batch + 1, rows, cols, ab.COL_MAJOR, out, None) == ab.ERR_SHAPE
    assert b"B200LSAP_NOALLOC_MAX_ELEMS" in lib.b200lsap_noalloc_last_error()
    torch.cuda.synchronize()
    assert int(out.min().item()) == -9
    # other argument errors
    assert _call(lib, big, 1, 33, 40, ab.ROW_MAJOR, out, None) == ab.ERR_SHAPE
    assert lib.b200lsap_solve_batch_noalloc_f32(0, None, None, 1, 4, 4, 1, None, C.c_void_p(out.data_ptr()), None) == ab.ERR_ARG


def test_noalloc_reports_a_failed_problem_through_the_device_status_and_takes_a_side_stream(handle):
    import torch
    import auctionalgorithm_b200 as ab
    lib = ab.lib()
    a = np.ones((3, 4, 8), np.float32)
    a[1, 2] = 0                                                  # a row without an edge: the reference never returns
    t = torch.from_numpy(a).cuda()
    out = torch.full((3, 4), -9, dtype=torch.int32, device="cuda:0")
    status = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        assert _call(lib, t, 3, 4, 8, ab.ROW_MAJOR, out, status, s.cuda_stream) == ab.OK
    s.synchronize()
    assert int(status.item()) == ab.ERR_INFEASIBLE
    got = out.cpu().numpy()
    assert (got[1] == -1).all() and sorted(got[0]) == [0, 1, 2, 3] and sorted(got[2]) == [0, 1, 2, 3]
    # the next call starts from a cleared control block
    t2 = torch.from_numpy(np.ones((3, 4, 8), np.float32)).cuda()
    torch.cuda.synchronize()
    assert _call(lib, t2, 3, 4, 8, ab.ROW_MAJOR, out, status) == ab.OK
    torch.cuda.synchronize()
    assert int(status.item()) == 0 and int(out.min().item()) >= 0
