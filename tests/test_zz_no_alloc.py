"""ccb_query_device_noalloc (include/cuclark_b200.h): the device-resident query on the library's fixed scratch array
instead of buffers the handle allocates.  No counterpart in the reference,
so the checker is the existing entry point: ccb_query_device on the same handle and inputs, bit for bit
(u16[5] per read, sparse rows, the statistics counters) -- the tolerance every integer test here uses -- on calls in a
row, which share the array and find the previous call's contents in it.

Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

K, G, LG = 31, 6, 60000


def _run(db, fn, batch, n, flags, rows):
    import torch
    from jn_cuclark_b200 import CCB_STATS, CCB_STATS_EX
    d_off, d_cont = batch
    out = torch.full((n * 5,), 0x5A5A, dtype=torch.int16, device="cuda")
    full = torch.full((n * db.row_stride,), 0x5A5A, dtype=torch.int16, device="cuda") if rows else None
    st = torch.zeros(4, dtype=torch.int64, device="cuda")
    fn(d_off, d_cont, n, out.data_ptr(), full.data_ptr() if rows else 0, st.data_ptr(), flags | CCB_STATS | CCB_STATS_EX,
       torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return (out.cpu().numpy().view(np.uint16).reshape(n, 5),
            full.cpu().numpy().view(np.uint16).reshape(n, db.row_stride) if rows else None,
            st.cpu().numpy().copy())


def _same(a, b):
    assert np.array_equal(a[0], b[0])
    assert (a[1] is None) == (b[1] is None) and (a[1] is None or np.array_equal(a[1], b[1]))
    assert np.array_equal(a[2], b[2]), (a[2], b[2])


def test_noalloc_entry_point_equals_query_device_on_calls_in_a_row():
    import torch
    from jn_cuclark_b200 import CuClarkDB, CCB_NO_ANCHOR, CCB_FORCE_DENSE, CCB_WARP_KERNEL
    from jn_cuclark_b200.lib import CcbError
    db = CuClarkDB(1, K, 1, G)
    db.synth_db(11, G, LG, G)
    assert db.anchor_synth(11, G, LG, G).ready          # so that the anchor pass's per-read records are in play
    NA, NB = 6000, 1500
    a = db.synth_reads_packed(22, 11, 0, NA, 150, G, LG)        # short reads: anchor scan + block kernel
    b = db.synth_reads_packed(23, 11, 0, NB, 300, G, LG)        # longer than the block kernel takes: redo list + warp kernel
    A, B = (a[0], a[1]), (b[0], b[1])
    cases = [(A, NA, 0, False), (B, NB, 0, True), (A, NA, 0, True), (A, NA, CCB_NO_ANCHOR, False),
             (B, NB, CCB_FORCE_DENSE, False), (A, NA, CCB_FORCE_DENSE, True), (A, NA, CCB_WARP_KERNEL, False), (A, NA, 0, False)]
    want = [_run(db, db.query_device, bt, n, fl, rows) for bt, n, fl, rows in cases]
    assert int((want[0][0][:, 1] != 0).sum()) > NA // 2         # the batches do classify
    assert 0.99 * NA * (150 - K + 1) < want[0][2][0] <= NA * (150 - K + 1)      # (one read in a thousand holds an N)
    # set-up call (looks the array's address up on this device), then calls back to back on the shared array:
    # each one finds what the one before left in it
    db.query_device_noalloc(0, 0, 0, 0)
    for _ in range(2):
        for (bt, n, fl, rows), w in zip(cases, want):
            _same(_run(db, db.query_device_noalloc, bt, n, fl, rows), w)
    # and it allocates nothing: the device's free memory does not move over further calls
    # (after one round of the same calls: the driver loads a kernel's code on its first launch -- here the variants without
    # statistics counters -- and that takes device memory once per process, whoever launches it)
    out = torch.zeros(NA * 5, dtype=torch.int16, device="cuda")
    db.query_device_noalloc(A[0], A[1], NA, out.data_ptr())
    db.query_device_noalloc(B[0], B[1], NB, out.data_ptr())
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        db.query_device_noalloc(A[0], A[1], NA, out.data_ptr())
        db.query_device_noalloc(B[0], B[1], NB, out.data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    assert np.array_equal(out.cpu().numpy().view(np.uint16)[:NB * 5].reshape(NB, 5), want[1][0])
    # sizes beyond the fixed scratch are refused before anything is launched
    with pytest.raises(CcbError, match="65537 reads in one call, the fixed scratch holds 65536"):
        db.query_device_noalloc(A[0], A[1], 65537, out.data_ptr())
    _same(_run(db, db.query_device_noalloc, A, NA, 0, False), want[0])      # the handle is still good
    for p in a[:2] + b[:2]:
        db.device_free(p)
    db.close()
    wide = CuClarkDB(1, K, 1, 8193)
    with pytest.raises(CcbError, match="8193 targets, the fixed scratch holds 8192"):
        wide.query_device_noalloc(0, 0, 1, 0)
    wide.close()


def test_noalloc_entry_point_at_its_limit():
    """CCB_NOALLOC_MAX_READS reads in one call: the lists and the per-read records fill the array to its last slot."""
    from jn_cuclark_b200 import CuClarkDB
    db = CuClarkDB(1, K, 1, G)
    db.synth_db(11, G, LG, G)
    assert db.anchor_synth(11, G, LG, G).ready
    N = 65536
    d_off, d_cont, _ = db.synth_reads_packed(24, 11, 0, N, 150, G, LG)
    want = _run(db, db.query_device, (d_off, d_cont), N, 0, True)
    for _ in range(2):
        _same(_run(db, db.query_device_noalloc, (d_off, d_cont), N, 0, True), want)
    db.device_free(d_off); db.device_free(d_cont); db.close()
