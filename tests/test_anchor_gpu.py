"""Anchor index (sequences + sampled k-mer positions in front of the table): results are identical with and
without it, and identical to the oracle; it only changes how many table sectors are read."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _classify(db, d_off, d_cont, n, flags):
    import torch
    out = torch.zeros(n * 5, dtype=torch.int16, device="cuda")
    st = torch.zeros(2, dtype=torch.int64, device="cuda")
    db.query_device(d_off, d_cont, n, out.data_ptr(), 0, st.data_ptr(), flags)
    torch.cuda.synchronize()
    return out.cpu().numpy().view(np.uint16).reshape(n, 5), int(st[0]), int(st[1])


@pytest.mark.parametrize("k,L,G,LG", [(31, 150, 6, 60000), (31, 100, 3, 30000), (21, 140, 5, 50000), (31, 151, 4, 40000), (31, 200, 4, 40000)])
def test_anchor_path_equals_plain_path_and_oracle(oracle, k, L, G, LG):
    import torch
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS, CCB_NO_ANCHOR
    NR = 6000
    db = CuClarkDB(1, k, 1, G)
    db.synth_db(11, G, LG, G)
    info = db.anchor_synth(11, G, LG, G)
    assert info.ready and info.entries > 0.08 * G * LG and info.dropped < 0.08 * info.entries
    d_off, d_cont, nc = db.synth_reads_packed(22, 11, 0, NR, L, G, LG)
    plain, l0, s0 = _classify(db, d_off, d_cont, NR, CCB_STATS | CCB_NO_ANCHOR)
    fast, l1, s1 = _classify(db, d_off, d_cont, NR, CCB_STATS)
    assert np.array_equal(fast, plain)
    assert l1 == l0                                   # every k-mer is still accounted for
    if L - k + 1 <= 128:
        assert s1 < 0.6 * s0, (s0, s1)                # and most of them no longer touch the table
    else:
        assert s1 == s0                               # longer reads take the plain path for now
    # against the oracle
    off = torch.empty(NR + 1, dtype=torch.int32, device="cuda"); cont = torch.empty(nc, dtype=torch.int16, device="cuda")
    import ctypes as C
    rt = C.CDLL("libcudart.so")
    rt.cudaMemcpy(C.c_void_p(off.data_ptr()), C.c_void_p(d_off), C.c_size_t((NR + 1) * 4), 3)
    rt.cudaMemcpy(C.c_void_p(cont.data_ptr()), C.c_void_p(d_cont), C.c_size_t(nc * 2), 3)
    odb = oracle.OracleDb.synth(11, G, LG, G, k, 1000003)
    want, _, nl = odb.query_packed(k, off.cpu().numpy().view(np.uint32), cont.cpu().numpy().view(np.uint16))
    assert np.array_equal(fast, want) and nl == l1
    # dropping the index restores the plain path
    db.anchor_drop()
    again, _, s2 = _classify(db, d_off, d_cont, NR, CCB_STATS)
    assert np.array_equal(again, plain) and s2 == s0
    db.device_free(d_off); db.device_free(d_cont); db.close()
