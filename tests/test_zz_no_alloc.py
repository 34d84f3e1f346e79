"""mpas_b200_halo_copy_no_alloc (include/mpas_b200.h; no counterpart in the reference, added for callers that cannot call
cudaMalloc): the same-process halo copy with the message buffer in a fixed device array of the library.  Its output must be
bitwise what the existing entry point mpas_b200_exch_halo_field2d gives for the same lists in the reference's copyList
form (endPoint = own rank, mpas_dmpar.F:5481-5504), on two calls in a row that share the buffer (the second one shorter
and with other slots, so stale contents of the first would show), and device memory in use must not change across the
calls.  Collected last on purpose (pytest -x): a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

MAX_REALS = 1 << 20  # MPAS_B200_NO_ALLOC_MAX_REALS


def _lib():
    L = common.host.lib()
    L.mpas_b200_halo_register_lists.argtypes = [C.c_int] * 3 + [C.c_void_p] * 5
    L.mpas_b200_exch_halo_field2d.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    L.mpas_b200_halo_copy_no_alloc.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 4
    L.mpas_b200_halo_copy_no_alloc.restype = C.c_int
    L.mpas_b200_register_field.argtypes = [C.c_void_p, C.c_size_t]
    for f in ("mpas_b200_upload", "mpas_b200_download", "mpas_b200_unregister_field"):
        getattr(L, f).argtypes = [C.c_void_p]
    return L


def _lists(rng, n_cols, n):
    """one send list (owned columns -> buffer slots) and one recv list (buffer slots -> halo columns), 1-based"""
    cols = rng.permutation(n_cols) + 1
    send_src = cols[:n].astype(np.int32)                  # columns packed
    recv_dst = cols[n:2 * n].astype(np.int32)             # columns overwritten (disjoint from the packed ones)
    send_dst = (rng.permutation(n) + 1).astype(np.int32)  # slot of each packed column
    recv_src = (rng.permutation(n) + 1).astype(np.int32)  # slot each halo column is read from
    return send_src, send_dst, recv_src, recv_dst


def _existing(L, a, dim1, lists):
    """the existing entry point: the lists registered as one copyList pair of halo layer 1, then exch_halo_field2d"""
    send_src, send_dst, recv_src, recv_dst = lists
    one = np.array([len(send_src)], dtype=np.int32)
    ep, lay = np.zeros(1, dtype=np.int32), np.ones(1, dtype=np.int32)
    assert L.mpas_b200_halo_comm_init(0, 1, None) == 0
    for which, (s, d) in ((0, (send_src, send_dst)), (1, (recv_src, recv_dst))):
        assert L.mpas_b200_halo_register_lists(1, which, 1, ep.ctypes.data, lay.ctypes.data, one.ctypes.data, s.ctypes.data, d.ctypes.data) == 0, common.host.last_error()
    out = a.copy()
    assert L.mpas_b200_register_field(out.ctypes.data, out.nbytes) == 0 and L.mpas_b200_upload(out.ctypes.data) == 0
    assert L.mpas_b200_exch_halo_field2d(out.ctypes.data, dim1, 1, 0) == 0, common.host.last_error()
    assert L.mpas_b200_download(out.ctypes.data) == 0 and L.mpas_b200_synchronize() == 0
    assert L.mpas_b200_unregister_field(out.ctypes.data) == 0
    assert L.mpas_b200_halo_finalize() == 0
    return out


class _Registered:
    """host arrays registered with the library (device mirrors made up front: the call under test allocates nothing)"""

    def __init__(self, L, arrays):
        self.L, self.arrays = L, arrays

    def __enter__(self):
        for x in self.arrays:
            assert self.L.mpas_b200_register_field(x.ctypes.data, x.nbytes) == 0, common.host.last_error()
            assert self.L.mpas_b200_upload(x.ctypes.data) == 0
        return self

    def __exit__(self, *exc):
        for x in self.arrays:
            self.L.mpas_b200_unregister_field(x.ctypes.data)


def _free_bytes():
    import torch
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def test_no_alloc_halo_copy_matches_the_exchange_entry_point_on_two_calls_in_a_row():
    L = _lib()
    rng = np.random.default_rng(2024)
    n_cols = 6000
    # call 1: 26 levels, 2500 columns; call 2: 55 levels, 900 columns, other slots -- both inside the fixed buffer
    cases = [(26, 2500), (55, 900)]
    inputs = [(d, rng.uniform(-1.0, 1.0, (n_cols + 1, d)), _lists(rng, n_cols, n)) for d, n in cases]
    want = [_existing(L, a, d, lists) for d, a, lists in inputs]
    got = [a.copy() for _, a, _ in inputs]
    flat = [x for g, (_, _, lists) in zip(got, inputs) for x in (g,) + lists]
    with _Registered(L, flat):
        # warm-up call on a scratch copy: the first call on a device looks the buffer's address up (cudaGetSymbolAddress)
        warm = inputs[0][1].copy()
        with _Registered(L, [warm]):
            assert L.mpas_b200_halo_copy_no_alloc(warm.ctypes.data, inputs[0][0], cases[0][1], *[x.ctypes.data for x in inputs[0][2]]) == 0, common.host.last_error()
            assert L.mpas_b200_synchronize() == 0
        free0 = _free_bytes()
        for g, (d, _, lists), (_, n) in zip(got, inputs, cases):  # two calls in a row, nothing in between
            assert L.mpas_b200_halo_copy_no_alloc(g.ctypes.data, d, n, *[x.ctypes.data for x in lists]) == 0, common.host.last_error()
        assert L.mpas_b200_synchronize() == 0
        free1 = _free_bytes()
        if free1 != free0:  # cudaMemGetInfo is device-wide: something else on the box may have moved; the calls must be neutral when repeated
            for (d, a, lists), (_, n) in zip(inputs, cases):
                w = a.copy()
                with _Registered(L, [w]):
                    free0 = _free_bytes()
                    assert L.mpas_b200_halo_copy_no_alloc(w.ctypes.data, d, n, *[x.ctypes.data for x in lists]) == 0, common.host.last_error()
                    assert L.mpas_b200_synchronize() == 0
                    free1 = _free_bytes()
                assert free1 == free0, "device memory in use changed across a no-alloc call"
        for g in got:
            assert L.mpas_b200_download(g.ctypes.data) == 0
        assert L.mpas_b200_synchronize() == 0
    for g, w, (_, a, _) in zip(got, want, inputs):
        assert np.array_equal(g, w)          # bitwise: a copy has no tolerance
        assert not np.array_equal(g, a)      # and something moved


def test_no_alloc_halo_copy_rejects_sizes_beyond_the_fixed_buffer():
    L = _lib()
    dim1, n = 64, MAX_REALS // 64 + 1
    a = np.zeros((n + 1, dim1))
    idx = np.arange(1, n + 1, dtype=np.int32)
    rc = L.mpas_b200_halo_copy_no_alloc(a.ctypes.data, dim1, n, idx.ctypes.data, idx.ctypes.data, idx.ctypes.data, idx.ctypes.data)
    assert rc != 0
    msg = common.host.last_error()
    assert "MPAS_B200_NO_ALLOC_MAX_REALS" in msg and str(MAX_REALS) in msg, msg
    # exactly the limit is accepted (n * dim1 == MAX_REALS), a zero-length list is a no-op
    n = MAX_REALS // dim1
    a = np.random.default_rng(5).uniform(0, 1, (2 * n + 1, dim1))
    src = np.arange(1, n + 1, dtype=np.int32)
    dst = np.arange(n + 1, 2 * n + 1, dtype=np.int32)
    before = a.copy()
    with _Registered(L, [a, src, dst]):
        assert L.mpas_b200_halo_copy_no_alloc(a.ctypes.data, dim1, 0, src.ctypes.data, src.ctypes.data, src.ctypes.data, dst.ctypes.data) == 0
        assert L.mpas_b200_halo_copy_no_alloc(a.ctypes.data, dim1, n, src.ctypes.data, src.ctypes.data, src.ctypes.data, dst.ctypes.data) == 0, common.host.last_error()
        assert L.mpas_b200_download(a.ctypes.data) == 0 and L.mpas_b200_synchronize() == 0
    assert np.array_equal(a[n:2 * n], before[:n]) and np.array_equal(a[:n], before[:n])
