"""Halo exchange on the GPU in the reference's own test idiom (src/core_test/mpas_test_core_halo_exch.F:631 fill, :1010
validate): every owned element is set to its global ID, every halo element to -1; after mpas_dmpar_exch_halo_field
(here mpas_b200_dycore_exchange: pack kernel -> message -> unpack kernel) every halo element of the selected layers must
hold the global ID of the element it mirrors, and the layers NOT selected must be untouched.  Cell, edge and vertex
fields, a 3-D field (scalars: num_scalars x nVertLevels per cell, mpas_dmpar.F:5565-5748), layer selections all / (/1/) /
(/2/) as atm_srk3 uses them (TI:642, 792, 845, 876-887).  Several blocks in one process (the reference's no-MPI copy
path, mpas_dmpar.F:5481-5504); the same check across processes over NCCL and over the peer-store transport is
tests/test_gpu_multirank.py."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

FIELDS = [("theta_m@1", "cell"), ("u@1", "edge"), ("vorticity", "vertex"), ("scalars@1", "cell"), ("w@1", "cell")]
IDNAME = {"cell": "indexToCellID", "edge": "indexToEdgeID", "vertex": "indexToVertexID"}
OWN = {"cell": "nCellsSolve", "edge": "nEdgesSolve", "vertex": "nVerticesSolve"}


def halo_layers(block, kind, n):
    """layer of every local element (0 = owned) from the recv lists (mpas_dmpar_get_exch_list, mpas_dmpar.F:1483-2038)"""
    lay = np.zeros(n, dtype=np.int32)
    for x in block.exchange_lists(kind, "recv"):
        lay[x["dest"] - 1] = x["layer"]
    return lay


def fill_exchange_validate(model, blocks, name, kind, layer, get, put, exchange):
    for ib, b in enumerate(blocks):
        a = np.full(tuple(reversed(b.info(name)[2])), -1.0)
        nown = int(b.cfg()[OWN[kind]])
        ids = b.field(IDNAME[kind])
        a[:nown] = ids[:nown].reshape((nown,) + (1,) * (a.ndim - 1))
        put(ib, name, a)
    exchange(name, layer)
    checked = 0
    for ib, b in enumerate(blocks):
        a = get(ib, name)
        ids = b.field(IDNAME[kind])[:-1].astype(np.float64)
        n = len(ids)
        lay = halo_layers(b, kind, n)
        nown = int(b.cfg()[OWN[kind]])
        assert (lay[:nown] == 0).all() and (lay[nown:] > 0).all(), "every halo element is in exactly one recv list"
        want = np.where((lay == 0) | (layer == 0) | (lay == layer), ids, -1.0)
        want = np.broadcast_to(want.reshape((n,) + (1,) * (a.ndim - 1)), a[:n].shape)
        assert np.array_equal(a[:n], want), (name, layer, ib)
        checked += int((lay > 0).sum())
    return checked


@pytest.fixture(scope="module")
def model():
    case, blocks = common.make_case(4, 4, nVertLevels=8, num_scalars=3, method="metis")
    gpu = common.GpuModel(blocks, common.level_config(4))
    return blocks, gpu


@pytest.mark.parametrize("name,kind", FIELDS)
@pytest.mark.parametrize("layer", [0, 1, 2])
def test_halo_exchange_reference_idiom_on_the_gpu(model, name, kind, layer):
    blocks, gpu = model
    n = fill_exchange_validate(gpu, blocks, name, kind, layer, gpu.field, gpu.set_field, gpu.exchange)
    assert n > 100


def test_edge_and_vertex_fields_have_a_third_halo_layer(model):
    blocks, gpu = model
    for kind in ("edge", "vertex"):
        layers = set()
        for b in blocks:
            layers |= {x["layer"] for x in b.exchange_lists(kind, "recv")}
        assert layers == {1, 2, 3}, (kind, layers)
    n = fill_exchange_validate(gpu, blocks, "u@1", "edge", 3, gpu.field, gpu.set_field, gpu.exchange)
    assert n > 100


def test_layer1_exchange_of_registered_host_arrays_with_host_built_lists():
    """mpas_b200_halo_register_lists + mpas_b200_exch_halo_field2d / 3d: the host's own send / recv lists (here the lists of
    block 0 of a 2-block decomposition, turned into the reference's copyList form: endPoint = own rank, so the message is a
    device copy) applied to arrays registered by HOST address.  What the receiver's halo column destList(i) must hold
    afterwards is the sender-side column that was packed into buffer slot srcList(i): the reference's buffer addressing
    (mpas_dmpar.F:5462 pack at (destList-1)*dim1, :5521 unpack from (srcList-1)*dim1)."""
    import ctypes as C
    case, blocks = common.make_case(4, 2, nVertLevels=8, num_scalars=3, method="metis")
    b = blocks[0]
    Lb = common.host.lib()
    Lb.mpas_b200_halo_register_lists.argtypes = [C.c_int] * 3 + [C.c_void_p] * 5
    Lb.mpas_b200_exch_halo_field2d.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    Lb.mpas_b200_exch_halo_field3d.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    assert Lb.mpas_b200_halo_comm_init(0, 1, None) == 0
    # fabricate a self exchange per layer: the recv lists of block 0 (from block 1) paired with send lists of the same length
    # that pack OWNED columns 1..n of block 0 itself
    for kind_i, kind in ((1, "cell"), (2, "edge")):
        recv = b.exchange_lists(kind, "recv")
        ep = np.zeros(len(recv), dtype=np.int32)
        lay = np.array([x["layer"] for x in recv], dtype=np.int32)
        nl = np.array([len(x["dest"]) for x in recv], dtype=np.int32)
        r_src = np.concatenate([x["src"] for x in recv]).astype(np.int32)
        r_dst = np.concatenate([x["dest"] for x in recv]).astype(np.int32)
        s_src = np.concatenate([np.arange(1, n + 1) for n in nl]).astype(np.int32)   # owned columns 1..n
        s_dst = np.concatenate([np.random.default_rng(3 + i).permutation(n) + 1 for i, n in enumerate(nl)]).astype(np.int32)  # buffer slots
        for which, (s, d) in ((0, (s_src, s_dst)), (1, (r_src, r_dst))):
            rc = Lb.mpas_b200_halo_register_lists(kind_i, which, len(recv), ep.ctypes.data, lay.ctypes.data, nl.ctypes.data, s.ctypes.data, d.ctypes.data)
            assert rc == 0, common.host.last_error()
        n_el = int(b.cfg()["nCells" if kind == "cell" else "nEdges"]) + 1
        rng = np.random.default_rng(11)
        for dims, mask in (((8,), 0), ((8,), 1), ((8,), 2), ((8, 3), 0), ((8, 3), 2)):
            a = rng.uniform(1, 2, (n_el,) + tuple(reversed(dims)))
            before = a.copy()
            assert Lb.mpas_b200_register_field(a.ctypes.data, a.nbytes) == 0
            assert Lb.mpas_b200_upload(a.ctypes.data) == 0
            if len(dims) == 1:
                rc = Lb.mpas_b200_exch_halo_field2d(a.ctypes.data, dims[0], kind_i, mask)
            else:
                rc = Lb.mpas_b200_exch_halo_field3d(a.ctypes.data, dims[1], dims[0], kind_i, mask)  # Fortran order: fastest first
            assert rc == 0, common.host.last_error()
            assert Lb.mpas_b200_download(a.ctypes.data) == 0 and Lb.mpas_b200_synchronize() == 0
            want = before.copy()
            off = 0
            for i, n in enumerate(nl):
                if mask == 0 or (mask >> (lay[i] - 1)) & 1:
                    buf = np.empty((n,) + before.shape[1:])
                    buf[s_dst[off:off + n] - 1] = before[s_src[off:off + n] - 1]
                    want[r_dst[off:off + n] - 1] = buf[r_src[off:off + n] - 1]
                off += n
            assert np.array_equal(a, want), (kind, dims, mask)
            assert not np.array_equal(a, before)
            assert Lb.mpas_b200_unregister_field(a.ctypes.data) == 0
    assert Lb.mpas_b200_halo_finalize() == 0
