"""Host-side decomposition and halo-list contract plus the reference's own
halo test idiom (src/core_test/mpas_test_core_halo_exch.F:631 setup, :1010 validate): fill owned elements
with their global ID, exchange, every halo element must then hold the global ID of the element it mirrors."""
import os

import numpy as np
import pytest

import common


@pytest.fixture(scope="module")
def case4():
    case, blocks = common.make_case(4, 4, nVertLevels=8, num_scalars=1)
    return case, blocks


def test_sfc_permutation_is_a_bijection_with_exact_inverse():
    case = common.host.HostCase(4, lloyd_iters=5, sfc=True)
    for kind, n in (("cell", case.nCells), ("edge", case.nEdges), ("vertex", case.nVertices)):
        p = case.permutation(kind)
        assert sorted(p.tolist()) == list(range(n))
        inv = np.empty_like(p)
        inv[p] = np.arange(n, dtype=p.dtype)
        assert np.array_equal(inv[p], np.arange(n))
        assert np.array_equal(p[inv], np.arange(n))


def test_sfc_mesh_equals_generator_mesh_up_to_permutation():
    a = common.host.HostCase(3, lloyd_iters=0, sfc=False)
    b = common.host.HostCase(3, lloyd_iters=0, sfc=True)
    ba = a.build_block(0, nVertLevels=4, full_fields=False)
    bb = b.build_block(0, nVertLevels=4, full_fields=False)
    p = b.permutation("cell")  # new -> old
    for name in ("areaCell", "latCell", "lonCell"):
        assert np.array_equal(bb.field(name)[:-1], ba.field(name)[:-1][p]), name
    # a block renumbers its edges in first-encounter order (mpas_block_decomp.F:369), so go through the global IDs
    pe = b.permutation("edge")
    ida, idb = ba.field("indexToEdgeID")[:-1], bb.field("indexToEdgeID")[:-1]
    for name in ("dvEdge", "dcEdge"):
        ga = np.empty(a.nEdges); ga[ida - 1] = ba.field(name)[:-1]  # generator numbering
        gb = np.empty(b.nEdges); gb[idb - 1] = bb.field(name)[:-1]  # SFC numbering
        assert np.array_equal(gb, ga[pe]), name


def test_local_ordering_rules(case4):
    case, blocks = case4
    part = np.zeros(case.nCells, dtype=np.int32)
    common.host.hostlib().mpas_host_case_get_partition(case.h, part.ctypes.data)
    for ib, b in enumerate(blocks):
        cfg = b.cfg()
        ids = b.field("indexToCellID")[:-1]
        nown = int(cfg["nCellsSolve"])
        # owned cells: exactly the cells of this block, ascending global ID (mpas_block_creator.F:99)
        assert np.array_equal(ids[:nown], np.nonzero(part == ib)[0] + 1)
        # halo cells belong to other blocks, no duplicates
        assert (part[ids[nown:] - 1] != ib).all()
        assert len(set(ids.tolist())) == len(ids)
        # an edge is owned iff its first cellsOnEdge entry is an owned cell (mpas_block_decomp.F:301-356)
        coe = b.field("cellsOnEdge")[:-1]
        nEown = int(cfg["nEdgesSolve"])
        assert (coe[:nEown, 0] <= nown).all()
        assert (coe[nEown:, 0] > nown).all()


def test_exchange_lists_are_in_ascending_global_id_order(case4):
    case, blocks = case4
    for b in blocks:
        for kind, idname in (("cell", "indexToCellID"), ("edge", "indexToEdgeID"), ("vertex", "indexToVertexID")):
            ids = b.field(idname)
            for x in b.exchange_lists(kind, "send"):
                order = np.argsort(x["dest"])
                g = ids[x["src"][order] - 1]
                assert (np.diff(g) > 0).all()
                assert sorted(x["dest"].tolist()) == list(range(1, len(g) + 1))
            for x in b.exchange_lists(kind, "recv"):
                order = np.argsort(x["src"])
                g = ids[x["dest"][order] - 1]
                assert (np.diff(g) > 0).all()


def test_send_and_recv_lists_pair_up(case4):
    case, blocks = case4
    for kind, idname in (("cell", "indexToCellID"), ("edge", "indexToEdgeID"), ("vertex", "indexToVertexID")):
        for ib, b in enumerate(blocks):
            for x in b.exchange_lists(kind, "send"):
                peer = blocks[x["endPoint"]]
                match = [y for y in peer.exchange_lists(kind, "recv") if y["endPoint"] == ib and y["layer"] == x["layer"]]
                assert len(match) == 1
                y = match[0]
                msg = np.zeros(len(x["src"]), dtype=np.int64)
                msg[x["dest"] - 1] = b.field(idname)[x["src"] - 1]
                assert np.array_equal(msg[y["src"] - 1], peer.field(idname)[y["dest"] - 1])


def test_halo_exchange_reference_idiom_oracle(case4):
    case, blocks = case4
    m = common.OracleModel(blocks, common.level_config(4))
    for name, idname, own in (("theta_m@1", "indexToCellID", "nCellsSolve"), ("u@1", "indexToEdgeID", "nEdgesSolve"),
                              ("vorticity", "indexToVertexID", "nVerticesSolve")):
        for ib, b in enumerate(blocks):
            a = m.field(ib, name)
            a[:] = -1.0
            nown = int(b.cfg()[own])
            a[:nown, :] = b.field(idname)[:nown, None]
        m.exchange(name)
        for ib, b in enumerate(blocks):
            a = m.field(ib, name)
            ids = b.field(idname)
            assert np.array_equal(a[:-1], np.broadcast_to(ids[:-1, None].astype(float), a[:-1].shape)), (name, ib)


def test_partition_file_round_trip(tmp_path):
    case = common.host.HostCase(3, lloyd_iters=0, sfc=True)
    prefix = str(tmp_path / "graph.info.part.")
    p1 = case.partition(3, "metis", prefix)
    assert os.path.exists(prefix + "3") and os.path.exists(str(tmp_path / "graph.info"))
    lines = open(prefix + "3").read().split()
    assert len(lines) == case.nCells and [int(x) for x in lines] == p1.tolist()
    head = open(str(tmp_path / "graph.info")).readline().split()
    assert int(head[0]) == case.nCells and int(head[1]) == case.nEdges
    case2 = common.host.HostCase(3, lloyd_iters=0, sfc=True)
    p2 = case2.partition(3, "file", prefix)
    assert np.array_equal(p1, p2)
    counts = np.bincount(p1, minlength=3)
    assert counts.min() > 0.8 * case.nCells / 3
