"""Local numbering of a block against a brute-force Python restatement written from the
reference's routines, not from csrc/host/decomp.cpp:

  cells     owned cells ascending global ID (mpas_block_creator.F:99), then halo layer h = the not-yet-present
            neighbours (cellsOnCell) of everything up to layer h-1, found in list order (mpas_block_decomp_add_halo,
            mpas_block_decomp.F:434-503) and SORTED per layer (mpas_block_creator.F:601)
  edges /   first-encounter order over the owned cells x edgesOnCell slots (mpas_block_decomp_all_edges_in_block,
  vertices  mpas_block_decomp.F:369-421); an element is owned iff its first non-zero cellsOnEdge / cellsOnVertex entry is
            an owned cell; owned elements keep the encounter order, the others are filled in from the END backwards, i.e.
            reversed (mpas_block_decomp_partitioned_edge_list, :301-356, ghostEdgeStart at :340-341); layer h+1 = the new
            elements met walking all cells up to halo layer h in local order (mpas_block_creator.F:834-876)
  indices   local connectivity = global connectivity renumbered, elements outside the block -> the garbage index n+1
            (mpas_block_creator.F:1445-1508)
"""
import numpy as np
import pytest

import common

NHALOS = 2


def _global(level, lloyd):
    case, blocks = common.make_case(level, 1, nVertLevels=4, num_scalars=1, lloyd=lloyd)
    b = blocks[0]
    g = {n: b.field(n)[:-1].copy() for n in ("cellsOnCell", "edgesOnCell", "verticesOnCell", "nEdgesOnCell", "cellsOnEdge", "cellsOnVertex",
                                             "verticesOnEdge", "edgesOnVertex")}
    for n, idn in (("cell", "indexToCellID"), ("edge", "indexToEdgeID"), ("vertex", "indexToVertexID")):
        ids = b.field(idn)[:-1]
        g["id_" + n] = ids.copy()
    return case, g


def _expected(g, part, ib):
    """brute force; everything in GLOBAL 1-based ids"""
    # a single block numbers its elements in its own local order: translate its local connectivity to global ids first
    idc, ide, idv = g["id_cell"], g["id_edge"], g["id_vertex"]
    nC = len(idc)
    pos_c = np.argsort(idc)  # global id - 1 -> local row of the one-block mesh

    def row_c(gid):
        return pos_c[gid - 1]

    def to_gid(ids, local):
        return int(ids[local - 1]) if 1 <= local <= len(ids) else 0

    owned = sorted(int(c) for c in np.nonzero(part == ib)[0] + 1)
    ownedset = set(owned)
    cells = list(owned)
    cell_end = [len(cells)]
    for h in range(1, NHALOS + 1):
        present = set(cells)
        new = []
        for c in cells:
            r = row_c(c)
            for i in range(g["nEdgesOnCell"][r]):
                n = to_gid(idc, int(g["cellsOnCell"][r, i]))
                if n != 0 and n not in present:
                    present.add(n)
                    new.append(n)
        cells += sorted(new)
        cell_end.append(len(cells))

    def dual(on_cell, ids, cells_on, deg):
        pos = np.argsort(ids)
        enc, seen = [], set()
        for c in cells[:cell_end[0]]:
            r = row_c(c)
            for i in range(g["nEdgesOnCell"][r]):
                e = to_gid(ids, int(on_cell[r, i]))
                if e not in seen:
                    seen.add(e)
                    enc.append(e)
        own, ghost = [], []
        for e in enc:
            first = 0
            for s in range(deg):
                first = to_gid(idc, int(cells_on[pos[e - 1], s]))
                if first != 0:
                    break
            (own if first in ownedset else ghost).append(e)
        out = own + ghost[::-1]
        ends = [len(own), len(out)]
        for h in range(2, NHALOS + 2):
            for c in cells[:cell_end[h - 1]]:
                r = row_c(c)
                for i in range(g["nEdgesOnCell"][r]):
                    e = to_gid(ids, int(on_cell[r, i]))
                    if e not in seen:
                        seen.add(e)
                        out.append(e)
            ends.append(len(out))
        return out, ends

    edges, edge_end = dual(g["edgesOnCell"], ide, g["cellsOnEdge"], 2)
    verts, vert_end = dual(g["verticesOnCell"], idv, g["cellsOnVertex"], 3)
    return cells, cell_end, edges, edge_end, verts, vert_end


@pytest.mark.parametrize("level,nblocks,method", [(3, 3, "sfc"), (4, 5, "metis")])
def test_local_numbering_matches_brute_force_restatement_of_the_reference_rules(level, nblocks, method):
    case1, g = _global(level, 10)
    case, blocks = common.make_case(level, nblocks, nVertLevels=4, num_scalars=1, method=method, lloyd=10)
    part = np.zeros(case.nCells, dtype=np.int32)
    common.host.hostlib().mpas_host_case_get_partition(case.h, part.ctypes.data)
    assert sorted(g["id_cell"].tolist()) == list(range(1, case.nCells + 1))
    n_reversed = 0
    for ib, b in enumerate(blocks):
        cells, cell_end, edges, edge_end, verts, vert_end = _expected(g, part, ib)
        cfg = b.cfg()
        assert b.field("indexToCellID")[:-1].tolist() == cells, ib
        assert b.field("indexToEdgeID")[:-1].tolist() == edges, ib
        assert b.field("indexToVertexID")[:-1].tolist() == verts, ib
        assert int(cfg["nCellsSolve"]) == cell_end[0] and int(cfg["nCells"]) == cell_end[-1]
        assert int(cfg["nEdgesSolve"]) == edge_end[0] and int(cfg["nEdges"]) == edge_end[-1]
        assert int(cfg["nVerticesSolve"]) == vert_end[0] and int(cfg["nVertices"]) == vert_end[-1]
        # the reversed halo-1 segment is really exercised: its encounter order is not monotone the other way round by accident
        n_reversed += edge_end[1] - edge_end[0]
        # per-layer sort of the cell halos: each layer ascending, layers not merged
        ids = np.array(cells)
        for a, z in zip(cell_end[:-1], cell_end[1:]):
            assert (np.diff(ids[a:z]) > 0).all()
        assert not (np.diff(ids[cell_end[0]:]) > 0).all()
    assert n_reversed > 20


@pytest.mark.parametrize("nblocks", [4])
def test_local_connectivity_is_the_global_one_renumbered_with_garbage_index_outside(nblocks):
    case1, g = _global(4, 10)
    case, blocks = common.make_case(4, nblocks, nVertLevels=4, num_scalars=1, method="metis", lloyd=10)
    G = {"cell": g["id_cell"], "edge": g["id_edge"], "vertex": g["id_vertex"]}
    rows = {k: np.argsort(v) for k, v in G.items()}
    spec = [("cellsOnEdge", "edge", "cell"), ("verticesOnEdge", "edge", "vertex"), ("edgesOnCell", "cell", "edge"), ("cellsOnCell", "cell", "cell"),
            ("verticesOnCell", "cell", "vertex"), ("cellsOnVertex", "vertex", "cell"), ("edgesOnVertex", "vertex", "edge")]
    idname = {"cell": "indexToCellID", "edge": "indexToEdgeID", "vertex": "indexToVertexID"}
    for b in blocks:
        ids = {k: b.field(v)[:-1] for k, v in idname.items()}
        local_of = {k: {int(gid): i + 1 for i, gid in enumerate(v)} for k, v in ids.items()}
        for name, row_kind, val_kind in spec:
            loc = b.field(name)[:-1]
            glob = g[name] if name in g else None
            if glob is None:
                continue
            n_garbage = len(ids[val_kind]) + 1
            nvalid = b.field("nEdgesOnCell")[:-1] if row_kind == "cell" else None
            bad = 0
            for i in range(loc.shape[0]):
                grow = glob[rows[row_kind][ids[row_kind][i] - 1]]
                width = int(nvalid[i]) if nvalid is not None else loc.shape[1]
                for s in range(width):
                    gl = int(grow[s])
                    gid = int(G[val_kind][gl - 1]) if gl >= 1 else 0
                    want = local_of[val_kind].get(gid, n_garbage) if gid else loc[i, s]
                    bad += int(loc[i, s] != want)
            assert bad == 0, (name, bad)
