#include "decomp.h"

#include <algorithm>
#include <cstdio>
#include <stdexcept>
#include <unordered_map>
#include <unordered_set>

namespace mpas {

#ifdef MPAS_B200_WITH_METIS
extern "C" int METIS_SetDefaultOptions(int64_t* options);
extern "C" int METIS_PartGraphKway(int64_t* nvtxs, int64_t* ncon, int64_t* xadj, int64_t* adjncy, int64_t* vwgt, int64_t* vsize,
                                   int64_t* adjwgt, int64_t* nparts, float* tpwgts, float* ubvec, int64_t* options, int64_t* edgecut,
                                   int64_t* part);
#endif

void partition_cells(const Mesh& G, int nBlocks, int method, std::vector<int>& part) {
    const int nC = G.nCells;
    part.assign(nC, 0);
    if (nBlocks <= 1) return;
#ifdef MPAS_B200_WITH_METIS
    if (method == 0) {
        std::vector<int64_t> xadj(nC + 1, 0), adj;
        adj.reserve((size_t)nC * 6);
        for (int c = 0; c < nC; ++c) {
            for (int i = 0; i < G.nEdgesOnCell[c]; ++i) {
                int n = G.cellsOnCell[(size_t)c * G.maxEdges + i];
                if (n >= 1 && n <= nC) adj.push_back(n - 1);
            }
            xadj[c + 1] = (int64_t)adj.size();
        }
        int64_t nv = nC, ncon = 1, np = nBlocks, cut = 0, options[40];
        METIS_SetDefaultOptions(options);
        options[8] = 20240611;  // METIS_OPTION_SEED
        std::vector<int64_t> p(nC);
        int rc = METIS_PartGraphKway(&nv, &ncon, xadj.data(), adj.data(), nullptr, nullptr, nullptr, &np, nullptr, nullptr, options, &cut, p.data());
        if (rc == 1) {
            for (int c = 0; c < nC; ++c) part[c] = (int)p[c];
            return;
        }
    }
#else
    (void)method;
#endif
    // contiguous, equal-sized chunks of the global (SFC) numbering
    for (int c = 0; c < nC; ++c) part[c] = (int)(((int64_t)c * nBlocks) / nC);
}

namespace {

// graph distance (<= maxd) from cell c (0-based) to the owned set of every nearby block
void block_distances(const Mesh& G, const std::vector<int>& part, int c, int maxd, std::vector<std::pair<int, int>>& out) {
    out.clear();
    auto note = [&](int b, int d) {
        for (auto& q : out)
            if (q.first == b) { q.second = std::min(q.second, d); return; }
        out.push_back({b, d});
    };
    int frontier[64], nf = 0, seen[128], ns = 0;
    frontier[nf++] = c;
    seen[ns++] = c;
    note(part[c], 0);
    for (int d = 1; d <= maxd; ++d) {
        int next[64], nn = 0;
        for (int f = 0; f < nf; ++f) {
            int cc = frontier[f];
            for (int i = 0; i < G.nEdgesOnCell[cc]; ++i) {
                int n = G.cellsOnCell[(size_t)cc * G.maxEdges + i] - 1;
                if (n < 0 || n >= G.nCells) continue;
                bool s = false;
                for (int k = 0; k < ns; ++k)
                    if (seen[k] == n) { s = true; break; }
                if (s) continue;
                if (ns < 128) seen[ns++] = n;
                if (nn < 64) next[nn++] = n;
                note(part[n], d);
            }
        }
        nf = nn;
        for (int k = 0; k < nn; ++k) frontier[k] = next[k];
    }
}
int dist_to(const std::vector<std::pair<int, int>>& d, int b) {
    for (auto& q : d)
        if (q.first == b) return q.second;
    return 1 << 20;
}

// finalize lists: sort by (endPoint, layer)
void sort_lists(std::vector<ExchangeList>& v) {
    std::sort(v.begin(), v.end(), [](const ExchangeList& a, const ExchangeList& b) {
        if (a.endPointID != b.endPointID) return a.endPointID < b.endPointID;
        return a.haloLayer < b.haloLayer;
    });
}

// Build send/recv lists of one entity kind.
//   localIDs: global id (1-based) of each local element; nOwned; layerEnd[h] = count incl. layer h
//   ownerOf(gid) -> owning block; neededBy(gid, out vector<(block, layer)>) for OWNED elements
template <class OwnerF, class NeededF>
void build_exchange(const std::vector<int>& localIDs, int nOwned, const std::vector<int>& layerEnd, int me, OwnerF ownerOf,
                    NeededF neededBy, ExchangeSet& ex) {
    ex.send.clear();
    ex.recv.clear();
    const int nLayers = (int)layerEnd.size() - 1;
    // ---- recv: halo elements grouped by (owner, layer); message order = ascending global id
    for (int h = 1; h <= nLayers; ++h) {
        std::unordered_map<int, std::vector<std::pair<int, int>>> byOwner;  // owner -> (gid, local index 1-based)
        for (int j = layerEnd[h - 1]; j < layerEnd[h]; ++j) byOwner[ownerOf(localIDs[j])].push_back({localIDs[j], j + 1});
        for (auto& kv : byOwner) {
            if (kv.first == me) throw std::runtime_error("decomp: halo element owned by self");
            std::vector<std::pair<int, int>> sorted = kv.second;
            std::sort(sorted.begin(), sorted.end());
            std::unordered_map<int, int> packing;  // gid -> position (1-based)
            for (size_t q = 0; q < sorted.size(); ++q) packing[sorted[q].first] = (int)q + 1;
            ExchangeList l;
            l.endPointID = kv.first;
            l.haloLayer = h;
            // enumerate in local halo-list order (mpas_dmpar.F:1911-1922)
            for (auto& gl : kv.second) { l.srcList.push_back(packing[gl.first]); l.destList.push_back(gl.second); }
            ex.recv.push_back(std::move(l));
        }
    }
    // ---- send: owned elements needed by other blocks
    std::unordered_map<long long, std::vector<std::pair<int, int>>> toSend;  // (block,layer) -> (gid, local 1-based)
    std::vector<std::pair<int, int>> need;
    for (int j = 0; j < nOwned; ++j) {
        neededBy(localIDs[j], need);
        for (auto& bl : need) {
            if (bl.first == me || bl.second < 1 || bl.second > nLayers) continue;
            toSend[((long long)bl.first << 8) | bl.second].push_back({localIDs[j], j + 1});
        }
    }
    for (auto& kv : toSend) {
        std::vector<std::pair<int, int>> sorted = kv.second;
        std::sort(sorted.begin(), sorted.end());
        ExchangeList l;
        l.endPointID = (int)(kv.first >> 8);
        l.haloLayer = (int)(kv.first & 0xff);
        for (size_t q = 0; q < sorted.size(); ++q) { l.srcList.push_back(sorted[q].second); l.destList.push_back((int)q + 1); }
        ex.send.push_back(std::move(l));
    }
    sort_lists(ex.send);
    sort_lists(ex.recv);
}

}  // namespace

void build_block(const Mesh& G, const std::vector<int>& part, int nBlocks, int iBlock, int nHalos, Block& out) {
    const int mE = G.maxEdges, mE2 = G.maxEdges2;
    out.blockID = iBlock;
    out.nBlocks = nBlocks;
    out.nHalos = nHalos;
    // ---- cells: owned sorted by global id (mpas_block_creator.F:99), halo layers each sorted (:601)
    std::vector<int> cells;  // global ids 1-based
    for (int c = 0; c < G.nCells; ++c)
        if (part[c] == iBlock) cells.push_back(c + 1);
    if (cells.empty()) throw std::runtime_error("decomp: block owns no cells");
    std::vector<int> cellEnd = {(int)cells.size()};
    std::unordered_map<int, int> cellLocal;  // gid -> local 1-based
    cellLocal.reserve(cells.size() * 2);
    for (size_t j = 0; j < cells.size(); ++j) cellLocal[cells[j]] = (int)j + 1;
    for (int h = 1; h <= nHalos; ++h) {
        std::vector<int> halo;
        std::unordered_set<int> added;
        for (size_t j = 0; j < cells.size(); ++j) {
            int g = cells[j] - 1;
            for (int i = 0; i < G.nEdgesOnCell[g]; ++i) {
                int n = G.cellsOnCell[(size_t)g * mE + i];
                if (n < 1 || n > G.nCells) continue;
                if (cellLocal.count(n) || added.count(n)) continue;
                added.insert(n);
                halo.push_back(n);
            }
        }
        std::sort(halo.begin(), halo.end());
        for (int n : halo) { cells.push_back(n); cellLocal[n] = (int)cells.size(); }
        cellEnd.push_back((int)cells.size());
    }
    // ---- edges / vertices: first-encounter order + owned/ghost partition (mpas_block_decomp.F:301-421)
    auto build_dual = [&](const std::vector<int>& onCell /*edgesOnCell or verticesOnCell*/, const std::vector<int>& cellsOn, int deg,
                          std::vector<int>& ids, std::vector<int>& layerEnd) {
        ids.clear();
        layerEnd.clear();
        std::unordered_set<int> seen;
        std::vector<int> enc;
        for (int j = 0; j < cellEnd[0]; ++j) {
            int g = cells[j] - 1;
            for (int i = 0; i < G.nEdgesOnCell[g]; ++i) {
                int e = onCell[(size_t)g * mE + i];
                if (seen.insert(e).second) enc.push_back(e);
            }
        }
        std::vector<int> owned, ghost;
        for (int e : enc) {
            int first = 0;
            for (int s = 0; s < deg; ++s) {
                first = cellsOn[(size_t)(e - 1) * deg + s];
                if (first != 0) break;
            }
            if (part[first - 1] == iBlock) owned.push_back(e);
            else ghost.push_back(e);
        }
        std::reverse(ghost.begin(), ghost.end());  // ghosts are filled from the end backwards (:340-341)
        ids = owned;
        layerEnd.push_back((int)ids.size());
        ids.insert(ids.end(), ghost.begin(), ghost.end());
        layerEnd.push_back((int)ids.size());
        for (int h = 2; h <= nHalos + 1; ++h) {  // mpas_block_creator.F:834-876
            for (int j = 0; j < cellEnd[h - 1]; ++j) {
                int g = cells[j] - 1;
                for (int i = 0; i < G.nEdgesOnCell[g]; ++i) {
                    int e = onCell[(size_t)g * mE + i];
                    if (seen.insert(e).second) ids.push_back(e);
                }
            }
            layerEnd.push_back((int)ids.size());
        }
    };
    std::vector<int> edges, edgeEnd, verts, vertEnd;
    build_dual(G.edgesOnCell, G.cellsOnEdge, 2, edges, edgeEnd);
    build_dual(G.verticesOnCell, G.cellsOnVertex, 3, verts, vertEnd);
    std::unordered_map<int, int> edgeLocal, vertLocal;
    edgeLocal.reserve(edges.size() * 2);
    vertLocal.reserve(verts.size() * 2);
    for (size_t j = 0; j < edges.size(); ++j) edgeLocal[edges[j]] = (int)j + 1;
    for (size_t j = 0; j < verts.size(); ++j) vertLocal[verts[j]] = (int)j + 1;

    // ---- local mesh with reindexed connectivity (mpas_block_creator.F:1376-1510)
    Mesh& m = out.mesh;
    m = Mesh();
    m.maxEdges = mE; m.maxEdges2 = mE2; m.on_a_sphere = G.on_a_sphere; m.sphere_radius = G.sphere_radius;
    m.planar_xmax = G.planar_xmax; m.planar_ymax = G.planar_ymax;
    const int nc = (int)cells.size(), ne = (int)edges.size(), nv = (int)verts.size();
    m.alloc(nc, ne, nv);
    m.nCellsSolve = cellEnd[0]; m.nEdgesSolve = edgeEnd[0]; m.nVerticesSolve = vertEnd[0];
    m.nCellsHalo = cellEnd; m.nEdgesHalo = edgeEnd; m.nVerticesHalo = vertEnd;
    auto LC = [&](int g) { auto it = cellLocal.find(g); return it == cellLocal.end() ? nc + 1 : it->second; };
    auto LE = [&](int g) { auto it = edgeLocal.find(g); return it == edgeLocal.end() ? ne + 1 : it->second; };
    auto LV = [&](int g) { auto it = vertLocal.find(g); return it == vertLocal.end() ? nv + 1 : it->second; };
    for (int j = 0; j < nc; ++j) {
        int g = cells[j] - 1;
        m.indexToCellID[j] = cells[j];
        m.xCell[j] = G.xCell[g]; m.yCell[j] = G.yCell[g]; m.zCell[j] = G.zCell[g];
        m.latCell[j] = G.latCell[g]; m.lonCell[j] = G.lonCell[g];
        m.areaCell[j] = G.areaCell[g]; m.meshDensity[j] = G.meshDensity[g];
        m.nEdgesOnCell[j] = G.nEdgesOnCell[g];
        for (int i = 0; i < mE; ++i) {
            bool used = i < G.nEdgesOnCell[g];
            m.cellsOnCell[(size_t)j * mE + i] = used ? LC(G.cellsOnCell[(size_t)g * mE + i]) : nc + 1;
            m.edgesOnCell[(size_t)j * mE + i] = used ? LE(G.edgesOnCell[(size_t)g * mE + i]) : ne + 1;
            m.verticesOnCell[(size_t)j * mE + i] = used ? LV(G.verticesOnCell[(size_t)g * mE + i]) : nv + 1;
        }
    }
    for (int j = 0; j < ne; ++j) {
        int g = edges[j] - 1;
        m.indexToEdgeID[j] = edges[j];
        m.xEdge[j] = G.xEdge[g]; m.yEdge[j] = G.yEdge[g]; m.zEdge[j] = G.zEdge[g];
        m.latEdge[j] = G.latEdge[g]; m.lonEdge[j] = G.lonEdge[g];
        m.dcEdge[j] = G.dcEdge[g]; m.dvEdge[j] = G.dvEdge[g]; m.angleEdge[j] = G.angleEdge[g];
        for (int s = 0; s < 2; ++s) {
            m.cellsOnEdge[2 * j + s] = LC(G.cellsOnEdge[2 * g + s]);
            m.verticesOnEdge[2 * j + s] = LV(G.verticesOnEdge[2 * g + s]);
        }
        m.nEdgesOnEdge[j] = G.nEdgesOnEdge[g];
        for (int k = 0; k < mE2; ++k) {
            bool used = k < G.nEdgesOnEdge[g];
            m.edgesOnEdge[(size_t)j * mE2 + k] = used ? LE(G.edgesOnEdge[(size_t)g * mE2 + k]) : ne + 1;
            m.weightsOnEdge[(size_t)j * mE2 + k] = G.weightsOnEdge[(size_t)g * mE2 + k];
        }
    }
    for (int j = 0; j < nv; ++j) {
        int g = verts[j] - 1;
        m.indexToVertexID[j] = verts[j];
        m.xVertex[j] = G.xVertex[g]; m.yVertex[j] = G.yVertex[g]; m.zVertex[j] = G.zVertex[g];
        m.latVertex[j] = G.latVertex[g]; m.lonVertex[j] = G.lonVertex[g];
        m.areaTriangle[j] = G.areaTriangle[g];
        for (int s = 0; s < 3; ++s) {
            m.cellsOnVertex[3 * j + s] = LC(G.cellsOnVertex[3 * g + s]);
            m.edgesOnVertex[3 * j + s] = LE(G.edgesOnVertex[3 * g + s]);
            m.kiteAreasOnVertex[3 * j + s] = G.kiteAreasOnVertex[3 * g + s];
        }
    }
    // garbage element
    m.cellsOnEdge[2 * ne] = m.cellsOnEdge[2 * ne + 1] = nc + 1;
    m.verticesOnEdge[2 * ne] = m.verticesOnEdge[2 * ne + 1] = nv + 1;
    for (int s = 0; s < 3; ++s) { m.cellsOnVertex[3 * nv + s] = nc + 1; m.edgesOnVertex[3 * nv + s] = ne + 1; }
    for (int i = 0; i < mE; ++i) {
        m.cellsOnCell[(size_t)nc * mE + i] = nc + 1; m.edgesOnCell[(size_t)nc * mE + i] = ne + 1;
        m.verticesOnCell[(size_t)nc * mE + i] = nv + 1;
    }
    for (int k = 0; k < mE2; ++k) m.edgesOnEdge[(size_t)ne * mE2 + k] = ne + 1;
    m.areaCell[nc] = 0.0;

    // ---- exchange lists
    if (nBlocks > 1) {
        std::vector<std::pair<int, int>> d0, d1, d2;
        auto cellOwner = [&](int gid) { return part[gid - 1]; };
        auto edgeOwner = [&](int gid) { return part[G.cellsOnEdge[2 * (gid - 1)] - 1]; };
        auto vertOwner = [&](int gid) { return part[G.cellsOnVertex[3 * (gid - 1)] - 1]; };
        auto cellNeeded = [&](int gid, std::vector<std::pair<int, int>>& need) {
            block_distances(G, part, gid - 1, nHalos, need);  // (block, dist) ; dist == halo layer
        };
        auto dualNeeded = [&](const int* cs, int deg, std::vector<std::pair<int, int>>& need) {
            need.clear();
            std::vector<std::pair<int, int>> dd[3];
            for (int s = 0; s < deg; ++s) block_distances(G, part, cs[s] - 1, nHalos, dd[s]);
            for (int s = 0; s < deg; ++s)
                for (auto& q : dd[s]) {
                    int b = q.first;
                    bool done = false;
                    for (auto& r : need)
                        if (r.first == b) done = true;
                    if (done) continue;
                    int mn = 1 << 20;
                    for (int t = 0; t < deg; ++t) mn = std::min(mn, dist_to(dd[t], b));
                    int layer;
                    if (mn == 0) layer = 1;  // touches an owned cell of b but is owned by us
                    else if (mn <= nHalos) layer = mn + 1;
                    else continue;
                    need.push_back({b, layer});
                }
        };
        auto edgeNeeded = [&](int gid, std::vector<std::pair<int, int>>& need) { dualNeeded(&G.cellsOnEdge[2 * (gid - 1)], 2, need); };
        auto vertNeeded = [&](int gid, std::vector<std::pair<int, int>>& need) { dualNeeded(&G.cellsOnVertex[3 * (gid - 1)], 3, need); };
        build_exchange(cells, cellEnd[0], cellEnd, iBlock, cellOwner, cellNeeded, out.cellExch);
        build_exchange(edges, edgeEnd[0], edgeEnd, iBlock, edgeOwner, edgeNeeded, out.edgeExch);
        build_exchange(verts, vertEnd[0], vertEnd, iBlock, vertOwner, vertNeeded, out.vertexExch);
    }
}

}  // namespace mpas
