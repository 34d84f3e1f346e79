// Horizontal MPAS mesh (global mesh or one decomposed block), reference names and shapes
// (src/core_atmosphere/Registry.xml:1127-1330).  All connectivity values are 1-based;
// arrays carry the trailing garbage element (index n+1) that off-block neighbours point to
// (src/framework/mpas_block_creator.F:1445-1508).  Geometry is on the UNIT sphere until
// scale_to_sphere() is called (the reference scales inside init_atm_case_jw,
// src/core_init_atmosphere/mpas_init_atm_cases.F:497-510).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace mpas {

struct Mesh {
    int nCells = 0, nEdges = 0, nVertices = 0;
    int nCellsSolve = 0, nEdgesSolve = 0, nVerticesSolve = 0;  // owned counts
    int maxEdges = 6, maxEdges2 = 12, vertexDegree = 3;
    bool on_a_sphere = true;
    double sphere_radius = 1.0;
    double planar_xmax = 0.0, planar_ymax = 0.0;  // planar meshes: GLOBAL maxima of xCell / yCell (the initial cases centre their bubble on them)
    // halo layer ends: cells: nCellsHalo[h] = count incl. halo layers <= h (h=0 owned)
    std::vector<int> nCellsHalo, nEdgesHalo, nVerticesHalo;

    // global IDs (1-based) of local elements
    std::vector<int> indexToCellID, indexToEdgeID, indexToVertexID;

    // geometry (size n+1)
    std::vector<double> xCell, yCell, zCell, latCell, lonCell;
    std::vector<double> xEdge, yEdge, zEdge, latEdge, lonEdge;
    std::vector<double> xVertex, yVertex, zVertex, latVertex, lonVertex;
    std::vector<double> dcEdge, dvEdge, angleEdge, areaCell, areaTriangle, meshDensity;
    std::vector<double> kiteAreasOnVertex;  // (3, nV+1)
    std::vector<double> weightsOnEdge;      // (maxEdges2, nE+1)

    // connectivity (1-based values)
    std::vector<int> nEdgesOnCell;    // (nC+1)
    std::vector<int> cellsOnCell;     // (maxEdges, nC+1)
    std::vector<int> edgesOnCell;     // (maxEdges, nC+1)
    std::vector<int> verticesOnCell;  // (maxEdges, nC+1)
    std::vector<int> cellsOnEdge;     // (2, nE+1)
    std::vector<int> verticesOnEdge;  // (2, nE+1)
    std::vector<int> nEdgesOnEdge;    // (nE+1)
    std::vector<int> edgesOnEdge;     // (maxEdges2, nE+1)
    std::vector<int> cellsOnVertex;   // (3, nV+1)
    std::vector<int> edgesOnVertex;   // (3, nV+1)

    void alloc(int nc, int ne, int nv);
};

// Icosahedral Voronoi mesh by recursive bisection, level n -> 10*4^n+2 cells.
// lloyd_iters > 0 applies fixed-topology centroidal relaxation.
void make_icosahedral_mesh(int level, Mesh& m, int lloyd_iters = 0);


// Doubly periodic planar mesh of nx x ny regular hexagons, cell spacing dc (m); ny even
void make_planar_hex_mesh(int nx, int ny, double dc, Mesh& m);
void compute_trisk_weights(Mesh& m);

// Recompute all metric arrays (lat/lon, dc, dv, angle, areas, kites, TRiSK weights)
// from positions + connectivity.
void compute_mesh_metrics(Mesh& m);

// Space-filling-curve renumbering of cells, edges and vertices. perm arrays map
// new index (0-based) -> old index (0-based); applying the inverse restores the
// original numbering bit-exactly.
struct MeshPermutation {
    std::vector<int> cellNewToOld, edgeNewToOld, vertexNewToOld;
    std::vector<int> cellOldToNew, edgeOldToNew, vertexOldToNew;
};
void sfc_reorder(const Mesh& in, Mesh& out, MeshPermutation& p);
void apply_permutation(const Mesh& in, const MeshPermutation& p, Mesh& out);
void invert_permutation(const MeshPermutation& p, MeshPermutation& inv);

// METIS graph file / partition file helpers (format consumed by
// src/framework/mpas_block_decomp.F:99-196: one 0-based block id per line).
void write_graph_info(const Mesh& m, const std::string& path);
void write_partition(const std::vector<int>& part, const std::string& path);
bool read_partition(const std::string& path, int nCells, std::vector<int>& part);

}  // namespace mpas
