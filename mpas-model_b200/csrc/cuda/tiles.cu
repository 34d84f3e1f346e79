// Edge-tile plans for the shared-memory staged stencil gathers (north_star: "gathered neighbour columns are staged in
// shared memory").
//
// The high-order horizontal fluxes (theta_m and w in atm_compute_dyn_tend_work TI:5046-5074, 5231-5252; the scalars in
// atm_advance_scalars_work TI:3330-3387 and atm_advance_scalars_mono_work TI:3916-3961) read, per edge and level, the
// columns of the ~10 cells of advCellsForEdge.  Edges are numbered along a space-filling curve, so consecutive edges
// share most of their stencil cells: a TILE of consecutive edges needs about 1.1 distinct cell columns per edge instead
// of 10.  A plan cuts the edge range into tiles (greedy: as many consecutive edges as fit `ncmax` distinct cells, at
// most `temax`), lists each tile's distinct cells and gives every (edge, stencil slot) the index of its cell in that
// list.  A kernel block then stages the tile's columns ONCE with bulk (TMA) copies and every gather becomes a
// shared-memory read.  The plan depends on the connectivity only; it is built on the host from a download of
// advCellsForEdge the first time a routine sees a mesh (never inside a stream capture) and cached by the device address
// of that array.
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "common.cuh"

namespace mpasb200 {

namespace {
struct PlanStore {
    EdgeTilePlan plan;
    void* d_all = nullptr;
};
typedef std::tuple<const void*, int, int, int, int> PlanKey;
std::map<PlanKey, PlanStore> g_plans;
std::mutex g_plan_mu;
}  // namespace

const EdgeTilePlan* edge_tile_plan(const int* d_advCellsForEdge, const int* d_nAdvCellsForEdge, int edgeStart, int nEdgesLoop, int nCells, int ncmax, int temax) {
    std::lock_guard<std::mutex> lk(g_plan_mu);
    const PlanKey key(d_advCellsForEdge, edgeStart, nEdgesLoop, ncmax, temax);
    auto it = g_plans.find(key);
    if (it != g_plans.end()) return &it->second.plan;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream(), &st) == cudaSuccess && st != cudaStreamCaptureStatusNone) return nullptr;
    if (nEdgesLoop <= 0 || ncmax < 17 || ncmax > 255 || temax < 1) return nullptr;
    std::vector<int> adv((size_t)nEdgesLoop * 15), nadv((size_t)nEdgesLoop);
    cudaStreamSynchronize(stream());
    if (cudaMemcpy(adv.data(), d_advCellsForEdge + (size_t)(edgeStart - 1) * 15, adv.size() * sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(nadv.data(), d_nAdvCellsForEdge + (edgeStart - 1), nadv.size() * sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    std::vector<int> first, ncs, cells;
    std::vector<unsigned char> lidx((size_t)nEdgesLoop * 16, 0);
    std::vector<int> stamp((size_t)nCells + 2, -1), local((size_t)nCells + 2, 0);
    first.push_back(0);
    int tile = 0, nc = 0, ne = 0;
    size_t base = 0;
    cells.resize((size_t)ncmax, 1);
    for (int e = 0; e < nEdgesLoop; ++e) {
        const int n = std::min(std::max(nadv[e], 0), 15);
        int fresh = 0;
        for (int j = 0; j < n; ++j) {
            int& c = adv[(size_t)e * 15 + j];
            if (c < 1) c = 1;
            if (c > nCells + 1) return nullptr;  // not a connectivity array we understand
            if (stamp[c] != tile) {
                bool dup = false;  // the same new cell twice in one stencil
                for (int q = 0; q < j; ++q) dup = dup || adv[(size_t)e * 15 + q] == c;
                if (!dup) ++fresh;
            }
        }
        if (ne > 0 && (nc + fresh > ncmax || ne == temax)) {  // close the tile
            ncs.push_back(nc);
            first.push_back(e);
            ++tile;
            nc = 0;
            ne = 0;
            base = cells.size();
            cells.resize(base + (size_t)ncmax, 1);
        }
        for (int j = 0; j < n; ++j) {
            const int c = adv[(size_t)e * 15 + j];
            if (stamp[c] != tile) {
                stamp[c] = tile;
                local[c] = nc;
                cells[base + nc] = c;
                ++nc;
            }
            lidx[(size_t)e * 16 + j] = (unsigned char)local[c];
        }
        ++ne;
    }
    ncs.push_back(nc);
    first.push_back(nEdgesLoop);
    const int nTiles = (int)ncs.size();
    // one allocation: [first (nTiles+1) | nc (nTiles) | cells (nTiles*ncmax) | lidx]
    const size_t o_first = 0, o_nc = o_first + (size_t)(nTiles + 1) * 4, o_cells = o_nc + (size_t)nTiles * 4;
    const size_t o_lidx = (o_cells + (size_t)nTiles * ncmax * 4 + 15) & ~(size_t)15, total = o_lidx + lidx.size();
    PlanStore ps;
    if (cudaMalloc(&ps.d_all, total) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    char* d = (char*)ps.d_all;
    cudaMemcpy(d + o_first, first.data(), (size_t)(nTiles + 1) * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d + o_nc, ncs.data(), (size_t)nTiles * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d + o_cells, cells.data(), (size_t)nTiles * ncmax * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d + o_lidx, lidx.data(), lidx.size(), cudaMemcpyHostToDevice);
    ps.plan.max_stencil = 0;
    for (int e = 0; e < nEdgesLoop; ++e) ps.plan.max_stencil = std::max(ps.plan.max_stencil, std::min(std::max(nadv[e], 0), 15));
    ps.plan.nTiles = nTiles;
    ps.plan.ncmax = ncmax;
    ps.plan.first = (const int*)(d + o_first);
    ps.plan.nc = (const int*)(d + o_nc);
    ps.plan.cells = (const int*)(d + o_cells);
    ps.plan.lidx = (const unsigned char*)(d + o_lidx);
    auto ins = g_plans.emplace(key, ps);
    return &ins.first->second.plan;
}

void release_edge_tile_plans(const void* d_advCellsForEdge) {
    std::lock_guard<std::mutex> lk(g_plan_mu);
    for (auto it = g_plans.begin(); it != g_plans.end();) {
        if (d_advCellsForEdge == nullptr || std::get<0>(it->first) == d_advCellsForEdge) {
            cudaFree(it->second.d_all);
            it = g_plans.erase(it);
        } else {
            ++it;
        }
    }
}

}  // namespace mpasb200
