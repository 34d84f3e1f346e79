// Block decomposition: partition -> owned/halo cell, edge, vertex lists -> exchange lists,
// with the reference's ordering rules:
//   src/framework/mpas_block_decomp.F:51-287 (partition file -> cells per block),
//   :301-356 (owned vs ghost edges), :369-421 (first-encounter edge order), :434-506 (add halo),
//   src/framework/mpas_block_creator.F:52-116, :266-470, :470-734, :734-940, :1376-1510,
//   src/framework/mpas_dmpar.F:1483-2038 (exchange lists: message order = ascending global ID).
#pragma once
#include <vector>

#include "mesh.h"
#include "pool.h"

namespace mpas {

// One exchange list, reference form (mpas_dmpar_types.inc mpas_exchange_list):
//   send: srcList = local owned index (1-based), destList = position in the message (1-based)
//   recv: srcList = position in the message (1-based), destList = local halo index (1-based)
struct ExchangeList {
    int endPointID = -1;  // neighbour block/rank
    int haloLayer = 0;    // 1-based halo layer
    std::vector<int> srcList, destList;
};
struct ExchangeSet {
    std::vector<ExchangeList> send, recv;  // sorted by (endPointID, haloLayer)
};

struct Block {
    int blockID = 0, nBlocks = 1, nHalos = 2;
    Mesh mesh;  // local mesh, local 1-based indices, garbage element n+1
    Pool pool;
    ExchangeSet cellExch, edgeExch, vertexExch;
};

// method 0: METIS k-way through libmetis when linked, else falls back to 1; 1: contiguous chunks of
// the (SFC-ordered) global numbering.
void partition_cells(const Mesh& global, int nBlocks, int method, std::vector<int>& part);

void build_block(const Mesh& global, const std::vector<int>& part, int nBlocks, int iBlock, int nHalos, Block& out);

}  // namespace mpas
