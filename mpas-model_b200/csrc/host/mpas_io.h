// MPAS grid / restart files on top of nc_classic (mpas_io.cpp)
#pragma once
#include <string>
#include <vector>

#include "decomp.h"
#include "mesh.h"

namespace mpas {
void write_grid_nc(const Mesh& m, const std::string& path, int version);
void read_grid_nc(const std::string& path, Mesh& m);
void create_restart_nc(const std::string& path, int version, const Mesh& global, const Block& proto, const std::vector<std::string>& names,
                       const std::string& xtime);
void write_restart_block(const std::string& path, const Block& b, const std::vector<std::string>& names);
int read_restart_block(const std::string& path, Block& b);
}  // namespace mpas
