// rt_io_vtk.hpp — in-memory legacy-VTK POLYDATA (see rt_io_vtk.cpp).
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace rtio {

struct VtkArray {
    size_t comps = 1;
    std::vector<double> v;  // tuples * comps
    size_t tuples() const { return comps ? v.size() / comps : 0; }
};

struct VtkPolyData {
    std::vector<double> points;  // 3 per point
    // vtkPolyData numbers its cells verts -> lines -> polys -> strips; CELL_DATA follows that order
    std::vector<std::vector<int64_t>> verts, lines, polys, strips;
    std::map<std::string, VtkArray> point_data, cell_data;
    size_t n_points() const { return points.size() / 3; }
    size_t n_cells() const { return verts.size() + lines.size() + polys.size() + strips.size(); }
};

void parse_polydata(const std::string &text, VtkPolyData &out);  // throws rtio::IoError
void load_polydata(const std::string &path, VtkPolyData &out);

}  // namespace rtio
