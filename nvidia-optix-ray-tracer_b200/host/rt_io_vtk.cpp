// rt_io_vtk.cpp — legacy-VTK POLYDATA reader (ASCII and BINARY) and the two particle loaders built on
// it.  Replaces vtkPolyDataReader + the field extraction of src/Util/VTKReaderImpl.cpp:24-251.
//
// Supported subset (everything the reference's inputs use, files/data_description.txt): DATASET
// POLYDATA with POINTS, VERTICES, LINES, POLYGONS, TRIANGLE_STRIPS (classic "n i0 i1 ..." layout and the
// OFFSETS/CONNECTIVITY layout of file version 5), POINT_DATA / CELL_DATA with SCALARS (+LOOKUP_TABLE),
// VECTORS, NORMALS, TEXTURE_COORDINATES, COLOR_SCALARS, TENSORS and FIELD arrays.  Values are kept
// as double, like vtkDataArray::GetComponent.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "rt_io_internal.hpp"
#include "rt_io_vtk.hpp"

namespace rtio {

namespace {

struct Cursor {
    const std::string &s;
    size_t p = 0;
    bool binary = false;
    explicit Cursor(const std::string &text) : s(text) {}

    void skip_ws() {
        while (p < s.size() && (s[p] == ' ' || s[p] == '\t' || s[p] == '\n' || s[p] == '\r')) p++;
    }
    bool eof() {
        skip_ws();
        return p >= s.size();
    }
    std::string line() {  // rest of the current line
        size_t e = s.find('\n', p);
        if (e == std::string::npos) e = s.size();
        std::string r = s.substr(p, e - p);
        p = std::min(e + 1, s.size());
        while (!r.empty() && (r.back() == '\r' || r.back() == ' ')) r.pop_back();
        return r;
    }
    std::string token() {
        skip_ws();
        const size_t b = p;
        while (p < s.size() && !(s[p] == ' ' || s[p] == '\t' || s[p] == '\n' || s[p] == '\r')) p++;
        if (b == p) raise(RT_IO_ERR_PARSE, "VTK: unexpected end of file");
        return s.substr(b, p - b);
    }
    std::string peek() {
        const size_t save = p;
        skip_ws();
        if (p >= s.size()) {
            p = save;
            return "";
        }
        std::string t = token();
        p = save;
        return t;
    }
    long long integer() {
        skip_ws();
        const char *b = s.c_str() + p;
        char *e = nullptr;
        const long long v = std::strtoll(b, &e, 10);
        if (e == b) raise(RT_IO_ERR_PARSE, "VTK: integer expected near '" + s.substr(p, 16) + "'");
        p += (size_t)(e - b);
        return v;
    }
    double number() {
        skip_ws();
        const char *b = s.c_str() + p;
        char *e = nullptr;
        const double v = std::strtod(b, &e);
        if (e == b) raise(RT_IO_ERR_PARSE, "VTK: number expected near '" + s.substr(p, 16) + "'");
        p += (size_t)(e - b);
        return v;
    }
    // BINARY payloads start right after the newline that ends the header line and are big-endian
    void begin_binary_block() {
        if (p < s.size() && s[p] == '\r') p++;
        if (p < s.size() && s[p] == '\n') p++;
    }
    const unsigned char *raw(size_t bytes) {
        if (p + bytes > s.size()) raise(RT_IO_ERR_PARSE, "VTK: truncated binary block");
        const unsigned char *r = reinterpret_cast<const unsigned char *>(s.data()) + p;
        p += bytes;
        return r;
    }
};

size_t type_size(const std::string &t) {
    if (t == "bit") raise(RT_IO_ERR_FORMAT, "VTK: bit arrays are not supported");
    if (t == "char" || t == "unsigned_char") return 1;
    if (t == "short" || t == "unsigned_short") return 2;
    if (t == "int" || t == "unsigned_int" || t == "float") return 4;
    if (t == "long" || t == "unsigned_long" || t == "double" || t == "vtktypeint64" || t == "vtktypeuint64" || t == "vtkidtype") return 8;
    if (t == "vtktypeint32" || t == "vtktypeuint32") return 4;
    raise(RT_IO_ERR_FORMAT, "VTK: unknown data type '" + t + "'");
}

std::string lower(std::string s) {
    for (auto &c : s) c = (char)std::tolower((unsigned char)c);
    return s;
}

double be_value(const unsigned char *b, const std::string &t) {
    unsigned char r[8];
    const size_t n = type_size(t);
    for (size_t i = 0; i < n; i++) r[i] = b[n - 1 - i];  // big-endian file -> little-endian host
    if (t == "float") { float v; std::memcpy(&v, r, 4); return v; }
    if (t == "double") { double v; std::memcpy(&v, r, 8); return v; }
    if (t == "char") { int8_t v; std::memcpy(&v, r, 1); return v; }
    if (t == "unsigned_char") { uint8_t v; std::memcpy(&v, r, 1); return v; }
    if (t == "short") { int16_t v; std::memcpy(&v, r, 2); return v; }
    if (t == "unsigned_short") { uint16_t v; std::memcpy(&v, r, 2); return v; }
    if (t == "int" || t == "vtktypeint32") { int32_t v; std::memcpy(&v, r, 4); return v; }
    if (t == "unsigned_int" || t == "vtktypeuint32") { uint32_t v; std::memcpy(&v, r, 4); return v; }
    if (t == "unsigned_long" || t == "vtktypeuint64") { uint64_t v; std::memcpy(&v, r, 8); return (double)v; }
    int64_t v;
    std::memcpy(&v, r, 8);
    return (double)v;
}

void read_values(Cursor &c, const std::string &type_in, size_t count, std::vector<double> &out) {
    const std::string type = lower(type_in);
    out.resize(count);
    if (!c.binary) {
        (void)type_size(type);
        for (size_t i = 0; i < count; i++) out[i] = c.number();
        return;
    }
    const size_t sz = type_size(type);
    c.begin_binary_block();
    const unsigned char *b = c.raw(sz * count);
    for (size_t i = 0; i < count; i++) out[i] = be_value(b + sz * i, type);
}

void read_cells(Cursor &c, std::vector<std::vector<int64_t>> &cells) {
    const long long n = c.integer();
    const long long size = c.integer();
    if (n < 0 || size < 0) raise(RT_IO_ERR_PARSE, "VTK: negative cell count");
    const std::string nx = c.peek();
    if (nx == "OFFSETS") {  // version 5: n = number of offsets, size = connectivity length
        c.token();
        const std::string ot = c.token();
        std::vector<double> off, con;
        read_values(c, ot, (size_t)n, off);
        if (c.token() != "CONNECTIVITY") raise(RT_IO_ERR_PARSE, "VTK: CONNECTIVITY expected");
        const std::string ct = c.token();
        read_values(c, ct, (size_t)size, con);
        for (size_t i = 0; i + 1 < off.size(); i++) {
            std::vector<int64_t> cell;
            for (size_t k = (size_t)off[i]; k < (size_t)off[i + 1] && k < con.size(); k++) cell.push_back((int64_t)con[k]);
            cells.push_back(std::move(cell));
        }
        return;
    }
    std::vector<double> flat;
    read_values(c, "int", (size_t)size, flat);
    size_t k = 0;
    for (long long i = 0; i < n; i++) {
        if (k >= flat.size()) raise(RT_IO_ERR_PARSE, "VTK: cell list shorter than declared");
        const size_t m = (size_t)flat[k++];
        if (k + m > flat.size()) raise(RT_IO_ERR_PARSE, "VTK: cell list shorter than declared");
        std::vector<int64_t> cell(m);
        for (size_t j = 0; j < m; j++) cell[j] = (int64_t)flat[k++];
        cells.push_back(std::move(cell));
    }
}

void read_attributes(Cursor &c, size_t n, std::map<std::string, VtkArray> &dst) {
    for (;;) {
        const std::string kw = c.peek();
        if (kw.empty() || kw == "POINT_DATA" || kw == "CELL_DATA" || kw == "POINTS" || kw == "VERTICES" || kw == "LINES" ||
            kw == "POLYGONS" || kw == "TRIANGLE_STRIPS")
            return;
        c.token();
        if (kw == "SCALARS") {
            VtkArray a;
            const std::string name = c.token();
            const std::string type = c.token();
            a.comps = 1;
            // optional component count, then LOOKUP_TABLE <name>
            std::string nx = c.peek();
            if (!nx.empty() && std::isdigit((unsigned char)nx[0])) {
                a.comps = (size_t)c.integer();
                nx = c.peek();
            }
            if (nx == "LOOKUP_TABLE") {
                c.token();
                c.token();
            }
            read_values(c, type, n * a.comps, a.v);
            dst[name] = std::move(a);
        } else if (kw == "VECTORS" || kw == "NORMALS") {
            VtkArray a;
            const std::string name = c.token();
            const std::string type = c.token();
            a.comps = 3;
            read_values(c, type, n * 3, a.v);
            dst[name] = std::move(a);
        } else if (kw == "TEXTURE_COORDINATES") {
            VtkArray a;
            const std::string name = c.token();
            a.comps = (size_t)c.integer();
            const std::string type = c.token();
            read_values(c, type, n * a.comps, a.v);
            dst[name] = std::move(a);
        } else if (kw == "TENSORS") {
            VtkArray a;
            const std::string name = c.token();
            const std::string type = c.token();
            a.comps = 9;
            read_values(c, type, n * 9, a.v);
            dst[name] = std::move(a);
        } else if (kw == "COLOR_SCALARS") {
            VtkArray a;
            const std::string name = c.token();
            a.comps = (size_t)c.integer();
            read_values(c, c.binary ? "unsigned_char" : "float", n * a.comps, a.v);
            dst[name] = std::move(a);
        } else if (kw == "LOOKUP_TABLE") {
            c.token();
            const size_t sz = (size_t)c.integer();
            std::vector<double> tmp;
            read_values(c, c.binary ? "unsigned_char" : "float", sz * 4, tmp);
        } else if (kw == "FIELD") {
            c.token();  // field name
            const long long na = c.integer();
            for (long long i = 0; i < na; i++) {
                VtkArray a;
                const std::string name = c.token();
                a.comps = (size_t)c.integer();
                const size_t tuples = (size_t)c.integer();
                const std::string type = c.token();
                read_values(c, type, a.comps * tuples, a.v);
                dst[name] = std::move(a);
            }
        } else if (kw == "METADATA") {  // written by newer VTK: skip to the blank line that ends it
            c.line();
            while (c.p < c.s.size()) {
                if (c.line().empty()) break;
            }
        } else {
            raise(RT_IO_ERR_FORMAT, "VTK: unsupported attribute keyword '" + kw + "'");
        }
    }
}

}  // namespace

void parse_polydata(const std::string &text, VtkPolyData &pd) {
    Cursor c(text);
    const std::string l0 = c.line();
    if (l0.rfind("# vtk DataFile", 0) != 0) raise(RT_IO_ERR_PARSE, "VTK: missing '# vtk DataFile' header");
    c.line();  // title
    const std::string fmt = c.token();
    if (fmt == "BINARY") c.binary = true;
    else if (fmt != "ASCII") raise(RT_IO_ERR_PARSE, "VTK: ASCII or BINARY expected, got '" + fmt + "'");
    if (c.token() != "DATASET") raise(RT_IO_ERR_PARSE, "VTK: DATASET expected");
    const std::string ds = c.token();
    if (ds != "POLYDATA") raise(RT_IO_ERR_FORMAT, "VTK: DATASET " + ds + " is not POLYDATA");
    while (!c.eof()) {
        const std::string kw = c.token();
        if (kw == "POINTS") {
            const size_t n = (size_t)c.integer();
            const std::string type = c.token();
            read_values(c, type, 3 * n, pd.points);
        } else if (kw == "VERTICES") {
            read_cells(c, pd.verts);
        } else if (kw == "LINES") {
            read_cells(c, pd.lines);
        } else if (kw == "POLYGONS") {
            read_cells(c, pd.polys);
        } else if (kw == "TRIANGLE_STRIPS") {
            read_cells(c, pd.strips);
        } else if (kw == "POINT_DATA") {
            const size_t n = (size_t)c.integer();
            read_attributes(c, n, pd.point_data);
        } else if (kw == "CELL_DATA") {
            const size_t n = (size_t)c.integer();
            read_attributes(c, n, pd.cell_data);
        } else if (kw == "METADATA") {
            c.line();
            while (c.p < c.s.size()) {
                if (c.line().empty()) break;
            }
        } else if (kw == "FIELD") {  // dataset-level field data
            std::map<std::string, VtkArray> tmp;
            c.p -= kw.size();
            read_attributes(c, 0, tmp);
        } else {
            raise(RT_IO_ERR_FORMAT, "VTK: unsupported section '" + kw + "'");
        }
    }
}

void load_polydata(const std::string &path, VtkPolyData &pd) {
    std::string text;
    if (!read_file(path, text)) raise(RT_IO_ERR_OPEN, "Failed to open file: " + path + "!");
    parse_polydata(text, pd);
}

}  // namespace rtio
