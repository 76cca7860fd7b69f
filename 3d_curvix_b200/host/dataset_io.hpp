// dataset_io.hpp -- on-disk formats either side of the edge-sketch path (host side of the drop-in).
//
// Inputs, as the reference's file_reader reads them (Edge_Reconst/file_reader.cpp):
//   <Dataset_Path><Dataset_Name>/<Scene_Name>/Edges/Edge_<i>_t<thr>.txt   x y theta strength per edge   (:50-78)
//   .../RnT/R_matrix.txt (or refined_R.txt)   3 doubles per row, 3 rows per view                         (:113-145)
//   .../RnT/T_matrix.txt (or refined_T.txt)   3 doubles per view                                         (:148-174)
//   .../RnT/K_matrix.txt                      3x3 per view, only with Use_Multiple_K                     (:177-212)
// Every reader is "whitespace separated doubles in stream order", exactly what `file >> double` loops do; a trailing
// incomplete record is dropped like the reference does.  Edge tables are returned as SoA columns because that is what
// cvx_set_view_edges takes (and what an Eigen column-major E x 4 matrix already is).
//
// Outputs, as EdgeSketch_Core::Stack_3D_Edges writes them (EdgeSketch_Core.cpp:750-763): `ofstream << Eigen::MatrixXd`,
// i.e. Eigen's default IOFormat: stream precision (6 significant digits), one space between coefficients, every
// coefficient right-aligned to the width of the widest one in the whole matrix, '\n' between rows, no trailing newline.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace curvix_host {

struct EdgeTable {   // one view: E edges, SoA
    std::vector<double> x, y, theta, strength;
    int size() const { return int(x.size()); }
};

// Reads a whole text file of whitespace separated doubles.  Returns false when the file cannot be opened.
inline bool read_doubles(const std::string& path, std::vector<double>* out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::string buf(size_t(sz > 0 ? sz : 0), '\0');
    const size_t got = sz > 0 ? std::fread(&buf[0], 1, size_t(sz), f) : 0;
    std::fclose(f);
    buf.resize(got);
    out->clear();
    out->reserve(got / 12 + 4);
    const char* p = buf.c_str();
    const char* end = p + buf.size();
    while (p < end) {
        char* q = nullptr;
        const double v = std::strtod(p, &q);
        if (q == p) break;   // `file >> double` stops at the first token that is not a number
        out->push_back(v);
        p = q;
    }
    return true;
}

inline std::string edge_file_path(const std::string& scene_dir, int view, int thresh) {
    return scene_dir + "/Edges/Edge_" + std::to_string(view) + "_t" + std::to_string(thresh) + ".txt";
}

inline bool read_edge_file(const std::string& path, EdgeTable* t) {
    std::vector<double> v;
    if (!read_doubles(path, &v)) return false;
    const size_t n = v.size() / 4;
    t->x.resize(n); t->y.resize(n); t->theta.resize(n); t->strength.resize(n);
    for (size_t i = 0; i < n; i++) { t->x[i] = v[4 * i]; t->y[i] = v[4 * i + 1]; t->theta[i] = v[4 * i + 2]; t->strength[i] = v[4 * i + 3]; }
    return true;
}

// Binary SoA cache of one parsed edge file: the reference re-parses all N text files twice per
// pass; here a file is parsed once per run and, when a cache directory is given, once ever.
// Layout: magic "CVXE", int32 version, int64 n, then x[n], y[n], theta[n], strength[n] (little-endian doubles).
inline bool read_edge_cache(const std::string& path, EdgeTable* t) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char magic[4]; int32_t ver = 0; int64_t n = -1;
    bool ok = std::fread(magic, 1, 4, f) == 4 && std::fread(&ver, 4, 1, f) == 1 && std::fread(&n, 8, 1, f) == 1 &&
              magic[0] == 'C' && magic[1] == 'V' && magic[2] == 'X' && magic[3] == 'E' && ver == 1 && n >= 0;
    if (ok) {
        std::vector<double>* cols[4] = {&t->x, &t->y, &t->theta, &t->strength};
        for (auto* c : cols) { c->resize(size_t(n)); if (n > 0 && std::fread(c->data(), 8, size_t(n), f) != size_t(n)) ok = false; }
    }
    std::fclose(f);
    return ok;
}
inline bool write_edge_cache(const std::string& path, const EdgeTable& t) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const int32_t ver = 1; const int64_t n = t.size();
    std::fwrite("CVXE", 1, 4, f); std::fwrite(&ver, 4, 1, f); std::fwrite(&n, 8, 1, f);
    const std::vector<double>* cols[4] = {&t.x, &t.y, &t.theta, &t.strength};
    for (auto* c : cols) if (n > 0) std::fwrite(c->data(), 8, size_t(n), f);
    return std::fclose(f) == 0;
}

// Rotations: n x 9 row-major; translations: n x 3.  A trailing incomplete matrix is dropped (file_reader.cpp:124-139).
inline bool read_R_file(const std::string& path, std::vector<double>* R) {
    std::vector<double> v;
    if (!read_doubles(path, &v)) return false;
    v.resize(v.size() / 9 * 9);
    R->swap(v);
    return true;
}
inline bool read_T_file(const std::string& path, std::vector<double>* T) {
    std::vector<double> v;
    if (!read_doubles(path, &v)) return false;
    v.resize(v.size() / 3 * 3);
    T->swap(v);
    return true;
}

// `ofstream << Eigen::MatrixXd` (row-major data in, rows x cols).
inline bool write_eigen_matrix(const std::string& path, const double* m, int64_t rows, int cols) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const int64_t n = rows * cols;
    std::vector<char> txt(size_t(n) * 32);
    std::vector<uint8_t> len(size_t(n) ? size_t(n) : 1);
    int width = 0;
    for (int64_t i = 0; i < n; i++) {
        const int l = std::snprintf(&txt[size_t(i) * 32], 32, "%.6g", m[i]);   // default ostream formatting at precision 6
        len[size_t(i)] = uint8_t(l);
        if (l > width) width = l;
    }
    std::string line;
    for (int64_t r = 0; r < rows; r++) {
        line.clear();
        for (int c = 0; c < cols; c++) {
            const int64_t i = r * cols + c;
            if (c) line.push_back(' ');
            line.append(size_t(width - len[size_t(i)]), ' ');
            line.append(&txt[size_t(i) * 32], len[size_t(i)]);
        }
        if (r + 1 < rows) line.push_back('\n');
        std::fwrite(line.data(), 1, line.size(), f);
    }
    return std::fclose(f) == 0;
}

}  // namespace curvix_host
