// MeshIO.h -- result writers and mesh readers of the driver layer at multi-million-element sizes.
//
// The reference writes its results with one `ostream << value << ", " ... << endl` per line (Mesh::outputPoints /
// outputSimplices / outputBoundaryNodes, src/Mesh.cpp:1049-1095: a flush per line) and reads FromFile meshes with a
// stringstream per line (utils::readTriangles, src/MeshUtils.h:669-733).  At SURVEY C5 (8.4 M points, 16.8 M triangles) that
// text IO takes far longer than the whole GPU stepping loop.  Here:
//   * write_text: the SAME bytes as the reference's loops (default ostream formatting of a double = printf "%g", 6 significant
//     digits; ", " separators; '\n' line ends), formatted by all host threads into per-chunk buffers and written with one
//     fwrite per chunk;
//   * write_bin / read_bin: a raw little-endian table, 32-byte header {"MMADMMB1", int64 rows, int64 cols, int32 dtype (0 = f64,
//     1 = i32), int32 0}, row-major data -- exact, and ~20x smaller to parse than text;
//   * read_text: whole-file read + strtod / strtol scan (commas, blanks and line ends are separators).
// utils::readTriangles accepts either format for each of its three files (the magic decides); Mesh::output* also writes the
// binary twin "<name>.bin" when MMADMM_BINARY_IO=1.
#ifndef MMADMM_MESH_IO_H
#define MMADMM_MESH_IO_H

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

namespace meshio {

static const char MAGIC[8] = {'M', 'M', 'A', 'D', 'M', 'M', 'B', '1'};

inline int host_threads() {
    unsigned n = std::thread::hardware_concurrency();
    return n ? (int)n : 1;
}

// rows x cols table, value(i, j) -> text exactly as `out << v(i,0) << ", " << ... << v(i,cols-1) << endl`
template <class T>
inline bool write_text(const char *fname, long rows, int cols, const std::function<T(long, int)> &value,
                       const std::function<bool(long)> &keepRow = nullptr) {
    FILE *f = fopen(fname, "wb");
    if (!f) return false;
    const int nthr = (rows > 20000) ? host_threads() : 1;
    const long per = (rows + nthr - 1) / nthr;
    std::vector<std::string> bufs(nthr);
    auto work = [&](int t) {
        const long r0 = t * per, r1 = std::min(rows, r0 + per);
        std::string &b = bufs[t];
        b.reserve((size_t)std::max<long>(r1 - r0, 0) * cols * 12);
        char tmp[64];
        for (long i = r0; i < r1; i++) {
            if (keepRow && !keepRow(i)) continue;
            for (int j = 0; j < cols; j++) {
                int n;
                if (sizeof(T) == sizeof(int) && (T)0.5 == (T)0) n = snprintf(tmp, sizeof(tmp), "%d", (int)value(i, j));
                else n = snprintf(tmp, sizeof(tmp), "%g", (double)value(i, j));
                b.append(tmp, n);
                if (j + 1 < cols) b.append(", ", 2);
            }
            b.push_back('\n');
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthr; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    bool ok = true;
    for (auto &b : bufs) ok = ok && fwrite(b.data(), 1, b.size(), f) == b.size();
    return fclose(f) == 0 && ok;
}

template <class T> inline bool write_bin(const char *fname, long rows, int cols, const std::function<T(long, int)> &value) {
    FILE *f = fopen(fname, "wb");
    if (!f) return false;
    int64_t hdr[2] = {rows, cols};
    int32_t tail[2] = {(sizeof(T) == sizeof(int) && (T)0.5 == (T)0) ? 1 : 0, 0};
    std::vector<T> row((size_t)rows * cols);
    for (long i = 0; i < rows; i++) for (int j = 0; j < cols; j++) row[(size_t)i * cols + j] = value(i, j);
    bool ok = fwrite(MAGIC, 1, 8, f) == 8 && fwrite(hdr, 8, 2, f) == 2 && fwrite(tail, 4, 2, f) == 2 &&
              fwrite(row.data(), sizeof(T), row.size(), f) == row.size();
    return fclose(f) == 0 && ok;
}

inline bool slurp(const char *fname, std::vector<char> &buf) {
    FILE *f = fopen(fname, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    buf.resize((size_t)n + 1);
    bool ok = fread(buf.data(), 1, (size_t)n, f) == (size_t)n;
    buf[n] = 0;
    fclose(f);
    return ok;
}
inline bool is_bin(const std::vector<char> &buf) { return buf.size() >= 32 && !memcmp(buf.data(), MAGIC, 8); }

// every number of a text or binary table, in file order
inline bool read_doubles(const char *fname, std::vector<double> &out) {
    std::vector<char> buf;
    if (!slurp(fname, buf)) return false;
    out.clear();
    if (is_bin(buf)) {
        int64_t hdr[2]; int32_t tail[2];
        memcpy(hdr, buf.data() + 8, 16); memcpy(tail, buf.data() + 24, 8);
        const size_t n = (size_t)hdr[0] * hdr[1];
        out.resize(n);
        if (tail[0] == 0) memcpy(out.data(), buf.data() + 32, n * 8);
        else { const int32_t *p = (const int32_t *)(buf.data() + 32); for (size_t i = 0; i < n; i++) out[i] = p[i]; }
        return true;
    }
    out.reserve(buf.size() / 8);
    const char *p = buf.data(), *end = buf.data() + buf.size() - 1;
    while (p < end) {
        while (p < end && (*p == ',' || *p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++;
        if (p >= end) break;
        char *q;
        const double v = strtod(p, &q);
        if (q == p) return false;
        out.push_back(v);
        p = q;
    }
    return true;
}
inline bool read_ints(const char *fname, std::vector<int> &out) {
    std::vector<char> buf;
    if (!slurp(fname, buf)) return false;
    out.clear();
    if (is_bin(buf)) {
        int64_t hdr[2]; int32_t tail[2];
        memcpy(hdr, buf.data() + 8, 16); memcpy(tail, buf.data() + 24, 8);
        const size_t n = (size_t)hdr[0] * hdr[1];
        out.resize(n);
        if (tail[0] == 1) memcpy(out.data(), buf.data() + 32, n * 4);
        else { const double *p = (const double *)(buf.data() + 32); for (size_t i = 0; i < n; i++) out[i] = (int)p[i]; }
        return true;
    }
    out.reserve(buf.size() / 4);
    const char *p = buf.data(), *end = buf.data() + buf.size() - 1;
    while (p < end) {
        while (p < end && (*p == ',' || *p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++;
        if (p >= end) break;
        char *q;
        const long v = strtol(p, &q, 10);
        if (q == p) return false;
        out.push_back((int)v);
        p = q;
    }
    return true;
}
inline bool binary_twin() { const char *e = getenv("MMADMM_BINARY_IO"); return e && e[0] == '1'; }

}  // namespace meshio
#endif
