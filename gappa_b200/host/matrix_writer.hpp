// Distance-matrix text output, byte-compatible with the reference.
//
// Formats and layout follow genesis utils/containers/matrix/writer.hpp:200-320 as driven by
// gappa src/options/matrix_output.cpp:76-111: tab separator, "matrix" / "list" / "triangular",
// optional row+column labels with a corner cell, values through the default ostream formatting
// (= printf "%g", 6 significant digits; writer.hpp:72,237).  Optional gzip (file_output.cpp:284-299).
// Rows are formatted in parallel: at N = 10k this is 1e8 cells, the reference's single-threaded
// operator<< loop takes minutes.
#pragma once
#include <zlib.h>

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace krdhost {

enum class MatrixFormat { Matrix, List, Triangular };

inline MatrixFormat parse_matrix_format(std::string v)
{
    for (auto& c : v) c = (char)std::tolower((unsigned char)c);   // CLI::ignore_case
    if (v == "matrix") return MatrixFormat::Matrix;
    if (v == "list") return MatrixFormat::List;
    if (v == "triangular") return MatrixFormat::Triangular;
    throw std::invalid_argument("--matrix-format: Invalid format '" + v + "'.");
}

inline void append_value(std::string& s, double v)
{
    char buf[40];
    int n = std::snprintf(buf, sizeof buf, "%g", v);
    s.append(buf, (size_t)n);
}

// One output row (or, for "list", the n lines of matrix row r).
inline void format_row(std::string& out, const double* m, size_t n, size_t r, MatrixFormat fmt,
                       const std::vector<std::string>* names)
{
    const double* row = m + r * n;
    if (fmt == MatrixFormat::List) {
        for (size_t c = 0; c < n; ++c) {
            if (names) { out += (*names)[r]; out += '\t'; out += (*names)[c]; out += '\t'; }
            append_value(out, row[c]);
            out += '\n';
        }
        return;
    }
    if (names) { out += (*names)[r]; out += '\t'; }
    if (fmt == MatrixFormat::Matrix) {
        for (size_t c = 0; c < n; ++c) { if (c) out += '\t'; append_value(out, row[c]); }
    } else {  // columns backwards, starting at the last one, down to col == row
        for (size_t cb = 0; cb < n - r; ++cb) { if (cb) out += '\t'; append_value(out, row[n - cb - 1]); }
    }
    out += '\n';
}

inline std::string format_header(size_t n, MatrixFormat fmt, const std::vector<std::string>& names, const std::string& corner)
{
    std::string h;
    if (fmt == MatrixFormat::List) return h;
    h += corner; h += '\t';
    for (size_t k = 0; k < n; ++k) {
        if (k) h += '\t';
        h += names[fmt == MatrixFormat::Triangular ? n - k - 1 : k];
    }
    h += '\n';
    return h;
}

// names == nullptr: --omit-matrix-labels.  Returns the number of (uncompressed) bytes written.
inline size_t write_matrix_file(const std::string& path, const double* m, size_t n, MatrixFormat fmt,
                                const std::vector<std::string>* names, const std::string& corner,
                                bool gzip, int threads)
{
    if (names && names->size() != n) throw std::invalid_argument("Number of row/column names differs from the matrix size.");
    gzFile gz = nullptr;
    FILE* fp = nullptr;
    if (gzip) { gz = gzopen(path.c_str(), "wb"); if (!gz) throw std::runtime_error("Cannot open output file '" + path + "'."); }
    else { fp = std::fopen(path.c_str(), "wb"); if (!fp) throw std::runtime_error("Cannot open output file '" + path + "'."); }
    size_t total = 0;
    auto emit = [&](const std::string& s) {
        if (s.empty()) return;
        total += s.size();
        if (gz) { if (gzwrite(gz, s.data(), (unsigned)s.size()) <= 0) throw std::runtime_error("Error writing '" + path + "'."); }
        else if (std::fwrite(s.data(), 1, s.size(), fp) != s.size()) throw std::runtime_error("Error writing '" + path + "'.");
    };
    if (names) emit(format_header(n, fmt, *names, corner));

    if (threads < 1) threads = 1;
    const size_t block = 256;                      // rows per batch and thread
    std::vector<std::string> bufs((size_t)threads);
    for (size_t r0 = 0; r0 < n; r0 += block * (size_t)threads) {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) {
            const size_t a = r0 + (size_t)t * block, b = std::min(n, a + block);
            bufs[(size_t)t].clear();
            if (a >= b) continue;
            auto work = [&, a, b, t]() { for (size_t r = a; r < b; ++r) format_row(bufs[(size_t)t], m, n, r, fmt, names); };
            if (threads == 1) work(); else pool.emplace_back(work);
        }
        for (auto& th : pool) th.join();
        for (int t = 0; t < threads; ++t) emit(bufs[(size_t)t]);
    }
    if (gz) gzclose(gz); else std::fclose(fp);
    return total;
}

}  // namespace krdhost
