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
#include <cmath>
#include <cstdint>
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

// "%g" (6 significant digits) without printf for the values a distance matrix holds.  The six digits come from ONE
// correctly rounded scaling of v by an exact power of ten (10^k is a double for |k| <= 22): scaled = v * 10^(5 - X) lies in
// [1e5, 1e6) with an absolute error below 1.2e-10, so its integer part and its distance to the rounding boundary .5 decide
// the digits with certainty unless the fraction lies within 1e-9 of .5 — those values (and anything outside
// [1e-17, 1e22), infinities, NaN) take the printf path.  Returns the length, 0 = not handled.
inline int format_g6_fast(char* out, double v)
{
    static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
                                   1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    char* p = out;
    if (std::signbit(v)) { *p++ = '-'; v = -v; }
    if (v == 0.0) { *p++ = '0'; return (int)(p - out); }
    if (!(v >= 1e-17 && v < 1e22)) return 0;                     // (also NaN)
    uint64_t bits;
    std::memcpy(&bits, &v, 8);
    const int e2 = (int)(bits >> 52) - 1022;                     // v = f * 2^e2, f in [0.5, 1)  (v is normal here)
    int X = ((e2 - 1) * 78913) >> 18;                            // floor((e2 - 1) * log10 2): floor(log10 v) or one less
    auto scale = [&](int k) { return k >= 0 ? v * P10[k] : v / P10[-k]; };
    if (X < -17) X = -17;
    double sc = scale(5 - X);
    if (sc >= 1e6) { ++X; sc = scale(5 - X); }
    else if (sc < 1e5) { --X; if (X < -17) return 0; sc = scale(5 - X); }
    if (!(sc >= 1e5 && sc < 1e6)) return 0;
    uint32_t n = (uint32_t)sc;                                    // (sc > 0: truncation = floor)
    const double frac = sc - (double)n;
    if (frac > 0.5 - 1e-9 && frac < 0.5 + 1e-9) return 0;         // too close to a tie to decide here
    n += frac > 0.5 ? 1u : 0u;
    if (n >= 1000000u) { n = 100000u; ++X; }
    char d[6];
    { const uint32_t hi = n / 1000u, lo = n - hi * 1000u;         // two three-digit halves
      d[0] = (char)('0' + hi / 100u); d[1] = (char)('0' + hi / 10u % 10u); d[2] = (char)('0' + hi % 10u);
      d[3] = (char)('0' + lo / 100u); d[4] = (char)('0' + lo / 10u % 10u); d[5] = (char)('0' + lo % 10u); }
    int nd = 6;
    while (nd > 1 && d[nd - 1] == '0') --nd;                     // %g drops trailing zeros
    if (X < -4 || X >= 6) {                                       // d.ddddde[+-]XX
        *p++ = d[0];
        if (nd > 1) { *p++ = '.'; for (int i = 1; i < nd; ++i) *p++ = d[i]; }
        *p++ = 'e';
        int ax = X;
        if (ax < 0) { *p++ = '-'; ax = -ax; } else *p++ = '+';
        if (ax >= 100) { *p++ = (char)('0' + ax / 100); ax %= 100; }
        *p++ = (char)('0' + ax / 10); *p++ = (char)('0' + ax % 10);
    } else if (X >= 0) {
        for (int i = 0; i <= X; ++i) *p++ = d[i];
        if (nd > X + 1) { *p++ = '.'; for (int i = X + 1; i < nd; ++i) *p++ = d[i]; }
    } else {
        *p++ = '0'; *p++ = '.';
        for (int i = 0; i < -X - 1; ++i) *p++ = '0';
        for (int i = 0; i < nd; ++i) *p++ = d[i];
    }
    return (int)(p - out);
}

constexpr size_t VALUE_MAX = 32;   // longest "%g" text of a double, with room to spare

// Writes "%g" of v at p (at least VALUE_MAX bytes free) and returns the end.
inline char* put_value(char* p, double v)
{
    int n = format_g6_fast(p, v);
    if (n == 0) n = std::snprintf(p, VALUE_MAX, "%g", v);
    return p + n;
}

inline void append_value(std::string& s, double v)
{
    char buf[VALUE_MAX + 8];
    s.append(buf, (size_t)(put_value(buf, v) - buf));
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
    // the values of the row are written through a pointer into a scratch row sized for the longest case (kept per thread:
    // growing the std::string itself would zero-fill four times the bytes the row ends up with), then appended once
    const size_t cells = (fmt == MatrixFormat::Matrix) ? n : n - r;
    thread_local std::vector<char> scratch;
    if (scratch.size() < cells * (VALUE_MAX + 1) + 1) scratch.resize(cells * (VALUE_MAX + 1) + 1);
    char* const p0 = scratch.data();
    char* p = p0;
    if (fmt == MatrixFormat::Matrix) {
        for (size_t c = 0; c < n; ++c) { if (c) *p++ = '\t'; p = put_value(p, row[c]); }
    } else {  // columns backwards, starting at the last one, down to col == row
        for (size_t cb = 0; cb < n - r; ++cb) { if (cb) *p++ = '\t'; p = put_value(p, row[n - cb - 1]); }
    }
    *p++ = '\n';
    out.append(p0, (size_t)(p - p0));
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

// One complete gzip member holding `in` (RFC 1952; members may simply be concatenated: gzip -d, zlib's gzread and the
// reference's GzipIStream read them as one stream).  This is what lets the batches of rows be compressed in parallel — one
// gzwrite stream runs at ~50 MB/s, which at N = 10k (1 GB of text) would dwarf everything else in the command.
inline std::string gzip_member(const std::string& in)
{
    z_stream zs{};
    if (deflateInit2(&zs, Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK)
        throw std::runtime_error("zlib: deflateInit2 failed.");
    std::string out;
    out.resize(deflateBound(&zs, (uLong)in.size()) + 64);
    zs.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(in.data()));
    zs.avail_in = (uInt)in.size();
    zs.next_out = reinterpret_cast<Bytef*>(&out[0]);
    zs.avail_out = (uInt)out.size();
    const int rc = deflate(&zs, Z_FINISH);
    const size_t produced = out.size() - zs.avail_out;
    deflateEnd(&zs);
    if (rc != Z_STREAM_END) throw std::runtime_error("zlib: deflate failed.");
    out.resize(produced);
    return out;
}

// names == nullptr: --omit-matrix-labels.  Returns the number of (uncompressed) bytes written.
inline size_t write_matrix_file(const std::string& path, const double* m, size_t n, MatrixFormat fmt,
                                const std::vector<std::string>* names, const std::string& corner,
                                bool gzip, int threads)
{
    if (names && names->size() != n) throw std::invalid_argument("Number of row/column names differs from the matrix size.");
    FILE* fp = std::fopen(path.c_str(), "wb");
    if (!fp) throw std::runtime_error("Cannot open output file '" + path + "'.");
    size_t total = 0;
    bool wrote = false;
    auto put = [&](const std::string& s) {
        wrote = wrote || !s.empty();
        if (!s.empty() && std::fwrite(s.data(), 1, s.size(), fp) != s.size()) { std::fclose(fp); throw std::runtime_error("Error writing '" + path + "'."); }
    };
    if (names) { const std::string h = format_header(n, fmt, *names, corner); total += h.size(); put(gzip && !h.empty() ? gzip_member(h) : h); }

    if (threads < 1) threads = 1;
    const size_t block = 256;                      // rows per batch and thread
    std::vector<std::string> bufs((size_t)threads), packed((size_t)threads);
    std::vector<size_t> plain((size_t)threads);
    std::vector<std::string> errs((size_t)threads);
    for (size_t r0 = 0; r0 < n; r0 += block * (size_t)threads) {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) {
            const size_t a = r0 + (size_t)t * block, b = std::min(n, a + block);
            bufs[(size_t)t].clear(); packed[(size_t)t].clear(); plain[(size_t)t] = 0;
            if (a >= b) continue;
            auto work = [&, a, b, t]() {
                try {
                    for (size_t r = a; r < b; ++r) format_row(bufs[(size_t)t], m, n, r, fmt, names);
                    plain[(size_t)t] = bufs[(size_t)t].size();
                    if (gzip && !bufs[(size_t)t].empty()) packed[(size_t)t] = gzip_member(bufs[(size_t)t]);   // one member per batch
                } catch (std::exception const& ex) { errs[(size_t)t] = ex.what(); }
            };
            if (threads == 1) work(); else pool.emplace_back(work);
        }
        for (auto& th : pool) th.join();
        for (int t = 0; t < threads; ++t) {
            if (!errs[(size_t)t].empty()) { std::fclose(fp); throw std::runtime_error(errs[(size_t)t]); }
            total += plain[(size_t)t];
            put(gzip ? packed[(size_t)t] : bufs[(size_t)t]);
        }
    }
    if (gzip && !wrote) put(gzip_member(std::string()));          // (an empty file is not a gzip file)
    if (std::fclose(fp) != 0) throw std::runtime_error("Error writing '" + path + "'.");
    return total;
}

}  // namespace krdhost
