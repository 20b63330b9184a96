// libkrd_host.so — C entry points of the host-side pieces (jplace ingest, matrix writer) so that
// they can be exercised without a GPU (tests/test_host_ingest.py) and reused from other languages.
// No CUDA in this file; the compute path lives in libkrd_b200.so (include/krd_b200.h).
#include <cstring>

#include "../../include/krd_host.h"
#include "ingest.hpp"
#include "matrix_writer.hpp"

using namespace krdhost;

namespace {
thread_local std::string g_err;
struct Loaded { MassSet ms; std::vector<std::string> files, labels; double diameter = 0.0; };
}  // namespace

extern "C" {

const char* krd_host_last_error() { return g_err.c_str(); }

krd_host_set* krd_host_load_jplace(int32_t n_paths, const char* const* paths, int32_t point_mass,
                                   int32_t ignore_multiplicities, int32_t threads)
{
    try {
        std::vector<std::string> raw(paths, paths + n_paths);
        auto L = new Loaded();
        L->files = resolve_jplace_paths(raw);
        for (const auto& f : L->files) L->labels.push_back(sample_label(f));
        IngestOptions opt;
        opt.point_mass = point_mass != 0; opt.ignore_multiplicities = ignore_multiplicities != 0; opt.threads = threads;
        L->ms = load_mass_set(L->files, opt);
        L->diameter = diameter(L->ms.tree);
        return reinterpret_cast<krd_host_set*>(L);
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return nullptr;
    }
}

void krd_host_free_set(krd_host_set* s) { delete reinterpret_cast<Loaded*>(s); }

void krd_host_set_sizes(const krd_host_set* s, int32_t* n_edges, int32_t* n_samples, int64_t* n_masses,
                        int32_t* root, double* diameter_out)
{
    auto L = reinterpret_cast<const Loaded*>(s);
    *n_edges = L->ms.tree.n_edges; *n_samples = L->ms.n_samples(); *n_masses = (int64_t)L->ms.edge.size();
    *root = L->ms.tree.root; *diameter_out = L->diameter;
}

void krd_host_set_copy(const krd_host_set* s, int32_t* parent, int32_t* child, int64_t* edge_num, double* bl,
                       int64_t* off, int32_t* edge, double* pos, double* mass, double* pendant)
{
    auto L = reinterpret_cast<const Loaded*>(s);
    const auto& t = L->ms.tree;
    auto cp = [](void* d, const void* src, size_t b) { if (d && b) std::memcpy(d, src, b); };
    cp(parent, t.parent.data(), t.parent.size() * 4); cp(child, t.child.data(), t.child.size() * 4);
    cp(edge_num, t.edge_num.data(), t.edge_num.size() * 8); cp(bl, t.bl.data(), t.bl.size() * 8);
    cp(off, L->ms.off.data(), L->ms.off.size() * 8); cp(edge, L->ms.edge.data(), L->ms.edge.size() * 4);
    cp(pos, L->ms.pos.data(), L->ms.pos.size() * 8); cp(mass, L->ms.mass.data(), L->ms.mass.size() * 8);
    cp(pendant, L->ms.pendant.data(), L->ms.pendant.size() * 8);
}

const char* krd_host_set_label(const krd_host_set* s, int32_t i)
{
    auto L = reinterpret_cast<const Loaded*>(s);
    return (i >= 0 && (size_t)i < L->labels.size()) ? L->labels[(size_t)i].c_str() : "";
}

int64_t krd_host_write_matrix(const char* path, const double* m, int32_t n, const char* format,
                              const char* const* labels, const char* corner, int32_t gzip, int32_t threads)
{
    try {
        std::vector<std::string> names;
        if (labels) for (int i = 0; i < n; ++i) names.emplace_back(labels[i]);
        return (int64_t)write_matrix_file(path, m, (size_t)n, parse_matrix_format(format ? format : "matrix"),
                                          labels ? &names : nullptr, corner ? corner : "", gzip != 0, threads);
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return -1;
    }
}

double krd_host_parse_number(const char* text)
{
    JsonScanner js(text, text + std::strlen(text), "text");
    try { return js.number().as_double(); } catch (...) { return std::nan(""); }
}

void krd_host_add_pendant_work(double* m, int32_t n, const double* pendant)
{
    // placement::earth_movers_distance( SampleSet, p, with_pendant_length = true ), placement/function/emd.cpp:118-126:
    // every cell (the diagonal too) gets the pendant work of both of its samples
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j < n; ++j) m[i * n + j] += pendant[i] + pendant[j];
}

}  // extern "C"
