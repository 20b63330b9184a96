// TEST INFRASTRUCTURE ONLY — never linked into, imported by or executed from the product path.
//
// Thin C-ABI driver around the UNMODIFIED genesis sources (compiled where they lie under
// /root/reference by oracle/Makefile) so that tests and bench.py's reference arm can call the
// reference's own implementation of the KR path on flat arrays:
//   * ref_emd_matrix    -> genesis::tree::earth_movers_distance(std::vector<MassTree>, p)
//                          (libs/genesis/lib/genesis/tree/mass_tree/emd.cpp:183-235), optionally
//                          after mass_tree_binify_masses (tree/mass_tree/functions.cpp:222-265)
//   * ref_jplace_load   -> the ingest gappa does in JplaceInputOptions::mass_tree_set()
//                          (src/options/jplace_input.cpp:129-171,295-332), dumped to flat arrays
//   * ref_write_matrix -> genesis::utils::MatrixWriter<double> as configured by gappa
//                          (src/options/matrix_output.cpp:76-111), to a file
//   * ref_squash       -> genesis::tree::SquashClustering::run (tree/mass_tree/squash_clustering.cpp:213-249)
//   * ref_kmeans       -> genesis::tree::MassTreeKmeans (tree/mass_tree/kmeans.cpp:59-208) via utils::Kmeans::run
//                          (utils/math/kmeans.hpp:118-168) as set up by gappa (src/commands/analyze/kmeans_phylogenetic.cpp:155-175)
//   * ref_diameter      -> genesis::tree::diameter (tree/common_tree/functions.cpp:178-205)
// This file is ours; it contains no reference code, it only calls it.

#include "genesis/placement/formats/jplace_reader.hpp"
#include "genesis/placement/formats/newick_reader.hpp"
#include "genesis/placement/function/functions.hpp"
#include "genesis/placement/function/masses.hpp"
#include "genesis/placement/function/operators.hpp"
#include "genesis/placement/sample.hpp"
#include "genesis/tree/common_tree/functions.hpp"
#include "genesis/tree/function/operators.hpp"
#include "genesis/tree/mass_tree/emd.hpp"
#include "genesis/tree/mass_tree/functions.hpp"
#include "genesis/tree/mass_tree/kmeans.hpp"
#include "genesis/tree/mass_tree/squash_clustering.hpp"
#include "genesis/tree/mass_tree/tree.hpp"
#include "genesis/utils/containers/matrix.hpp"
#include "genesis/utils/containers/matrix/writer.hpp"
#include "genesis/utils/io/input_source.hpp"
#include "genesis/utils/io/output_target.hpp"

#include <omp.h>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

using namespace genesis;

namespace {

thread_local std::string g_err;

// Newick with {edge} tags from flat parent/child arrays; children in listing order.
std::string flat_to_newick(int n_edges, const int* parent, const int* child, const double* bl, int root)
{
    int n_nodes = n_edges + 1;
    std::vector<std::vector<int>> kids(n_nodes);
    for (int e = 0; e < n_edges; ++e) kids[parent[e]].push_back(e);
    std::ostringstream os;
    os.precision(17);
    // iterative emit
    struct Fr { int node; int up_edge; size_t next; };
    std::vector<Fr> st;
    st.push_back({root, -1, 0});
    while (!st.empty()) {
        Fr& f = st.back();
        if (f.next == 0 && !kids[f.node].empty()) os << "(";
        if (f.next < kids[f.node].size()) {
            if (f.next > 0) os << ",";
            int e = kids[f.node][f.next++];
            st.push_back({child[e], e, 0});
            continue;
        }
        if (!kids[f.node].empty()) os << ")";
        os << "n" << f.node;
        if (f.up_edge >= 0) os << ":" << bl[f.up_edge] << "{" << f.up_edge << "}";
        st.pop_back();
    }
    os << ";";
    return os.str();
}

struct Dump {
    std::vector<int> parent, child, edge_num;
    std::vector<double> bl;
    int root = 0;
    std::vector<int64_t> off;
    std::vector<int> edge;
    std::vector<double> pos, mass;
    std::vector<double> pendant;
    double diameter = 0.0;
};

} // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

int ref_max_threads() { return omp_get_max_threads(); }

// Build N mass trees on the given topology, fill masses[pos] += mass (std::map semantics as in
// tree/mass_tree/tree.hpp:199), optionally binify, run the reference all-pairs EMD.
int ref_emd_matrix(int n_edges, const int* parent, const int* child, const double* bl, int root,
                   int n_samples, const int64_t* off, const int* edge, const double* pos,
                   const double* mass, int bins, double p, int threads, double* out,
                   double* seconds)
{
    try {
        if (threads > 0) omp_set_num_threads(threads);
        auto nw = flat_to_newick(n_edges, parent, child, bl, root);
        auto ptree = placement::PlacementTreeNewickReader().read(utils::from_string(nw));
        if ((int)ptree.edge_count() != n_edges) throw std::runtime_error("edge count mismatch");
        std::vector<size_t> e2g(n_edges);
        for (auto const& e : ptree.edges()) {
            e2g[e.data<placement::PlacementEdgeData>().edge_num()] = e.index();
        }
        auto base = tree::convert_common_tree_to_mass_tree(ptree);
        // restore exact branch lengths (the newick text round-trips %.17g, but be safe)
        for (int e = 0; e < n_edges; ++e)
            base.edge_at(e2g[e]).data<tree::MassTreeEdgeData>().branch_length = bl[e];
        std::vector<tree::MassTree> trees(n_samples);
        #pragma omp parallel for schedule(dynamic)
        for (int i = 0; i < n_samples; ++i) {
            trees[i] = base;
            for (int64_t m = off[i]; m < off[i + 1]; ++m) {
                trees[i].edge_at(e2g[edge[m]]).data<tree::MassTreeEdgeData>().masses[pos[m]] += mass[m];
            }
        }
        if (bins > 0) {
            for (auto& t : trees) tree::mass_tree_binify_masses(t, (size_t)bins);
        }
        auto t0 = std::chrono::steady_clock::now();
        auto mat = tree::earth_movers_distance(trees, p);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        std::memcpy(out, mat.data().data(), sizeof(double) * (size_t)n_samples * n_samples);
        return 0;
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return 1;
    }
}

// Same inputs, but dump the (optionally binified) mass trees back as flat sorted arrays
// (edge asc in OUR numbering, pos asc). Caller sizes the outputs >= off[n_samples].
int ref_binify_flat(int n_edges, const int* parent, const int* child, const double* bl, int root,
                    int n_samples, const int64_t* off, const int* edge, const double* pos,
                    const double* mass, int bins, int64_t* out_off, int* out_edge,
                    double* out_pos, double* out_mass)
{
    try {
        auto nw = flat_to_newick(n_edges, parent, child, bl, root);
        auto ptree = placement::PlacementTreeNewickReader().read(utils::from_string(nw));
        std::vector<size_t> e2g(n_edges);
        for (auto const& e : ptree.edges())
            e2g[e.data<placement::PlacementEdgeData>().edge_num()] = e.index();
        auto base = tree::convert_common_tree_to_mass_tree(ptree);
        for (int e = 0; e < n_edges; ++e)
            base.edge_at(e2g[e]).data<tree::MassTreeEdgeData>().branch_length = bl[e];
        int64_t w = 0;
        out_off[0] = 0;
        for (int i = 0; i < n_samples; ++i) {
            auto t = base;
            for (int64_t m = off[i]; m < off[i + 1]; ++m)
                t.edge_at(e2g[edge[m]]).data<tree::MassTreeEdgeData>().masses[pos[m]] += mass[m];
            if (bins > 0) tree::mass_tree_binify_masses(t, (size_t)bins);
            for (int e = 0; e < n_edges; ++e) {
                for (auto const& kv : t.edge_at(e2g[e]).data<tree::MassTreeEdgeData>().masses) {
                    out_edge[w] = e; out_pos[w] = kv.first; out_mass[w] = kv.second; ++w;
                }
            }
            out_off[i + 1] = w;
        }
        return 0;
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return 1;
    }
}

// ---- jplace ingest exactly as gappa's mass_tree_set() ---------------------------------------

void* ref_jplace_load(int n_files, const char** paths, int point_mass, int ignore_mult)
{
    try {
        using namespace genesis::placement;
        auto d = new Dump();
        std::vector<tree::MassTree> mts(n_files);
        d->pendant.resize(n_files);
        JplaceReader reader;
        for (int fi = 0; fi < n_files; ++fi) {
            auto smp = reader.read(utils::from_file(paths[fi]));
            if (point_mass) { filter_n_max_weight_placements(smp); normalize_weight_ratios(smp); }
            if (ignore_mult) {
                for (auto& pq : smp) {
                    auto const tm = total_multiplicity(pq);
                    for (auto& nm : pq.names()) nm.multiplicity /= tm;
                }
            }
            auto r = convert_sample_to_mass_tree(smp, true);
            mts[fi] = std::move(r.first);
            d->pendant[fi] = r.second;
            if (fi == 0) {
                for (auto const& e : smp.tree().edges())
                    d->edge_num.push_back(e.data<PlacementEdgeData>().edge_num());
            }
        }
        if (!tree::identical_topology(mts))
            throw std::runtime_error("Sample reference trees do not have identical topology.");
        tree::mass_trees_make_average_branch_lengths(mts);
        auto const& t0 = mts[0];
        d->root = (int)t0.root_node().index();
        for (auto const& e : t0.edges()) {
            d->parent.push_back((int)e.primary_node().index());
            d->child.push_back((int)e.secondary_node().index());
            d->bl.push_back(e.data<tree::MassTreeEdgeData>().branch_length);
        }
        d->diameter = tree::diameter(t0);
        d->off.push_back(0);
        for (auto const& t : mts) {
            for (auto const& e : t.edges()) {
                for (auto const& kv : e.data<tree::MassTreeEdgeData>().masses) {
                    d->edge.push_back((int)e.index());
                    d->pos.push_back(kv.first);
                    d->mass.push_back(kv.second);
                }
            }
            d->off.push_back((int64_t)d->edge.size());
        }
        return d;
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return nullptr;
    }
}

void ref_dump_sizes(void* h, int* n_edges, int* n_samples, int64_t* n_masses, int* root, double* diameter)
{
    auto d = static_cast<Dump*>(h);
    *n_edges = (int)d->parent.size();
    *n_samples = (int)d->off.size() - 1;
    *n_masses = (int64_t)d->edge.size();
    *root = d->root;
    *diameter = d->diameter;
}

void ref_dump_copy(void* h, int* parent, int* child, int* edge_num, double* bl, int64_t* off,
                   int* edge, double* pos, double* mass, double* pendant)
{
    auto d = static_cast<Dump*>(h);
    std::memcpy(parent, d->parent.data(), d->parent.size() * sizeof(int));
    std::memcpy(child, d->child.data(), d->child.size() * sizeof(int));
    std::memcpy(edge_num, d->edge_num.data(), d->edge_num.size() * sizeof(int));
    std::memcpy(bl, d->bl.data(), d->bl.size() * sizeof(double));
    std::memcpy(off, d->off.data(), d->off.size() * sizeof(int64_t));
    std::memcpy(edge, d->edge.data(), d->edge.size() * sizeof(int));
    std::memcpy(pos, d->pos.data(), d->pos.size() * sizeof(double));
    std::memcpy(mass, d->mass.data(), d->mass.size() * sizeof(double));
    std::memcpy(pendant, d->pendant.data(), d->pendant.size() * sizeof(double));
}

void ref_dump_free(void* h) { delete static_cast<Dump*>(h); }

// gappa's MatrixOutputOptions::write_matrix (src/options/matrix_output.cpp:76-111) on a flat matrix:
// format 0 = matrix, 1 = list, 2 = triangular; labels == nullptr -> --omit-matrix-labels.
int ref_write_matrix(const char* path, const double* m, int n, int format, const char** labels, const char* corner)
{
    try {
        utils::Matrix<double> mat((size_t)n, (size_t)n);
        for (size_t r = 0; r < (size_t)n; ++r) for (size_t c = 0; c < (size_t)n; ++c) mat(r, c) = m[r * n + c];
        auto writer = utils::MatrixWriter<double>();
        writer.format(format == 0 ? utils::MatrixWriter<double>::Format::kMatrix
                      : format == 1 ? utils::MatrixWriter<double>::Format::kList
                                    : utils::MatrixWriter<double>::Format::kTriangular);
        auto target = utils::to_file(path);
        if (labels) {
            std::vector<std::string> names(labels, labels + n);
            writer.write(mat, target, names, names, corner);
        } else {
            writer.write(mat, target);
        }
        return 0;
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return 1;
    }
}

// SquashClustering on the same flat inputs; mergers out as (index_a, index_b, distance_a, distance_b) per merge,
// tree_out (may be null) receives tree_string(labels "0".."n-1"), at most tree_cap bytes.
int ref_squash(int n_edges, const int* parent, const int* child, const double* bl, int root,
               int n_samples, const int64_t* off, const int* edge, const double* pos, const double* mass,
               int bins, double p, int64_t* out_index /* 2*(n-1) */, double* out_dist /* 2*(n-1) */,
               char* tree_out, int64_t tree_cap, double* seconds)
{
    try {
        auto nw = flat_to_newick(n_edges, parent, child, bl, root);
        auto ptree = placement::PlacementTreeNewickReader().read(utils::from_string(nw));
        std::vector<size_t> e2g(n_edges);
        for (auto const& e : ptree.edges()) e2g[e.data<placement::PlacementEdgeData>().edge_num()] = e.index();
        auto base = tree::convert_common_tree_to_mass_tree(ptree);
        for (int e = 0; e < n_edges; ++e) base.edge_at(e2g[e]).data<tree::MassTreeEdgeData>().branch_length = bl[e];
        std::vector<tree::MassTree> trees(n_samples);
        for (int i = 0; i < n_samples; ++i) {
            trees[i] = base;
            for (int64_t m = off[i]; m < off[i + 1]; ++m)
                trees[i].edge_at(e2g[edge[m]]).data<tree::MassTreeEdgeData>().masses[pos[m]] += mass[m];
            if (bins > 0) tree::mass_tree_binify_masses(trees[i], (size_t)bins);
        }
        tree::SquashClustering sc;
        sc.p(p);
        auto t0 = std::chrono::steady_clock::now();
        sc.run(std::move(trees));
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        auto const& mg = sc.mergers();
        for (size_t k = 0; k < mg.size(); ++k) {
            out_index[2 * k] = (int64_t)mg[k].index_a; out_index[2 * k + 1] = (int64_t)mg[k].index_b;
            out_dist[2 * k] = mg[k].distance_a; out_dist[2 * k + 1] = mg[k].distance_b;
        }
        if (tree_out && tree_cap > 0) {
            std::vector<std::string> labels;
            for (int i = 0; i < n_samples; ++i) labels.push_back(std::to_string(i));
            auto ts = sc.tree_string(labels);
            std::strncpy(tree_out, ts.c_str(), (size_t)tree_cap - 1);
            tree_out[tree_cap - 1] = 0;
        }
        return 0;
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return 1;
    }
}

// MassTreeKmeans on the same flat inputs, started from given data points as centroids (what both of the
// reference's initialisation strategies produce).  bins as in gappa: binify the data and accumulate_centroid_masses.
int ref_kmeans(int n_edges, const int* parent, const int* child, const double* bl, int root,
               int n_samples, const int64_t* off, const int* edge, const double* pos, const double* mass,
               int bins, int k, int max_iterations, const int* init_centroids,
               int64_t* out_assign /* n */, double* out_dist /* n */, double* out_var /* k */, int64_t* out_iterations, double* seconds)
{
    try {
        auto nw = flat_to_newick(n_edges, parent, child, bl, root);
        auto ptree = placement::PlacementTreeNewickReader().read(utils::from_string(nw));
        std::vector<size_t> e2g(n_edges);
        for (auto const& e : ptree.edges()) e2g[e.data<placement::PlacementEdgeData>().edge_num()] = e.index();
        auto base = tree::convert_common_tree_to_mass_tree(ptree);
        for (int e = 0; e < n_edges; ++e) base.edge_at(e2g[e]).data<tree::MassTreeEdgeData>().branch_length = bl[e];
        std::vector<tree::MassTree> trees(n_samples);
        for (int i = 0; i < n_samples; ++i) {
            trees[i] = base;
            for (int64_t m = off[i]; m < off[i + 1]; ++m)
                trees[i].edge_at(e2g[edge[m]]).data<tree::MassTreeEdgeData>().masses[pos[m]] += mass[m];
            if (bins > 0) tree::mass_tree_binify_masses(trees[i], (size_t)bins);
        }
        tree::MassTreeKmeans km;
        if (bins > 0) km.accumulate_centroid_masses((size_t)bins);
        km.max_iterations((size_t)max_iterations);
        km.initialization_strategy(utils::KmeansInitializationStrategy::kNone);
        std::vector<tree::MassTree> cents;
        for (int c = 0; c < k; ++c) cents.push_back(trees[init_centroids[c]]);
        km.centroids(cents);
        auto t0 = std::chrono::steady_clock::now();
        auto const it = km.run(trees, (size_t)k);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        auto const info = km.cluster_info(trees);
        for (int i = 0; i < n_samples; ++i) { out_assign[i] = (int64_t)km.assignments()[i]; out_dist[i] = info.distances[i]; }
        for (int c = 0; c < k; ++c) out_var[c] = info.variances[c];
        *out_iterations = (int64_t)it;
        return 0;
    } catch (std::exception const& ex) {
        g_err = ex.what();
        return 1;
    }
}

} // extern "C"
