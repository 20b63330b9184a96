// gappa-b200 — `analyze krd` (and `analyze squash`, cluster tree only) with the flag surface of the reference command
// (gappa src/commands/analyze/krd.cpp:58-158) on top of the B200 library (include/krd_b200.h).
//
//   gappa-b200 analyze krd --jplace-path <files|dirs>... [--exponent p] [--normalize]
//        [--point-mass] [--ignore-multiplicities] [--out-dir d] [--file-prefix s] [--file-suffix s]
//        [--compress] [--matrix-format matrix|list|triangular] [--omit-matrix-labels]
//        [--allow-file-overwriting] [--verbose] [--threads n] [--log-file f]
//   gappa-b200 analyze squash --jplace-path ... [--exponent p] [--point-mass] [--ignore-multiplicities] [--out-dir d]
//        [--file-prefix s] [--file-suffix s]       -> cluster.newick  (gappa src/commands/analyze/squash.cpp:94-175;
//        the per-cluster coloured tree drawings of the reference are visualisation and out of scope)
//   gappa-b200 analyze phylogenetic-kmeans --jplace-path ... --k "1-5,8" --bins B [--write-overview-file] [--point-mass]
//        [--ignore-multiplicities] [--out-dir d] [--file-prefix s] [--file-suffix s]  -> k_<k>_assignments.csv, overview.csv
//        (gappa src/commands/analyze/kmeans_phylogenetic.cpp:134-199; additive: --seed n, --init-centroids i,j,...)
//   additive: --krd-exponent (alias), --bins B (mass_tree_binify_masses first), --device d, --gpus G (analyze krd: tiles
//             sharded over devices d..d+G-1, one host thread per GPU; --exchange nccl (default: NCCL all-gather of the
//             result tiles over NVLink, every GPU then copies its row block of the matrix to the host) | host (no
//             collective: every GPU copies its own packed tiles to the host, unpacked there)),
//             --write-binary f (raw row-major doubles, for full-precision comparisons)
//
// Exit code 0 on success; on error prints "Error: <message>" like src/main.cpp:167-173 and exits 1.
#include <sys/stat.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <thread>
#include <limits>
#include <random>
#include <sstream>
#include <algorithm>

#include "../../include/krd_b200.h"
#include "ingest.hpp"
#include "matrix_writer.hpp"

using namespace krdhost;

namespace {

struct KrdOptions {
    bool squash = false;      // `analyze squash` instead of `analyze krd`
    bool kmeans = false;      // `analyze phylogenetic-kmeans`
    std::string ks, init_centroids;   // --k range list; --init-centroids (additive): explicit sample indices instead of k-means++
    bool overview_file = false;
    long long seed = -1;      // --seed (additive): k-means++ random seed; default: clock, like the reference's Options::random_engine
    std::vector<std::string> jplace_paths;
    double exponent = 1.0;
    bool normalize = false, point_mass = false, ignore_multiplicities = false;
    std::string out_dir = ".", prefix, suffix, matrix_format = "matrix", log_file, write_binary;
    bool compress = false, omit_labels = false, allow_overwrite = false, verbose = false;
    int threads = 0, bins = 0, device = 0, gpus = 1;
    std::string exchange = "nccl";
};

const char* kUsage =
    "Usage: gappa-b200 analyze krd | squash | phylogenetic-kmeans [OPTIONS]\n"
    "Calculate the pairwise Kantorovich-Rubinstein (KR) distance matrix between samples.\n\n"
    "Input:\n  --jplace-path TEXT ... REQUIRED   List of jplace files or directories to process.\n"
    "Settings:\n  --exponent FLOAT=1                Exponent for KR integration.\n"
    "  --normalize                       Divide the KR distance by the tree length to get normalized values.\n"
    "  --point-mass                      Treat every pquery as a point mass on its highest-weight placement.\n"
    "  --ignore-multiplicities           Set the multiplicity of each pquery to 1.0.\n"
    "  --bins INT=0                      (additive) binify masses into this many bins per edge first.\n"
    "  --k TEXT  --write-overview-file   (phylogenetic-kmeans) values of k, e.g. 1-5,8.\n"
    "  --seed INT  --init-centroids LIST (phylogenetic-kmeans, additive) k-means++ seed / explicit start samples.\n"
    "Matrix Output:\n  --out-dir TEXT=.  --file-prefix TEXT  --file-suffix TEXT  --compress\n"
    "  --matrix-format TEXT=matrix       matrix | list | triangular\n  --omit-matrix-labels\n"
    "Global Options:\n  --allow-file-overwriting  --verbose  --threads UINT  --log-file TEXT  --device INT  --gpus INT  --exchange nccl|host  --help\n";

[[noreturn]] void usage_error(const std::string& msg)
{
    std::cerr << msg << "\nRun with --help for more information.\n";
    std::exit(106);   // CLI11 ExitCodes::ExtrasError-like; any non-zero
}

bool valid_filefix(const std::string& s)
{
    if (s.empty()) return true;
    if (std::isspace((unsigned char)s.front()) || std::isspace((unsigned char)s.back())) return false;
    for (char c : s) if (!std::isprint((unsigned char)c) || std::string("<>:\"\\/|?*").find(c) != std::string::npos) return false;
    return true;
}

KrdOptions parse_args(int argc, char** argv)
{
    KrdOptions o;
    int i = 1;
    if (argc >= 2 && (std::string(argv[1]) == "--help" || std::string(argv[1]) == "-h")) { std::cout << kUsage; std::exit(0); }
    if (argc < 3 || std::string(argv[1]) != "analyze" || (std::string(argv[2]) != "krd" && std::string(argv[2]) != "squash" && std::string(argv[2]) != "phylogenetic-kmeans"))
        usage_error("gappa-b200 implements `analyze krd`, `analyze squash` and `analyze phylogenetic-kmeans`.");
    o.squash = std::string(argv[2]) == "squash";
    o.kmeans = std::string(argv[2]) == "phylogenetic-kmeans";
    i = 3;
    std::vector<std::string> args;
    for (; i < argc; ++i) {
        std::string a = argv[i];
        const size_t eq = a.find('=');
        if (a.rfind("--", 0) == 0 && eq != std::string::npos) { args.push_back(a.substr(0, eq)); args.push_back(a.substr(eq + 1)); }
        else args.push_back(a);
    }
    auto is_opt = [](const std::string& s) { return s.rfind("--", 0) == 0; };
    for (size_t k = 0; k < args.size(); ++k) {
        const std::string& a = args[k];
        auto value = [&]() -> std::string {
            if (k + 1 >= args.size() || is_opt(args[k + 1])) usage_error(a + ": 1 required TEXT missing");
            return args[++k];
        };
        auto number = [&](auto conv) {
            const std::string v = value();
            try { size_t used = 0; auto r = conv(v, &used); if (used != v.size()) throw 0; return r; }
            catch (...) { usage_error("Could not convert: " + a + " = " + v); }
        };
        if (a == "--jplace-path") {
            size_t before = o.jplace_paths.size();
            while (k + 1 < args.size() && !is_opt(args[k + 1])) o.jplace_paths.push_back(args[++k]);
            if (o.jplace_paths.size() == before) usage_error("--jplace-path: At least 1 required");
        } else if (a == "--exponent" || a == "--krd-exponent") o.exponent = number([](const std::string& s, size_t* u) { return std::stod(s, u); });
        else if (a == "--normalize") o.normalize = true;
        else if (a == "--point-mass") o.point_mass = true;
        else if (a == "--ignore-multiplicities") o.ignore_multiplicities = true;
        else if (a == "--out-dir") o.out_dir = value();
        else if (a == "--file-prefix") o.prefix = value();
        else if (a == "--file-suffix") o.suffix = value();
        else if (a == "--compress") o.compress = true;
        else if (a == "--matrix-format") o.matrix_format = value();
        else if (a == "--omit-matrix-labels") o.omit_labels = true;
        else if (a == "--allow-file-overwriting") o.allow_overwrite = true;
        else if (a == "--verbose") o.verbose = true;
        else if (a == "--threads") o.threads = number([](const std::string& s, size_t* u) { return std::stoi(s, u); });
        else if (a == "--log-file") o.log_file = value();
        else if (a == "--bins") o.bins = number([](const std::string& s, size_t* u) { return std::stoi(s, u); });
        else if (a == "--device") o.device = number([](const std::string& s, size_t* u) { return std::stoi(s, u); });
        else if (a == "--gpus") o.gpus = number([](const std::string& s, size_t* u) { return std::stoi(s, u); });
        else if (a == "--exchange") o.exchange = value();
        else if (a == "--write-binary") o.write_binary = value();
        else if (a == "--k") o.ks = value();
        else if (a == "--write-overview-file") o.overview_file = true;
        else if (a == "--init-centroids") o.init_centroids = value();
        else if (a == "--seed") o.seed = number([](const std::string& s, size_t* u) { return std::stoll(s, u); });
        else if (a == "--help" || a == "-h") { std::cout << kUsage; std::exit(0); }
        else usage_error("The following argument was not expected: " + a);
    }
    if (o.jplace_paths.empty()) usage_error("--jplace-path is required");
    for (const auto& p : o.jplace_paths)
        if (!is_file_path(p) && !is_dir_path(p)) usage_error("--jplace-path: Path does not exist: " + p);
    if (!valid_filefix(o.prefix)) usage_error("--file-prefix: File prefix contains invalid characters (`<>:\"\\/|?*`), non-printable characters, or surrounding whitespace.");
    if (!valid_filefix(o.suffix)) usage_error("--file-suffix: File suffix contains invalid characters (`<>:\"\\/|?*`), non-printable characters, or surrounding whitespace.");
    parse_matrix_format(o.matrix_format);   // validates
    if (o.bins < 0) usage_error("--bins: must be >= 0");
    if (o.gpus < 1 || o.gpus > 64) usage_error("--gpus: must be in 1..64");
    if (o.exchange != "nccl" && o.exchange != "host") usage_error("--exchange: must be nccl or host");
    if (o.threads <= 0) {
        o.threads = (int)std::thread::hardware_concurrency();
        if (o.threads <= 0) o.threads = 1;
    }
    return o;
}

struct Log {
    std::ofstream file;
    void msg(const std::string& s) { std::cout << s << std::endl; if (file.is_open()) file << s << std::endl; }
};

void mkdirs(const std::string& dir)
{
    std::string cur;
    for (size_t i = 0; i <= dir.size(); ++i) {
        if (i == dir.size() || dir[i] == '/') {
            if (!cur.empty() && !is_dir_path(cur) && ::mkdir(cur.c_str(), 0775) != 0 && !is_dir_path(cur))
                throw std::runtime_error("Cannot create directory: " + cur);
        }
        if (i < dir.size()) cur += dir[i];
    }
}

void check_krd(krd_handle* h, krd_status st)
{
    if (st == KRD_OK) return;
    const std::string m = krd_last_error(h);
    if (st == KRD_ERR_INCOMPATIBLE_TREES) throw std::invalid_argument(m);   // emd.cpp:78,109,123,172
    throw std::runtime_error(m);
}

// Masses -> device.  With --bins B the entries are binned HERE (the reference's expression, mass_tree/functions.cpp:231-237,
// evaluated exactly as the library's device pass evaluates it) and cross PCIe in the compact form of krd_set_masses_binned*:
// cell = edge * B + bin as uint16 (trees with n_edges * B <= 65536) or uint32, plus the mass — 10 / 12 bytes per entry
// instead of 20.  Equal cells of a sample stay separate entries; the library sums them in input order, as it does for the
// full form, so both routes give the same bits.
template <typename CT>
void binned_cells(const MassSet& ms, int bins, int threads, std::vector<CT>& cell)
{
    const size_t M = ms.edge.size();
    cell.resize(M);
    const double nb = (double)bins, below = std::nextafter(nb, 0.0);
    auto work = [&](size_t m0, size_t m1) {
        for (size_t m = m0; m < m1; ++m) {
            const double v = ms.pos[m] / ms.tree.bl[(size_t)ms.edge[m]] * nb;
            double pn = (0.0 < v) ? v : 0.0;          // (NaN -> 0, like std::max(0.0, v))
            pn = (below < pn) ? below : pn;
            cell[m] = (CT)((uint32_t)ms.edge[m] * (uint32_t)bins + (uint32_t)std::floor(pn));
        }
    };
    const int T = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(threads, 1), M / 65536 + 1));
    std::vector<std::thread> pool;
    for (int t = 1; t < T; ++t) pool.emplace_back(work, M * (size_t)t / (size_t)T, M * (size_t)(t + 1) / (size_t)T);
    work(0, M / (size_t)T);
    for (auto& w : pool) w.join();
}

krd_status set_masses(krd_handle* h, const MassSet& ms, int bins, int threads)
{
    const int32_t n = ms.n_samples();
    const int64_t cells = (int64_t)ms.tree.n_edges * bins;
    if (bins < 1 || cells > 0xFFFFFFF0ll)
        return krd_set_masses(h, n, ms.off.data(), ms.edge.data(), ms.pos.data(), ms.mass.data());
    for (int32_t e : ms.edge)                          // (the ingest has checked edge numbers; a foreign MassSet may not have)
        if (e < 0 || e >= ms.tree.n_edges) return krd_set_masses(h, n, ms.off.data(), ms.edge.data(), ms.pos.data(), ms.mass.data());
    if (cells <= 65536) {
        std::vector<uint16_t> cell;
        binned_cells(ms, bins, threads, cell);
        return krd_set_masses_binned16(h, n, ms.off.data(), cell.data(), ms.mass.data());
    }
    std::vector<uint32_t> cell;
    binned_cells(ms, bins, threads, cell);
    return krd_set_masses_binned(h, n, ms.off.data(), cell.data(), ms.mass.data());
}

int run_krd(const KrdOptions& o, Log& log)
{
    using clk = std::chrono::steady_clock;
    // krd.cpp:119-121 / file_output.cpp:229-277: fail early if the output exists
    std::string dir = o.out_dir;
    while (dir.size() > 1 && dir.back() == '/') dir.pop_back();
    const std::string out_path = dir + "/" + o.prefix + "krd_matrix" + o.suffix + ".csv" + (o.compress ? ".gz" : "");
    if (is_file_path(out_path)) {
        if (!o.allow_overwrite)
            throw std::runtime_error("Output file already exists: " + out_path + "\nUse --allow-file-overwriting to allow gappa to overwrite the file.");
        log.msg("Warning: Output file already exists and will be overwritten: " + out_path);
    }

    const auto files = resolve_jplace_paths(o.jplace_paths);
    log.msg("Found " + std::to_string(files.size()) + " jplace file" + (files.size() > 1 ? "s" : ""));   // file_input.cpp:168-170
    if (files.size() < 2) throw std::runtime_error("Cannot run krd with fewer than 2 samples.");       // krd.cpp:131-133

    auto t0 = clk::now();
    IngestOptions io;
    io.point_mass = o.point_mass; io.ignore_multiplicities = o.ignore_multiplicities; io.threads = o.threads;
    MassSet ms = load_mass_set(files, io);
    auto t1 = clk::now();

    log.msg("Calculating pairwise KR distances.");
    const int32_t n = ms.n_samples();
    std::vector<double> matrix((size_t)n * (size_t)n, 0.0);
    krd_stats st{};
    // krd.cpp:144-150: --normalize divides by the DIAMETER of the (averaged) tree
    const double divisor = o.normalize ? diameter(ms.tree) : 1.0;
    if (o.gpus <= 1) {
        krd_config cfg{};
        cfg.device = o.device; cfg.bins = o.bins;
        krd_handle* h = nullptr;
        if (krd_create(&cfg, &h) != KRD_OK) throw std::runtime_error(krd_last_error(nullptr));
        try {
            check_krd(h, krd_set_tree(h, ms.tree.n_edges, ms.tree.parent.data(), ms.tree.child.data(), ms.tree.bl.data(), ms.tree.root));
            check_krd(h, set_masses(h, ms, o.bins, o.threads > 0 ? o.threads : (int)std::thread::hardware_concurrency()));
            check_krd(h, krd_compute(h, o.exponent, divisor, matrix.data()));
            krd_get_stats(h, &st);
        } catch (...) { krd_destroy(h); throw; }
        krd_destroy(h);
    } else {
        // --gpus G (additive): one host thread per GPU (devices o.device .. o.device+G-1), every GPU
        // gets the replicated inputs over its own PCIe link, computes its contiguous share of the tile list and
        // copies its packed tiles back; the host scatters them into the matrix.  No collective: only the host
        // needs the result (the NCCL all-gather variant is what bench.py / one-process-per-GPU callers use).
        const int G = o.gpus;
        std::vector<std::string> errors((size_t)G);
        std::vector<krd_stats> stats((size_t)G);
        std::vector<double> packed;
        std::vector<std::thread> workers;
        std::vector<krd_handle*> hs((size_t)G, nullptr);
        int64_t shard = -1;
        // phase 1: create + upload on all GPUs concurrently
        for (int g = 0; g < G; ++g) workers.emplace_back([&, g]() {
            try {
                krd_config cfg{};
                cfg.device = o.device + g; cfg.bins = o.bins;
                if (krd_create(&cfg, &hs[(size_t)g]) != KRD_OK) throw std::runtime_error(krd_last_error(nullptr));
                krd_handle* h = hs[(size_t)g];
                check_krd(h, krd_set_tree(h, ms.tree.n_edges, ms.tree.parent.data(), ms.tree.child.data(), ms.tree.bl.data(), ms.tree.root));
                check_krd(h, set_masses(h, ms, o.bins, o.threads > 0 ? o.threads : (int)std::thread::hardware_concurrency()));
            } catch (std::exception const& ex) { errors[(size_t)g] = ex.what(); }
        });
        for (auto& w : workers) w.join();
        workers.clear();
        auto cleanup = [&]() { for (auto* h : hs) if (h) krd_destroy(h); };
        for (const auto& e : errors) if (!e.empty()) { cleanup(); throw std::runtime_error(e); }
        if (o.exchange == "nccl") {
            // phase 2 (default): every GPU computes its share, the packed tiles are all-gathered with NCCL over NVLink and
            // every GPU copies ITS row block of the finished matrix to the host over its own PCIe link
            std::vector<void*> comms((size_t)G, nullptr);
            std::vector<int32_t> devs((size_t)G);
            for (int g = 0; g < G; ++g) devs[(size_t)g] = o.device + g;
            if (krd_nccl_init_all(G, devs.data(), comms.data()) != KRD_OK) {
                const std::string msg = krd_last_error(nullptr);
                cleanup();
                throw std::runtime_error(msg + " (use --exchange host to run without NCCL)");
            }
            for (int g = 0; g < G; ++g) workers.emplace_back([&, g]() {
                try {
                    const int32_t r0 = (int32_t)((int64_t)g * n / G), r1 = (int32_t)((int64_t)(g + 1) * n / G);
                    check_krd(hs[(size_t)g], krd_compute_allgather_host(hs[(size_t)g], o.exponent, divisor, g, G, comms[(size_t)g], r0, r1,
                                                                        matrix.data() + (size_t)r0 * (size_t)n));
                    krd_get_stats(hs[(size_t)g], &stats[(size_t)g]);
                } catch (std::exception const& ex) { errors[(size_t)g] = ex.what(); }
            });
            for (auto& w : workers) w.join();
            krd_nccl_destroy(G, comms.data());
            cleanup();
            for (const auto& e : errors) if (!e.empty()) throw std::runtime_error(e);
        } else {
            shard = krd_shard_doubles(hs[0], G);
            packed.assign((size_t)shard * (size_t)G, 0.0);
            // phase 2 (--exchange host): every GPU computes and returns its own packed tiles; no collective
            for (int g = 0; g < G; ++g) workers.emplace_back([&, g]() {
                try {
                    check_krd(hs[(size_t)g], krd_compute_shard_host(hs[(size_t)g], o.exponent, divisor, g, G, packed.data() + (size_t)g * (size_t)shard));
                    krd_get_stats(hs[(size_t)g], &stats[(size_t)g]);
                } catch (std::exception const& ex) { errors[(size_t)g] = ex.what(); }
            });
            for (auto& w : workers) w.join();
            cleanup();
            for (const auto& e : errors) if (!e.empty()) throw std::runtime_error(e);
            if (krd_unpack_shards_host(n, G, packed.data(), matrix.data()) != KRD_OK) throw std::runtime_error("krd_unpack_shards_host failed");
        }
        st = stats[0];
        for (const auto& q : stats) { st.ms_pairs = std::max(st.ms_pairs, q.ms_pairs); st.ms_build = std::max(st.ms_build, q.ms_build); }
    }
    auto t2 = clk::now();

    log.msg("Writing distance matrix.");
    mkdirs(dir);
    std::vector<std::string> names;
    for (const auto& f : files) names.push_back(sample_label(f));
    write_matrix_file(out_path, matrix.data(), (size_t)n, parse_matrix_format(o.matrix_format),
                      o.omit_labels ? nullptr : &names, "Sample", o.compress, o.threads);
    if (!o.write_binary.empty()) {
        std::ofstream b(o.write_binary, std::ios::binary);
        b.write(reinterpret_cast<const char*>(matrix.data()), (std::streamsize)(matrix.size() * sizeof(double)));
    }
    auto t3 = clk::now();
    if (o.verbose) {
        auto sec = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
        char buf[512];
        std::snprintf(buf, sizeof buf,
                      "samples %d, edges %lld, masses %lld, segments %lld | ingest %.3f s, device path %.3f s "
                      "(F build %.3f ms, pair kernel %.3f ms, epilogue %.3f ms), write %.3f s",
                      n, (long long)st.n_edges, (long long)st.n_masses, (long long)st.n_segments, sec(t0, t1), sec(t1, t2),
                      st.ms_build, st.ms_pairs, st.ms_epilogue, sec(t2, t3));
        log.msg(buf);
    }
    log.msg("Finished");
    return 0;
}

// SquashClustering::tree_string (genesis tree/mass_tree/squash_clustering.cpp:251-274)
std::string squash_tree_string(std::vector<std::string> list, const std::vector<krd_merger>& mergers)
{
    const size_t n = list.size();
    for (size_t i = 0; i < mergers.size(); ++i) {
        const auto& cm = mergers[i];
        const std::string a = list[(size_t)cm.index_a] + ":" + std::to_string(cm.distance_a);
        const std::string b = list[(size_t)cm.index_b] + ":" + std::to_string(cm.distance_b);
        list[(size_t)cm.index_a].clear(); list[(size_t)cm.index_b].clear();
        list.push_back("(" + a + "," + b + ")" + std::to_string(i + n));
    }
    return list.back() + ";";
}

int run_squash(const KrdOptions& o, Log& log)
{
    if (o.exponent <= 0.0)   // squash.cpp:98-103
        throw std::runtime_error("--exponent (" + std::to_string(o.exponent) + "): Invalid exponent value for KR distance calculation. Has to be > 0.0.");
    std::string dir = o.out_dir;
    while (dir.size() > 1 && dir.back() == '/') dir.pop_back();
    const std::string out_path = dir + "/" + o.prefix + "cluster" + o.suffix + ".newick";
    if (is_file_path(out_path)) {
        if (!o.allow_overwrite)
            throw std::runtime_error("Output file already exists: " + out_path + "\nUse --allow-file-overwriting to allow gappa to overwrite the file.");
        log.msg("Warning: Output file already exists and will be overwritten: " + out_path);
    }
    const auto files = resolve_jplace_paths(o.jplace_paths);
    log.msg("Found " + std::to_string(files.size()) + " jplace file" + (files.size() > 1 ? "s" : ""));
    IngestOptions io;
    io.point_mass = o.point_mass; io.ignore_multiplicities = o.ignore_multiplicities; io.threads = o.threads;
    MassSet ms = load_mass_set(files, io);
    log.msg("Running Squash Clustering");
    const int32_t n = ms.n_samples();
    krd_config cfg{};
    cfg.device = o.device; cfg.bins = o.bins;
    krd_handle* h = nullptr;
    if (krd_create(&cfg, &h) != KRD_OK) throw std::runtime_error(krd_last_error(nullptr));
    std::vector<krd_merger> mergers((size_t)std::max(n - 1, 0));
    try {
        check_krd(h, krd_set_tree(h, ms.tree.n_edges, ms.tree.parent.data(), ms.tree.child.data(), ms.tree.bl.data(), ms.tree.root));
        check_krd(h, set_masses(h, ms, o.bins, o.threads > 0 ? o.threads : (int)std::thread::hardware_concurrency()));
        check_krd(h, krd_squash(h, o.exponent, mergers.data()));
    } catch (...) { krd_destroy(h); throw; }
    krd_destroy(h);
    log.msg("Writing output files");
    mkdirs(dir);
    std::vector<std::string> names;
    for (const auto& f : files) names.push_back(sample_label(f));
    std::ofstream out(out_path, std::ios::binary);
    if (!out) throw std::runtime_error("Cannot open output file '" + out_path + "'.");
    out << squash_tree_string(names, mergers);
    log.msg("Finished");
    return 0;
}

// genesis::utils::split_range_list ("1-5,8,10,12" -> sorted unique numbers), as used by get_k_values (kmeans.cpp:138-161)
std::vector<int> parse_k_values(const std::string& ks)
{
    const std::string msg = "--k (" + ks + "): Invalid list of values for k. Needs to be a comma-separated list of positive numbers or ranges, e.g., 5-10,12,15";
    std::vector<int> out;
    size_t i = 0;
    auto num = [&](size_t& q) -> long {
        while (q < ks.size() && ks[q] == ' ') ++q;
        size_t b = q; long v = 0;
        while (q < ks.size() && std::isdigit((unsigned char)ks[q])) { v = v * 10 + (ks[q] - '0'); if (v > 1000000) throw std::runtime_error(msg); ++q; }
        if (q == b) throw std::runtime_error(msg);
        while (q < ks.size() && ks[q] == ' ') ++q;
        return v;
    };
    while (i < ks.size()) {
        long a = num(i), b = a;
        if (i < ks.size() && ks[i] == '-') { ++i; b = num(i); }
        if (b < a) throw std::runtime_error(msg);
        for (long v = a; v <= b; ++v) out.push_back((int)v);
        if (i < ks.size()) { if (ks[i] != ',') throw std::runtime_error(msg); ++i; if (i == ks.size()) throw std::runtime_error(msg); }
    }
    std::sort(out.begin(), out.end());
    out.erase(std::unique(out.begin(), out.end()), out.end());
    if (out.empty() || out[0] == 0) throw std::runtime_error(msg);
    return out;
}

std::string fmt_g(double v) { char b[64]; std::snprintf(b, sizeof b, "%g", v); return b; }   // default ostream << double

// `analyze phylogenetic-kmeans` (src/commands/analyze/kmeans_phylogenetic.cpp:134-199, kmeans.cpp:163-269): assignment
// tables k_<k>_assignments.csv and, with --write-overview-file, overview.csv.  The centroid tree drawings
// (write_pkmeans_cluster_trees: svg / nexus / phyloxml colour trees) are not produced.
int run_kmeans(const KrdOptions& o, Log& log)
{
    const std::vector<int> ks = parse_k_values(o.ks);
    std::string dir = o.out_dir;
    while (dir.size() > 1 && dir.back() == '/') dir.pop_back();
    auto out_path = [&](const std::string& infix) { return dir + "/" + o.prefix + infix + o.suffix + ".csv"; };
    std::vector<std::string> outs;
    for (int k : ks) outs.push_back(out_path("k_" + std::to_string(k) + "_assignments"));
    if (o.overview_file) outs.push_back(out_path("overview"));
    for (const auto& f : outs) if (is_file_path(f)) {
        if (!o.allow_overwrite)
            throw std::runtime_error("Output file already exists: " + f + "\nUse --allow-file-overwriting to allow gappa to overwrite the file.");
        log.msg("Warning: Output file already exists and will be overwritten: " + f);
    }
    const auto files = resolve_jplace_paths(o.jplace_paths);
    log.msg("Found " + std::to_string(files.size()) + " jplace file" + (files.size() > 1 ? "s" : ""));
    if (files.size() < 2) throw std::runtime_error("Cannot run k-means with fewer than 2 samples.");   // kmeans_phylogenetic.cpp:146-148
    log.msg("Reading samples.");
    IngestOptions io;
    io.point_mass = o.point_mass; io.ignore_multiplicities = o.ignore_multiplicities; io.threads = o.threads;
    MassSet ms = load_mass_set(files, io);
    const int32_t n = ms.n_samples();
    std::vector<int32_t> fixed_init;
    if (!o.init_centroids.empty()) {
        std::stringstream ss(o.init_centroids); std::string tok;
        while (std::getline(ss, tok, ',')) fixed_init.push_back((int32_t)std::stol(tok));
    }
    const unsigned long long seed = o.seed >= 0 ? (unsigned long long)o.seed
                                                : (unsigned long long)std::chrono::system_clock::now().time_since_epoch().count();
    std::mt19937_64 engine(seed);
    if (fixed_init.empty()) log.msg("Random seed for k-means++: " + std::to_string(seed));

    krd_config cfg{};
    cfg.device = o.device; cfg.bins = o.bins;
    krd_handle* h = nullptr;
    if (krd_create(&cfg, &h) != KRD_OK) throw std::runtime_error(krd_last_error(nullptr));
    struct Row { int k; double dist, var; };
    std::vector<Row> overview;
    mkdirs(dir);
    try {
        check_krd(h, krd_set_tree(h, ms.tree.n_edges, ms.tree.parent.data(), ms.tree.child.data(), ms.tree.bl.data(), ms.tree.root));
        check_krd(h, set_masses(h, ms, o.bins, o.threads > 0 ? o.threads : (int)std::thread::hardware_concurrency()));
        std::vector<double> D;   // pairwise KR matrix: the data-to-data-point distances k-means++ draws from
        std::vector<int32_t> assign((size_t)n), prev_assign;
        std::vector<double> dist((size_t)n);
        for (size_t ki = 0; ki < ks.size(); ++ki) {
            const int k = ks[ki];
            log.msg("\nRunning Phylogenetic Kmeans with k=" + std::to_string(k));
            if (k > n) throw std::runtime_error("Cannot run Kmeans with more clusters (k == " + std::to_string(k) + ") than data points (" + std::to_string(n) + ")");
            std::vector<int32_t> init;
            if (!fixed_init.empty()) {
                if ((int)fixed_init.size() != k) throw std::runtime_error("--init-centroids: needs exactly k sample indices");
                init = fixed_init;
            } else {
                // k-means++ (utils/math/kmeans.hpp:614-659): first centroid uniform, the others with probability
                // proportional to the squared distance to the nearest centroid chosen so far
                if (D.empty()) { D.resize((size_t)n * n); check_krd(h, krd_compute(h, 1.0, 1.0, D.data())); }
                init.push_back((int32_t)std::uniform_int_distribution<size_t>(0, (size_t)n - 1)(engine));
                std::vector<double> probs((size_t)n);
                for (int c = 1; c < k; ++c) {
                    for (int d = 0; d < n; ++d) {
                        double m = std::numeric_limits<double>::max();
                        for (int32_t q : init) m = std::min(m, D[(size_t)d * n + q]);
                        probs[(size_t)d] = m * m;
                    }
                    init.push_back((int32_t)std::discrete_distribution<size_t>(probs.begin(), probs.end())(engine));
                }
            }
            int32_t iterations = 0;
            // gappa runs every k on ONE MassTreeKmeans object: the assignment vector of the previous k is still there when
            // the next run starts (kmeans.hpp:296-306), and runtime_checks_ (kmeans.hpp:569-576) rejects it if it does not fit
            const int32_t* carry = nullptr;
            if (ki > 0) {
                for (int d = 0; d < n; ++d)
                    if (assign[(size_t)d] >= k)
                        throw std::runtime_error("Invalid assignment " + std::to_string(assign[(size_t)d]) + " >= k = " + std::to_string(k) + ".");
                prev_assign = assign;
                carry = prev_assign.data();
            }
            check_krd(h, krd_kmeans_ex(h, k, 100, init.data(), carry, assign.data(), dist.data(), &iterations));
            log.msg("Finished after " + std::to_string(iterations) + " iterations");
            std::ofstream out(outs[ki], std::ios::binary);
            if (!out) throw std::runtime_error("Cannot open output file '" + outs[ki] + "'.");
            out << "Sample\tCluster\tDistance\n";
            for (int d = 0; d < n; ++d) out << sample_label(files[(size_t)d]) << "\t" << assign[(size_t)d] << "\t" << fmt_g(dist[(size_t)d]) << "\n";
            // print_cluster_info (kmeans.cpp:194-248)
            double avg_dst = 0.0, avg_var = 0.0;
            for (int c = 0; c < k; ++c) {
                double cd = 0.0, cv = 0.0; size_t cnt = 0;
                for (int d = 0; d < n; ++d) if (assign[(size_t)d] == c) { avg_dst += dist[(size_t)d]; avg_var += dist[(size_t)d] * dist[(size_t)d]; cd += dist[(size_t)d]; cv += dist[(size_t)d] * dist[(size_t)d]; ++cnt; }
                log.msg("Cluster " + std::to_string(c) + ": " + std::to_string(cnt) + " samples, with a variance of " + fmt_g(cnt ? cv / (double)cnt : 0.0) +
                        " and average distance " + fmt_g(cd / (double)cnt));
            }
            avg_dst /= (double)n; avg_var /= (double)n;
            log.msg("Total average distance: " + fmt_g(avg_dst));
            log.msg("Total average variance: " + fmt_g(avg_var));
            overview.push_back({k, avg_dst, avg_var});
        }
    } catch (...) { krd_destroy(h); throw; }
    krd_destroy(h);
    if (o.overview_file) {
        std::ofstream out(outs.back(), std::ios::binary);
        if (!out) throw std::runtime_error("Cannot open output file '" + outs.back() + "'.");
        out << "k\tDistance\tVariance\n";
        for (const auto& r : overview) out << r.k << "\t" << fmt_g(r.dist) << "\t" << fmt_g(r.var) << "\n";
    }
    log.msg("Finished");
    return 0;
}

}  // namespace

int main(int argc, char** argv)
{
    KrdOptions o = parse_args(argc, argv);
    Log log;
    if (!o.log_file.empty()) log.file.open(o.log_file);
    try {
        return o.kmeans ? run_kmeans(o, log) : o.squash ? run_squash(o, log) : run_krd(o, log);
    } catch (std::exception const& ex) {
        std::cerr << "Error: " << ex.what() << std::endl;
        if (log.file.is_open()) log.file << "Error: " << ex.what() << std::endl;
        return 1;
    }
}
